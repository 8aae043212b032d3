// ORACLE — TEST INFRASTRUCTURE ONLY.  CPU restatement of the EQP-cluster compute-counts path.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// build, link or execute anything under oracle/.  The shipped library never calls into it.
//
// PARITY UNPINNED: the reference holds no golden vectors for this path (its samples.tgz blob is
// missing) and cannot run here (no JVM / python2 / bedtools).  Fidelity is argued
// function by function against the cited reference lines; tests/golden holds hand-traced cases
//  and the literal SAM lines of SamRecord.java:1253-1268.
//
// Citation shorthands: J/ = exon-pipeline-scripts/java/source/
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <set>
#include <stdexcept>
#include <string>
#include <string_view>
#include <unordered_map>
#include <unordered_set>
#include <vector>

namespace eqp {

using sv = std::string_view;

struct Error : std::runtime_error { using std::runtime_error::runtime_error; };

inline std::string read_file(const std::string& path) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) throw Error("ERROR: file " + path + " not found.");
  std::string s; char buf[1 << 16]; size_t n;
  while ((n = fread(buf, 1, sizeof buf, f)) > 0) s.append(buf, n);
  fclose(f);
  return s;
}
inline void write_file(const std::string& path, const std::string& s) {
  FILE* f = fopen(path.c_str(), "wb");
  if (!f) throw Error("ERROR: file " + path + " cannot be written to.");
  fwrite(s.data(), 1, s.size(), f); fclose(f);
}

// Iterate '\n'-terminated lines (BufferedReader.readLine semantics; a trailing '\r' is kept out).
struct Lines {
  sv text; size_t pos = 0;
  explicit Lines(sv t) : text(t) {}
  bool next(sv& line) {
    if (pos >= text.size()) return false;
    size_t e = text.find('\n', pos);
    if (e == sv::npos) e = text.size();
    line = text.substr(pos, e - pos);
    if (!line.empty() && line.back() == '\r') line.remove_suffix(1);
    pos = e + 1;
    return true;
  }
};

inline std::vector<sv> split(sv s, char sep) {
  std::vector<sv> out; size_t p = 0;
  while (true) {
    size_t e = s.find(sep, p);
    if (e == sv::npos) { out.push_back(s.substr(p)); break; }
    out.push_back(s.substr(p, e - p)); p = e + 1;
  }
  return out;
}
// java.util.StringTokenizer: runs of separators collapse, no empty tokens.
inline std::vector<sv> tokenize(sv s, char sep) {
  std::vector<sv> out; size_t p = 0;
  while (p < s.size()) {
    while (p < s.size() && s[p] == sep) p++;
    if (p >= s.size()) break;
    size_t e = s.find(sep, p);
    if (e == sv::npos) e = s.size();
    out.push_back(s.substr(p, e - p)); p = e;
  }
  return out;
}
inline int to_int(sv s) {                               // Integer.parseInt
  if (s.empty()) throw Error("String  cannot be converted to an integer.");
  size_t i = 0; bool neg = false;
  if (s[0] == '-' || s[0] == '+') { neg = s[0] == '-'; i = 1; }
  if (i >= s.size()) throw Error("String " + std::string(s) + " cannot be converted to an integer.");
  long long v = 0;
  for (; i < s.size(); i++) {
    if (s[i] < '0' || s[i] > '9') throw Error("String " + std::string(s) + " cannot be converted to an integer.");
    v = v * 10 + (s[i] - '0');
    if (v > 2147483648LL) throw Error("String " + std::string(s) + " cannot be converted to an integer.");
  }
  return (int)(neg ? -v : v);
}
inline bool ends_with(sv s, sv suf) { return s.size() >= suf.size() && s.substr(s.size() - suf.size()) == suf; }
inline bool starts_with(sv s, sv pre) { return s.size() >= pre.size() && s.substr(0, pre.size()) == pre; }

// java.lang.Double.toString for finite non-negative doubles (JDK >= 19 shortest-repr digits;
// older JDKs can print one extra digit for a few values — documented in DESIGN.md).
inline std::string java_double(double v) {
  if (v == 0) return "0.0";
  char buf[64]; int prec;
  for (prec = 1; prec <= 17; prec++) {
    snprintf(buf, sizeof buf, "%.*e", prec - 1, v);
    if (strtod(buf, nullptr) == v) break;
  }
  std::string digits; int exp10 = 0;
  { char* e = strchr(buf, 'e'); exp10 = atoi(e + 1); for (char* p = buf; p < e; p++) if (*p >= '0' && *p <= '9') digits += *p; }
  while (digits.size() > 1 && digits.back() == '0') digits.pop_back();
  std::string out;
  if (v >= 1e-3 && v < 1e7) {
    if (exp10 >= 0) {
      std::string ip = digits.substr(0, std::min<size_t>(digits.size(), exp10 + 1));
      while ((int)ip.size() < exp10 + 1) ip += '0';
      std::string fp = digits.size() > (size_t)exp10 + 1 ? digits.substr(exp10 + 1) : "0";
      out = ip + "." + fp;
    } else {
      out = "0." + std::string(-exp10 - 1, '0') + digits;
    }
  } else {
    out = digits.substr(0, 1) + "." + (digits.size() > 1 ? digits.substr(1) : "0") + "E" + std::to_string(exp10);
  }
  return out;
}

// Counter(5).getZeroFilledCount — J/util-source/Counter.java:41-83 ("-A%05d" alignment ids)
inline std::string zero_filled(int counter, int digits) {
  char buf[32]; snprintf(buf, sizeof buf, "%0*d", digits, counter); return buf;
}

}  // namespace eqp
