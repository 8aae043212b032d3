// Host-side helpers of the compute-counts drop-in: gz-aware file reading and the reference's number formats.
#pragma once
#include <zlib.h>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>

namespace eqphost {

// Reads a plain or gzip file completely (UtilLib.getBufferedReader: ".gz" suffix => GZIPInputStream, UtilLib.java:322-383;
// gzread also passes plain files through).
inline std::string read_maybe_gz(const std::string& path) {
  gzFile f = gzopen(path.c_str(), "rb");
  if (!f) throw std::runtime_error("File " + path + " not found.");
  gzbuffer(f, 1 << 20);
  std::string out; std::string buf(1 << 22, '\0'); int n;
  while ((n = gzread(f, &buf[0], (unsigned)buf.size())) > 0) out.append(buf.data(), (size_t)n);
  bool bad = n < 0; gzclose(f);
  if (bad) throw std::runtime_error("Problem reading " + path);
  return out;
}

// Same, piecewise: fn(buf, n) is called with consecutive pieces of the inflated text; it returns false to stop.
template <class F>
inline void stream_maybe_gz(const std::string& path, F&& fn, size_t piece = (size_t)8 << 20) {
  gzFile f = gzopen(path.c_str(), "rb");
  if (!f) throw std::runtime_error("File " + path + " not found.");
  gzbuffer(f, 1 << 20);
  std::string buf(piece, '\0'); int n;
  while ((n = gzread(f, &buf[0], (unsigned)buf.size())) > 0) if (!fn(buf.data(), (size_t)n)) break;
  bool bad = n < 0; gzclose(f);
  if (bad) throw std::runtime_error("Problem reading " + path);
}

// java.lang.Double.toString: shortest digits that round-trip (JDK >= 19 exactly; older JDKs print the same digits except in
// rare cases where they emit one digit more), decimal notation for 1e-3 <= |d| < 1e7, else computerised scientific notation.
inline std::string java_double(double d) {
  if (std::isnan(d)) return "NaN";
  if (std::isinf(d)) return d > 0 ? "Infinity" : "-Infinity";
  if (d == 0) return std::signbit(d) ? "-0.0" : "0.0";
  char buf[64]; int prec = 1;
  for (; prec <= 17; prec++) { snprintf(buf, sizeof buf, "%.*e", prec - 1, d); if (strtod(buf, nullptr) == d) break; }
  std::string s = buf; bool neg = s[0] == '-'; if (neg) s.erase(0, 1);
  size_t e = s.find('e'); int exp10 = atoi(s.c_str() + e + 1); std::string digits = s.substr(0, e);
  digits.erase(digits.find('.') == std::string::npos ? digits.size() : digits.find('.'), 1);   // d.ddd -> dddd
  while (digits.size() > 1 && digits.back() == '0') digits.pop_back();
  std::string out;
  double a = std::fabs(d);
  if (a >= 1e-3 && a < 1e7) {
    if (exp10 >= 0) {
      std::string ip = digits.substr(0, std::min(digits.size(), (size_t)exp10 + 1));
      while ((int)ip.size() < exp10 + 1) ip += '0';
      std::string fp = digits.size() > (size_t)exp10 + 1 ? digits.substr(exp10 + 1) : "0";
      out = ip + "." + fp;
    } else out = "0." + std::string((size_t)(-exp10 - 1), '0') + digits;
  } else {
    out = digits.substr(0, 1) + "." + (digits.size() > 1 ? digits.substr(1) : "0") + "E" + std::to_string(exp10);
  }
  return neg ? "-" + out : out;
}

}  // namespace eqphost
