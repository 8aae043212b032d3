// pgzip — bench / test tooling: the concatenation of the input files as ONE gzip member, compressed by several threads the way pigz
// does it (raw deflate per 8 MB piece, sync-flushed, concatenated; CRC-32 combined).  Any gzip reader takes the result.
//   pgzip <out.gz> <level> <threads> <in1> [in2 ...]
#include <zlib.h>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

struct Piece { int file; uint64_t off, n; std::vector<unsigned char> z; uLong crc; };

int main(int argc, char** argv) {
  if (argc < 5) { fprintf(stderr, "usage: pgzip <out.gz> <level> <threads> <in1> [in2 ...]\n"); return 2; }
  const int level = atoi(argv[2]); int threads = atoi(argv[3]); if (threads < 1) threads = 1;
  const uint64_t PIECE = 8u << 20;
  std::vector<std::string> in(argv + 4, argv + argc); std::vector<Piece> pieces;
  for (size_t f = 0; f < in.size(); f++) {
    FILE* fp = fopen(in[f].c_str(), "rb"); if (!fp) { fprintf(stderr, "pgzip: cannot open %s\n", in[f].c_str()); return 1; }
    fseeko(fp, 0, SEEK_END); uint64_t size = (uint64_t)ftello(fp); fclose(fp);
    for (uint64_t off = 0; off < size; off += PIECE) pieces.push_back(Piece{(int)f, off, std::min(PIECE, size - off), {}, 0});
  }
  std::atomic<size_t> next{0}; std::atomic<int> failed{0};
  auto work = [&]() {
    std::vector<unsigned char> buf;
    for (size_t k; (k = next.fetch_add(1)) < pieces.size();) {
      Piece& p = pieces[k]; buf.resize(p.n);
      FILE* fp = fopen(in[p.file].c_str(), "rb"); if (!fp || fseeko(fp, (off_t)p.off, SEEK_SET) || fread(buf.data(), 1, p.n, fp) != p.n) { failed = 1; if (fp) fclose(fp); return; }
      fclose(fp);
      z_stream s; memset(&s, 0, sizeof s);
      if (deflateInit2(&s, level, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) { failed = 1; return; }
      p.z.resize(deflateBound(&s, p.n) + 16);
      s.next_in = buf.data(); s.avail_in = (uInt)p.n; s.next_out = p.z.data(); s.avail_out = (uInt)p.z.size();
      if (deflate(&s, Z_SYNC_FLUSH) != Z_OK || s.avail_in) { failed = 1; deflateEnd(&s); return; }
      p.z.resize(p.z.size() - s.avail_out); deflateEnd(&s);
      p.crc = crc32(crc32(0L, Z_NULL, 0), buf.data(), (uInt)p.n);
    }
  };
  std::vector<std::thread> pool; for (int t = 0; t < threads; t++) pool.emplace_back(work);
  for (auto& t : pool) t.join();
  if (failed) { fprintf(stderr, "pgzip: read or deflate failed\n"); return 1; }
  FILE* out = fopen(argv[1], "wb"); if (!out) { fprintf(stderr, "pgzip: cannot write %s\n", argv[1]); return 1; }
  const unsigned char hdr[10] = {0x1f, 0x8b, 8, 0, 0, 0, 0, 0, 0, 3}; fwrite(hdr, 1, 10, out);
  uLong crc = crc32(0L, Z_NULL, 0); uint64_t total = 0;
  for (Piece& p : pieces) { fwrite(p.z.data(), 1, p.z.size(), out); crc = crc32_combine(crc, p.crc, (z_off_t)p.n); total += p.n; }
  const unsigned char fin[2] = {0x03, 0x00}; fwrite(fin, 1, 2, out);                      // final empty fixed block
  unsigned char tr[8]; for (int i = 0; i < 4; i++) { tr[i] = (unsigned char)(crc >> (8 * i)); tr[4 + i] = (unsigned char)(total >> (8 * i)); }
  fwrite(tr, 1, 8, out);
  if (fclose(out)) { fprintf(stderr, "pgzip: write failed\n"); return 1; }
  return 0;
}
