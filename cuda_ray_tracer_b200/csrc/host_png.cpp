// host_png.cpp — RGBA8 PNG output/input on zlib alone (libpng is not on the box image).
// Replaces save_image / load_image (include/libpng.h:53-63, src/libpng.cpp:15-107): 8 bits per
// channel, colour type 6 (RGBA), non-interlaced.  Pixel bytes are what matters for parity; the file
// bytes differ from libpng's (filter choice, deflate level).
#include <zlib.h>

#include <algorithm>
#include <atomic>
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/rt_api.h"

void rt_set_error(const std::string& msg);

namespace {
void put32(std::vector<unsigned char>& v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }
void chunk(FILE* f, const char* type, const unsigned char* data, size_t n) {
  unsigned char len[4] = {(unsigned char)(n >> 24), (unsigned char)(n >> 16), (unsigned char)(n >> 8), (unsigned char)n};
  fwrite(len, 1, 4, f);
  fwrite(type, 1, 4, f);
  if (n) fwrite(data, 1, n, f);
  uLong c = crc32(0L, (const Bytef*)type, 4);
  if (n) c = crc32(c, data, (uInt)n);
  unsigned char cb[4] = {(unsigned char)(c >> 24), (unsigned char)(c >> 16), (unsigned char)(c >> 8), (unsigned char)c};
  fwrite(cb, 1, 4, f);
}
inline int paeth(int a, int b, int c) {
  int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}
}  // namespace

// Row-block-parallel encoder (the pigz construction).  The image is cut into blocks of rows; each
// worker filters its rows (Sub: every row depends only on itself) and deflates them as a RAW deflate stream that
// ends on a byte boundary (Z_SYNC_FLUSH; the last block ends with Z_FINISH).  Concatenated behind one zlib header
// and followed by the Adler-32 of the whole filtered image (adler32_combine of the per-block sums) they form one
// valid zlib stream = one IDAT chunk, whose CRC is likewise combined from per-block CRCs.  A block starts with an
// empty window, which costs a fraction of a percent of file size at >= 64 rows per block.
// RT_PNG_THREADS / RT_PNG_LEVEL override the thread count and the deflate level (default 4).
extern "C" int rt_png_write(const char* path, int w, int h, const uint8_t* rgba) {
  if (!path || !rgba || w <= 0 || h <= 0) { rt_set_error("rt_png_write: bad argument"); return RT_ERR_INVALID; }
  const size_t stride = (size_t)w * 4;
  unsigned threads = std::thread::hardware_concurrency();
  if (const char* e = getenv("RT_PNG_THREADS")) { const int v = atoi(e); if (v >= 1) threads = (unsigned)v; }
  if (threads < 1) threads = 1;
  int level = 4;
  if (const char* e = getenv("RT_PNG_LEVEL")) { const int v = atoi(e); if (v >= 0 && v <= 9) level = v; }
  const size_t min_rows = std::max<size_t>(1, ((size_t)1 << 18) / (stride + 1));  // >= 256 KiB of filtered bytes per block
  const size_t nblocks = std::max<size_t>(1, std::min<size_t>((size_t)h / min_rows, (size_t)threads * 4));
  struct Block { size_t y0 = 0, y1 = 0; std::vector<unsigned char> z; uLong adler = 1, crc = 0; size_t raw_len = 0; bool ok = false; };
  std::vector<Block> blocks(nblocks);
  for (size_t i = 0; i < nblocks; i++) { blocks[i].y0 = (size_t)h * i / nblocks; blocks[i].y1 = (size_t)h * (i + 1) / nblocks; }
  auto work = [&](size_t i) {
    Block& b = blocks[i];
    const size_t rows = b.y1 - b.y0;
    b.raw_len = (stride + 1) * rows;
    std::vector<unsigned char> raw(b.raw_len);
    for (size_t y = b.y0; y < b.y1; y++) {  // filter type 1 (Sub): cheap and compresses smooth renders well
      unsigned char* row = raw.data() + (stride + 1) * (y - b.y0);
      const uint8_t* src = rgba + stride * y;
      row[0] = 1;
      for (size_t x = 0; x < 4 && x < stride; x++) row[1 + x] = src[x];
      for (size_t x = 4; x < stride; x++) row[1 + x] = (unsigned char)(src[x] - src[x - 4]);
    }
    b.adler = adler32(adler32(0L, Z_NULL, 0), raw.data(), (uInt)raw.size());
    z_stream zs; memset(&zs, 0, sizeof(zs));
    if (deflateInit2(&zs, level, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) return;
    b.z.resize(deflateBound(&zs, (uLong)raw.size()) + 16);
    zs.next_in = raw.data(); zs.avail_in = (uInt)raw.size();
    zs.next_out = b.z.data(); zs.avail_out = (uInt)b.z.size();
    const bool last = i + 1 == nblocks;
    const int rc = deflate(&zs, last ? Z_FINISH : Z_SYNC_FLUSH);
    b.ok = (last ? rc == Z_STREAM_END : rc == Z_OK) && zs.avail_in == 0 && zs.avail_out > 0;
    b.z.resize(b.z.size() - zs.avail_out);
    deflateEnd(&zs);
    b.crc = crc32(crc32(0L, Z_NULL, 0), b.z.data(), (uInt)b.z.size());
  };
  if ((stride + 1) * (size_t)h / nblocks > 0x7FFFFFF0ull) { rt_set_error("rt_png_write: image too large"); return RT_ERR_INVALID; }
  {
    std::atomic<size_t> next{0};
    auto body = [&]() { for (size_t i; (i = next.fetch_add(1)) < nblocks;) work(i); };
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < std::min<size_t>(threads, nblocks); t++) pool.emplace_back(body);
    body();
    for (auto& t : pool) t.join();
  }
  size_t zlen = 2 + 4;
  for (auto& b : blocks) { if (!b.ok) { rt_set_error("rt_png_write: deflate failed"); return RT_ERR_IO; } zlen += b.z.size(); }
  if (zlen > 0x7FFFFFFFull) { rt_set_error("rt_png_write: compressed image exceeds one IDAT chunk"); return RT_ERR_INVALID; }
  const unsigned char zhdr[2] = {0x78, 0x5E};  // deflate, 32K window, FLEVEL 1, check bits: 0x785E % 31 == 0
  uLong adler = blocks[0].adler;
  for (size_t i = 1; i < nblocks; i++) adler = adler32_combine(adler, blocks[i].adler, (z_off_t)blocks[i].raw_len);
  const unsigned char ztail[4] = {(unsigned char)(adler >> 24), (unsigned char)(adler >> 16), (unsigned char)(adler >> 8), (unsigned char)adler};
  uLong crc = crc32(crc32(0L, (const Bytef*)"IDAT", 4), zhdr, 2);
  for (auto& b : blocks) crc = crc32_combine(crc, b.crc, (z_off_t)b.z.size());
  crc = crc32(crc, ztail, 4);

  FILE* f = fopen(path, "wb");
  if (!f) { rt_set_error(std::string("rt_png_write: cannot open ") + path); return RT_ERR_IO; }
  static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
  fwrite(sig, 1, 8, f);
  std::vector<unsigned char> ihdr;
  put32(ihdr, (uint32_t)w); put32(ihdr, (uint32_t)h);
  ihdr.push_back(8); ihdr.push_back(6); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
  chunk(f, "IHDR", ihdr.data(), ihdr.size());
  {
    std::vector<unsigned char> be; put32(be, (uint32_t)zlen);
    fwrite(be.data(), 1, 4, f); fwrite("IDAT", 1, 4, f); fwrite(zhdr, 1, 2, f);
    for (auto& b : blocks) fwrite(b.z.data(), 1, b.z.size(), f);
    fwrite(ztail, 1, 4, f);
    be.clear(); put32(be, (uint32_t)crc); fwrite(be.data(), 1, 4, f);
  }
  chunk(f, "IEND", nullptr, 0);
  const bool bad = ferror(f) != 0;
  if (fclose(f) != 0 || bad) { rt_set_error(std::string("rt_png_write: write failed: ") + path); return RT_ERR_IO; }
  return RT_OK;
}

extern "C" int rt_png_read(const char* path, int* w_out, int* h_out, uint8_t** rgba_out) {
  if (!path || !w_out || !h_out || !rgba_out) { rt_set_error("rt_png_read: bad argument"); return RT_ERR_INVALID; }
  FILE* f = fopen(path, "rb");
  if (!f) { rt_set_error(std::string("rt_png_read: cannot open ") + path); return RT_ERR_IO; }
  std::vector<unsigned char> file;
  unsigned char tmp[65536]; size_t k;
  while ((k = fread(tmp, 1, sizeof(tmp), f)) > 0) file.insert(file.end(), tmp, tmp + k);
  fclose(f);
  if (file.size() < 8 || memcmp(file.data(), "\x89PNG\r\n\x1a\n", 8) != 0) { rt_set_error("rt_png_read: not a PNG"); return RT_ERR_IO; }
  size_t p = 8; uint32_t W = 0, H = 0; int depth = 0, ctype = 0, interlace = 0;
  std::vector<unsigned char> idat;
  while (p + 12 <= file.size()) {
    uint32_t len = (file[p] << 24) | (file[p + 1] << 16) | (file[p + 2] << 8) | file[p + 3];
    const char* type = (const char*)&file[p + 4];
    const unsigned char* data = &file[p + 8];
    if (p + 12 + len > file.size()) break;
    if (!memcmp(type, "IHDR", 4)) {
      W = (data[0] << 24) | (data[1] << 16) | (data[2] << 8) | data[3]; H = (data[4] << 24) | (data[5] << 16) | (data[6] << 8) | data[7];
      depth = data[8]; ctype = data[9]; interlace = data[12];
    } else if (!memcmp(type, "IDAT", 4)) idat.insert(idat.end(), data, data + len);
    else if (!memcmp(type, "IEND", 4)) break;
    p += 12 + len;
  }
  if (depth != 8 || (ctype != 6 && ctype != 2) || interlace != 0 || W == 0 || H == 0) {
    rt_set_error("rt_png_read: only 8-bit non-interlaced RGB/RGBA is supported"); return RT_ERR_IO;
  }
  const int bpp = ctype == 6 ? 4 : 3;
  const size_t stride = (size_t)W * bpp;
  std::vector<unsigned char> raw((stride + 1) * H);
  uLongf rawlen = (uLongf)raw.size();
  if (uncompress(raw.data(), &rawlen, idat.data(), (uLong)idat.size()) != Z_OK || rawlen != raw.size()) {
    rt_set_error("rt_png_read: inflate failed"); return RT_ERR_IO;
  }
  std::vector<unsigned char> img(stride * H);
  for (uint32_t y = 0; y < H; y++) {
    const unsigned char* row = &raw[(stride + 1) * y];
    unsigned char* out = &img[stride * y];
    const unsigned char* up = y ? &img[stride * (y - 1)] : nullptr;
    for (size_t x = 0; x < stride; x++) {
      int a = x >= (size_t)bpp ? out[x - bpp] : 0, b = up ? up[x] : 0, c = (up && x >= (size_t)bpp) ? up[x - bpp] : 0, v = row[1 + x];
      switch (row[0]) {
        case 0: break; case 1: v += a; break; case 2: v += b; break; case 3: v += (a + b) / 2; break; case 4: v += paeth(a, b, c); break;
        default: rt_set_error("rt_png_read: bad filter"); return RT_ERR_IO;
      }
      out[x] = (unsigned char)v;
    }
  }
  uint8_t* res = (uint8_t*)malloc((size_t)W * H * 4);
  if (!res) { rt_set_error("rt_png_read: out of memory"); return RT_ERR_NOMEM; }
  for (size_t i = 0; i < (size_t)W * H; i++) {
    res[4 * i] = img[bpp * i]; res[4 * i + 1] = img[bpp * i + 1]; res[4 * i + 2] = img[bpp * i + 2];
    res[4 * i + 3] = bpp == 4 ? img[4 * i + 3] : 0xFF;  // missing alpha = opaque, libpng.cpp:50-51
  }
  *w_out = (int)W; *h_out = (int)H; *rgba_out = res;
  return RT_OK;
}

extern "C" void rt_free(void* p) { free(p); }
