// host_png.cpp — RGBA8 PNG output/input on zlib alone (libpng is not on the box image).
// Replaces save_image / load_image (include/libpng.h:53-63, src/libpng.cpp:15-107): 8 bits per
// channel, colour type 6 (RGBA), non-interlaced.  Pixel bytes are what matters for parity; the file
// bytes differ from libpng's (filter choice, deflate level).
#include <zlib.h>

#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
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

extern "C" int rt_png_write(const char* path, int w, int h, const uint8_t* rgba) {
  if (!path || !rgba || w <= 0 || h <= 0) { rt_set_error("rt_png_write: bad argument"); return RT_ERR_INVALID; }
  const size_t stride = (size_t)w * 4;
  std::vector<unsigned char> raw((stride + 1) * (size_t)h);
  for (int y = 0; y < h; y++) {  // filter type 1 (Sub): cheap and compresses smooth renders well
    unsigned char* row = raw.data() + (stride + 1) * (size_t)y;
    const uint8_t* src = rgba + stride * (size_t)y;
    row[0] = 1;
    for (size_t x = 0; x < stride; x++) row[1 + x] = (unsigned char)(src[x] - (x >= 4 ? src[x - 4] : 0));
  }
  uLongf bound = compressBound((uLong)raw.size());
  std::vector<unsigned char> z(bound);
  if (compress2(z.data(), &bound, raw.data(), (uLong)raw.size(), 6) != Z_OK) { rt_set_error("rt_png_write: deflate failed"); return RT_ERR_IO; }
  FILE* f = fopen(path, "wb");
  if (!f) { rt_set_error(std::string("rt_png_write: cannot open ") + path); return RT_ERR_IO; }
  static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
  fwrite(sig, 1, 8, f);
  std::vector<unsigned char> ihdr;
  put32(ihdr, (uint32_t)w); put32(ihdr, (uint32_t)h);
  ihdr.push_back(8); ihdr.push_back(6); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
  chunk(f, "IHDR", ihdr.data(), ihdr.size());
  chunk(f, "IDAT", z.data(), bound);
  chunk(f, "IEND", nullptr, 0);
  fclose(f);
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
