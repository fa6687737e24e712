// png.cpp — PNG decode to RGB8 / encode from RGB8 or gray8, on zlib. Stands in for the reference's use of
// lodepng (src/textures/texture.cpp:75-86 decode as LCT_RGB 8-bit; src/fb/framebuffer.cpp:109-143 encode).
// Non-interlaced images of colour types 0, 2, 3, 4, 6 at 8 or 16 bits (palette also at 1/2/4 bits).
#include "png.h"

#include <zlib.h>

#include <cstdio>
#include <cstring>

namespace qb {

static uint32_t be32(const uint8_t *p) { return ((uint32_t) p[0] << 24) | ((uint32_t) p[1] << 16) | ((uint32_t) p[2] << 8) | p[3]; }

static int paeth(int a, int b, int c)
{
  int p = a + b - c, pa = p > a ? p - a : a - p, pb = p > b ? p - b : b - p, pc = p > c ? p - c : c - p;
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

bool png_decode_rgb8(const char *path, std::vector<uint8_t> *rgb, int *width, int *height, std::string *err)
{
  FILE *fp = fopen(path, "rb");
  if (!fp) { if (err) *err = std::string("cannot open ") + path; return false; }
  std::vector<uint8_t> buf;
  uint8_t tmp[1 << 16];
  size_t n;
  while ((n = fread(tmp, 1, sizeof tmp, fp)) > 0) buf.insert(buf.end(), tmp, tmp + n);
  fclose(fp);
  static const uint8_t sig[8] = {137, 80, 78, 71, 13, 10, 26, 10};
  if (buf.size() < 33 || memcmp(buf.data(), sig, 8) != 0) { if (err) *err = "not a PNG"; return false; }
  uint32_t W = 0, H = 0;
  int depth = 0, ctype = 0, interlace = 0;
  std::vector<uint8_t> idat, plte;
  size_t pos = 8;
  while (pos + 12 <= buf.size()) {
    const uint32_t len = be32(&buf[pos]);
    const char *tag = (const char *) &buf[pos + 4];
    const uint8_t *data = &buf[pos + 8];
    if (pos + 12 + (size_t) len > buf.size()) { if (err) *err = "truncated PNG chunk"; return false; }
    if (!memcmp(tag, "IHDR", 4)) {
      W = be32(data); H = be32(data + 4); depth = data[8]; ctype = data[9]; interlace = data[12];
    } else if (!memcmp(tag, "PLTE", 4)) {
      plte.assign(data, data + len);
    } else if (!memcmp(tag, "IDAT", 4)) {
      idat.insert(idat.end(), data, data + len);
    } else if (!memcmp(tag, "IEND", 4)) {
      break;
    }
    pos += 12 + (size_t) len;
  }
  if (W == 0 || H == 0 || interlace != 0) { if (err) *err = "unsupported PNG (empty or interlaced)"; return false; }
  int channels = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
  if (!channels || !(depth == 8 || depth == 16 || (ctype == 3 && (depth == 1 || depth == 2 || depth == 4)) ||
                     (ctype == 0 && (depth == 1 || depth == 2 || depth == 4)))) {
    if (err) *err = "unsupported PNG colour type / depth";
    return false;
  }
  const size_t bpp_bits = (size_t) channels * depth;
  const size_t stride = (W * bpp_bits + 7) / 8;
  const size_t bpp = bpp_bits >= 8 ? bpp_bits / 8 : 1;
  std::vector<uint8_t> raw((stride + 1) * (size_t) H);
  uLongf rawlen = (uLongf) raw.size();
  if (uncompress(raw.data(), &rawlen, idat.data(), (uLong) idat.size()) != Z_OK || rawlen != raw.size()) {
    if (err) *err = "PNG inflate failed";
    return false;
  }
  std::vector<uint8_t> img(stride * (size_t) H);
  for (size_t y = 0; y < H; ++y) {
    const uint8_t ft = raw[y * (stride + 1)];
    const uint8_t *src = &raw[y * (stride + 1) + 1];
    uint8_t *dst = &img[y * stride];
    const uint8_t *up = y ? &img[(y - 1) * stride] : nullptr;
    for (size_t x = 0; x < stride; ++x) {
      const int a = x >= bpp ? dst[x - bpp] : 0, b = up ? up[x] : 0, c = (up && x >= bpp) ? up[x - bpp] : 0;
      int v = src[x];
      switch (ft) {
        case 0: break;
        case 1: v += a; break;
        case 2: v += b; break;
        case 3: v += (a + b) >> 1; break;
        case 4: v += paeth(a, b, c); break;
        default: if (err) *err = "bad PNG filter"; return false;
      }
      dst[x] = (uint8_t) v;
    }
  }
  rgb->assign((size_t) W * H * 3, 0);
  for (size_t y = 0; y < H; ++y) {
    const uint8_t *row = &img[y * stride];
    for (size_t x = 0; x < W; ++x) {
      uint8_t *o = &(*rgb)[(y * W + x) * 3];
      if (depth < 8) {
        const size_t bit = x * depth;
        const int v = (row[bit >> 3] >> (8 - depth - (bit & 7))) & ((1 << depth) - 1);
        if (ctype == 3) {
          if ((size_t) v * 3 + 2 < plte.size()) { o[0] = plte[v * 3]; o[1] = plte[v * 3 + 1]; o[2] = plte[v * 3 + 2]; }
        } else {
          o[0] = o[1] = o[2] = (uint8_t) ((v * 255) / ((1 << depth) - 1));
        }
        continue;
      }
      const size_t bs = depth / 8;  // 16-bit samples keep their most significant byte
      const uint8_t *px = row + x * channels * bs;
      switch (ctype) {
        case 0: case 4: o[0] = o[1] = o[2] = px[0]; break;
        case 2: case 6: o[0] = px[0]; o[1] = px[bs]; o[2] = px[2 * bs]; break;
        case 3:
          if ((size_t) px[0] * 3 + 2 < plte.size()) { o[0] = plte[px[0] * 3]; o[1] = plte[px[0] * 3 + 1]; o[2] = plte[px[0] * 3 + 2]; }
          break;
      }
    }
  }
  *width = (int) W;
  *height = (int) H;
  return true;
}

static void put_chunk(std::vector<uint8_t> *out, const char *tag, const uint8_t *data, size_t len)
{
  const uint8_t l[4] = {(uint8_t) (len >> 24), (uint8_t) (len >> 16), (uint8_t) (len >> 8), (uint8_t) len};
  out->insert(out->end(), l, l + 4);
  const size_t start = out->size();
  out->insert(out->end(), tag, tag + 4);
  if (len) out->insert(out->end(), data, data + len);
  const uint32_t crc = (uint32_t) crc32(0L, out->data() + start, (uInt) (len + 4));
  const uint8_t c[4] = {(uint8_t) (crc >> 24), (uint8_t) (crc >> 16), (uint8_t) (crc >> 8), (uint8_t) crc};
  out->insert(out->end(), c, c + 4);
}

bool png_encode(const char *path, const uint8_t *pixels, int width, int height, int channels)
{
  if (channels != 1 && channels != 3) return false;
  const size_t stride = (size_t) width * channels;
  std::vector<uint8_t> raw((stride + 1) * (size_t) height);
  for (int y = 0; y < height; ++y) {
    raw[y * (stride + 1)] = 0;
    memcpy(&raw[y * (stride + 1) + 1], pixels + (size_t) y * stride, stride);
  }
  uLongf clen = compressBound((uLong) raw.size());
  std::vector<uint8_t> comp(clen);
  if (compress2(comp.data(), &clen, raw.data(), (uLong) raw.size(), 6) != Z_OK) return false;
  std::vector<uint8_t> out = {137, 80, 78, 71, 13, 10, 26, 10};
  uint8_t ihdr[13] = {(uint8_t) (width >> 24), (uint8_t) (width >> 16), (uint8_t) (width >> 8), (uint8_t) width,
                      (uint8_t) (height >> 24), (uint8_t) (height >> 16), (uint8_t) (height >> 8), (uint8_t) height,
                      8, (uint8_t) (channels == 3 ? 2 : 0), 0, 0, 0};
  put_chunk(&out, "IHDR", ihdr, 13);
  put_chunk(&out, "IDAT", comp.data(), clen);
  put_chunk(&out, "IEND", nullptr, 0);
  FILE *fp = fopen(path, "wb");
  if (!fp) return false;
  const bool ok = fwrite(out.data(), 1, out.size(), fp) == out.size();
  fclose(fp);
  return ok;
}

}  // namespace qb
