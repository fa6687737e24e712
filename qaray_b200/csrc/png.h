// png.h — PNG <-> raw 8-bit pixels (zlib-backed); see png.cpp.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace qb {
// decode to tightly packed RGB8, row 0 first
bool png_decode_rgb8(const char *path, std::vector<uint8_t> *rgb, int *width, int *height, std::string *err);
// encode RGB8 (channels=3) or gray8 (channels=1)
bool png_encode(const char *path, const uint8_t *pixels, int width, int height, int channels);
}  // namespace qb
