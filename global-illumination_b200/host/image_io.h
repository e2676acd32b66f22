// host/image_io.h -- writes the 8-bit RGB frame gi_render returns (row 0 = bottom row, like R2Image) to a file chosen
// by extension, as R2Image::Write does (src/R2Shapes/R2Image.cpp:317-340).
#pragma once
#include <cstdint>
#include <string>
namespace gi {
// .png (8-bit RGB, stored deflate blocks: no zlib dependency), .ppm (ASCII P3, the reference's layout), .bmp (24-bit).
// Other extensions the reference accepts (.jpg .tif .raw .pgm) return false with a message in `err`.
bool write_image(const std::string &path, const uint8_t *rgb, int width, int height, std::string &err);
}  // namespace gi
