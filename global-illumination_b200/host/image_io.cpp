// host/image_io.cpp -- see image_io.h
#include "image_io.h"
#include <cstdio>
#include <cstring>
#include <vector>

namespace gi {
namespace {

uint32_t crc_table[256];
bool crc_ready = false;
uint32_t crc32(const uint8_t *p, size_t n, uint32_t c = 0) {
  if (!crc_ready) {
    for (uint32_t i = 0; i < 256; i++) { uint32_t v = i; for (int k = 0; k < 8; k++) v = (v & 1) ? 0xEDB88320u ^ (v >> 1) : v >> 1; crc_table[i] = v; }
    crc_ready = true;
  }
  c = ~c;
  for (size_t i = 0; i < n; i++) c = crc_table[(c ^ p[i]) & 255] ^ (c >> 8);
  return ~c;
}
void be32(std::vector<uint8_t> &v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }
void chunk(std::vector<uint8_t> &out, const char *type, const std::vector<uint8_t> &data) {
  be32(out, (uint32_t)data.size());
  std::vector<uint8_t> body(type, type + 4);
  body.insert(body.end(), data.begin(), data.end());
  out.insert(out.end(), body.begin(), body.end());
  be32(out, crc32(body.data(), body.size()));
}

// WritePNG (R2Image.cpp:1390-1455): the top PNG row is the image's last row
bool write_png(FILE *f, const uint8_t *rgb, int w, int h) {
  std::vector<uint8_t> raw;
  raw.reserve((size_t)h * (3 * (size_t)w + 1));
  for (int y = h - 1; y >= 0; y--) { raw.push_back(0); raw.insert(raw.end(), rgb + (size_t)y * w * 3, rgb + (size_t)(y + 1) * w * 3); }
  std::vector<uint8_t> z = {0x78, 0x01};
  uint32_t a = 1, b = 0;
  for (size_t i = 0; i < raw.size(); i++) { a = (a + raw[i]) % 65521; b = (b + a) % 65521; }
  for (size_t off = 0; off < raw.size() || off == 0; off += 65535) {
    const size_t n = raw.size() - off < 65535 ? raw.size() - off : 65535;
    z.push_back(off + n >= raw.size() ? 1 : 0);
    z.push_back(n & 255); z.push_back(n >> 8); z.push_back(~n & 255); z.push_back((~n >> 8) & 255);
    z.insert(z.end(), raw.begin() + off, raw.begin() + off + n);
    if (raw.empty()) break;
  }
  be32(z, (b << 16) | a);
  std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1a, '\n'}, ihdr;
  be32(ihdr, (uint32_t)w); be32(ihdr, (uint32_t)h);
  ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
  chunk(out, "IHDR", ihdr); chunk(out, "IDAT", z); chunk(out, "IEND", {});
  return fwrite(out.data(), 1, out.size(), f) == out.size();
}

// WritePPM(filename, ascii = 1) (R2Image.cpp:768-806)
bool write_ppm(FILE *f, const uint8_t *rgb, int w, int h) {
  fprintf(f, "P3\n%d %d\n255\n", w, h);
  for (int j = h - 1; j >= 0; j--) {
    for (int i = 0; i < w; i++) {
      const uint8_t *p = rgb + ((size_t)j * w + i) * 3;
      fprintf(f, "%-3d %-3d %-3d  ", p[0], p[1], p[2]);
      if (((i + 1) % 4) == 0) fprintf(f, "\n");
    }
    if ((w % 4) != 0) fprintf(f, "\n");
  }
  fprintf(f, "\n");
  return !ferror(f);
}

// WriteBMP (R2Image.cpp:463-540): 24-bit, bottom-up, BGR, rows padded to 4 bytes
bool write_bmp(FILE *f, const uint8_t *rgb, int w, int h) {
  const int rowsize = (3 * w + 3) & ~3;
  uint8_t hdr[54]; memset(hdr, 0, sizeof(hdr));
  auto le32 = [&](int o, uint32_t v) { hdr[o] = v; hdr[o + 1] = v >> 8; hdr[o + 2] = v >> 16; hdr[o + 3] = v >> 24; };
  hdr[0] = 'B'; hdr[1] = 'M'; le32(2, 54 + (uint32_t)rowsize * h); le32(10, 54); le32(14, 40); le32(18, w); le32(22, h);
  hdr[26] = 1; hdr[28] = 24; le32(34, (uint32_t)rowsize * h); le32(38, 2925); le32(42, 2925);
  if (fwrite(hdr, 1, 54, f) != 54) return false;
  std::vector<uint8_t> row(rowsize, 0);
  for (int j = 0; j < h; j++) {
    for (int i = 0; i < w; i++) { const uint8_t *p = rgb + ((size_t)j * w + i) * 3; row[3 * i] = p[2]; row[3 * i + 1] = p[1]; row[3 * i + 2] = p[0]; }
    if (fwrite(row.data(), 1, rowsize, f) != (size_t)rowsize) return false;
  }
  return true;
}

}  // namespace

bool write_image(const std::string &path, const uint8_t *rgb, int width, int height, std::string &err) {
  const size_t dot = path.rfind('.');
  if (dot == std::string::npos) { err = "Output file has no extension (e.g., .jpg)."; return false; }
  const std::string ext = path.substr(dot);
  const bool png = !strncmp(ext.c_str(), ".png", 4), ppm = !strncmp(ext.c_str(), ".ppm", 4), bmp = !strncmp(ext.c_str(), ".bmp", 4);
  if (!png && !ppm && !bmp) { err = "Unsupported image file extension " + ext + " (this build writes .png, .ppm, .bmp)"; return false; }
  FILE *f = fopen(path.c_str(), ppm ? "w" : "wb");
  if (!f) { err = "Unable to open image file: " + path; return false; }
  const bool ok = png ? write_png(f, rgb, width, height) : (ppm ? write_ppm(f, rgb, width, height) : write_bmp(f, rgb, width, height));
  fclose(f);
  if (!ok) err = "Unable to write image file: " + path;
  return ok;
}

}  // namespace gi
