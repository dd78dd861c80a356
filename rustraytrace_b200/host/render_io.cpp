// render_io.cpp — host mirror of src/render_io.rs:3-30 (write_ppm_from_accum): RGBA f32 sums ->
// text "P3".  The per-pixel arithmetic here is the reference's own, in f32, so the CLI can write
// an image straight from the accumulation buffer; the device-side twin is rrt_resolve_rgb8.
#include <cmath>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/rrt_host.h"

static inline unsigned quantise(float x, float scale) {
  float r = x * scale;
  if (!std::isfinite(r)) r = 0.0f;          // :16-18
  r = std::sqrt(r > 0.0f ? r : 0.0f);       // :20-22  r.max(0.0).sqrt()
  if (r < 0.0f) r = 0.0f;
  if (r > 0.999f) r = 0.999f;               // :24-26  clamp(0, 0.999) * 256 as u8
  return (unsigned)(r * 256.0f);
}

static int write_text(const char* path, int width, int height, const std::vector<unsigned char>& rgb) {
  FILE* f = path ? std::fopen(path, "w") : stdout;
  if (!f) return -1;
  std::string buf;
  buf.reserve((size_t)width * height * 12 + 32);
  buf += "P3\n" + std::to_string(width) + " " + std::to_string(height) + "\n255\n";  // :6
  for (size_t i = 0; i < (size_t)width * height; ++i) {
    buf += std::to_string(rgb[3 * i]); buf += ' ';
    buf += std::to_string(rgb[3 * i + 1]); buf += ' ';
    buf += std::to_string(rgb[3 * i + 2]); buf += '\n';
  }
  size_t n = std::fwrite(buf.data(), 1, buf.size(), f);
  if (path) std::fclose(f); else std::fflush(f);
  return n == buf.size() ? 0 : -1;
}

extern "C" int rrth_write_ppm_from_accum(const char* path, int width, int height, const float* accum,
                                         uint32_t samples_per_pixel) {
  if (!accum || width <= 0 || height <= 0) return -1;
  float scale = samples_per_pixel > 0 ? 1.0f / (float)samples_per_pixel : 0.0f;  // :8
  std::vector<unsigned char> rgb((size_t)width * height * 3);
  for (size_t i = 0; i < (size_t)width * height; ++i)
    for (int c = 0; c < 3; ++c) rgb[3 * i + c] = (unsigned char)quantise(accum[4 * i + c], scale);
  return write_text(path, width, height, rgb);
}

extern "C" int rrth_write_ppm_rgb8(const char* path, int width, int height, const uint8_t* rgb) {
  if (!rgb || width <= 0 || height <= 0) return -1;
  std::vector<unsigned char> v(rgb, rgb + (size_t)width * height * 3);
  return write_text(path, width, height, v);
}
