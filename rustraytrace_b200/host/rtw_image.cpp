// rtw_image.cpp — RtwImage::new's search order and load (the_rest_of_your_life/rtw_image.rs:10-67).
//
// Search order kept from the reference: $RTW_IMAGES/<file>, <file>, then images/<file> in the
// current directory and up to six parents (:14-32).  An extra directory handed in by the caller
// is tried first (the Python package passes the repo's images/ so tests do not depend on cwd).
//
// Decoding: the reference uses image 0.25.9 -> zune-jpeg 0.5.12 (Cargo.lock).  Here a candidate file
// is decoded natively: baseline JPEG by jpeg_decode.cpp, binary PPM (P6) by the reader below; a
// decoded sibling "<stem>.ppm" is accepted as well.  Texel parity against zune-jpeg is unpinned
// (+-1 LSB from IDCT rounding).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "scene_builder.hpp"

namespace rrt_host {

static bool read_ppm_p6(const std::string& path, LoadedImage& out) {
  FILE* f = std::fopen(path.c_str(), "rb");
  if (!f) return false;
  auto next_int = [&](int& v) -> bool {
    int c = std::fgetc(f);
    for (;;) {
      while (c == ' ' || c == '\n' || c == '\r' || c == '\t') c = std::fgetc(f);
      if (c == '#') { while (c != '\n' && c != EOF) c = std::fgetc(f); continue; }
      break;
    }
    if (c < '0' || c > '9') return false;
    v = 0;
    while (c >= '0' && c <= '9') { v = v * 10 + (c - '0'); c = std::fgetc(f); }
    return true;  // consumed exactly one whitespace char after the number
  };
  bool ok = std::fgetc(f) == 'P' && std::fgetc(f) == '6';
  int w = 0, h = 0, mx = 0;
  ok = ok && next_int(w) && next_int(h) && next_int(mx) && mx == 255 && w > 0 && h > 0;
  if (ok) {
    out.rgb.resize((size_t)w * h * 3);
    ok = std::fread(out.rgb.data(), 1, out.rgb.size(), f) == out.rgb.size();
  }
  std::fclose(f);
  if (!ok) { out.rgb.clear(); return false; }
  out.width = w; out.height = h;
  return true;
}

static bool read_file(const std::string& path, std::vector<uint8_t>& bytes) {
  FILE* f = std::fopen(path.c_str(), "rb");
  if (!f) return false;
  std::fseek(f, 0, SEEK_END);
  long n = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  bool ok = n > 0;
  if (ok) { bytes.resize((size_t)n); ok = std::fread(bytes.data(), 1, bytes.size(), f) == bytes.size(); }
  std::fclose(f);
  return ok;
}

// RtwImage::load (:57-67) for one candidate path.
static bool load_candidate(const std::string& path, LoadedImage& out) {
  std::vector<uint8_t> bytes;
  if (read_file(path, bytes) && bytes.size() > 3) {
    if (bytes[0] == 0xFF && bytes[1] == 0xD8) {
      std::string err;
      if (decode_jpeg(bytes, out, err)) return true;
      std::fprintf(stderr, "ERROR: '%s': %s\n", path.c_str(), err.c_str());
    } else if (read_ppm_p6(path, out)) {
      return true;
    }
  }
  size_t dot = path.find_last_of('.');
  size_t slash = path.find_last_of('/');
  std::string stem = (dot != std::string::npos && (slash == std::string::npos || dot > slash))
                         ? path.substr(0, dot) : path;
  return stem + ".ppm" != path && read_ppm_p6(stem + ".ppm", out);
}

bool load_rtw_image(const std::string& filename, const char* extra_dir, LoadedImage& out) {
  out.width = out.height = 0; out.rgb.clear();
  if (extra_dir && *extra_dir) {
    if (load_candidate(std::string(extra_dir) + "/" + filename, out)) return true;
  }
  if (const char* dir = std::getenv("RTW_IMAGES")) {  // :14-19
    if (load_candidate(std::string(dir) + "/" + filename, out)) return true;
  }
  if (load_candidate(filename, out)) return true;  // :21-23
  std::string prefix;
  for (int i = 0; i < 7; ++i) {  // :25-32
    if (load_candidate(prefix + "images/" + filename, out)) return true;
    prefix += "../";
  }
  std::fprintf(stderr, "ERROR: Could not load image file '%s'.\n", filename.c_str());  // :34
  return false;
}

}  // namespace rrt_host

// C export (include/rrt_host.h): the loader by itself, for tests and tools.
extern "C" int rrth_load_image(const char* filename, const char* extra_dir, int32_t* width, int32_t* height,
                               uint8_t* rgb, uint64_t rgb_capacity) {
  rrt_host::LoadedImage im;
  if (!filename || !rrt_host::load_rtw_image(filename, extra_dir, im)) return -1;
  if (width) *width = im.width;
  if (height) *height = im.height;
  if (rgb) {
    if (rgb_capacity < im.rgb.size()) return -2;
    std::memcpy(rgb, im.rgb.data(), im.rgb.size());
  }
  return 0;
}
