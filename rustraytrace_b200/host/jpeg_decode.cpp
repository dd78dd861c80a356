// jpeg_decode.cpp — baseline (SOF0, 8-bit, Huffman) JPEG decoder for image textures.
//
// Stands in for `image::open(path).to_rgb8()` (the_rest_of_your_life/rtw_image.rs:57-67), which in the
// reference resolves to image 0.25.9 -> zune-jpeg 0.5.12 (Cargo.lock).  That crate is not in the
// reference tree, so this is a from-scratch restatement of the published baseline algorithm
// (ITU-T T.81): Huffman-coded 8x8 DCT blocks, dequantisation, inverse DCT, YCbCr -> RGB (JFIF).
// Scope: what the reference's only asset needs and a little more — 1 or 3 components, sampling
// factors 1 or 2 (chroma replicated, no smoothing filter), restart intervals.  Progressive,
// arithmetic-coded and 12-bit files are rejected.  Texel parity with zune-jpeg is unpinned
// (IDCT rounding may differ by +-1 LSB); tests compare against libjpeg-turbo via PIL.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "scene_builder.hpp"

namespace rrt_host {
namespace {

struct Huff {
  uint8_t bits[17] = {0};
  uint8_t vals[256] = {0};
  int mincode[17], maxcode[18], valptr[17];
  bool present = false;
  void build() {
    int code = 0, k = 0;
    for (int l = 1; l <= 16; ++l) {
      valptr[l] = k;
      mincode[l] = code;
      code += bits[l];
      k += bits[l];
      maxcode[l] = bits[l] ? code - 1 : -1;
      code <<= 1;
    }
    maxcode[17] = 0x7fffffff;
    present = true;
  }
};

struct Component {
  int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0;
  int pred = 0;
  int bw = 0, bh = 0;  // blocks per row / column (padded to whole MCUs)
  std::vector<uint8_t> plane;  // bw*8 x bh*8 samples
};

struct BitReader {
  const uint8_t* p;
  const uint8_t* end;
  uint32_t acc = 0;
  int n = 0;
  bool hit_marker = false;
  int bit() {
    if (n == 0) {
      uint8_t b = 0;
      if (p < end && !hit_marker) {
        b = *p++;
        if (b == 0xFF) {
          uint8_t b2 = p < end ? *p : 0;
          if (b2 == 0) ++p;                       // stuffed zero
          else { hit_marker = true; --p; b = 0; }  // a marker: feed zeros, leave it for the caller
        }
      }
      acc = b; n = 8;
    }
    --n;
    return (acc >> n) & 1;
  }
  int receive(int s) { int v = 0; for (int i = 0; i < s; ++i) v = (v << 1) | bit(); return v; }
  void reset() { n = 0; acc = 0; hit_marker = false; }
};

inline int extend(int v, int t) { return (t && v < (1 << (t - 1))) ? v - (1 << t) + 1 : v; }

int decode_symbol(BitReader& br, const Huff& h) {
  int code = 0;
  for (int l = 1; l <= 16; ++l) {
    code = (code << 1) | br.bit();
    if (h.maxcode[l] >= 0 && code <= h.maxcode[l] && code >= h.mincode[l]) return h.vals[h.valptr[l] + code - h.mincode[l]];
  }
  return -1;
}

const int kZigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48,
                         41, 34, 27, 20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22,
                         15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// separable inverse DCT in double precision, then level shift, round, clamp
void idct8x8(const int* coef, uint8_t* out, int stride) {
  static double c[8][8];
  static bool init = false;
  if (!init) {
    for (int x = 0; x < 8; ++x)
      for (int u = 0; u < 8; ++u) c[x][u] = (u == 0 ? std::sqrt(0.125) : 0.5) * std::cos((2 * x + 1) * u * M_PI / 16.0);
    init = true;
  }
  double tmp[64];
  for (int y = 0; y < 8; ++y)      // rows of coefficients -> horizontal pass
    for (int x = 0; x < 8; ++x) {
      double s = 0;
      for (int u = 0; u < 8; ++u) s += c[x][u] * coef[y * 8 + u];
      tmp[y * 8 + x] = s;
    }
  for (int x = 0; x < 8; ++x)
    for (int y = 0; y < 8; ++y) {
      double s = 0;
      for (int v = 0; v < 8; ++v) s += c[y][v] * tmp[v * 8 + x];
      int val = (int)std::floor(s + 128.5);
      out[y * stride + x] = (uint8_t)(val < 0 ? 0 : (val > 255 ? 255 : val));
    }
}

inline int rd16(const uint8_t* p) { return (p[0] << 8) | p[1]; }

}  // namespace

bool decode_jpeg(const std::vector<uint8_t>& file, LoadedImage& out, std::string& err) {
  const uint8_t* d = file.data();
  const size_t n = file.size();
  if (n < 4 || d[0] != 0xFF || d[1] != 0xD8) { err = "not a JPEG (no SOI)"; return false; }
  uint16_t qt[4][64] = {{0}};
  Huff dc[4], ac[4];
  std::vector<Component> comp;
  int width = 0, height = 0, restart = 0, hmax = 1, vmax = 1;
  size_t i = 2;
  while (i + 4 <= n) {
    if (d[i] != 0xFF) { err = "marker expected"; return false; }
    while (i < n && d[i] == 0xFF) ++i;  // fill bytes
    const uint8_t m = d[i++];
    if (m == 0xD9) break;
    if (m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;
    if (i + 2 > n) break;
    const int L = rd16(d + i);
    if (L < 2 || i + L > n) { err = "segment runs past the end of the file"; return false; }
    const uint8_t* s = d + i + 2;
    const uint8_t* e = d + i + L;
    if (m == 0xDB) {  // DQT
      while (s < e) {
        int pq = s[0] >> 4, tq = s[0] & 15; ++s;
        if (tq > 3 || s + (pq ? 128 : 64) > e) { err = "bad DQT"; return false; }
        for (int k = 0; k < 64; ++k) { qt[tq][kZigzag[k]] = pq ? (uint16_t)rd16(s) : *s; s += pq ? 2 : 1; }
      }
    } else if (m == 0xC4) {  // DHT
      while (s < e) {
        int tc = s[0] >> 4, th = s[0] & 15; ++s;
        if (th > 3 || tc > 1 || s + 16 > e) { err = "bad DHT"; return false; }
        Huff& h = tc ? ac[th] : dc[th];
        int total = 0;
        for (int l = 1; l <= 16; ++l) { h.bits[l] = s[l - 1]; total += h.bits[l]; }
        s += 16;
        if (total > 256 || s + total > e) { err = "bad DHT"; return false; }
        std::memcpy(h.vals, s, total);
        s += total;
        h.build();
      }
    } else if (m == 0xC0 || m == 0xC1) {  // SOF0 / SOF1 (extended sequential, Huffman)
      if (L < 8) { err = "bad SOF"; return false; }
      if (s[0] != 8) { err = "only 8-bit JPEG is supported"; return false; }
      height = rd16(s + 1); width = rd16(s + 3);
      int nc = s[5];
      if ((nc != 1 && nc != 3) || width < 1 || height < 1) { err = "unsupported component count"; return false; }
      if (L < 8 + 3 * nc) { err = "bad SOF"; return false; }   // the component table must fit its segment
      if ((long long)width * height > (1LL << 28)) { err = "image too large"; return false; }
      comp.resize(nc);
      for (int k = 0; k < nc; ++k) {
        comp[k].id = s[6 + 3 * k]; comp[k].h = s[7 + 3 * k] >> 4; comp[k].v = s[7 + 3 * k] & 15; comp[k].tq = s[8 + 3 * k] & 3;
        if (comp[k].h < 1 || comp[k].h > 2 || comp[k].v < 1 || comp[k].v > 2) { err = "unsupported sampling factor"; return false; }
        hmax = std::max(hmax, comp[k].h); vmax = std::max(vmax, comp[k].v);
      }
    } else if (m == 0xC2 || (m >= 0xC5 && m <= 0xCF && m != 0xC8 && m != 0xCC)) {
      err = "progressive / lossless / arithmetic JPEG is not supported (baseline only)"; return false;
    } else if (m == 0xDD) {
      if (L < 4) { err = "bad DRI"; return false; }
      restart = rd16(s);
    } else if (m == 0xDA) {  // SOS: one interleaved scan holding every component
      if (comp.empty()) { err = "SOS before SOF"; return false; }
      if (L < 3) { err = "bad SOS"; return false; }
      int ns = s[0];
      if (ns != (int)comp.size()) { err = "non-interleaved scans are not supported"; return false; }
      if (L < 6 + 2 * ns) { err = "bad SOS"; return false; }
      for (int k = 0; k < ns; ++k) {
        int cid = s[1 + 2 * k];
        const int td = s[2 + 2 * k] >> 4, ta = s[2 + 2 * k] & 15;
        if (td > 3 || ta > 3) { err = "bad SOS"; return false; }   // dc[4], ac[4]
        for (auto& c : comp) if (c.id == cid) { c.td = td; c.ta = ta; }
      }
      i += L;
      const int mcux = (width + 8 * hmax - 1) / (8 * hmax), mcuy = (height + 8 * vmax - 1) / (8 * vmax);
      for (auto& c : comp) {
        c.bw = mcux * c.h; c.bh = mcuy * c.v;
        c.plane.assign((size_t)c.bw * 8 * c.bh * 8, 0);
        if (!dc[c.td].present || !ac[c.ta].present) { err = "scan uses an undefined Huffman table"; return false; }
      }
      BitReader br; br.p = d + i; br.end = d + n;
      int until_restart = restart;
      for (int my = 0; my < mcuy; ++my)
        for (int mx = 0; mx < mcux; ++mx) {
          if (restart && until_restart == 0) {  // RSTn: byte-align, skip the marker, reset predictors
            br.reset();
            while (br.p + 1 < br.end && !(br.p[0] == 0xFF && br.p[1] >= 0xD0 && br.p[1] <= 0xD7)) ++br.p;
            if (br.p + 1 < br.end) br.p += 2;
            for (auto& c : comp) c.pred = 0;
            until_restart = restart;
          }
          for (auto& c : comp)
            for (int by = 0; by < c.v; ++by)
              for (int bx = 0; bx < c.h; ++bx) {
                int coef[64] = {0};
                int t = decode_symbol(br, dc[c.td]);
                if (t < 0 || t > 11) { err = "corrupt DC code"; return false; }
                c.pred += extend(br.receive(t), t);
                coef[0] = c.pred * qt[c.tq][0];
                for (int k = 1; k < 64;) {
                  int rs = decode_symbol(br, ac[c.ta]);
                  if (rs < 0) { err = "corrupt AC code"; return false; }
                  int r = rs >> 4, sz = rs & 15;
                  if (sz == 0) { if (r == 15) { k += 16; continue; } break; }
                  k += r;
                  if (k > 63) { err = "AC run past the block"; return false; }
                  coef[kZigzag[k]] = extend(br.receive(sz), sz) * qt[c.tq][kZigzag[k]];
                  ++k;
                }
                const int X = (mx * c.h + bx) * 8, Y = (my * c.v + by) * 8;
                idct8x8(coef, c.plane.data() + (size_t)Y * c.bw * 8 + X, c.bw * 8);
              }
          if (restart) --until_restart;
        }
      // ---- colour: replicate subsampled planes, YCbCr -> RGB (JFIF), grey -> RGB
      out.width = width; out.height = height;
      out.rgb.resize((size_t)width * height * 3);
      for (int y = 0; y < height; ++y)
        for (int x = 0; x < width; ++x) {
          auto sample = [&](const Component& c) {
            int sx = x * c.h / hmax, sy = y * c.v / vmax;
            return (int)c.plane[(size_t)sy * c.bw * 8 + sx];
          };
          uint8_t* px = &out.rgb[((size_t)y * width + x) * 3];
          if (comp.size() == 1) { px[0] = px[1] = px[2] = (uint8_t)sample(comp[0]); continue; }
          double Y = sample(comp[0]), cb = sample(comp[1]) - 128.0, cr = sample(comp[2]) - 128.0;
          int r = (int)std::floor(Y + 1.402 * cr + 0.5);
          int g = (int)std::floor(Y - 0.344136 * cb - 0.714136 * cr + 0.5);
          int b = (int)std::floor(Y + 1.772 * cb + 0.5);
          px[0] = (uint8_t)(r < 0 ? 0 : (r > 255 ? 255 : r));
          px[1] = (uint8_t)(g < 0 ? 0 : (g > 255 ? 255 : g));
          px[2] = (uint8_t)(b < 0 ? 0 : (b > 255 ? 255 : b));
        }
      return true;
    }
    i += L;
  }
  err = "no scan found";
  return false;
}

}  // namespace rrt_host
