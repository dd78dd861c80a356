// warp_sim.cpp — CPU model of the warp-level cost of the render kernel's BVH traversal.
//
// The render kernel (rustraytrace_b200/csrc/render_kernel.cuh + intersect.cuh) is bound by how its 32
// lanes re-converge, not by the instructions of one lane: a warp pays, per path segment, for the front
// run, for the LONGEST descent of any lane between two leaf rounds, and for every leaf round
// (profiles/README.md r01i/r01j).  This tool replays that schedule on the host: it traces paths through
// a book scene with the flattener's own tree (flatten.cpp is compiled into it), groups them the way
// the kernel does (8x4-pixel tile per warp, lane = pixel, a finished lane starts its next sample at
// once), and charges every warp iteration with the instruction counts read off the ncu SASS profile.
// A different tree (other RRT_* constants of flatten.cpp, set with -D when building this file) is
// scored on the SAME rays, so tree shapes can be compared without a GPU:
//
//   make warp_sim [SIMFLAGS="-DRRT_FRONT_MAX=8 ..."]      (g++ on this file + rustraytrace_b200/csrc/flatten.cpp, -> build/warp_sim)
//   build/warp_sim <book 1|2|3|4> <scene_no|-1> <image_width> [tiles=1500] [samples_per_lane=8]
//   RRT_NO_FRONT=1 build/warp_sim 1 -1 1200       # the same rays against the tree without a front run
// Scheduling variants (environment): SIM_NO_CHAIN / SIM_CHAIN (consecutive leaves in one round off / on),
// SIM_DEFER=k (a lane puts k leaves aside before it waits), SIM_MAX_LEAF, SIM_SINGLE_LEAF_MAX,
// SIM_TILE=WxH (pixels of a warp's tile), SIM_REGEN_MIN=k (dead lanes start new paths k at a time),
// SIM_QUEUE=1 [SIM_SORT=w] [SIM_POOL=n] (decoupled ray queue), SIM_GRID=lambda (uniform grid instead of the tree).
//
// Not part of the product and not an oracle: sphere/quad hits and the three book-1 materials in f32,
// no textures (albedo from the material record), no media, no light sampling.  Validation against the
// measured A/B runs is in profiles/README.md.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../include/rrt_host.h"
#include "../rustraytrace_b200/csrc/flatten.hpp"

using namespace rrt;

namespace {

struct V3 { float x, y, z; };
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 norm(V3 a) { return (1.0f / std::sqrt(dot(a, a))) * a; }
inline int as_int(float f) { int i; std::memcpy(&i, &f, 4); return i; }

struct Rng {  // xorshift64*: the model needs plausible paths, not the kernel's Philox keys
  uint64_t s;
  explicit Rng(uint64_t seed) : s(seed * 0x9E3779B97F4A7C15ull + 0x1234567ull) { next(); next(); }
  uint64_t next() { s ^= s >> 12; s ^= s << 25; s ^= s >> 27; return s * 0x2545F4914F6CDD1Dull; }
  float u() { return (float)(next() >> 40) * (1.0f / 16777216.0f); }
};
V3 uniform_sphere(Rng& g) {
  float z = 1.0f - 2.0f * g.u(), r = std::sqrt(std::max(0.0f, 1.0f - z * z)), p = 6.2831853f * g.u();
  return {r * std::cos(p), r * std::sin(p), z};
}

struct Ray { V3 o, d; float tm; };
struct Hit { float t; int prim; };       // prim = flat index (| QUAD bit), -1 none
const int QUAD_BIT = 0x40000000;

// What one lane does in one world_hit32 call, as the kernel's while-while loop sees it: rounds of
// (node steps until the lane holds a leaf or has finished, leaves tested in that round).
struct Trace {
  std::vector<uint16_t> steps, leaves;  // per round
  int front_tests = 0;
};

struct Tree {
  FlatScene F;
  std::vector<int> id_to_flat;           // construction id -> flat prim index (for `skip`)
  bool chain = true;                     // sphere-only instances run consecutive leaves in one round
  bool inline_box = false;
  bool chain_spheres_only = false;
  int defer = 1;                         // leaves a lane puts aside before it waits for a leaf round (1 = the kernel)
};

int prim_id(const FlatScene& F, int flat, bool quad) { return as_int(quad ? F.prims[4 * flat + 3].y : F.prims[4 * flat + 2].y); }

void sphere_hit(const FlatScene& F, int i, const Ray& r, float a, float inv_a, float tmin, int skip, Hit& h) {
  const float4 p0 = F.prims[4 * i], p1 = F.prims[4 * i + 1];
  V3 c{p0.x + r.tm * p1.x, p0.y + r.tm * p1.y, p0.z + r.tm * p1.z};
  V3 oc = c - r.o;
  float hh = dot(r.d, oc), k = hh * inv_a;
  V3 perp = oc - k * r.d;
  float disc = p0.w * p0.w - dot(perp, perp);
  if (disc < 0.0f) return;
  float sq = std::sqrt(disc * a), root;
  if (i == skip) {
    if (hh <= 0.0f) return;
    root = (hh + sq) * inv_a;
    if (!(root > tmin && root <= h.t)) return;
  } else {
    root = (hh - sq) * inv_a;
    if (!(root > tmin && root <= h.t)) { root = (hh + sq) * inv_a; if (!(root > tmin && root <= h.t)) return; }
  }
  if (root == h.t && h.prim >= 0) return;  // tie rules do not matter to the cost model
  h.t = root; h.prim = i;
}
void quad_hit(const FlatScene& F, int i, const Ray& r, float tmin, int skip, Hit& h) {
  if (i == skip) return;
  const float4 p0 = F.prims[4 * i], p1 = F.prims[4 * i + 1], p2 = F.prims[4 * i + 2];
  V3 n{p0.x, p0.y, p0.z};
  float denom = dot(n, r.d);
  if (std::fabs(denom) < 1e-8f) return;
  float t = (p0.w - dot(n, r.o)) / denom;
  if (!(t >= tmin && t <= h.t)) return;
  V3 p = r.o + t * r.d;
  float al = dot(p, V3{p1.x, p1.y, p1.z}) + p1.w, be = dot(p, V3{p2.x, p2.y, p2.z}) + p2.w;
  if (!(al >= 0.0f && al <= 1.0f && be >= 0.0f && be <= 1.0f)) return;
  h.t = t; h.prim = i | QUAD_BIT;
}
void leaf_hit(const FlatScene& F, int leaf, const Ray& r, float a, float inv_a, float tmin, int skip, Hit& h) {
  if ((leaf & 0xFF) == LEAF_BOX_MARK) {  // box leaf: the model tests the six quads
    int first = as_int(F.boxes[2 * (leaf >> 8)].w);
    for (int k = 0; k < 6; ++k) quad_hit(F, first + k, r, tmin, skip, h);
    return;
  }
  int first = leaf >> 8, ns = (leaf >> 4) & 15, nq = leaf & 15;
  for (int k = 0; k < ns; ++k) sphere_hit(F, first + k, r, a, inv_a, tmin, skip, h);
  for (int k = 0; k < nq; ++k) quad_hit(F, first + ns + k, r, tmin, skip, h);
}

// intersect.cuh world_hit32 on the host, recording the lane's rounds.
Hit world_hit(const Tree& T, const Ray& r, float tmin, int skip, Trace* tr) {
  const FlatScene& F = T.F;
  Hit h{INFINITY, -1};
  const float a = dot(r.d, r.d), inv_a = 1.0f / a;
  int cur = F.root_ref;
  if (tr) { tr->steps.clear(); tr->leaves.clear(); tr->front_tests = 0; }
  if (cur < 0) {  // brute-force root leaf: all lanes together; a quad costs about two spheres (C5: 50 instructions per quad)
    leaf_hit(F, ~cur, r, a, inv_a, tmin, skip, h);
    if (tr) tr->front_tests = ((~cur >> 4) & 15) + 2 * (~cur & 15);
    return h;
  }
  if (F.front_ref != -1) {
    leaf_hit(F, ~F.front_ref, r, a, inv_a, tmin, skip, h);
    if (tr) tr->front_tests = ((~F.front_ref >> 4) & 15) + (~F.front_ref & 15);
  }
  auto rcp = [](float d) { return 1.0f / (std::fabs(d) > 1e-30f ? d : std::copysign(1e-30f, d)); };
  const float id[3] = {rcp(r.d.x), rcp(r.d.y), rcp(r.d.z)};
  const float oo[3] = {r.o.x * id[0], r.o.y * id[1], r.o.z * id[2]};
  int stack[64], sp = 0;
  auto count = [](int ref) { const int lf = ~ref; return (lf & 0xFF) == LEAF_BOX_MARK ? 3 : ((lf >> 4) & 15) + (lf & 15); };  // (a packed sphere pair: 2 here, priced by SIM_C_PAIR below)  // primitives tested (a box leaf ~ 3: slab test, entry face, fall-back)
  for (;;) {
    bool done = false;
    int steps = 0, leaves = 0, buf[8], nb = 0;
    // descend; with T.defer > 1 a found leaf is put aside and the lane goes on with its stack
    // (without the leaf's hit distance to cull by) until it holds T.defer leaves
    for (;;) {
      while (cur >= 0) {
        ++steps;
        const float* f = reinterpret_cast<const float*>(&F.nodes[4 * cur]);
        float nr[2], fr[2];
        for (int c = 0; c < 2; ++c) {
          float tn = tmin, tf = h.t;
          for (int ax = 0; ax < 3; ++ax) {
            float tc = f[2 * ax + c] * id[ax] - oo[ax], e = f[6 + 2 * ax + c] * std::fabs(id[ax]);
            tn = std::max(tn, tc - e); tf = std::min(tf, tc + e);
          }
          nr[c] = tn; fr[c] = tf;
        }
        bool h0 = nr[0] <= fr[0] * 1.0000004f, h1 = nr[1] <= fr[1] * 1.0000004f;
        int r0 = as_int(f[12]), r1 = as_int(f[13]);
        if (h0 && h1) {
          if (nr[1] < nr[0]) std::swap(r0, r1);
          if (sp < 64) stack[sp++] = r1;
          cur = r0;
        } else if (h0) cur = r0;
        else if (h1) cur = r1;
        else { if (sp == 0) { done = true; break; } cur = stack[--sp]; }
      }
      if (done) break;
      if (T.inline_box && ((~cur) & 0xFF) == LEAF_BOX_MARK) {   // SIM_INLINE_BOX: a box leaf is settled by its parent's slab test, no leaf round
        leaf_hit(F, ~cur, r, a, inv_a, tmin, skip, h);
        if (sp == 0) { done = true; break; }
        cur = stack[--sp];
        continue;
      }
      buf[nb++] = cur;
      if (nb >= T.defer) break;
      if (sp == 0) { done = true; break; }
      cur = stack[--sp];
    }
    for (int k = 0; k < nb; ++k) { leaves += count(buf[k]); leaf_hit(F, ~buf[k], r, a, inv_a, tmin, skip, h); }
    if (!done) {
      for (;;) {
        if (sp == 0) { done = true; break; }
        cur = stack[--sp];
        if (!(T.chain && T.defer == 1 && cur < 0)) break;
        if (T.chain_spheres_only && ((~cur) & 0xFF) == LEAF_BOX_MARK) break;   // SIM_CHAIN=s: only sphere/quad leaves are chained
        leaves += count(cur); leaf_hit(F, ~cur, r, a, inv_a, tmin, skip, h);   // consecutive leaves in one round
      }
    }
    if (tr && (steps || leaves)) { tr->steps.push_back((uint16_t)steps); tr->leaves.push_back((uint16_t)leaves); }
    if (done) break;
  }
  return h;
}

struct Cam {
  V3 c, p00, du, dv, ddu, ddv;
  bool defocus;
  int W, H, max_depth;
};

struct Seg { Ray r; int skip_id; };               // skip as construction id: the same for every tree
struct Iter { std::vector<Seg> lanes; };          // one pass of the warp's main loop: the alive lanes' segments

// Scatter with the book-1/2 rules (render_kernel.cuh shade_step without textures, media and PDFs).
bool scatter(const FlatScene& F, const Ray& r, const Hit& h, int bounce, int max_depth, Rng& g, Ray& out) {
  const bool quad = h.prim & QUAD_BIT;
  const int i = h.prim & ~QUAD_BIT;
  V3 p = r.o + h.t * r.d, n;
  int mat;
  if (quad) { const float4 p0 = F.prims[4 * i]; n = norm(V3{p0.x, p0.y, p0.z}); mat = as_int(F.prims[4 * i + 3].x); }
  else {
    const float4 p0 = F.prims[4 * i], p1 = F.prims[4 * i + 1];
    n = norm(p - V3{p0.x + r.tm * p1.x, p0.y + r.tm * p1.y, p0.z + r.tm * p1.z});
    if (p0.w < 0) n = -1.0f * n;
    mat = as_int(F.prims[4 * i + 2].x);
  }
  bool front = dot(r.d, n) < 0.0f;
  if (!front) n = -1.0f * n;
  const DevMaterial& M = F.materials[mat];
  V3 att{M.albedo.x, M.albedo.y, M.albedo.z};
  if (M.tex >= 0) att = {0.5f, 0.5f, 0.5f};  // textured: a mid grey stands in for the roulette
  V3 dir;
  switch (M.kind) {
    case RRT_M_LAMBERTIAN: case RRT_M_ISOTROPIC: dir = n + uniform_sphere(g); break;
    case RRT_M_METAL: {
      V3 ud = r.d; dir = norm(ud - (2.0f * dot(ud, n)) * n) + M.fuzz * uniform_sphere(g);
      if (!(dot(dir, n) > 0.0f)) return false;
      break;
    }
    case RRT_M_DIELECTRIC: {
      float ri = front ? 1.0f / M.ior : M.ior;
      V3 ud = norm(r.d);
      float ct = std::min(-dot(ud, n), 1.0f), st = std::sqrt(std::max(0.0f, 1.0f - ct * ct));
      float r0 = (1.0f - ri) / (1.0f + ri); r0 *= r0;
      float refl = r0 + (1.0f - r0) * std::pow(1.0f - ct, 5.0f);
      if (ri * st > 1.0f || refl > g.u()) dir = ud - (2.0f * dot(ud, n)) * n;
      else { V3 perp = ri * (ud + ct * n); dir = perp + (-std::sqrt(std::fabs(1.0f - dot(perp, perp)))) * n; }
      att = {1.0f, 1.0f, 1.0f};
      break;
    }
    default: return false;  // lights, empty
  }
  if (bounce >= 5) {
    float rr = std::min(std::max(std::max(att.x, std::max(att.y, att.z)), 0.05f), 0.95f);
    if (g.u() > rr) return false;
  }
  if (bounce + 1 >= max_depth) return false;
  out.o = p; out.d = dir; out.tm = r.tm;
  return true;
}

}  // namespace

int main(int argc, char** argv) {
  if (argc < 4) { std::fprintf(stderr, "usage: warp_sim <book 1|2|3|4> <scene_no|-1> <image_width> [tiles] [samples_per_lane]\n"); return 2; }
  const int book = std::atoi(argv[1]), scene_no = std::atoi(argv[2]), width = std::atoi(argv[3]);
  const int n_tiles = argc > 4 ? std::atoi(argv[4]) : 1500, spl = argc > 5 ? std::atoi(argv[5]) : 8;
  rrth_overrides ov; rrth_overrides_none(&ov);
  ov.has = RRTH_O_IMAGE_WIDTH | RRTH_O_MAX_DEPTH; ov.image_width = width; ov.max_depth = 50;
  rrth_scene* sc = nullptr;
  if (rrth_scene_build(book, scene_no, 0x5EED1234ull, &ov, nullptr, &sc) != 0) { std::fprintf(stderr, "scene: %s\n", rrth_last_error()); return 1; }
  const rrt_scene_desc* desc = rrth_scene_desc(sc);
  const rrt_camera* cam = rrth_scene_camera(sc);

  // the tree under test (this build's constants, RRT_NO_FRONT / RRT_NO_BOX_LEAVES from the environment)
  Tree T;
  std::string err;
  const int max_leaf = std::getenv("SIM_MAX_LEAF") ? std::atoi(std::getenv("SIM_MAX_LEAF")) : 1;  // the kernel's loop assumes 1
  const int single_max = std::getenv("SIM_SINGLE_LEAF_MAX") ? std::atoi(std::getenv("SIM_SINGLE_LEAF_MAX")) : RRT_SINGLE_LEAF_MAX;
  if (flatten_scene(desc, T.F, err, single_max, max_leaf) != RRT_OK) { std::fprintf(stderr, "flatten: %s\n", err.c_str()); return 1; }
  T.chain = (!(T.F.features & SF_QUADS) || std::getenv("SIM_CHAIN")) && !std::getenv("SIM_NO_CHAIN");
  T.inline_box = std::getenv("SIM_INLINE_BOX") != nullptr;
  T.chain_spheres_only = std::getenv("SIM_CHAIN") && std::getenv("SIM_CHAIN")[0] == 's';
  if (std::getenv("SIM_DEFER")) T.defer = std::min(8, std::max(1, std::atoi(std::getenv("SIM_DEFER"))));
  // the rays come from a fixed tree, so that every variant is scored on the same segments
  Tree B;
  {
    const char* nf = std::getenv("RRT_NO_FRONT");
    std::string keep = nf ? nf : "";
    setenv("RRT_NO_FRONT", "1", 1);
    if (flatten_scene(desc, B.F, err) != RRT_OK) return 1;
    if (nf) setenv("RRT_NO_FRONT", keep.c_str(), 1); else unsetenv("RRT_NO_FRONT");
  }
  auto index_ids = [](Tree& t) {
    int maxid = 0;
    for (int i = 0; i < t.F.n_prims; ++i) maxid = std::max(maxid, std::max(as_int(t.F.prims[4 * i + 2].y), as_int(t.F.prims[4 * i + 3].y)));
    t.id_to_flat.assign((size_t)maxid + 2, -1);
    // a record is a quad iff its p3 holds the int block; the flattener emits spheres with p3 = (uvcos, uvsin, 0, 0):
    // take the kind from the leaf references instead
    std::vector<char> is_quad(t.F.n_prims, 0);
    auto mark = [&](int ref) {
      int leaf = ~ref;
      if ((leaf & 0xFF) == LEAF_BOX_MARK) { int f = as_int(t.F.boxes[2 * (leaf >> 8)].w); for (int k = 0; k < 6; ++k) is_quad[f + k] = 1; return; }
      int first = leaf >> 8, ns = (leaf >> 4) & 15, nq = leaf & 15;
      for (int k = 0; k < nq; ++k) is_quad[first + ns + k] = 1;
    };
    if (t.F.front_ref != -1) mark(t.F.front_ref);
    if (t.F.root_ref < 0) mark(t.F.root_ref);
    for (size_t n = 0; n < t.F.nodes.size() / 4; ++n) {
      int r0 = as_int(t.F.nodes[4 * n + 3].x), r1 = as_int(t.F.nodes[4 * n + 3].y);
      if (r0 < 0) mark(r0);
      if (r1 < 0) mark(r1);
    }
    for (int i = 0; i < t.F.n_prims; ++i) t.id_to_flat[prim_id(t.F, i, is_quad[i])] = i;
  };
  index_ids(T); index_ids(B);

  DevCamera c32; DevCamera64 c64;
  const int H = camera_initialize(cam, c32, c64);
  Cam C{{c32.cx, c32.cy, c32.cz}, {c32.p00x, c32.p00y, c32.p00z}, {c32.dux, c32.duy, c32.duz}, {c32.dvx, c32.dvy, c32.dvz},
        {c32.ddux, c32.dduy, c32.dduz}, {c32.ddvx, c32.ddvy, c32.ddvz}, c32.defocus != 0, cam->image_width, H, cam->max_depth};

  // ---- paths, grouped as the kernel groups them -------------------------------------------------
  std::vector<Iter> iters;
  Rng pick(7);
  // tile of a warp: SIM_TILE=WxH pixels with W*H <= 32 (the kernel: 8x4, lane = pixel); smaller tiles put
  // several lanes on one pixel, each tile pixel then gets 32/(W*H) times the samples
  int TW = 8, TH = 4;
  if (const char* e = std::getenv("SIM_TILE")) std::sscanf(e, "%dx%d", &TW, &TH);
  const int NP = TW * TH, spp_pixel = spl * 32 / NP;
  const int tiles_x = (C.W + TW - 1) / TW, tiles_y = (C.H + TH - 1) / TH;
  double n_paths = 0, n_segs = 0;
  // the kernel takes over neighbours' samples except in the Perlin instances and scenes of <= 2 primitives
  const bool steal = !std::getenv("SIM_NO_STEAL") && !(T.F.features & SF_NOISE) && T.F.n_prims > 2;
  const int regen_min = std::getenv("SIM_REGEN_MIN") ? std::atoi(std::getenv("SIM_REGEN_MIN")) : 1;
  for (int t = 0; t < n_tiles; ++t) {
    const int tx = (int)(pick.next() % tiles_x), ty = (int)(pick.next() % tiles_y);
    int next_sample[32], tgt[32], bounce[32], skip_id[32];
    bool alive[32];
    Ray ray[32];
    std::vector<Rng> g;
    for (int l = 0; l < 32; ++l) {
      g.emplace_back((uint64_t)t * 64 + l);
      const int px = tx * TW + (l % NP) % TW, py = ty * TH + (l % NP) / TW;
      next_sample[l] = (l < NP && px < C.W && py < C.H) ? 0 : spp_pixel;
      tgt[l] = l % NP; alive[l] = false; bounce[l] = 0; skip_id[l] = -1;
    }
    for (;;) {
      Iter it;
      bool any = false;
      int n_dead = 0;
      for (int l = 0; l < 32; ++l) n_dead += !alive[l];
      const bool regen = n_dead >= regen_min || n_dead == 32;   // SIM_REGEN_MIN: dead lanes wait until that many can start together
      for (int l = 0; l < 32; ++l) {
        if (!alive[l] && !regen) continue;
        if (!alive[l]) {  // next sample of its own pixel, then of a neighbour's (sample stealing)
          int tries = 0;
          for (; tries < (steal ? NP : 1); ++tries) { if (next_sample[tgt[l]] < spp_pixel) break; tgt[l] = (tgt[l] + 1) % NP; }
          if (!steal && next_sample[tgt[l]] >= spp_pixel) continue;
          if (tries == NP) continue;
          ++next_sample[tgt[l]];
          const int px = tx * TW + tgt[l] % TW, py = ty * TH + tgt[l] / TW;
          const float fx = px + g[l].u() - 0.5f, fy = py + g[l].u() - 0.5f;
          V3 rel = C.p00 + fx * C.du + fy * C.dv, o = C.c;
          if (C.defocus) {
            float rr = std::sqrt(g[l].u()), ph = 6.2831853f * g[l].u();
            V3 lens = (rr * std::cos(ph)) * C.ddu + (rr * std::sin(ph)) * C.ddv;
            o = o + lens; rel = rel - lens;
          }
          ray[l] = {o, rel, T.F.has_motion ? g[l].u() : 0.0f};
          alive[l] = true; bounce[l] = 0; skip_id[l] = -1; n_paths += 1;
        }
        any = true;
        it.lanes.push_back({ray[l], skip_id[l]});
        const int skip = skip_id[l] >= 0 ? B.id_to_flat[skip_id[l]] : -1;
        Hit h = world_hit(B, ray[l], 0.001f, skip, nullptr);
        n_segs += 1;
        Ray nr;
        if (h.prim < 0 || !scatter(B.F, ray[l], h, bounce[l], C.max_depth, g[l], nr)) { alive[l] = false; continue; }
        skip_id[l] = prim_id(B.F, h.prim & ~QUAD_BIT, (h.prim & QUAD_BIT) != 0);
        ray[l] = nr; ++bounce[l];
      }
      if (!any) break;
      iters.push_back(std::move(it));
    }
  }

  // ---- score the tree under test -----------------------------------------------------------------
  // instruction counts per event, from the SASS profile of the r01j kernel (profiles/groups_r01j.txt)
  // the fixed part of a segment: ray set-up, Philox, shading, regeneration (C2 445); media add two boundary
  // tests per medium and the free-flight draw, the noise instance its texture code (C4: ~1000 in all)
  const double C_NODE = 43, C_LEAF_ROUND = 12, C_LEAF = 34, C_FRONT = 23, C_SETUP = 45;
  const double C_FIXED = 445 + 130.0 * T.F.media.size() + ((T.F.features & SF_NOISE) ? 200 : 0) + ((T.F.features & SF_TEXTURES) ? 60 : 0);
  double longest = 0, warp_steps = 0, lane_steps = 0, rounds = 0, leaf_rounds = 0, lane_leaves = 0, lanes = 0, cost = 0, front = 0;
  Trace tr;
  std::vector<Trace> lane_tr;
  for (const Iter& it : iters) {
    lane_tr.assign(it.lanes.size(), Trace());
    size_t R = 0;
    for (size_t k = 0; k < it.lanes.size(); ++k) {
      const Seg& s = it.lanes[k];
      const int skip = s.skip_id >= 0 ? T.id_to_flat[s.skip_id] : -1;
      world_hit(T, s.r, 0.001f, skip, &lane_tr[k]);
      R = std::max(R, lane_tr[k].steps.size());
      for (size_t q = 0; q < lane_tr[k].steps.size(); ++q) { lane_steps += lane_tr[k].steps[q]; lane_leaves += lane_tr[k].leaves[q]; }
    }
    {  // the longest lane of the iteration: what the warp would pay without any wait at the leaf rounds
      int m = 0;
      for (const Trace& t2 : lane_tr) { int tot = 0; for (uint16_t v : t2.steps) tot += v; m = std::max(m, tot); }
      longest += m;
    }
    lanes += it.lanes.size();
    double c = C_SETUP + C_FIXED + C_FRONT * lane_tr[0].front_tests;
    front += lane_tr[0].front_tests;
    for (size_t q = 0; q < R; ++q) {
      int ms = 0, ml = 0;
      for (const Trace& t2 : lane_tr) if (q < t2.steps.size()) { ms = std::max<int>(ms, t2.steps[q]); ml = std::max<int>(ml, t2.leaves[q]); }
      warp_steps += ms;
      c += C_NODE * ms;
      if (ml) { leaf_rounds += 1; c += C_LEAF_ROUND + C_LEAF * ml; }
    }
    rounds += R;
    cost += c;
  }
  // ---- SIM_QUEUE=1: the same rays through a decoupled traversal (a pool of rays, any lane takes the next
  // one as soon as its own is finished; rounds as in the kernel: descend to a leaf, leaf round) -----------
  if (std::getenv("SIM_QUEUE")) {
    const double C_FETCH = std::getenv("SIM_C_FETCH") ? std::atof(std::getenv("SIM_C_FETCH")) : 20, C_DONE = 8;
    const double C_SHADE = C_SETUP + C_FIXED + 40;   // per 32 segments, converged, + state traffic
    std::vector<Trace> all;
    double fronts = 0;
    // SIM_SORT=w: rays sorted inside windows of w by (direction octant, origin cell) before they are handed out
    std::vector<Seg> segs;
    for (const Iter& it : iters) for (const Seg& s : it.lanes) segs.push_back(s);
    if (const char* e = std::getenv("SIM_SORT")) {
      const size_t w = (size_t)std::atoi(e);
      float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
      for (const Seg& s : segs) { const float o[3] = {s.r.o.x, s.r.o.y, s.r.o.z}; for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], o[a]); hi[a] = std::max(hi[a], o[a]); } }
      auto key = [&](const Seg& s) {
        const float o[3] = {s.r.o.x, s.r.o.y, s.r.o.z};
        unsigned k = (s.r.d.x < 0) | ((s.r.d.y < 0) << 1) | ((s.r.d.z < 0) << 2);
        unsigned c[3];
        for (int a = 0; a < 3; ++a) c[a] = (unsigned)std::min(15.0f, 16.0f * (o[a] - lo[a]) / std::max(1e-9f, hi[a] - lo[a]));
        unsigned m = 0;
        for (int b = 3; b >= 0; --b) for (int a = 0; a < 3; ++a) m = (m << 1) | ((c[a] >> b) & 1);
        return (k << 12) | m;
      };
      for (size_t i = 0; i < segs.size(); i += w)
        std::stable_sort(segs.begin() + i, segs.begin() + std::min(segs.size(), i + w), [&](const Seg& a, const Seg& b) { return key(a) < key(b); });
    }
    for (int once = 0; once < 1; ++once)
      for (const Seg& s : segs) {
        Trace t2;
        const int skip = s.skip_id >= 0 ? T.id_to_flat[s.skip_id] : -1;
        world_hit(T, s.r, 0.001f, skip, &t2);
        fronts += t2.front_tests;
        all.push_back(std::move(t2));
      }
    const int pool = std::getenv("SIM_POOL") ? std::atoi(std::getenv("SIM_POOL")) : 1 << 30;   // rays per phase (tail at the end of each)
    double qcost = 0, qsteps = 0, lane_active = 0;
    size_t next = 0;
    while (next < all.size()) {
      const size_t end = std::min(all.size(), next + (size_t)pool);
      int cur[32], rnd[32];
      for (int l = 0; l < 32; ++l) cur[l] = -1;
      for (;;) {
        bool fetched = false, any = false;
        for (int l = 0; l < 32; ++l) {
          while (cur[l] < 0 && next < end) {   // rays with no rounds at all (front run settles them) finish at once
            cur[l] = (int)next++; rnd[l] = 0; fetched = true;
            if (all[cur[l]].steps.empty()) cur[l] = -1;
          }
          any |= cur[l] >= 0;
        }
        if (fetched) qcost += C_FETCH;
        if (!any) break;
        int ms = 0, ml = 0, act = 0; bool fin = false;
        for (int l = 0; l < 32; ++l) if (cur[l] >= 0) {
          const Trace& t2 = all[cur[l]];
          ms = std::max<int>(ms, t2.steps[rnd[l]]); ml = std::max<int>(ml, t2.leaves[rnd[l]]);
          lane_active += t2.steps[rnd[l]]; ++act;
          if (++rnd[l] >= (int)t2.steps.size()) { cur[l] = -1; fin = true; }
        }
        qsteps += ms;
        qcost += C_NODE * ms + (ml ? C_LEAF_ROUND + C_LEAF * ml : 0) + (fin ? C_DONE : 0);
      }
    }
    const double nseg = (double)all.size();
    const double front_cost = C_FRONT * (fronts / nseg) * nseg / 32.0, shade_cost = C_SHADE * nseg / 32.0;
    std::printf("QUEUE model (pool %d): traversal %.0f + front %.0f + shade/regen %.0f = %.0f warp instructions per 32 segments; node steps per 32 segments %.2f (lane occupancy %.0f %%)\n",
                pool, 32 * qcost / nseg, 32 * front_cost / nseg, 32 * shade_cost / nseg, 32 * (qcost + front_cost + shade_cost) / nseg, 32 * qsteps / nseg, 100.0 * lane_active / (32.0 * qsteps));
    std::printf("QUEUE vs kernel, warp instructions per segment-lane: %.1f vs %.1f\n", (qcost + front_cost + shade_cost) / nseg, cost / lanes);
  }
  // ---- SIM_GRID=lambda: the tree replaced by a uniform grid over its spheres (what flatten.cpp choose_grids / emit_grids build) ----
  if (const char* ge = std::getenv("SIM_GRID")) {
    const double lambda = std::atof(ge);
    struct Box { float lo[3], hi[3]; int prim; };
    std::vector<Box> items;
    std::vector<char> in_front(T.F.n_prims, 0);
    if (T.F.front_ref != -1) { int lf = ~T.F.front_ref, first = lf >> 8, n = ((lf >> 4) & 15) + (lf & 15); for (int k = 0; k < n; ++k) in_front[first + k] = 1; }
    for (int i = 0; i < T.F.n_prims; ++i) {
      if (in_front[i]) continue;
      const float4 p0 = T.F.prims[4 * i], p1 = T.F.prims[4 * i + 1];
      Box b; b.prim = i;
      const float c[3] = {p0.x, p0.y, p0.z}, d[3] = {p1.x, p1.y, p1.z};
      for (int a = 0; a < 3; ++a) { b.lo[a] = std::min(c[a], c[a] + d[a]) - std::fabs(p0.w); b.hi[a] = std::max(c[a], c[a] + d[a]) + std::fabs(p0.w); }
      items.push_back(b);
    }
    float glo[3] = {1e30f, 1e30f, 1e30f}, ghi[3] = {-1e30f, -1e30f, -1e30f};
    double avg_ext[3] = {0, 0, 0};
    for (const Box& b : items) for (int a = 0; a < 3; ++a) { glo[a] = std::min(glo[a], b.lo[a]); ghi[a] = std::max(ghi[a], b.hi[a]); avg_ext[a] += (b.hi[a] - b.lo[a]) / items.size(); }
    double ext[3], V = 1;
    for (int a = 0; a < 3; ++a) { ext[a] = std::max<double>(ghi[a] - glo[a], avg_ext[a]); V *= ext[a]; }
    const double dens = std::cbrt(lambda * items.size() / V);
    int res[3]; float cell[3], inv_cell[3];
    for (int a = 0; a < 3; ++a) { res[a] = std::max(1, std::min(256, (int)std::lround(ext[a] * dens))); cell[a] = (ghi[a] - glo[a]) / res[a]; inv_cell[a] = res[a] / std::max(1e-20f, ghi[a] - glo[a]); }
    std::vector<std::vector<int>> cells((size_t)res[0] * res[1] * res[2]);
    size_t refs = 0;
    for (const Box& b : items) {
      int c0[3], c1[3];
      for (int a = 0; a < 3; ++a) { c0[a] = std::max(0, std::min(res[a] - 1, (int)((b.lo[a] - glo[a]) * inv_cell[a]))); c1[a] = std::max(0, std::min(res[a] - 1, (int)((b.hi[a] - glo[a]) * inv_cell[a]))); }
      for (int z = c0[2]; z <= c1[2]; ++z) for (int y = c0[1]; y <= c1[1]; ++y) for (int x = c0[0]; x <= c1[0]; ++x) { cells[((size_t)z * res[1] + y) * res[0] + x].push_back(b.prim); ++refs; }
    }
    std::printf("GRID %dx%dx%d over %zu items, %.2f references per item, %.2f per cell\n", res[0], res[1], res[2], items.size(), (double)refs / items.size(), (double)refs / cells.size());
    const double G_SETUP = 45, G_CELL = 24, G_SPH = 30;   // instruction prices: clip + set-up, one cell step, one sphere test
    double gcost = 0, gl = 0, tot_cells = 0, tot_tests = 0, warp_cells = 0, warp_tests = 0, wrong = 0;
    for (const Iter& it : iters) {
      std::vector<std::vector<int>> per_cell(it.lanes.size());
      size_t R = 0;
      int nfront = 0;
      for (size_t k = 0; k < it.lanes.size(); ++k) {
        const Seg& sg = it.lanes[k];
        const Ray& r = sg.r;
        const int skip = sg.skip_id >= 0 ? T.id_to_flat[sg.skip_id] : -1;
        Hit h{INFINITY, -1};
        const float a = dot(r.d, r.d), inv_a = 1.0f / a;
        if (T.F.front_ref != -1) { leaf_hit(T.F, ~T.F.front_ref, r, a, inv_a, 0.001f, skip, h); nfront = ((~T.F.front_ref >> 4) & 15) + (~T.F.front_ref & 15); }
        const float o[3] = {r.o.x, r.o.y, r.o.z}, d[3] = {r.d.x, r.d.y, r.d.z};
        float t0 = 0.001f, t1 = h.t;
        for (int ax = 0; ax < 3; ++ax) {
          const float id = 1.0f / (std::fabs(d[ax]) > 1e-30f ? d[ax] : std::copysign(1e-30f, d[ax]));
          float ta = (glo[ax] - o[ax]) * id, tb = (ghi[ax] - o[ax]) * id;
          if (ta > tb) std::swap(ta, tb);
          t0 = std::max(t0, ta); t1 = std::min(t1, tb);
        }
        std::vector<int>& pc = per_cell[k];
        if (t0 <= t1) {
          int ic[3], stp[3]; float tmax[3], tdel[3];
          for (int ax = 0; ax < 3; ++ax) {
            const float p = o[ax] + t0 * d[ax];
            ic[ax] = std::max(0, std::min(res[ax] - 1, (int)((p - glo[ax]) * inv_cell[ax])));
            const float id = 1.0f / (std::fabs(d[ax]) > 1e-30f ? d[ax] : std::copysign(1e-30f, d[ax]));
            stp[ax] = d[ax] >= 0 ? 1 : -1;
            tmax[ax] = (glo[ax] + (ic[ax] + (d[ax] >= 0 ? 1 : 0)) * cell[ax] - o[ax]) * id; tdel[ax] = cell[ax] * std::fabs(id);
          }
          for (;;) {
            const std::vector<int>& c = cells[((size_t)ic[2] * res[1] + ic[1]) * res[0] + ic[0]];
            pc.push_back((int)c.size());
            for (int pi : c) sphere_hit(T.F, pi, r, a, inv_a, 0.001f, skip, h);
            const int ax = tmax[0] < tmax[1] ? (tmax[0] < tmax[2] ? 0 : 2) : (tmax[1] < tmax[2] ? 1 : 2);
            if (tmax[ax] > std::min(t1, h.t)) break;
            ic[ax] += stp[ax]; tmax[ax] += tdel[ax];
            if (ic[ax] < 0 || ic[ax] >= res[ax]) break;
          }
        }
        if (world_hit(T, r, 0.001f, skip, nullptr).prim != h.prim) wrong += 1;   // same closest hit as the tree?
        R = std::max(R, pc.size());
        tot_cells += pc.size(); for (int v : pc) tot_tests += v;
      }
      double c = C_SETUP + C_FIXED + C_FRONT * nfront + G_SETUP;
      for (size_t q = 0; q < R; ++q) {
        int mt = 0;
        for (const auto& pc : per_cell) if (q < pc.size()) mt = std::max(mt, pc[q]);
        c += G_CELL + G_SPH * mt; warp_cells += 1; warp_tests += mt;
      }
      gcost += c; gl += it.lanes.size();
    }
    const double nI = (double)iters.size();
    std::printf("GRID per warp and segment: %.2f cell rounds (%.2f per lane), %.2f sphere tests (%.2f per lane) | modelled %.0f warp instructions; %.0f of %.0f closest hits differ from the tree's\n",
                warp_cells / nI, tot_cells / gl, warp_tests / nI, tot_tests / gl, gcost / nI, wrong, gl);
  }
  const double ni = (double)iters.size();
  std::printf("scene book %d/%d width %d: %d prims, %zu nodes, front run %d, chain %d | %.0f paths, %.2f segments per path, %.1f lanes per warp iteration\n",
              book, scene_no, width, T.F.n_prims, T.F.nodes.size() / 4, (int)(front / ni + 0.5), (int)T.chain, n_paths, n_segs / n_paths, lanes / ni);
  std::printf("per warp and segment: %.2f node steps (%.2f per lane, %.0f %%), %.2f rounds, %.2f leaf rounds, %.2f leaf tests per lane | modelled %.0f warp instructions\n",
              warp_steps / ni, lane_steps / lanes, 100.0 * (lane_steps / lanes) / (warp_steps / ni), rounds / ni, leaf_rounds / ni, lane_leaves / lanes, cost / ni);
  std::printf("modelled warp instructions per path: %.1f\n", cost / n_paths);
  std::printf("longest lane of a segment: %.2f node steps: of the warp's %.2f, %.2f are the spread between lanes and %.2f the waits at leaf rounds\n",
              longest / ni, warp_steps / ni, longest / ni - lane_steps / lanes, (warp_steps - longest) / ni);
  rrth_scene_free(sc);
  return 0;
}
