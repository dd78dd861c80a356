// flatten.cpp — see flatten.hpp.  Host-only C++17, f64 throughout; f32 conversion happens at the
// very end and is conservative for boxes.
#include "flatten.hpp"

#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <map>
#include <queue>

namespace rrt {
namespace {

struct D3 {
  double x, y, z;
  D3() : x(0), y(0), z(0) {}
  D3(double a, double b, double c) : x(a), y(b), z(c) {}
  explicit D3(const double* p) : x(p[0]), y(p[1]), z(p[2]) {}
  double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
inline D3 operator+(D3 a, D3 b) { return D3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline D3 operator-(D3 a, D3 b) { return D3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline D3 operator*(double s, D3 a) { return D3(s * a.x, s * a.y, s * a.z); }
inline double dot(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline D3 cross(D3 a, D3 b) { return D3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }

// object -> world: p_w = R p + t with R a rotation about Y (hittable.rs:158-170: x = c x' + s z',
// z = -s x' + c z').
struct Xform {
  double c = 1.0, s = 0.0;
  D3 t;
  bool rotated = false;
  D3 rot(D3 p) const { return D3(c * p.x + s * p.z, p.y, -s * p.x + c * p.z); }
  D3 point(D3 p) const { return rot(p) + t; }
};

struct BakedPrim {
  int kind = 0, id = -1, order = 0, mat = -1, flags = 0;
  D3 c0, dc; double r = 0;            // sphere
  D3 Q, u, v, n, w; double d = 0, area = 0;  // quad
  double uvcos = 1, uvsin = 0;
  double lo[3], hi[3];
  int box = -1, face = -1;   // member of box group `box`, face slot 0..5 = -x,+x,-y,+y,-z,+z
  int grid = -1;             // build item: member of uniform grid `grid` (0 spheres, 1 boxes), see build_grids
};

// Six quads that are exactly the faces of one axis-aligned box (quad.rs box_() under translations only).
struct BoxGroup {
  int member[6];      // index into the world prim list, by face slot
  double lo[3], hi[3];
};

// Axis of an axis-aligned rectangle (quad whose edges u, v each run along one coordinate axis: the walls, floor,
// ceiling and light of the Cornell boxes, the_rest_of_your_life/mod.rs:82-130), or -1.
int rect_axis(const BakedPrim& p) {
  auto along = [](const D3& e) {
    const int nz = (e.x != 0.0) + (e.y != 0.0) + (e.z != 0.0);
    return nz == 1 ? (e.x != 0.0 ? 0 : (e.y != 0.0 ? 1 : 2)) : -1;
  };
  if (p.kind != PRIM_QUAD) return -1;
  const int au = along(p.u), av = along(p.v);
  if (au < 0 || av < 0 || au == av) return -1;
  return 3 - au - av;
}

void pad_axis(double& lo, double& hi) {  // Aabb::pad_to_minimums, aabb.rs:107-118
  const double delta = 0.0001;
  if (hi - lo < delta) { lo -= delta / 2; hi += delta / 2; }
}

void bake_sphere(const rrt_hittable& h, int id, const Xform& X, BakedPrim& p) {
  p.kind = PRIM_SPHERE; p.id = id; p.mat = h.material;
  p.c0 = X.point(D3(h.a));
  p.dc = X.rot(D3(h.b));
  p.r = std::max(h.c[0], 0.0);  // sphere.rs:23 radius.max(0.0)
  p.uvcos = X.c; p.uvsin = X.s;
  if (X.rotated) p.flags |= PF_UV_ROTATED;
  double rr = std::fabs(h.c[0]);
  D3 c1 = p.c0 + p.dc;
  for (int a = 0; a < 3; ++a) {
    p.lo[a] = std::min(p.c0[a], c1[a]) - rr;
    p.hi[a] = std::max(p.c0[a], c1[a]) + rr;
    pad_axis(p.lo[a], p.hi[a]);
  }
}

void quad_derive(BakedPrim& p) {  // Quad::new, quad.rs:24-48
  D3 n = cross(p.u, p.v);
  double len = std::sqrt(dot(n, n));
  p.n = D3(n.x / len, n.y / len, n.z / len);
  p.d = dot(p.n, p.Q);
  double nn = dot(n, n);
  p.w = D3(n.x / nn, n.y / nn, n.z / nn);
  p.area = len;
}

void bake_quad(const rrt_hittable& h, int id, const Xform& X, BakedPrim& p) {
  p.kind = PRIM_QUAD; p.id = id; p.mat = h.material;
  p.Q = X.point(D3(h.a));
  p.u = X.rot(D3(h.b));
  p.v = X.rot(D3(h.c));
  quad_derive(p);
  D3 corners[4] = {p.Q, p.Q + p.u, p.Q + p.v, p.Q + p.u + p.v};  // quad.rs:50-54
  for (int a = 0; a < 3; ++a) {
    p.lo[a] = p.hi[a] = corners[0][a];
    for (int k = 1; k < 4; ++k) { p.lo[a] = std::min(p.lo[a], corners[k][a]); p.hi[a] = std::max(p.hi[a], corners[k][a]); }
    pad_axis(p.lo[a], p.hi[a]);
  }
}

struct BakedMedium { int first, count, mat, id; double density; };

// ------------------------------------------------------------------------------------------------
// Reference traversal order.  At an exact equal-t tie the reference's answer depends on the order
// in which it visits primitives: list order (hittable_list.rs:46-51), and inside a BvhNode the
// left-then-right order of the tree that BvhNode::build produced (bvh.rs:159-172).  To give every
// primitive the same rank the reference would, the flattener replays the reference's bounding
// boxes (aabb.rs, sphere.rs:19-41, quad.rs:50-54, hittable.rs:79-82, 107-139) and its 12-bucket
// SAH partition (bvh.rs:21-156) — only to obtain the visiting order; the tree itself is discarded.
struct RBox {
  double lo[3], hi[3];
  static RBox empty() {
    RBox b;
    for (int a = 0; a < 3; ++a) { b.lo[a] = std::numeric_limits<double>::infinity(); b.hi[a] = -b.lo[a]; }
    return b;
  }
  void pad() {  // Aabb::new -> pad_to_minimums, aabb.rs:24-28, 107-118
    const double delta = 0.0001;
    for (int a = 0; a < 3; ++a)
      if (hi[a] - lo[a] < delta) { double p = delta / 2.0; lo[a] -= p; hi[a] += p; }
  }
  static RBox from_points(D3 p, D3 q) {  // aabb.rs:30-35
    RBox b;
    for (int a = 0; a < 3; ++a) { if (p[a] <= q[a]) { b.lo[a] = p[a]; b.hi[a] = q[a]; } else { b.lo[a] = q[a]; b.hi[a] = p[a]; } }
    b.pad();
    return b;
  }
  static RBox from_boxes(const RBox& x, const RBox& y) {  // aabb.rs:37-43, interval.rs:19-24
    RBox b;
    for (int a = 0; a < 3; ++a) { b.lo[a] = x.lo[a] <= y.lo[a] ? x.lo[a] : y.lo[a]; b.hi[a] = x.hi[a] >= y.hi[a] ? x.hi[a] : y.hi[a]; }
    b.pad();
    return b;
  }
  int longest_axis() const {  // aabb.rs:86-98
    double sx = hi[0] - lo[0], sy = hi[1] - lo[1], sz = hi[2] - lo[2];
    if (sx > sy) return sx > sz ? 0 : 2;
    return sy > sz ? 1 : 2;
  }
  double surface_area() const {  // aabb.rs:100-105
    double a = hi[0] - lo[0], b = hi[1] - lo[1], c = hi[2] - lo[2];
    return 2.0 * (a * b + a * c + b * c);
  }
};

struct RefOrder {
  const rrt_scene_desc* d;
  std::vector<RBox> memo;
  std::vector<char> have;
  explicit RefOrder(const rrt_scene_desc* desc) : d(desc), memo(desc->n_hittables), have(desc->n_hittables, 0) {}

  // Hittable::bounding_box of the reference object behind desc index idx (indices already validated)
  RBox bbox(int idx, int depth = 0) {
    if (idx < 0 || idx >= d->n_hittables || depth > 64) return RBox::empty();
    if (have[idx]) return memo[idx];
    const rrt_hittable& h = d->hittables[idx];
    RBox b = RBox::empty();
    switch (h.kind) {
      case RRT_H_SPHERE: {
        D3 c0(h.a), c1 = D3(h.a) + D3(h.b), rv(h.c[0], h.c[0], h.c[0]);
        b = RBox::from_points(c0 - rv, c0 + rv);
        if (h.b[0] != 0 || h.b[1] != 0 || h.b[2] != 0) b = RBox::from_boxes(b, RBox::from_points(c1 - rv, c1 + rv));
        break;
      }
      case RRT_H_QUAD: {
        D3 q(h.a), u(h.b), v(h.c);
        b = RBox::from_boxes(RBox::from_points(q, q + u + v), RBox::from_points(q + u, q + v));
        break;
      }
      case RRT_H_LIST:  // hittable_list.rs:25-33
        for (int k = 0; k < h.n_children; ++k) {
          RBox cb = bbox(d->children[h.first_child + k], depth + 1);
          b = k == 0 ? cb : RBox::from_boxes(b, cb);
        }
        break;
      case RRT_H_BVH:  // bvh.rs:24-27
        for (int k = 0; k < h.n_children; ++k) b = RBox::from_boxes(b, bbox(d->children[h.first_child + k], depth + 1));
        break;
      case RRT_H_TRANSLATE: {  // hittable.rs:79-82, aabb.rs:121-127
        RBox cb = bbox(h.first_child, depth + 1);
        for (int a = 0; a < 3; ++a) { b.lo[a] = cb.lo[a] + h.a[a]; b.hi[a] = cb.hi[a] + h.a[a]; }
        b.pad();
        break;
      }
      case RRT_H_ROTATE_Y: {  // hittable.rs:107-139
        RBox cb = bbox(h.first_child, depth + 1);
        double radians = h.a[0] * M_PI / 180.0, st = std::sin(radians), ct = std::cos(radians);
        double inf = std::numeric_limits<double>::infinity();
        D3 mn(inf, inf, inf), mx(-inf, -inf, -inf);
        for (int i = 0; i < 2; ++i) for (int j = 0; j < 2; ++j) for (int k = 0; k < 2; ++k) {
          double x = i ? cb.hi[0] : cb.lo[0], y = j ? cb.hi[1] : cb.lo[1], z = k ? cb.hi[2] : cb.lo[2];
          double t[3] = {ct * x + st * z, y, -st * x + ct * z};
          mn = D3(std::fmin(mn.x, t[0]), std::fmin(mn.y, t[1]), std::fmin(mn.z, t[2]));
          mx = D3(std::fmax(mx.x, t[0]), std::fmax(mx.y, t[1]), std::fmax(mx.z, t[2]));
        }
        b = RBox::from_points(mn, mx);
        break;
      }
      case RRT_H_CONSTANT_MEDIUM: b = bbox(h.first_child, depth + 1); break;  // constant_medium.rs:81-83
      default: break;
    }
    memo[idx] = b; have[idx] = 1;
    return b;
  }

  // BvhNode::build (bvh.rs:21-156) on objs[first, first+span): appends the left-to-right visiting
  // order of the resulting tree to `out`.
  void build(std::vector<int>& objs, size_t first, size_t span, std::vector<int>& out) {
    if (span == 0) return;
    if (span == 1) { out.push_back(objs[first]); return; }  // left == right: same object twice, one rank
    if (span == 2) { out.push_back(objs[first]); out.push_back(objs[first + 1]); return; }
    RBox box = RBox::empty();
    for (size_t i = 0; i < span; ++i) box = RBox::from_boxes(box, bbox(objs[first + i]));
    const size_t NB = 12;
    const int axis = box.longest_axis();
    auto median = [&]() {
      std::stable_sort(objs.begin() + first, objs.begin() + first + span,
                       [&](int a, int b) { return bbox(a).lo[axis] < bbox(b).lo[axis]; });
      size_t mid = span / 2;
      build(objs, first, mid, out);
      build(objs, first + mid, span - mid, out);
    };
    double cmin = std::numeric_limits<double>::infinity(), cmax = -cmin;
    for (size_t i = 0; i < span; ++i) {
      RBox b = bbox(objs[first + i]);
      double c = 0.5 * (b.lo[axis] + b.hi[axis]);
      if (c < cmin) cmin = c;
      if (c > cmax) cmax = c;
    }
    if (std::fabs(cmax - cmin) < 1e-12) { median(); return; }
    auto bucket_of = [&](int obj) {
      RBox b = bbox(obj);
      double c = 0.5 * (b.lo[axis] + b.hi[axis]);
      double f = (c - cmin) / (cmax - cmin) * (double)NB;
      size_t idx = f > 0.0 ? (size_t)f : 0;  // `as usize` saturates
      if (idx >= NB) idx = NB - 1;
      return idx;
    };
    size_t count[NB] = {0};
    RBox bb[NB];
    for (size_t k = 0; k < NB; ++k) bb[k] = RBox::empty();
    for (size_t i = 0; i < span; ++i) {
      size_t k = bucket_of(objs[first + i]);
      count[k] += 1;
      bb[k] = RBox::from_boxes(bb[k], bbox(objs[first + i]));
    }
    RBox rbox[NB];
    size_t rcount[NB];
    RBox acc = RBox::empty();
    size_t accn = 0;
    for (size_t k = NB; k-- > 0;) { accn += count[k]; acc = RBox::from_boxes(acc, bb[k]); rbox[k] = acc; rcount[k] = accn; }
    RBox lbox = RBox::empty();
    size_t lcount = 0, best_split = 0;
    double best_cost = std::numeric_limits<double>::infinity();
    for (size_t k = 0; k + 1 < NB; ++k) {
      lcount += count[k];
      lbox = RBox::from_boxes(lbox, bb[k]);
      if (lcount == 0 || rcount[k + 1] == 0) continue;
      double cost = lbox.surface_area() * (double)lcount + rbox[k + 1].surface_area() * (double)rcount[k + 1];
      if (cost < best_cost) { best_cost = cost; best_split = k; }
    }
    if (!std::isfinite(best_cost)) { median(); return; }
    size_t mid = 0;
    for (size_t i = 0; i < span; ++i)
      if (bucket_of(objs[first + i]) <= best_split) { std::swap(objs[first + i], objs[first + mid]); mid += 1; }
    if (mid == 0 || mid == span) { median(); return; }
    build(objs, first, mid, out);
    build(objs, first + mid, span - mid, out);
  }
};

struct Walker {
  const rrt_scene_desc* d;
  RefOrder ref;
  explicit Walker(const rrt_scene_desc* desc) : d(desc), ref(desc) {}
  std::vector<BakedPrim> world;
  std::vector<BakedPrim> boundary;
  std::vector<BakedMedium> media;
  std::vector<BoxGroup> boxes;   // groups found in `world`
  std::string err;
  int status = RRT_OK;

  bool fail(int st, const std::string& m) { if (status == RRT_OK) { status = st; err = m; } return false; }

  // Are out[first..first+6) the six faces of one axis-aligned box?  Then they become one BVH leaf.
  void find_box(std::vector<BakedPrim>& out, size_t first) {
    BoxGroup g;
    for (int a = 0; a < 3; ++a) { g.lo[a] = std::numeric_limits<double>::infinity(); g.hi[a] = -g.lo[a]; }
    int axis_of[6];
    double plane[6], rlo[6][3], rhi[6][3];
    for (int k = 0; k < 6; ++k) {
      const BakedPrim& p = out[first + k];
      if (p.kind != PRIM_QUAD) return;
      int ua = -1, va = -1;
      for (int a = 0; a < 3; ++a) {
        if (p.u[a] != 0.0) { if (ua >= 0) return; ua = a; }
        if (p.v[a] != 0.0) { if (va >= 0) return; va = a; }
      }
      if (ua < 0 || va < 0 || ua == va) return;
      axis_of[k] = 3 - ua - va;
      plane[k] = p.Q[axis_of[k]];
      D3 far = p.Q + p.u + p.v;
      for (int a = 0; a < 3; ++a) {
        rlo[k][a] = std::min(p.Q[a], far[a]); rhi[k][a] = std::max(p.Q[a], far[a]);
        g.lo[a] = std::min(g.lo[a], rlo[k][a]); g.hi[a] = std::max(g.hi[a], rhi[k][a]);
      }
    }
    for (int a = 0; a < 3; ++a) if (!(g.hi[a] > g.lo[a])) return;
    for (int s = 0; s < 6; ++s) g.member[s] = -1;
    for (int k = 0; k < 6; ++k) {
      const int ax = axis_of[k];
      const double tol = 1e-9 * std::max(1.0, std::fabs(g.hi[ax]) + std::fabs(g.lo[ax]));
      int side;
      if (std::fabs(plane[k] - g.lo[ax]) <= tol) side = 0;
      else if (std::fabs(plane[k] - g.hi[ax]) <= tol) side = 1;
      else return;
      for (int a = 0; a < 3; ++a) {   // the face must span the whole box in its own two axes
        if (a == ax) continue;
        const double t2 = 1e-9 * std::max(1.0, std::fabs(g.hi[a]) + std::fabs(g.lo[a]));
        if (std::fabs(rlo[k][a] - g.lo[a]) > t2 || std::fabs(rhi[k][a] - g.hi[a]) > t2) return;
      }
      if (g.member[2 * ax + side] >= 0) return;
      g.member[2 * ax + side] = (int)(first + k);
    }
    for (int s = 0; s < 6; ++s) {
      out[g.member[s]].box = (int)boxes.size();
      out[g.member[s]].face = s;
    }
    boxes.push_back(g);
  }

  bool walk(int idx, const Xform& X, std::vector<BakedPrim>& out, int depth, bool in_boundary) {
    if (idx < 0 || idx >= d->n_hittables) return fail(RRT_ERR_INVALID, "hittable index out of range");
    if (depth > 64) return fail(RRT_ERR_INVALID, "scene graph deeper than 64 (cycle?)");
    const rrt_hittable& h = d->hittables[idx];
    switch (h.kind) {
      case RRT_H_SPHERE: case RRT_H_QUAD: {
        if (h.material < 0 || h.material >= d->n_materials)
          return fail(RRT_ERR_INVALID, "primitive material index out of range");
        BakedPrim p;
        if (h.kind == RRT_H_SPHERE) bake_sphere(h, idx, X, p); else bake_quad(h, idx, X, p);
        p.order = (int)out.size();
        for (int a = 0; a < 3; ++a)
          if (!std::isfinite(p.lo[a]) || !std::isfinite(p.hi[a]))
            return fail(RRT_ERR_INVALID, "primitive with non-finite bounds (degenerate quad?)");
        out.push_back(p);
        return true;
      }
      case RRT_H_LIST: case RRT_H_BVH: {
        if (h.first_child < 0 || h.n_children < 0 || h.first_child + h.n_children > d->n_children)
          return fail(RRT_ERR_INVALID, "list children range out of bounds");
        std::vector<int> kids(d->children + h.first_child, d->children + h.first_child + h.n_children);
        for (int k : kids)
          if (k < 0 || k >= d->n_hittables) return fail(RRT_ERR_INVALID, "hittable index out of range");
        if (h.kind == RRT_H_BVH && kids.size() > 2) {  // visit in the order BvhNode::hit would
          std::vector<int> ordered;
          ref.build(kids, 0, kids.size(), ordered);
          kids.swap(ordered);
        }
        const size_t before = out.size();
        for (int k : kids)
          if (!walk(k, X, out, depth + 1, in_boundary)) return false;
        if (!in_boundary && kids.size() == 6 && out.size() == before + 6) find_box(out, before);
        return true;
      }
      case RRT_H_TRANSLATE: {  // hittable.rs:86-93: p_parent = p_child + offset
        Xform Y = X;
        Y.t = X.point(D3(h.a));
        return walk(h.first_child, Y, out, depth + 1, in_boundary);
      }
      case RRT_H_ROTATE_Y: {  // hittable.rs:107-112
        double radians = h.a[0] * M_PI / 180.0;
        double s2 = std::sin(radians), c2 = std::cos(radians);
        Xform Y = X;
        Y.c = X.c * c2 - X.s * s2;
        Y.s = X.c * s2 + X.s * c2;
        Y.rotated = true;
        return walk(h.first_child, Y, out, depth + 1, in_boundary);
      }
      case RRT_H_CONSTANT_MEDIUM: {
        if (in_boundary) return fail(RRT_ERR_UNSUPPORTED, "constant medium nested inside a medium boundary");
        if (h.material < 0 || h.material >= d->n_materials)
          return fail(RRT_ERR_INVALID, "medium phase material index out of range");
        BakedMedium m;
        m.first = (int)boundary.size(); m.mat = h.material; m.id = idx; m.density = h.a[0];
        if (!walk(h.first_child, X, boundary, depth + 1, true)) return false;
        m.count = (int)boundary.size() - m.first;
        media.push_back(m);
        return true;
      }
      default:
        return fail(RRT_ERR_INVALID, "unknown hittable kind");
    }
  }
};

// ------------------------------------------------------------------------------------------------
// BVH: binary, full-sweep SAH over the three axes (host, once per scene).  Not the reference's
// 12-bucket builder (bvh.rs:34-156) — the closest hit does not depend on the tree.
// Tree-shape constants; overridable at build time by scripts/warp_sim.cpp (the CPU model of the
// warp-level traversal cost), which compiles this file with other values to compare trees.
#ifndef RRT_SAH_NODE_COST
#define RRT_SAH_NODE_COST 0.3            // price of a node step against one primitive test
#endif
#ifndef RRT_FRONT_MAX
#define RRT_FRONT_MAX 4                  // primitives the front run may hold
#endif
#ifndef RRT_FRONT_SHARE_SPHERES
#define RRT_FRONT_SHARE_SPHERES 0.01     // box-area share of the remaining bounds, sphere-only scenes
#endif
#ifndef RRT_FRONT_SHARE_OTHER
#define RRT_FRONT_SHARE_OTHER 0.5        // the same with quads or boxes in the tree
#endif
#ifndef RRT_FRONT_MEDIAN_FACTOR
#define RRT_FRONT_MEDIAN_FACTOR 10.0     // and at least this many times the median item's box area
#endif
enum { ITEM_BOX = 2, FRONT_MAX = RRT_FRONT_MAX };  // build item standing for a BoxGroup (BakedPrim::kind beyond PRIM_SPHERE / PRIM_QUAD)

struct BNode {
  double lo[3], hi[3];
  int left = -1, right = -1;   // children (BNode indices) or -1
  int first = 0, count = 0;    // leaf range in `order`
};

struct Builder {
  const std::vector<BakedPrim>& P;
  std::vector<int> order;
  std::vector<BNode> nodes;
  int max_leaf;
  int max_depth_seen = 0;
  bool pair_leaves = false;   // two spheres left: one leaf (tested as a packed pair), whatever SAH says

  Builder(const std::vector<BakedPrim>& p, int ml) : P(p), max_leaf(ml) {
    order.resize(P.size());
    for (size_t i = 0; i < P.size(); ++i) order[i] = (int)i;
  }
  static double area(const double* lo, const double* hi) {
    double a = hi[0] - lo[0], b = hi[1] - lo[1], c = hi[2] - lo[2];
    return 2.0 * (a * b + a * c + b * c);
  }
  void bounds(int first, int count, double* lo, double* hi) const {
    for (int a = 0; a < 3; ++a) { lo[a] = std::numeric_limits<double>::infinity(); hi[a] = -lo[a]; }
    for (int i = first; i < first + count; ++i) {
      const BakedPrim& p = P[order[i]];
      for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], p.lo[a]); hi[a] = std::max(hi[a], p.hi[a]); }
    }
  }
  bool leaf_ok(int first, int count) const {
    int ns = 0, nq = 0;
    for (int i = first; i < first + count; ++i) {
      if (P[order[i]].kind == ITEM_BOX && count > 1) return false;  // a box group is a leaf of its own
      (P[order[i]].kind == PRIM_SPHERE ? ns : nq)++;
    }
    return ns <= LEAF_MAX_PER_KIND && nq <= LEAF_MAX_PER_KIND;
  }
  int build(int first, int count, int depth) {
    max_depth_seen = std::max(max_depth_seen, depth);
    int me = (int)nodes.size();
    nodes.push_back(BNode());
    bounds(first, count, nodes[me].lo, nodes[me].hi);
    nodes[me].first = first; nodes[me].count = count;
    if (count <= 1 || (count <= max_leaf && depth >= 40 && leaf_ok(first, count))) return me;
    if (pair_leaves && count == 2 && P[order[first]].kind == PRIM_SPHERE && P[order[first + 1]].kind == PRIM_SPHERE) return me;
    double parent_area = area(nodes[me].lo, nodes[me].hi);
    double best_cost = std::numeric_limits<double>::infinity();
    int best_axis = -1, best_split = -1;
    std::vector<double> right_area(count);
    std::vector<int> tmp(order.begin() + first, order.begin() + first + count);
    std::vector<int> best_order;
    std::vector<int> sorted_by[3];
    if (depth < 40) {
      for (int axis = 0; axis < 3; ++axis) {
        std::vector<int>& idx = sorted_by[axis];
        idx = tmp;
        std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) {
          return P[a].lo[axis] + P[a].hi[axis] < P[b].lo[axis] + P[b].hi[axis];
        });
        double lo[3], hi[3];
        for (int a = 0; a < 3; ++a) { lo[a] = std::numeric_limits<double>::infinity(); hi[a] = -lo[a]; }
        for (int i = count - 1; i > 0; --i) {
          const BakedPrim& p = P[idx[i]];
          for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], p.lo[a]); hi[a] = std::max(hi[a], p.hi[a]); }
          right_area[i] = area(lo, hi);
        }
        for (int a = 0; a < 3; ++a) { lo[a] = std::numeric_limits<double>::infinity(); hi[a] = -lo[a]; }
        for (int i = 1; i < count; ++i) {
          const BakedPrim& p = P[idx[i - 1]];
          for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], p.lo[a]); hi[a] = std::max(hi[a], p.hi[a]); }
          double cost = area(lo, hi) * i + right_area[i] * (count - i);
          if (cost < best_cost) { best_cost = cost; best_axis = axis; best_split = i; }
        }
      }
      if (best_axis >= 0) best_order.swap(sorted_by[best_axis]);
    }
    // leaf cost = count intersections; split cost = 1 traversal step + SAH-weighted children
    bool can_leaf = count <= max_leaf && leaf_ok(first, count);
    if (best_axis >= 0 && parent_area > 0) {
      // a leaf round costs the whole warp a reconvergence, so a node step is priced well below one
      // intersection: single-primitive leaves measured 1-2 % faster than SAH's usual 1.0 (profiles/README.md)
      double split_cost = RRT_SAH_NODE_COST + best_cost / parent_area;
      if (can_leaf && (double)count <= split_cost) return me;
    } else if (can_leaf) {
      return me;
    }
    int mid;
    if (best_axis < 0) {  // too deep or degenerate: median split on the longest axis
      int axis = 0;
      double ext = -1;
      for (int a = 0; a < 3; ++a) if (nodes[me].hi[a] - nodes[me].lo[a] > ext) { ext = nodes[me].hi[a] - nodes[me].lo[a]; axis = a; }
      best_order = tmp;
      std::stable_sort(best_order.begin(), best_order.end(), [&](int a, int b) {
        return P[a].lo[axis] + P[a].hi[axis] < P[b].lo[axis] + P[b].hi[axis];
      });
      mid = count / 2;
    } else {
      mid = best_split;
    }
    std::copy(best_order.begin(), best_order.end(), order.begin() + first);
    int l = build(first, mid, depth + 1);
    int r = build(first + mid, count - mid, depth + 1);
    nodes[me].left = l; nodes[me].right = r;
    return me;
  }
};

inline float round_up(double v) {
  float f = (float)v;
  if ((double)f < v) f = std::nextafterf(f, std::numeric_limits<float>::infinity());
  return f;
}
inline float as_float(int i) { float f; std::memcpy(&f, &i, 4); return f; }
inline float4 f4(double a, double b, double c, double d) { float4 r; r.x = (float)a; r.y = (float)b; r.z = (float)c; r.w = (float)d; return r; }
inline float4 i4(int a, int b, int c, int d) { float4 r; r.x = as_float(a); r.y = as_float(b); r.z = as_float(c); r.w = as_float(d); return r; }

void emit_prim32(const BakedPrim& p, std::vector<float4>& out) {
  if (p.kind == PRIM_SPHERE) {
    out.push_back(f4(p.c0.x, p.c0.y, p.c0.z, p.r));
    out.push_back(f4(p.dc.x, p.dc.y, p.dc.z, p.r > 0 ? 1.0 / p.r : 0.0));
    out.push_back(i4(p.mat, p.id, p.order, p.flags));
    out.push_back(f4(p.uvcos, p.uvsin, 0, 0));
  } else {
    D3 A = cross(p.v, p.w), B = cross(p.w, p.u);
    out.push_back(f4(p.n.x, p.n.y, p.n.z, p.d));
    out.push_back(f4(A.x, A.y, A.z, -dot(p.Q, A)));
    out.push_back(f4(B.x, B.y, B.z, -dot(p.Q, B)));
    out.push_back(i4(p.mat, p.id, p.order, p.flags));
  }
}
void emit_prim64(const BakedPrim& p, std::vector<double>& out) {
  double q[16] = {0};
  if (p.kind == PRIM_SPHERE) {
    q[0] = p.c0.x; q[1] = p.c0.y; q[2] = p.c0.z; q[3] = p.r; q[4] = p.dc.x; q[5] = p.dc.y; q[6] = p.dc.z;
  } else {
    q[0] = p.n.x; q[1] = p.n.y; q[2] = p.n.z; q[3] = p.d;
    q[4] = p.Q.x; q[5] = p.Q.y; q[6] = p.Q.z;
    q[7] = p.u.x; q[8] = p.u.y; q[9] = p.u.z;
    q[10] = p.v.x; q[11] = p.v.y; q[12] = p.v.z;
    q[13] = p.w.x; q[14] = p.w.y; q[15] = p.w.z;
  }
  out.insert(out.end(), q, q + 16);
}

// conservative f32 box as (center, half-extent): the half-extent is rounded up from the padded
// f64 box, the pad being relative so that it covers the f32 slab arithmetic
void box32(const double* lo, const double* hi, float* fc, float* fh) {
  for (int a = 0; a < 3; ++a) {
    double m = std::max(std::fabs(lo[a]), std::fabs(hi[a]));
    double pad = 4e-7 * m + 1e-7 * (hi[a] - lo[a]);
    float c = (float)(0.5 * (lo[a] + hi[a]));
    double need = std::max((hi[a] + pad) - (double)c, (double)c - (lo[a] - pad));
    fc[a] = c;
    fh[a] = round_up(need);
  }
}

bool tex_reads_uv(const rrt_scene_desc* d, int t, int depth) {  // only ImageTexture reads (u,v)
  if (t < 0 || t >= d->n_textures || depth > 8) return false;
  const rrt_texture& x = d->textures[t];
  if (x.kind == RRT_T_IMAGE) return true;
  if (x.kind == RRT_T_CHECKER) return tex_reads_uv(d, x.even, depth + 1) || tex_reads_uv(d, x.odd, depth + 1);
  return false;
}

bool tex_needs_ok(const rrt_scene_desc* d, int t, int depth) {
  if (t < 0) return true;
  if (t >= d->n_textures || depth > 8) return false;
  const rrt_texture& x = d->textures[t];
  if (x.kind == RRT_T_CHECKER) return tex_needs_ok(d, x.even, depth + 1) && tex_needs_ok(d, x.odd, depth + 1) && x.even >= 0 && x.odd >= 0;
  if (x.kind == RRT_T_IMAGE) return x.image >= 0 && x.image < d->n_images;
  if (x.kind == RRT_T_NOISE) return x.perlin >= 0 && x.perlin < d->n_perlins;
  return x.kind == RRT_T_SOLID;
}

// Uniform grids over even fields of equal items (device_types.h DevScene::grids).  Items of one kind and one
// size class (spheres by radius, boxes by their longest edge, classes a factor of two wide) that number at
// least RRT_GRID_MIN_ITEMS are a candidate field; its grid takes its resolution from the item density
// (RRT_GRID_LAMBDA cells per item) and references every item from each cell its (padded) bounds overlap.
// Accepted when balanced — few references per item, no crowded cell — because that is when "read the cell
// you stand in" beats "descend the hierarchy" (profiles/README.md r02k).  At most one grid of spheres and one
// of boxes.  RRT_NO_GRID=1 keeps the BVH, RRT_FORCE_GRID=1 skips the balance test, RRT_GRID_LAMBDA=x
// overrides the density.
#ifndef RRT_GRID_LAMBDA
#define RRT_GRID_LAMBDA 0.5
#endif
#ifndef RRT_GRID_LAMBDA_BOX
#define RRT_GRID_LAMBDA_BOX 0.5
#endif
#ifndef RRT_GRID_MIN_ITEMS
#define RRT_GRID_MIN_ITEMS 64
#endif
struct GridGeom {
  double lo[3], hi[3], cell[3];
  double lo0[3], hi0[3], lo1[3], hi1[3];   // the members' bounds at shutter time 0 and 1, padded like lo / hi
  int res[3];
  std::vector<std::vector<int>> lists;   // per cell: item indices
  size_t n_refs = 0, crowded = 0;
};
// members: indices into items.  Returns false when the field is not worth a grid.
bool grid_geometry(const std::vector<BakedPrim>& items, const std::vector<int>& members, bool force, GridGeom& G) {
  if (members.empty()) return false;
  double mean_ext[3] = {0, 0, 0};
  for (int a = 0; a < 3; ++a) { G.lo[a] = std::numeric_limits<double>::infinity(); G.hi[a] = -G.lo[a]; }
  for (int i : members)
    for (int a = 0; a < 3; ++a) {
      G.lo[a] = std::min(G.lo[a], items[i].lo[a]); G.hi[a] = std::max(G.hi[a], items[i].hi[a]);
      mean_ext[a] += (items[i].hi[a] - items[i].lo[a]) / (double)members.size();
    }
  const bool of_boxes = items[members[0]].kind == ITEM_BOX;
  const char* le = std::getenv(of_boxes ? "RRT_GRID_LAMBDA_BOX" : "RRT_GRID_LAMBDA");
  const double lambda = le ? std::atof(le) : (of_boxes ? RRT_GRID_LAMBDA_BOX : RRT_GRID_LAMBDA);
  // a flat field is one cell thick: an axis no longer than the items themselves gets the items' extent
  double ext[3], vol = 1;
  for (int a = 0; a < 3; ++a) { ext[a] = std::max(G.hi[a] - G.lo[a], std::max(mean_ext[a], 1e-9)); vol *= ext[a]; }
  // A dense cloud (the 1 000 overlapping spheres of the book-2 final scene fill 40 % of their bounds) stops a ray
  // within a cell or two, so what counts is how many items a cell lists: four times the cells.  Measured, Mpaths/s
  // at 500 spp: C4 2 998 / 3 099 / 3 139 at 0.5 / 1 / 2 cells per item, C2 (a sparse layer) 9 831 / 9 727 / 9 539 at 0.35 / 0.5 / 1.
  double fill = 0;
  if (!of_boxes) {
    for (int i : members) fill += 4.18879 * items[i].r * items[i].r * items[i].r;
    fill /= vol;
  }
  const double lambda_eff = (!le && fill > 0.25) ? 4.0 * lambda : lambda;
  const double dens = std::cbrt(lambda_eff * (double)members.size() / vol);
  // Boxes that tile an axis exactly (the 20 x 20 ground boxes of the book-2 final scene, the_next_week/mod.rs:309-327:
  // equal widths at whole multiples of the width) get cells that ARE the tiles along that axis: one box per cell
  // instead of the nine a padded membership would list.  The walk can then differ from the tree only for a ray that
  // passes within f32 rounding of a vertical box edge, where either answer is a rounding of the same silhouette.
  bool tiled[3] = {false, false, false};
  double tile_lo[3] = {0, 0, 0}, tile_w[3] = {0, 0, 0};
  if (of_boxes && !std::getenv("RRT_GRID_NO_TILING")) {
    for (int a = 0; a < 3; ++a) {
      const BakedPrim& f = items[members[0]];
      const double wd = f.dc[a] - f.c0[a];
      double flo = std::numeric_limits<double>::infinity(), fhi = -flo;
      bool same = wd > 0;
      for (int i : members) {
        same = same && std::fabs((items[i].dc[a] - items[i].c0[a]) - wd) <= 1e-9 * wd;
        flo = std::min(flo, items[i].c0[a]); fhi = std::max(fhi, items[i].dc[a]);
      }
      if (!same) continue;
      const double n = (fhi - flo) / wd;
      if (n < 2 || n > 128 || std::fabs(n - std::round(n)) > 1e-6) continue;
      bool on_lattice = true;
      for (int i : members) { const double k = (items[i].c0[a] - flo) / wd; on_lattice = on_lattice && std::fabs(k - std::round(k)) <= 1e-6; }
      if (!on_lattice) continue;
      tiled[a] = true; tile_lo[a] = flo; tile_w[a] = wd;
      G.res[a] = (int)std::lround(n); G.lo[a] = flo; G.hi[a] = fhi; G.cell[a] = wd;
    }
  }
  for (int a = 0; a < 3; ++a) {
    if (tiled[a]) continue;
    G.res[a] = std::max(1, std::min(128, (int)std::lround(ext[a] * dens)));
    const double pad = 1e-3 * ext[a] / G.res[a] + 1e-6 * std::max(std::fabs(G.lo[a]), std::fabs(G.hi[a]));   // the bounds, outwards
    G.lo[a] -= pad; G.hi[a] += pad;
    G.cell[a] = (G.hi[a] - G.lo[a]) / G.res[a];
  }
  for (int a = 0; a < 3; ++a) {   // bounds at the two ends of the shutter (moving spheres: c0 and c0 + dc), inside lo / hi
    G.lo0[a] = G.lo1[a] = std::numeric_limits<double>::infinity(); G.hi0[a] = G.hi1[a] = -G.lo0[a];
    for (int i : members) {
      const BakedPrim& p = items[i];
      if (p.kind == PRIM_SPHERE) {
        const double c0 = p.c0[a], c1 = p.c0[a] + p.dc[a];
        G.lo0[a] = std::min(G.lo0[a], c0 - p.r); G.hi0[a] = std::max(G.hi0[a], c0 + p.r);
        G.lo1[a] = std::min(G.lo1[a], c1 - p.r); G.hi1[a] = std::max(G.hi1[a], c1 + p.r);
      } else {
        G.lo0[a] = std::min(G.lo0[a], p.lo[a]); G.hi0[a] = std::max(G.hi0[a], p.hi[a]);
        G.lo1[a] = G.lo0[a]; G.hi1[a] = G.hi0[a];
      }
    }
    const double pad = tiled[a] ? 0.0 : 1e-3 * G.cell[a] + 1e-6 * std::max(std::fabs(G.lo[a]), std::fabs(G.hi[a]));
    G.lo0[a] = std::max(G.lo[a], G.lo0[a] - pad); G.lo1[a] = std::max(G.lo[a], G.lo1[a] - pad);
    G.hi0[a] = std::min(G.hi[a], G.hi0[a] + pad); G.hi1[a] = std::min(G.hi[a], G.hi1[a] + pad);
  }
  G.lists.assign((size_t)G.res[0] * G.res[1] * G.res[2], std::vector<int>());
  G.n_refs = 0; G.crowded = 0;
  for (int i : members) {
    int c0[3], c1[3];
    for (int a = 0; a < 3; ++a) {
      if (tiled[a]) { c0[a] = c1[a] = std::max(0, std::min(G.res[a] - 1, (int)std::lround((items[i].c0[a] - tile_lo[a]) / tile_w[a]))); continue; }
      // padded by a thousandth of a cell: the f32 walk may leave a cell that far from where f64 would
      const double pad = 1e-3 * G.cell[a] + 4e-6 * std::max(std::fabs(items[i].lo[a]), std::fabs(items[i].hi[a]));
      c0[a] = std::max(0, std::min(G.res[a] - 1, (int)std::floor((items[i].lo[a] - pad - G.lo[a]) / G.cell[a])));
      c1[a] = std::max(0, std::min(G.res[a] - 1, (int)std::floor((items[i].hi[a] + pad - G.lo[a]) / G.cell[a])));
    }
    // a sphere is listed where the SPHERE reaches, not where its box does: the cells at the corners of the box go
    // (a sixth of the references of the book-2 cloud, whose spheres are wider than its cells).  A moving sphere is
    // measured from the box of its centre's path, per axis — exact for motion along an axis, conservative otherwise.
    const BakedPrim& it = items[i];
    const bool ball = it.kind == PRIM_SPHERE && !std::getenv("RRT_GRID_BOX_MEMBERSHIP");
    for (int z = c0[2]; z <= c1[2]; ++z)
      for (int y = c0[1]; y <= c1[1]; ++y)
        for (int x = c0[0]; x <= c1[0]; ++x) {
          if (ball) {
            const int c[3] = {x, y, z};
            double d2 = 0;
            for (int a = 0; a < 3; ++a) {
              const double pad = 1e-3 * G.cell[a] + 4e-6 * std::fabs(it.c0[a]);
              const double lo = G.lo[a] + c[a] * G.cell[a] - pad, hi = G.lo[a] + (c[a] + 1) * G.cell[a] + pad;
              const double ca = std::min(it.c0[a], it.c0[a] + it.dc[a]), cb = std::max(it.c0[a], it.c0[a] + it.dc[a]);
              const double d = cb < lo ? lo - cb : (ca > hi ? ca - hi : 0.0);
              d2 += d * d;
            }
            if (d2 > it.r * it.r) continue;
          }
          G.lists[((size_t)z * G.res[1] + y) * G.res[0] + x].push_back(i); ++G.n_refs;
        }
  }
  for (const std::vector<int>& l : G.lists) G.crowded = std::max(G.crowded, l.size());
  if (std::getenv("RRT_GRID_DEBUG"))
    std::fprintf(stderr, "grid candidate: %zu items, %dx%dx%d cells, %.2f references per item, most in a cell %zu\n", members.size(),
                 G.res[0], G.res[1], G.res[2], (double)G.n_refs / (double)members.size(), G.crowded);
  if (G.crowded > 255 || G.n_refs >= (1u << 22)) return false;
  if (!force && ((double)G.n_refs > 16.0 * (double)members.size() || G.crowded > 64)) return false;   // not an even field
  return true;
}

// Tags the members of the chosen fields (BakedPrim::grid = 0 for the sphere grid, 1 for the box grid).  Grids are
// only adopted when everything else fits the front run (the kernel then walks no tree): returns the number adopted.
int choose_grids(std::vector<BakedPrim>& items) {
  if (std::getenv("RRT_NO_GRID") || std::getenv("RRT_NO_FRONT")) return 0;   // (what is not in a grid goes to the front run)
  const bool force = std::getenv("RRT_FORCE_GRID") != nullptr;
  int adopted = 0;
  for (int kind = 0; kind < 2; ++kind) {   // 0: spheres, 1: boxes
    std::map<int, std::vector<int>> classes;
    for (size_t i = 0; i < items.size(); ++i) {
      const BakedPrim& p = items[i];
      if (kind == 0 ? p.kind != PRIM_SPHERE : p.kind != ITEM_BOX) continue;
      double size = kind == 0 ? p.r : std::max(p.hi[0] - p.lo[0], std::max(p.hi[1] - p.lo[1], p.hi[2] - p.lo[2]));
      if (!(size > 0)) continue;
      classes[force ? 0 : (int)std::floor(std::log2(size))].push_back((int)i);
    }
    const std::vector<int>* best = nullptr;
    for (const auto& kv : classes) if (!best || kv.second.size() > best->size()) best = &kv.second;
    if (!best || (best->size() < (size_t)RRT_GRID_MIN_ITEMS && !force)) continue;
    GridGeom G;
    if (!grid_geometry(items, *best, force, G)) continue;
    for (int i : *best) items[i].grid = kind;
    ++adopted;
  }
  if (!adopted) return 0;
  int ns = 0, nq = 0, nb = 0;
  for (const BakedPrim& p : items) if (p.grid < 0) (p.kind == PRIM_SPHERE ? ns : p.kind == ITEM_BOX ? nb : nq)++;
  if (ns > LEAF_MAX_PER_KIND || nq > LEAF_MAX_PER_KIND || nb > 0) {   // the rest does not fit the front run
    for (BakedPrim& p : items) p.grid = -1;
    return 0;
  }
  return adopted;
}

// After the primitives are emitted: the chosen grids with their cell tables (item_ref: build item -> flat sphere
// index or box record).
void emit_grids(const std::vector<BakedPrim>& items, const std::vector<int>& item_ref, FlatScene& out) {
  out.grid_cells.clear(); out.grid_refs.clear(); out.grids.clear();
  const bool force = std::getenv("RRT_FORCE_GRID") != nullptr;
  std::vector<DevGrid> slots(RRT_MAX_GRIDS);
  int present = 0;
  for (int kind = 0; kind < RRT_MAX_GRIDS; ++kind) {
    DevGrid& D = slots[kind];
    std::memset(&D, 0, sizeof D);
    D.kind = kind;
    std::vector<int> members;
    for (size_t i = 0; i < items.size(); ++i) if (items[i].grid == kind) members.push_back((int)i);
    GridGeom G;
    if (members.empty() || !grid_geometry(items, members, force, G)) continue;
    D.first_cell = (int)out.grid_cells.size();
    D.n_items = (int)members.size();
    for (const std::vector<int>& l : G.lists) {
      out.grid_cells.push_back(((unsigned)out.grid_refs.size() << 8) | (unsigned)l.size());
      for (int i : l) out.grid_refs.push_back(item_ref[i]);
    }
    for (int a = 0; a < 3; ++a) {
      D.lo[a] = (float)G.lo[a]; D.hi[a] = (float)G.hi[a];
      // clip bounds rounded outwards (towards lo / hi, which they never leave)
      D.clo[a] = std::max(D.lo[a], std::nextafterf((float)G.lo0[a], -INFINITY)); D.chi[a] = std::min(D.hi[a], std::nextafterf((float)G.hi0[a], INFINITY));
      const float l1 = std::max(D.lo[a], std::nextafterf((float)G.lo1[a], -INFINITY)), h1 = std::min(D.hi[a], std::nextafterf((float)G.hi1[a], INFINITY));
      D.dclo[a] = l1 - D.clo[a]; D.dchi[a] = h1 - D.chi[a];
      if (D.dclo[a] != 0.0f || D.dchi[a] != 0.0f) { D.moving = 1; out.features |= SF_GRID_MOVING; }
      D.cell[a] = (float)G.cell[a]; D.inv[a] = (float)(1.0 / G.cell[a]);
      D.res[a] = G.res[a];
    }
    ++present;
  }
  if (!present) { out.grid_cells.clear(); out.grid_refs.clear(); return; }
  if (out.grid_refs.empty()) out.grid_refs.push_back(0);
  out.grids = slots;   // both slots, an absent one has n_items == 0
  out.features |= SF_GRID;
}

}  // namespace

int flatten_scene(const rrt_scene_desc* desc, FlatScene& out, std::string& err, int single_leaf_max, int max_leaf, bool tree_optional) {
  if (!desc) { err = "scene description is null"; return RRT_ERR_INVALID; }
  if (desc->n_hittables < 0 || (desc->n_hittables > 0 && !desc->hittables) ||
      (desc->n_children > 0 && !desc->children) || (desc->n_materials > 0 && !desc->materials) ||
      (desc->n_textures > 0 && !desc->textures) || (desc->n_images > 0 && !desc->images) ||
      (desc->n_perlins > 0 && !desc->perlins)) {
    err = "scene description has a null array with a positive count"; return RRT_ERR_INVALID;
  }
  Walker w(desc);
  Xform I;
  if (desc->n_hittables > 0 || desc->world >= 0) {
    if (!w.walk(desc->world, I, w.world, 0, false)) { err = w.err; return w.status; }
  }
  // ---- materials / textures / images / perlin
  for (int i = 0; i < desc->n_materials; ++i) {
    const rrt_material& m = desc->materials[i];
    if (m.kind < RRT_M_LAMBERTIAN || m.kind > RRT_M_EMPTY) { err = "unknown material kind"; return RRT_ERR_INVALID; }
    bool textured = m.kind == RRT_M_LAMBERTIAN || m.kind == RRT_M_DIFFUSE_LIGHT || m.kind == RRT_M_ISOTROPIC;
    if (textured && (m.texture < 0 || !tex_needs_ok(desc, m.texture, 0))) {
      err = "material texture index invalid"; return RRT_ERR_INVALID;
    }
    DevMaterial dm; std::memset(&dm, 0, sizeof dm);
    dm.kind = m.kind; dm.tex = textured ? m.texture : -1;
    dm.fuzz = (float)m.fuzz; dm.ior = (float)m.refraction_index;
    dm.albedo = f4(m.albedo[0], m.albedo[1], m.albedo[2], 0);
    if (textured && desc->textures[m.texture].kind == RRT_T_SOLID) {  // SolidColor fast path
      const rrt_texture& t = desc->textures[m.texture];
      dm.albedo = f4(t.color[0], t.color[1], t.color[2], 0);
      dm.tex = -1;
    }
    dm.albedo.w = (dm.tex >= 0 && tex_reads_uv(desc, dm.tex, 0)) ? 1.0f : 0.0f;  // needs sphere uv
    out.materials.push_back(dm);
  }
  for (int i = 0; i < desc->n_textures; ++i) {
    const rrt_texture& t = desc->textures[i];
    DevTexture dt; std::memset(&dt, 0, sizeof dt);
    dt.kind = t.kind;
    switch (t.kind) {
      case RRT_T_SOLID: dt.v = f4(t.color[0], t.color[1], t.color[2], 0); break;
      case RRT_T_CHECKER: dt.a = t.even; dt.b = t.odd; dt.v = f4(0, 0, 0, 1.0 / t.scale); break;  // texture.rs:51
      case RRT_T_IMAGE: dt.a = t.image; break;
      case RRT_T_NOISE: dt.a = t.perlin; dt.v = f4(0, 0, 0, t.scale); break;
      default: err = "unknown texture kind"; return RRT_ERR_INVALID;
    }
    out.textures.push_back(dt);
  }
  for (int i = 0; i < desc->n_images; ++i) {
    const rrt_image& im = desc->images[i];
    DevImage di; di.width = im.rgb ? im.width : 0; di.height = im.rgb ? im.height : 0;
    di.offset = (int)(out.texels.size() / 4); di.pad = 0;
    if (im.rgb) {
      size_t n = (size_t)im.width * im.height;
      size_t base = out.texels.size();
      out.texels.resize(base + n * 4);
      for (size_t k = 0; k < n; ++k) {
        out.texels[base + 4 * k + 0] = im.rgb[3 * k + 0];
        out.texels[base + 4 * k + 1] = im.rgb[3 * k + 1];
        out.texels[base + 4 * k + 2] = im.rgb[3 * k + 2];
        out.texels[base + 4 * k + 3] = 255;
      }
    }
    out.images.push_back(di);
  }
  for (int i = 0; i < desc->n_perlins; ++i) {
    const rrt_perlin& p = desc->perlins[i];
    out.perlins.emplace_back();
    DevPerlin& dp = out.perlins.back();
    for (int k = 0; k < 256; ++k) {
      dp.randvec[k] = f4(p.randvec[k][0], p.randvec[k][1], p.randvec[k][2], 0);
      dp.perm[0][k] = (uint8_t)(p.perm_x[k] & 255);
      dp.perm[1][k] = (uint8_t)(p.perm_y[k] & 255);
      dp.perm[2][k] = (uint8_t)(p.perm_z[k] & 255);
    }
  }
  // ---- media boundaries (outside the BVH; spheres first, then quads, per medium)
  for (const BakedMedium& m : w.media) {
    DevMedium dm; std::memset(&dm, 0, sizeof dm);
    dm.first = (int)(out.bprims.size() / 4);
    dm.mat = m.mat; dm.id = m.id;
    dm.neg_inv_density = (float)(-1.0 / m.density);  // constant_medium.rs:26
    for (int pass = 0; pass < 2; ++pass)
      for (int k = m.first; k < m.first + m.count; ++k)
        if (w.boundary[k].kind == pass) { emit_prim32(w.boundary[k], out.bprims); (pass == 0 ? dm.nsph : dm.nquad)++; }
    out.media.push_back(dm);
  }
  // ---- lights (book 3): a list of spheres/quads, or a single one
  if (desc->lights >= 0) {
    if (desc->lights >= desc->n_hittables) { err = "lights index out of range"; return RRT_ERR_INVALID; }
    std::vector<int> ids;
    const rrt_hittable& L = desc->hittables[desc->lights];
    if (L.kind == RRT_H_LIST) {
      if (L.first_child < 0 || L.first_child + L.n_children > desc->n_children) { err = "lights children out of range"; return RRT_ERR_INVALID; }
      for (int k = 0; k < L.n_children; ++k) ids.push_back(desc->children[L.first_child + k]);
    } else {
      ids.push_back(desc->lights);
    }
    for (int id : ids) {
      if (id < 0 || id >= desc->n_hittables) { err = "light index out of range"; return RRT_ERR_INVALID; }
      const rrt_hittable& h = desc->hittables[id];
      DevLight dl; std::memset(&dl, 0, sizeof dl);
      BakedPrim p;
      if (h.kind == RRT_H_SPHERE) {
        bake_sphere(h, id, I, p);
        dl.kind = 0; dl.a = f4(p.c0.x, p.c0.y, p.c0.z, p.r);
      } else if (h.kind == RRT_H_QUAD) {
        bake_quad(h, id, I, p);
        D3 A = cross(p.v, p.w), B = cross(p.w, p.u);
        dl.kind = 1;
        dl.a = f4(p.n.x, p.n.y, p.n.z, p.d);
        dl.b = f4(A.x, A.y, A.z, -dot(p.Q, A));
        dl.c = f4(B.x, B.y, B.z, -dot(p.Q, B));
        dl.q = f4(p.Q.x, p.Q.y, p.Q.z, p.area);
        dl.u = f4(p.u.x, p.u.y, p.u.z, 0);
        dl.v = f4(p.v.x, p.v.y, p.v.z, 0);
      } else {
        dl.kind = 2;  // trait defaults: pdf_value 0, random (1,0,0)  (hittable.rs:58-64)
      }
      out.lights.push_back(dl);
    }
  }
  // ---- BVH over the world prims
  const std::vector<BakedPrim>& P = w.world;
  out.n_prims = (int)P.size();
  for (const BakedPrim& p : P) if (p.kind == PRIM_SPHERE && (p.dc.x != 0 || p.dc.y != 0 || p.dc.z != 0)) out.has_motion = 1;
  for (const BakedPrim& p : w.boundary) if (p.kind == PRIM_SPHERE && (p.dc.x != 0 || p.dc.y != 0 || p.dc.z != 0)) out.has_motion = 1;
  if (out.has_motion) out.features |= SF_MOTION;
  for (const BakedPrim& p : P) if (p.kind != PRIM_SPHERE) out.features |= SF_QUADS;
  if (!out.media.empty()) out.features |= SF_MEDIA;
  for (const DevMaterial& m : out.materials) if (m.tex >= 0) out.features |= SF_TEXTURES;  // solid colours were folded into albedo
  if (out.features & SF_TEXTURES)
    for (const DevTexture& t : out.textures) if (t.kind == RRT_T_NOISE) out.features |= SF_NOISE;
  if (P.empty()) { out.root_ref = make_leaf_ref(0, 0, 0); return RRT_OK; }
  int ns = 0, nq = 0;
  for (const BakedPrim& p : P) (p.kind == PRIM_SPHERE ? ns : nq)++;
  bool single = (int)P.size() <= single_leaf_max && ns <= LEAF_MAX_PER_KIND && nq <= LEAF_MAX_PER_KIND;
  // build items: every primitive, except that the six quads of a box group are one item
  const bool use_boxes = !single && !w.boxes.empty() && !std::getenv("RRT_NO_BOX_LEAVES");
  std::vector<BakedPrim> items;
  std::vector<int> item_src;   // index into P, or -1 - box group index
  for (size_t i = 0; i < P.size(); ++i)
    if (!use_boxes || P[i].box < 0) { items.push_back(P[i]); item_src.push_back((int)i); }
  if (use_boxes) {
    out.features |= SF_BOXES;
    for (size_t g = 0; g < w.boxes.size(); ++g) {
      BakedPrim it;
      it.kind = ITEM_BOX; it.box = (int)g;
      for (int a = 0; a < 3; ++a) { it.lo[a] = w.boxes[g].lo[a]; it.hi[a] = w.boxes[g].hi[a]; }
      it.c0 = D3(w.boxes[g].lo[0], w.boxes[g].lo[1], w.boxes[g].lo[2]); it.dc = D3(w.boxes[g].hi[0], w.boxes[g].hi[1], w.boxes[g].hi[2]);   // the box itself (grid tiling test)
      for (int s = 0; s < 6; ++s)   // the faces' own (padded) bounds, as the reference's list box has them
        for (int a = 0; a < 3; ++a) { it.lo[a] = std::min(it.lo[a], P[w.boxes[g].member[s]].lo[a]); it.hi[a] = std::max(it.hi[a], P[w.boxes[g].member[s]].hi[a]); }
      items.push_back(it); item_src.push_back(-1 - (int)g);
    }
  }
  // front run: a primitive far larger than its neighbours (the r = 1000 ground sphere of the book-1/2
  // sphere scenes, mod.rs:59-63, and their three r = 1 spheres among 480 of r = 0.2) sits high in any
  // BVH as a leaf of its own and is reached by a large share of the rays, each time costing the warp a
  // leaf round.  Such primitives leave the tree and are tested ahead of the traversal by all lanes
  // together (DevScene::front_ref).  Criterion: box area >= 10x the median item's and >= a share of
  // the bounds of what is left (the surface-area estimate of "hit by a ray that enters the scene"):
  // 1 % in sphere scenes, where it measured +8-10 % on top of the ground sphere alone (C1, C2, C3a);
  // 50 % with quads or boxes in the tree, where taking the four largest spheres and the light out of
  // the book-2 final scene measured 1.6 % slower.  Repeated on what is left, because without the
  // ground sphere the bounds are those of the rest.  At most FRONT_MAX primitives.
  int n_front = 0;
  const int n_grids_chosen = single ? 0 : choose_grids(items);
  if (n_grids_chosen > 0) {   // everything outside the grids is the front run: the render kernel walks no tree
    std::vector<BakedPrim> it2; std::vector<int> src2;
    for (int pass = 0; pass < 2; ++pass)
      for (size_t i = 0; i < items.size(); ++i)
        if ((items[i].grid < 0) == (pass == 0)) { it2.push_back(items[i]); src2.push_back(item_src[i]); if (pass == 0) ++n_front; }
    items.swap(it2); item_src.swap(src2);
  } else if (!single && !std::getenv("RRT_NO_FRONT")) {
    bool spheres_only = true;
    for (const BakedPrim& p : items) if (p.kind != PRIM_SPHERE) spheres_only = false;
    const double share = spheres_only ? RRT_FRONT_SHARE_SPHERES : RRT_FRONT_SHARE_OTHER;
    std::vector<char> is_front(items.size(), 0);
    for (bool again = true; again && n_front < FRONT_MAX;) {
      again = false;
      double lo[3], hi[3];
      for (int a = 0; a < 3; ++a) { lo[a] = std::numeric_limits<double>::infinity(); hi[a] = -lo[a]; }
      std::vector<double> areas;
      for (size_t i = 0; i < items.size(); ++i) {
        if (is_front[i]) continue;
        areas.push_back(Builder::area(items[i].lo, items[i].hi));
        for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], items[i].lo[a]); hi[a] = std::max(hi[a], items[i].hi[a]); }
      }
      if (areas.size() <= 2) break;
      const double total = Builder::area(lo, hi);
      std::nth_element(areas.begin(), areas.begin() + areas.size() / 2, areas.end());
      const double median = areas[areas.size() / 2];
      for (size_t i = 0; i < items.size() && n_front < FRONT_MAX; ++i) {
        const double ar = Builder::area(items[i].lo, items[i].hi);
        if (!is_front[i] && items[i].kind != ITEM_BOX && ar >= share * total && ar >= RRT_FRONT_MEDIAN_FACTOR * median) { is_front[i] = 1; ++n_front; again = true; }
      }
    }
    if (n_front > 0) {  // front items first, construction order kept inside both groups
      std::vector<BakedPrim> it2; std::vector<int> src2;
      for (int pass = 1; pass >= 0; --pass)
        for (size_t i = 0; i < items.size(); ++i)
          if (is_front[i] == pass) { it2.push_back(items[i]); src2.push_back(item_src[i]); }
      items.swap(it2); item_src.swap(src2);
    }
  }
  std::vector<BakedPrim> front(items.begin(), items.begin() + n_front);
  items.erase(items.begin(), items.begin() + n_front);
  Builder B(items, single ? (int)items.size() : max_leaf);
  // (scripts/warp_sim.cpp only, with SIM_MAX_LEAF=2: the render kernel tests one primitive per leaf, so never at max_leaf 1)
  B.pair_leaves = max_leaf >= 2 && std::getenv("RRT_PAIR_LEAVES") != nullptr;
  int root = 0;
  if (single) {
    B.nodes.push_back(BNode());
    B.bounds(0, (int)items.size(), B.nodes[0].lo, B.nodes[0].hi);
    B.nodes[0].first = 0; B.nodes[0].count = (int)items.size();
  } else if (tree_optional && n_grids_chosen > 0) {
    // The caller will only render (rrt_render_frame: upload, render, free) and the scene is walked through grids: the
    // tree would serve the f64 probe alone.  Every item a leaf of its own, in construction order; no inner nodes.
    for (size_t i = 0; i < items.size(); ++i) { BNode n; n.first = (int)i; n.count = 1; B.nodes.push_back(n); }
    root = -1;
  } else {
    root = B.build(0, (int)items.size(), 0);
  }
  out.bvh_depth = B.max_depth_seen;
  // leaf -> contiguous prim run (spheres first, then quads); prims emitted in `order` sequence
  std::vector<int> leaf_ref(B.nodes.size(), 0);
  std::vector<int> item_flat(items.size(), -1);   // tree item -> flat primitive index (box groups: their box record)
  {
    // emit prims leaf by leaf in B.order sequence (DFS), reordering within the leaf by kind
    int emitted = 0;
    {
      int s = 0, q = 0;
      for (int pass = 0; pass < 2; ++pass)
        for (const BakedPrim& p : front)
          if (p.kind == pass) { emit_prim32(p, out.prims); emit_prim64(p, out.prims64); (pass == 0 ? s : q)++; }
      out.front_ref = make_leaf_ref(0, s, q);
      emitted = n_front;
    }
    // B.order ranges of leaves are disjoint and cover [0,n); walk nodes, leaves own [first,count)
    std::vector<int> leaves;
    for (size_t i = 0; i < B.nodes.size(); ++i) if (B.nodes[i].left < 0) leaves.push_back((int)i);
    std::sort(leaves.begin(), leaves.end(), [&](int a, int b) { return B.nodes[a].first < B.nodes[b].first; });
    for (int li : leaves) {
      const BNode& n = B.nodes[li];
      if (n.count == 1 && items[B.order[n.first]].kind == ITEM_BOX) {  // box leaf: six quads in face order + a box record
        const BoxGroup& g = w.boxes[items[B.order[n.first]].box];
        float lo[3], hi[3], fc[3], fh[3];
        box32(items[B.order[n.first]].lo, items[B.order[n.first]].hi, fc, fh);   // padded, rounded outward
        for (int a = 0; a < 3; ++a) { lo[a] = fc[a] - fh[a]; hi[a] = fc[a] + fh[a]; }
        float4 b0 = f4(lo[0], lo[1], lo[2], 0), b1 = f4(hi[0], hi[1], hi[2], 0);
        std::memcpy(&b0.w, &emitted, 4);
        leaf_ref[li] = ~(((int)(out.boxes.size() / 2) << 8) | LEAF_BOX_MARK);
        item_flat[B.order[n.first]] = (int)(out.boxes.size() / 2);   // a box item is known by its record
        out.boxes.push_back(b0); out.boxes.push_back(b1);
        for (int s = 0; s < 6; ++s) { emit_prim32(P[g.member[s]], out.prims); emit_prim64(P[g.member[s]], out.prims64); }
        emitted += 6;
        continue;
      }
      int s = 0, q = 0;
      for (int pass = 0; pass < 2; ++pass) {
        std::vector<int> run;
        for (int k = n.first; k < n.first + n.count; ++k)
          if (items[B.order[k]].kind == pass) run.push_back(B.order[k]);
        // quads in the reference's visiting order: among quads a later one wins an exact-t tie (quad.rs:72
        // `contains`), so a leaf that tests them in that order with `t <= closest` needs no tie lookup
        // (quad_pairs_hit32 relies on it; quad_hit32's explicit rule agrees with it)
        if (pass == 1) std::stable_sort(run.begin(), run.end(), [&](int a, int b) { return items[a].order < items[b].order; });
        // Brute-force root leaf: axis-aligned rectangles first, grouped by axis (rect_quads_hit32 tests one in a dozen
        // instructions), the other quads after them; inside a group the reference's order stays.  Two quads that share
        // a plane (a box standing on the floor) keep the reference's equal-t rule — the later one wins — as long as
        // they are tested in that order: not when a rectangle comes after a coplanar quad that is none.
        if (pass == 1 && single && !std::getenv("RRT_NO_RECT_QUADS")) {
          bool coplanar = false;   // ... in the one order the grouping would reverse
          for (size_t x = 0; x < run.size() && !coplanar; ++x)
            for (size_t y = x + 1; y < run.size() && !coplanar; ++y) {
              const BakedPrim& A = items[run[x]]; const BakedPrim& Bq = items[run[y]];   // A before Bq in the reference
              const double c = dot(A.n, Bq.n), scale = 1.0 + std::fabs(A.d) + std::fabs(Bq.d);
              const bool same_plane = std::fabs(std::fabs(c) - 1.0) < 1e-12 && std::fabs(A.d - (c > 0 ? Bq.d : -Bq.d)) < 1e-9 * scale;
              if (same_plane && rect_axis(A) < 0 && rect_axis(Bq) >= 0) coplanar = true;
            }
          if (!coplanar) {
            std::stable_sort(run.begin(), run.end(), [&](int a, int b) { return (unsigned)rect_axis(items[a]) < (unsigned)rect_axis(items[b]); });
            for (int i : run) {
              const int ax = rect_axis(items[i]);
              if (ax < 0) break;
              const BakedPrim& q = items[i];
              const int b = (ax + 1) % 3, c = (ax + 2) % 3;
              const double e_b = q.u[b] + q.v[b], e_c = q.u[c] + q.v[c];   // one of u, v runs along b, the other along c
              out.rects.push_back(f4(q.Q[ax], q.Q[b] + 0.5 * e_b, q.Q[c] + 0.5 * e_c, 0));
              out.rects.push_back(f4(0.5 * std::fabs(e_b), 0.5 * std::fabs(e_c), 0, 0));
              out.n_rects[ax]++;
            }
          }
        }
        for (int i : run) { item_flat[i] = (int)(out.prims.size() / 4); emit_prim32(items[i], out.prims); emit_prim64(items[i], out.prims64); (pass == 0 ? s : q)++; }
      }
      leaf_ref[li] = make_leaf_ref(emitted, s, q);
      emitted += n.count;
    }
  }
  if (root < 0) {   // no tree (tree_optional): grids and front run only
    out.root_ref = 0;
    emit_grids(items, item_flat, out);
    return out.grids.empty() ? RRT_ERR_INVALID : RRT_OK;
  }
  if (B.nodes[root].left < 0) {
    out.root_ref = leaf_ref[root];
    // Brute-force root leaf: its quads again as pair records for the packed plane tests of
    // quad_pairs_hit32 (device_types.h); an odd count is padded with a quad of normal 0, which the
    // |denom| >= 1e-8 test of quad.rs:67 rejects.
    const int leaf = ~out.root_ref, first = leaf >> 8, ns = (leaf >> 4) & 15, nq = leaf & 15;
    const int n_rect = out.n_rects[0] + out.n_rects[1] + out.n_rects[2];   // the first n_rect quads are rectangles (out.rects)
    if ((leaf & 0xFF) != LEAF_BOX_MARK && nq > 0 && !std::getenv("RRT_NO_QUAD_PAIRS")) {
      for (int k = n_rect; k < nq; k += 2) {
        float4 a[3], b[3];
        for (int j = 0; j < 3; ++j) {
          a[j] = out.prims[4 * (first + ns + k) + j];
          b[j] = k + 1 < nq ? out.prims[4 * (first + ns + k + 1) + j] : f4(0, 0, 0, 0);
        }
        for (int j = 0; j < 3; ++j) {   // (x0, x1, y0, y1) / (z0, z1, w0, w1) of plane, alpha and beta rows
          out.qpairs.push_back(f4(a[j].x, b[j].x, a[j].y, b[j].y));
          out.qpairs.push_back(f4(a[j].z, b[j].z, a[j].w, b[j].w));
        }
      }
    }
    return RRT_OK;
  }
  // inner nodes breadth-first
  std::vector<int> inner_index(B.nodes.size(), -1);
  std::vector<int> bfs;
  {
    std::queue<int> q; q.push(root);
    while (!q.empty()) {
      int n = q.front(); q.pop();
      inner_index[n] = (int)bfs.size(); bfs.push_back(n);
      int ch[2] = {B.nodes[n].left, B.nodes[n].right};
      for (int c : ch) if (B.nodes[c].left >= 0) q.push(c);
    }
  }
  for (int n : bfs) {
    const BNode& l = B.nodes[B.nodes[n].left];
    const BNode& r = B.nodes[B.nodes[n].right];
    float lc[3], lh[3], rc[3], rh[3];
    box32(l.lo, l.hi, lc, lh); box32(r.lo, r.hi, rc, rh);
    int ref0 = l.left >= 0 ? inner_index[B.nodes[n].left] : leaf_ref[B.nodes[n].left];
    int ref1 = r.left >= 0 ? inner_index[B.nodes[n].right] : leaf_ref[B.nodes[n].right];
    float4 rec[3];
    node_set_child(rec, 0, lc, lh); node_set_child(rec, 1, rc, rh);
    out.nodes.push_back(rec[0]); out.nodes.push_back(rec[1]); out.nodes.push_back(rec[2]);
    out.nodes.push_back(i4(ref0, ref1, 0, 0));
  }
  out.root_ref = 0;
  if (n_grids_chosen > 0) emit_grids(items, item_flat, out);
#ifdef RRT_BVH4
  // Experiment (profiles/README.md, r02): the binary tree collapsed to 4-wide nodes — a node's children
  // are its grandchildren (a leaf child stays one child) — so that a traversal step tests four boxes with
  // two packed slab tests.  Same leaves, same boxes; breadth-first like the binary array.
  {
    std::vector<int> index4(B.nodes.size(), -1), bfs4;
    auto kids = [&](int n, int* k) {   // up to four BNode indices
      int c = 0;
      const int ch[2] = {B.nodes[n].left, B.nodes[n].right};
      for (int s = 0; s < 2; ++s) {
        if (B.nodes[ch[s]].left < 0) k[c++] = ch[s];
        else { k[c++] = B.nodes[ch[s]].left; k[c++] = B.nodes[ch[s]].right; }
      }
      return c;
    };
    std::queue<int> q; q.push(root);
    while (!q.empty()) {
      int n = q.front(); q.pop();
      index4[n] = (int)bfs4.size(); bfs4.push_back(n);
      int k[4]; int c = kids(n, k);
      for (int i = 0; i < c; ++i) if (B.nodes[k[i]].left >= 0) q.push(k[i]);
    }
    for (int n : bfs4) {
      int k[4]; const int c = kids(n, k);
      float4 rec[8];
      std::memset(rec, 0, sizeof rec);
      int refs[4] = {0, 0, 0, 0};
      for (int i = 0; i < 4; ++i) {
        float cc[3] = {0, 0, 0}, hh[3] = {-1e30f, -1e30f, -1e30f};   // an empty slot: never hit
        if (i < c) {
          box32(B.nodes[k[i]].lo, B.nodes[k[i]].hi, cc, hh);
          refs[i] = B.nodes[k[i]].left >= 0 ? index4[k[i]] : leaf_ref[k[i]];
        }
        node_set_child(rec + (i < 2 ? 0 : 3), i & 1, cc, hh);
      }
      rec[6] = i4(refs[0], refs[1], refs[2], refs[3]);
      for (int i = 0; i < 8; ++i) out.nodes4.push_back(rec[i]);
    }
  }
#endif
  return RRT_OK;
}

// Camera::initialize, the_rest_of_your_life/camera.rs:108-154 (book 1/2 copies are the same apart
// from sqrt_spp).  All f64; the f32 block keeps pixel00 relative to the centre.
int camera_initialize(const rrt_camera* cam, DevCamera& c32, DevCamera64& c64) {
  int image_height = (int)((double)cam->image_width / cam->aspect_ratio);
  if (image_height < 1) image_height = 1;
  int sqrt_spp = (int)std::sqrt((double)cam->samples_per_pixel);
  double recip_sqrt_spp = 1.0 / (double)sqrt_spp;
  D3 center(cam->lookfrom);
  double theta = cam->vfov * M_PI / 180.0;
  double h = std::tan(theta / 2.0);
  double viewport_height = 2.0 * h * cam->focus_dist;
  double viewport_width = viewport_height * ((double)cam->image_width / (double)image_height);
  D3 wv = D3(cam->lookfrom) - D3(cam->lookat);
  double wl = std::sqrt(dot(wv, wv));
  D3 w(wv.x / wl, wv.y / wl, wv.z / wl);
  D3 uc = cross(D3(cam->vup), w);
  double ul = std::sqrt(dot(uc, uc));
  D3 u(uc.x / ul, uc.y / ul, uc.z / ul);
  D3 v = cross(w, u);
  D3 viewport_u = viewport_width * u;
  D3 viewport_v = viewport_height * D3(-v.x, -v.y, -v.z);
  D3 du(viewport_u.x / cam->image_width, viewport_u.y / cam->image_width, viewport_u.z / cam->image_width);
  D3 dv(viewport_v.x / image_height, viewport_v.y / image_height, viewport_v.z / image_height);
  D3 half_u(viewport_u.x / 2.0, viewport_u.y / 2.0, viewport_u.z / 2.0);
  D3 half_v(viewport_v.x / 2.0, viewport_v.y / 2.0, viewport_v.z / 2.0);
  D3 upper_left = center - (cam->focus_dist * w) - half_u - half_v;
  D3 pixel00 = upper_left + 0.5 * (du + dv);
  double defocus_radius = cam->focus_dist * std::tan((cam->defocus_angle / 2.0) * M_PI / 180.0);
  D3 ddu = defocus_radius * u, ddv = defocus_radius * v;  // u * r in the reference; commutative
  D3 rel = pixel00 - center;
  c32.cx = (float)center.x; c32.cy = (float)center.y; c32.cz = (float)center.z;
  c32.p00x = (float)rel.x; c32.p00y = (float)rel.y; c32.p00z = (float)rel.z;
  c32.dux = (float)du.x; c32.duy = (float)du.y; c32.duz = (float)du.z;
  c32.dvx = (float)dv.x; c32.dvy = (float)dv.y; c32.dvz = (float)dv.z;
  c32.ddux = (float)ddu.x; c32.dduy = (float)ddu.y; c32.dduz = (float)ddu.z;
  c32.ddvx = (float)ddv.x; c32.ddvy = (float)ddv.y; c32.ddvz = (float)ddv.z;
  c32.bgx = (float)cam->background[0]; c32.bgy = (float)cam->background[1]; c32.bgz = (float)cam->background[2];
  c32.recip_sqrt_spp = (float)recip_sqrt_spp;
  c32.sqrt_spp = sqrt_spp;
  c32.defocus = cam->defocus_angle > 0.0 ? 1 : 0;
  for (int a = 0; a < 3; ++a) { c64.c[a] = center[a]; c64.p00[a] = pixel00[a]; c64.du[a] = du[a]; c64.dv[a] = dv[a]; }
  return image_height;
}

}  // namespace rrt
