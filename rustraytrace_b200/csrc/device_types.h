// device_types.h — HBM layout of a flattened scene (shared by the host flattener and the kernels).
//
// Everything the reference keeps as an Arc<enum> pointer tree (hittable.rs:180-188) is baked into
// flat arrays:
//   prims   : one 64-byte record (4 x float4) per world-space leaf primitive, stored in BVH-leaf
//             order so a leaf is a contiguous run; within a leaf spheres come first, then quads.
//   nodes   : one 64-byte record per inner BVH node holding BOTH children's boxes, breadth-first
//             (the hot top of the tree is contiguous and stays L1-resident).
//   bprims  : boundary primitives of constant media (same record format), outside the BVH.
//   prims64 : f64 twin of prims (16 doubles per record) for the probe kernel.
#pragma once
#include <stdint.h>


#if defined(__CUDACC__)
#include <cuda_runtime.h>
#else
struct float4 { float x, y, z, w; };
#endif

namespace rrt {

// ---- prim record (float4 p[4]) -------------------------------------------------------------
// sphere: p0 = (c0.x, c0.y, c0.z, r)       p1 = (dc.x, dc.y, dc.z, 1/r)
//         p2 = (mat, id, order, flags) as int bits   p3 = (uv_cos, uv_sin, 0, 0)
// quad:   p0 = (n.x, n.y, n.z, d)           p1 = (A.x, A.y, A.z, a0)   alpha = p.A + a0
//         p2 = (B.x, B.y, B.z, b0)          p3 = (mat, id, order, flags) as int bits
//   with A = v x w, B = w x u, a0 = -Q.A, b0 = -Q.B  (quad.rs:76-79: alpha = w.(hp x v) = hp.(v x w))
// f64 twin (double q[16]):
// sphere: q0..2 = c0, q3 = r, q4..6 = dc
// quad:   q0..2 = n, q3 = d, q4..6 = Q, q7..9 = u, q10..12 = v, q13..15 = w   (quad.rs formulas verbatim)
// quad pair record (float4 q[6]), only for the quads of a brute-force root leaf (scenes of <= 16
// primitives: the Cornell boxes): two quads side by side in every 64-bit pair, so that the plane,
// alpha and beta equations of both are packed FMAs (fma.rn.f32x2) and the whole leaf is straight-line
// code with predicated accepts instead of three divergent early-outs per quad.
//   q0 = (n0.x, n1.x, n0.y, n1.y)  q1 = (n0.z, n1.z, d0, d1)
//   q2 = (A0.x, A1.x, A0.y, A1.y)  q3 = (A0.z, A1.z, a0_0, a0_1)
//   q4 = (B0.x, B1.x, B0.y, B1.y)  q5 = (B0.z, B1.z, b0_0, b0_1)
// Pair k holds quads first+2k and first+2k+1 of the leaf's quad run; a missing partner has n = 0.
enum PrimKind { PRIM_SPHERE = 0, PRIM_QUAD = 1 };

// Numbering of the 32x32-pixel super tiles for tile sharding (RRT_SHARD_TILE): tile k belongs to rank k % world and
// is that rank's (k / world)-th.  Row sy of the frame is numbered from column sy on (wrapping), so that a rank's
// tiles run along diagonals: numbered straight across, a frame whose tile row is a multiple of the rank count (3840
// pixels = 120 tiles, 8 ranks) gave every rank whole columns, and the Cornell box 5 % of imbalance (profiles/README.md r02r).
#if defined(__CUDACC__)
#define RRT_HD __host__ __device__
#else
#define RRT_HD
#endif
RRT_HD inline int super_tile_index(int sx, int sy, int stiles_x) {
  int c = sx - sy % stiles_x;
  if (c < 0) c += stiles_x;
  return sy * stiles_x + c;
}
RRT_HD inline void super_tile_xy(int k, int stiles_x, int& sx, int& sy) {
  sy = k / stiles_x;
  sx = (k % stiles_x + sy % stiles_x) % stiles_x;
}
enum PrimFlags { PF_UV_ROTATED = 1 };

// ---- BVH inner node (float4 n[4]) ------------------------------------------------------------
// boxes are (center c, half-extent e): the slab planes are c*(1/d) - o*(1/d) -+ e*|1/d|, three FMAs
// per axis and no min/max pairs (the ALU pipe, not the FMA pipe, was the busy one: profiles/).
// The two children sit side by side in every 64-bit pair so one packed FMA (fma.rn.f32x2, SASS
// FFMA2) serves both: as 12 floats, f[2*axis + child] = centre, f[6 + 2*axis + child] = half-extent.
// n0 = (c0.c.x, c1.c.x, c0.c.y, c1.c.y)   n1 = (c0.c.z, c1.c.z, c0.e.x, c1.e.x)
// n2 = (c0.e.y, c1.e.y, c0.e.z, c1.e.z)   n3 = (ref0, ref1, 0, 0) as int bits
// ref >= 0: inner node index.  ref < 0: leaf, ~ref = (first << 8) | (n_spheres << 4) | n_quads.
// Box leaf: ~ref = (box << 8) | 0xFF.  The six quads of an axis-aligned box (quad.rs box_()) are one leaf;
// boxes[2*box] = (lo.xyz, first as int bits), boxes[2*box+1] = (hi.xyz, 0); the quads sit at
// first .. first+5 in face order -x, +x, -y, +y, -z, +z.  The f32 kernel slab-tests the box, runs the
// real quad test on the entry face only, and falls back to all six whenever that is not conclusive.
// Traversal stack entries per lane (intersect.cuh).  The flattener switches to median splits at depth 40
// (flatten.cpp), so a tree over fewer than 2^24 primitives is shallower than this; rrt_scene_create
// refuses a deeper one (RRT_ERR_UNSUPPORTED) instead of letting the kernel drop a subtree.
#define RRT_STACK 64
static const int LEAF_BOX_MARK = 0xFF;
static const int LEAF_MAX_PER_KIND = 15;
inline
#if defined(__CUDACC__)
__host__ __device__
#endif
int make_leaf_ref(int first, int nsph, int nquad) { return ~((first << 8) | (nsph << 4) | nquad); }

// accessors for the host flattener, the flatten check and the f64 probe (the render kernel reads pairs)
inline
#if defined(__CUDACC__)
__host__ __device__
#endif
void node_set_child(float4* n, int child, const float c[3], const float e[3]) {
  float* f = reinterpret_cast<float*>(n);
  for (int a = 0; a < 3; ++a) { f[2 * a + child] = c[a]; f[6 + 2 * a + child] = e[a]; }
}

// ---- scene features ------------------------------------------------------------------------------
// What a flattened scene needs from the render kernel.  render_kernel<INTEGRATOR, FEAT> is compiled
// for a few feature sets and a scene runs on the smallest one covering it (kernels.cu).
enum SceneFeature {
  SF_MOTION = 1,    // a sphere with center1 != center0 (ray time, sphere.rs:70)
  SF_QUADS = 2,     // quads in the BVH
  SF_MEDIA = 4,     // constant media
  SF_TEXTURES = 8,  // a material whose texture is not a solid colour (checker / image / noise)
  SF_NOISE = 16,    // a Perlin noise texture: its turbulence loop alone costs 16 registers (80 instead of 64)
  SF_BOXES = 32,    // box leaves: the six quads of an axis-aligned box_() behind one slab test
  SF_ALL = 63,
  // Not a capability but the way the scene is walked (DevScene::grids): even fields in uniform grids and the rest in
  // the front run, no tree.  An instance is compiled for one or the other, so that neither carries the other's code
  // (the general book-2 instance with both was 75 KB of SASS, and its warps waited for instruction fetches 1.8 times per
  // issue: profiles/README.md r02l).  A grid scene runs on an instance with this bit, any other on one without.
  SF_GRID = 64,
  // ... and a grid's members move (DevGrid::dclo / dchi): rays are clipped to the field's bounds at their own time.
  // A bit of its own because the code costs the book-2 final scene, whose fields stand still, 2.8 % (profiles/README.md r02u).
  SF_GRID_MOVING = 128
};

struct DevMaterial {  // 32 B
  int kind;           // rrt_material_kind
  int tex;            // texture index or -1
  float fuzz;
  float ior;
  float4 albedo;      // metal albedo; for textured kinds a copy of a solid colour (fast path flag in .w)
};

struct DevTexture {   // 32 B
  int kind;           // rrt_texture_kind
  int a;              // checker: even   image: image index   noise: perlin index
  int b;              // checker: odd
  int pad;
  float4 v;           // solid: rgb    checker: .w = inv_scale    noise: .w = scale
};

struct DevImage {
  int width, height;
  int offset;         // first texel in the uchar4 texel pool
  int pad;
};

// Perlin tables (perlin.rs:4-9): 256 gradient vectors + three permutations
struct DevPerlin {
  float4 randvec[256];
  uint8_t perm[3][256];
};

struct DevMedium {    // constant_medium.rs:10-14
  int first, nsph, nquad;  // boundary run in bprims
  int mat;                 // Isotropic phase function
  float neg_inv_density;
  int id;                  // desc index of the medium
  int pad0, pad1;
};

// light list entry (lights of camera.rs:63; hittable_list.rs:60-75)
// kind 0 sphere: a = (c, r)
// kind 1 quad:   a = (n, d), b = (A, a0), c = (B, b0), q = (Q, area), u = (u, 0), v = (v, 0)
// kind 2 other:  pdf_value 0, random (1,0,0)  (trait defaults, hittable.rs:58-64)
struct DevLight {
  float4 a, b, c, q, u, v;
  int kind, pad0, pad1, pad2;
};

// CameraInternals (camera.rs:49-60), f32, direction kept relative to the centre
struct DevCamera {
  float cx, cy, cz;          // center
  float p00x, p00y, p00z;    // pixel00_loc - center
  float dux, duy, duz;       // pixel_delta_u
  float dvx, dvy, dvz;       // pixel_delta_v
  float ddux, dduy, dduz;    // defocus_disk_u
  float ddvx, ddvy, ddvz;    // defocus_disk_v
  float bgx, bgy, bgz;       // background
  float recip_sqrt_spp;
  int sqrt_spp;
  int defocus;               // defocus_angle > 0
};
struct DevCamera64 {
  double c[3], p00[3], du[3], dv[3];
};

#define RRT_MAX_GRIDS 2
struct DevGrid {
  float lo[3], hi[3];     // bounds (of the whole shutter interval): the cells are laid over these
  // What a ray is clipped to before it walks: the bounds at ray time tm, lo(tm) = clo + tm dclo, hi(tm) = chi + tm dchi.
  // Linear motion makes the items' extremes at tm at most the interpolated extremes of tm = 0 and tm = 1, so this is
  // conservative; for the bouncing spheres (the_next_week/mod.rs:96-149: centres rise by up to 0.5) the slab a ray has
  // to cross is 0.4 + 0.5 tm high instead of 0.9.  Static fields: dclo = dchi = 0.
  float clo[3], chi[3], dclo[3], dchi[3];
  float cell[3], inv[3];  // cell size and its reciprocal per axis
  int res[3];
  int first_cell;         // offset of this grid's cells in DevScene::grid_cells
  int kind;               // 0: cells list spheres (flat primitive indices), 1: boxes (records of DevScene::boxes)
  int n_items;
  int moving;             // any of dclo / dchi is not zero
  int pad_;
};

struct DevScene {
  const float4* prims;
  const float4* nodes;
  const float4* bprims;
  const float4* boxes;   // two float4 per box leaf
  const float4* qpairs;  // quad pair records of a brute-force root leaf (its quads that are not in `rects`), or null
  // Axis-aligned rectangles of a brute-force root leaf (walls, floor, ceiling and light of the Cornell boxes): the leaf's
  // first n_rects[0] + n_rects[1] + n_rects[2] quads, grouped by the axis a of their normal.  Two float4 each:
  // (plane coordinate along a, centre along b, centre along c, -) (half extent along b, along c, -, -), b = a+1, c = a+2 mod 3.
  const float4* rects;
  int n_rects[3];
  const float4* nodes4;  // experiment (-DRRT_BVH4, profiles/README.md r02): the same tree collapsed to 4-wide nodes,
                         //   8 float4 each: children 0,1 and 2,3 as two slab2 pair records, then the four refs; else null
  const double* prims64;
  const DevMaterial* materials;
  const DevTexture* textures;
  const DevImage* images;
  const unsigned char* texels;  // uchar4 pool (r,g,b,255)
  const DevPerlin* perlins;
  const DevMedium* media;
  const DevLight* lights;
  int root_ref;
  int front_ref;     // leaf-format run tested ahead of the traversal (primitives spanning most of the scene); ~0 = none
  int n_nodes, n_prims, n_media, n_lights;
  int has_motion;    // any sphere with dc != 0
  int features;      // SceneFeature bits
  // Uniform grids over even fields of equal primitives (flatten.cpp build_grids), n_grids == 0: none.
  // A field — the 480 small spheres of the book-1 final scene, the 400 ground boxes and the 1 000-sphere
  // cluster of the book-2 final scene — is many primitives of one size spread evenly: a ray has to descend
  // ~10 levels of any BVH to learn that its neighbourhood is empty, where a grid reads the cell it stands in.
  // With grids the render kernel walks no tree at all: whatever is not in a grid is in the front run
  // (front_ref), tested by all lanes together.  The BVH over the grids' primitives stays for the f64 probe
  // and for RRT_NO_GRID=1.  grid_cells[DevGrid::first_cell + (z ry + y) rx + x] = first reference << 8 |
  // count; grid_refs[] = flat sphere indices (kind 0) or box records (kind 1).
  const unsigned* grid_cells;
  const int* grid_refs;
  DevGrid grids[RRT_MAX_GRIDS];
  int n_grids;
};

// One launch's share of the frame.
struct DevWork {
  int width, height;
  int tiles_x;          // 8x4-pixel warp tiles per row
  int n_tiles;          // warp tiles owned by this rank
  int chunks;           // sample chunks per tile
  int s_begin, s_end;   // sample range of this rank
  int shard_mode, rank, world;
  int stiles_x, stiles_y;   // 32x32 super tiles (tile sharding)
  int integrator;
  int max_depth;
  int spp_eff;
  unsigned seed_lo, seed_hi;
  unsigned rkeys[20];   // Philox round keys of the seed (rng.cuh philox_round_keys; 2 per round, up to 10 rounds)
  int atomic_accum;     // chunks > 1: accumulate with red.add
  int compact;          // tile sharding for a gather: accum holds this rank's super tiles back to back
                        //   (1024 float4 each, row-major inside the tile) instead of a full frame
};

}  // namespace rrt
