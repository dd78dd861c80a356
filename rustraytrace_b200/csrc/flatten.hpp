// flatten.hpp — host side of rrt_scene_create: rrt_scene_desc (pointer-tree shaped, f64) ->
// flat world-space primitives + our own BVH + SoA tables.  Replaces, for the GPU, what the
// reference does with Arc trees: BvhNode::new (bvh.rs:16-156), Translate/RotateY wrappers
// (hittable.rs:73-178), HittableList nesting (hittable_list.rs), make_box lists (quad.rs:112-134).
#pragma once
#include <string>
#include <vector>

#include "../../include/rrt.h"
#include "device_types.h"

namespace rrt {

struct FlatScene {
  std::vector<float4> prims;      // 4 per prim
  std::vector<double> prims64;    // 16 per prim
  std::vector<float4> boxes;     // box leaves, two float4 each (device_types.h)
  std::vector<float4> nodes;      // 4 per inner node
  std::vector<float4> nodes4;     // -DRRT_BVH4 experiment only: 8 per 4-wide node
  std::vector<float4> qpairs;     // root-leaf quads again, 6 float4 per pair (device_types.h)
  std::vector<float4> rects;      // axis-aligned rectangles of a root leaf, 2 float4 each (device_types.h)
  int n_rects[3] = {0, 0, 0};     // how many with normal x, y, z: the first quads of the root leaf, in that order
  std::vector<unsigned> grid_cells;  // uniform grids over even fields of spheres / boxes (device_types.h DevScene), pooled; or empty
  std::vector<int> grid_refs;
  std::vector<DevGrid> grids;        // at most RRT_MAX_GRIDS
  std::vector<float4> bprims;     // 4 per medium-boundary prim
  std::vector<DevMaterial> materials;
  std::vector<DevTexture> textures;
  std::vector<DevImage> images;
  std::vector<unsigned char> texels;  // uchar4 pool
  std::vector<DevPerlin> perlins;
  std::vector<DevMedium> media;
  std::vector<DevLight> lights;
  int root_ref = -1;
  int front_ref = -1;  // leaf-format run of primitives tested ahead of the traversal (none: ~0)
  int n_prims = 0;
  int has_motion = 0;
  int features = 0;   // SceneFeature bits (device_types.h): which render kernel instance the scene needs
  int bvh_depth = 0;
};

// Returns RRT_OK or a negative rrt_status with `err` filled.
#ifndef RRT_SINGLE_LEAF_MAX
#define RRT_SINGLE_LEAF_MAX 16
#endif
// single_leaf_max: scenes with at most this many primitives become one root leaf (brute force).
// max_leaf: primitives per leaf below the root; the render kernel's traversal loop relies on 1.
// tree_optional: the caller will never probe the scene — where it is walked through grids no BVH is built (FlatScene::nodes
// stays empty; rrt_render_frame, 1 ms of host time on the book-1 final scene).
int flatten_scene(const rrt_scene_desc* desc, FlatScene& out, std::string& err,
                  int single_leaf_max = RRT_SINGLE_LEAF_MAX, int max_leaf = 1, bool tree_optional = false);

// Camera::initialize (camera.rs:108-154) -> device camera blocks.  Returns image height.
int camera_initialize(const rrt_camera* cam, DevCamera& c32, DevCamera64& c64);

}  // namespace rrt
