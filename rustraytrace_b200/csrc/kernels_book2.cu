// kernels_book2.cu — render_kernel instances for the book-2 integrator, smallest feature set first
// (select_render_variant takes the first that covers the scene).
#include "render_kernel_pool.cuh"

namespace rrt {

const RenderVariant* render_variants_book2(int* n) {
  static const RenderVariant table[] = {
    {0, render_kernel<RRT_INT_BOOK2, 0, true>, render_kernel<RRT_INT_BOOK2, 0, false>, nullptr},
    {SF_TEXTURES, render_kernel<RRT_INT_BOOK2, SF_TEXTURES, true>, render_kernel<RRT_INT_BOOK2, SF_TEXTURES, false>, nullptr},
    {SF_MOTION | SF_TEXTURES, render_kernel<RRT_INT_BOOK2, SF_MOTION | SF_TEXTURES, true>, render_kernel<RRT_INT_BOOK2, SF_MOTION | SF_TEXTURES, false>, render_kernel_pool<RRT_INT_BOOK2, SF_MOTION | SF_TEXTURES>},
    {SF_QUADS, render_kernel<RRT_INT_BOOK2, SF_QUADS, true>, render_kernel<RRT_INT_BOOK2, SF_QUADS, false>, render_kernel_pool<RRT_INT_BOOK2, SF_QUADS>},
    {SF_QUADS | SF_MEDIA, render_kernel<RRT_INT_BOOK2, SF_QUADS | SF_MEDIA, true>, render_kernel<RRT_INT_BOOK2, SF_QUADS | SF_MEDIA, false>, nullptr},
    {SF_TEXTURES | SF_NOISE, render_kernel<RRT_INT_BOOK2, SF_TEXTURES | SF_NOISE, true>, render_kernel<RRT_INT_BOOK2, SF_TEXTURES | SF_NOISE, false>, nullptr},
    {SF_QUADS | SF_TEXTURES | SF_NOISE, render_kernel<RRT_INT_BOOK2, SF_QUADS | SF_TEXTURES | SF_NOISE, true>, render_kernel<RRT_INT_BOOK2, SF_QUADS | SF_TEXTURES | SF_NOISE, false>, nullptr},
    {SF_ALL & ~SF_NOISE, render_kernel<RRT_INT_BOOK2, SF_ALL & ~SF_NOISE, true>, render_kernel<RRT_INT_BOOK2, SF_ALL & ~SF_NOISE, false>, nullptr},
    {SF_ALL, render_kernel<RRT_INT_BOOK2, SF_ALL, true>, render_kernel<RRT_INT_BOOK2, SF_ALL, false>, render_kernel_pool<RRT_INT_BOOK2, SF_ALL>},
    // scenes walked through grids (device_types.h SF_GRID): static spheres, the bouncing spheres, and the general instance (the final scene)
    {SF_GRID, render_kernel<RRT_INT_BOOK2, SF_GRID, true>, render_kernel<RRT_INT_BOOK2, SF_GRID, false>, nullptr},
    {SF_GRID | SF_GRID_MOVING | SF_MOTION | SF_TEXTURES, render_kernel<RRT_INT_BOOK2, SF_GRID | SF_GRID_MOVING | SF_MOTION | SF_TEXTURES, true>, render_kernel<RRT_INT_BOOK2, SF_GRID | SF_GRID_MOVING | SF_MOTION | SF_TEXTURES, false>, nullptr},
    {SF_GRID | SF_ALL, render_kernel<RRT_INT_BOOK2, SF_GRID | SF_ALL, true>, render_kernel<RRT_INT_BOOK2, SF_GRID | SF_ALL, false>, render_kernel_pool<RRT_INT_BOOK2, SF_GRID | SF_ALL>},
    {SF_GRID | SF_GRID_MOVING | SF_ALL, render_kernel<RRT_INT_BOOK2, SF_GRID | SF_GRID_MOVING | SF_ALL, true>, render_kernel<RRT_INT_BOOK2, SF_GRID | SF_GRID_MOVING | SF_ALL, false>, nullptr},
  };
  *n = (int)(sizeof table / sizeof table[0]);
  return table;
}

}  // namespace rrt
