// kernels_book3.cu — render_kernel instances for the book-3 integrator, smallest feature set first
// (select_render_variant takes the first that covers the scene).
#include "render_kernel_pool.cuh"

namespace rrt {

const RenderVariant* render_variants_book3(int* n) {
  static const RenderVariant table[] = {
    {SF_QUADS, render_kernel<RRT_INT_BOOK3, SF_QUADS, true>, render_kernel<RRT_INT_BOOK3, SF_QUADS, false>, render_kernel_pool<RRT_INT_BOOK3, SF_QUADS>},
    {SF_QUADS | SF_MEDIA, render_kernel<RRT_INT_BOOK3, SF_QUADS | SF_MEDIA, true>, render_kernel<RRT_INT_BOOK3, SF_QUADS | SF_MEDIA, false>, nullptr},
    {SF_ALL, render_kernel<RRT_INT_BOOK3, SF_ALL, true>, render_kernel<RRT_INT_BOOK3, SF_ALL, false>, nullptr},
    {SF_GRID | SF_ALL, render_kernel<RRT_INT_BOOK3, SF_GRID | SF_ALL, true>, render_kernel<RRT_INT_BOOK3, SF_GRID | SF_ALL, false>, nullptr},   // scenes walked through grids
    {SF_GRID | SF_GRID_MOVING | SF_ALL, render_kernel<RRT_INT_BOOK3, SF_GRID | SF_GRID_MOVING | SF_ALL, true>, render_kernel<RRT_INT_BOOK3, SF_GRID | SF_GRID_MOVING | SF_ALL, false>, nullptr},
  };
  *n = (int)(sizeof table / sizeof table[0]);
  return table;
}

}  // namespace rrt
