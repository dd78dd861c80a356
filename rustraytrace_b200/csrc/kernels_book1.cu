// kernels_book1.cu — render_kernel instances for the book-1 integrator, smallest feature set first
// (select_render_variant takes the first that covers the scene).
#include "render_kernel_pool.cuh"

namespace rrt {

const RenderVariant* render_variants_book1(int* n) {
  static const RenderVariant table[] = {
    {0, render_kernel<RRT_INT_BOOK1, 0, true>, render_kernel<RRT_INT_BOOK1, 0, false>, render_kernel_pool<RRT_INT_BOOK1, 0>},
    {SF_ALL, render_kernel<RRT_INT_BOOK1, SF_ALL, true>, render_kernel<RRT_INT_BOOK1, SF_ALL, false>, nullptr},
    // scenes walked through grids (device_types.h SF_GRID): the book-1 final scene, and the general instance
    {SF_GRID, render_kernel<RRT_INT_BOOK1, SF_GRID, true>, render_kernel<RRT_INT_BOOK1, SF_GRID, false>, nullptr},
    {SF_GRID | SF_ALL, render_kernel<RRT_INT_BOOK1, SF_GRID | SF_ALL, true>, render_kernel<RRT_INT_BOOK1, SF_GRID | SF_ALL, false>, nullptr},
  };
  *n = (int)(sizeof table / sizeof table[0]);
  return table;
}

}  // namespace rrt
