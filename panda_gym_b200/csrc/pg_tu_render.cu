// pg_tu_render.cu -- headless render("rgb_array"): the ray-caster kernel (pg_render.cuh) and its launcher
#include "pg_render.cuh"
#include "pg_launch.h"

namespace pg {

void launch_render(const Params& P, const Camera& cam, int env, uint8_t* out, cudaStream_t s) {
  dim3 block(16, 16), grid((cam.width + 15) / 16, (cam.height + 15) / 16);
  k_render<<<grid, block, 0, s>>>(P, cam, env, out);
}

}  // namespace pg
