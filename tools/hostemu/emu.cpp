// DEVELOPMENT TOOL ONLY -- see cuda_runtime.h in this directory.
#include "cuda_runtime.h"

namespace emu {

thread_local uint3e t_threadIdx, t_blockIdx;
thread_local dim3 t_blockDim, t_gridDim;
thread_local BlockCtx* t_block = nullptr;
thread_local WarpCtx* t_warp = nullptr;
thread_local int t_lane = 0;

static thread_local unsigned char* t_dyn = nullptr;
unsigned char* dyn_smem() { return t_dyn; }

void launch(dim3 grid, dim3 block, size_t smem, const std::function<void()>& body) {
  const int nthreads = (int)(block.x * block.y * block.z);
  const int nwarps = (nthreads + 31) / 32;
  for (unsigned bz = 0; bz < grid.z; ++bz)
    for (unsigned by = 0; by < grid.y; ++by)
      for (unsigned bx = 0; bx < grid.x; ++bx) {
        BlockCtx ctx;
        ctx.bar.reset(nthreads);
        ctx.warps = std::vector<WarpCtx>(nwarps);
        for (int w = 0; w < nwarps; ++w) ctx.warps[w].bar.reset(std::min(32, nthreads - 32 * w));
        ctx.dyn.assign(smem + 64, 0);
        std::vector<std::thread> th;
        th.reserve(nthreads);
        for (int t = 0; t < nthreads; ++t) {
          th.emplace_back([&, t] {
            t_threadIdx = {(unsigned)(t % block.x), (unsigned)((t / block.x) % block.y), (unsigned)(t / (block.x * block.y))};
            t_blockIdx = {bx, by, bz};
            t_blockDim = block;
            t_gridDim = grid;
            t_block = &ctx;
            t_warp = &ctx.warps[t / 32];
            t_lane = t % 32;
            uintptr_t a = (reinterpret_cast<uintptr_t>(ctx.dyn.data()) + 15) & ~(uintptr_t)15;
            t_dyn = reinterpret_cast<unsigned char*>(a);
            body();
            t_warp->bar.drop();
            ctx.bar.drop();
          });
        }
        for (auto& x : th) x.join();
      }
}

}  // namespace emu
