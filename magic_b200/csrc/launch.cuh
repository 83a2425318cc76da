// Kernel launches of the training step: programmatic dependent launch (PDL).
//
// A step is ~35 short-to-medium kernels on one stream; with plain stream order each boundary costs the grid drain,
// the launch and the next kernel's prologue (barrier init, TMEM allocation, tensor-map prefetch).  Launched with
// cudaLaunchAttributeProgrammaticStreamSerialization, a kernel's CTAs become resident as soon as the previous
// kernel has triggered (ptx::grid_dep_launch(): the GEMM does it when a CTA has issued its last MMA; the other
// kernels trigger implicitly at exit) and an SM has room; they run their prologue and then block in
// ptx::grid_dep_wait() until the previous grid has completed and its writes are visible.  EVERY kernel launched through launch_chain() must execute grid_dep_wait() before it touches global
// memory -- also when it does not consume its predecessor's output -- because completion of a kernel is what its
// successor waits for, and that must imply completion of everything before it.
// MAGIC_PDL=0 falls back to plain stream order (the attribute is simply not set; the PTX is then a no-op).
#pragma once
#include <cuda_runtime.h>
#include <cstdlib>

namespace magic {

inline bool pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("MAGIC_PDL");
    v = (e == nullptr || atoi(e) != 0) ? 1 : 0;
  }
  return v == 1;
}

template <typename... P, typename... A>
inline cudaError_t launch_chain(void (*kernel)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                A... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof cfg);
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, P(args)...);
}

}  // namespace magic
