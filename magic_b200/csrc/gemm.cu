// Host-side planning and launch of the tcgen05 GEMM: tensor maps, split-K choice, finishing pass.
#include "gemm.h"
#include "kernels.h"
#include "launch.cuh"
#include <cstdio>
#include <cstring>

namespace magic {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;

// split-K partial sums (fp32 workspace) -> fused epilogue.  One thread per (row, 32-column chunk).
template <int EPI>
__global__ void gemm_finish_kernel(EpiArgs e, const float* __restrict__ ws, long long ws_ld, int M, int N) {
  ptx::grid_dep_wait();
  const int chunks = (N + 31) >> 5;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int row = (int)(idx / chunks);
  const int c = (int)(idx % chunks);
  if (row >= M) return;
  const float4* src = reinterpret_cast<const float4*>(ws + (long long)row * ws_ld + c * 32);
  uint32_t r[32];
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    float4 v = src[q];
    r[4 * q] = __float_as_uint(v.x);
    r[4 * q + 1] = __float_as_uint(v.y);
    r[4 * q + 2] = __float_as_uint(v.z);
    r[4 * q + 3] = __float_as_uint(v.w);
  }
  float dummy = 0.f;
  epilogue_chunk<EPI, false>(e, TileOut{}, r, row, c * 32, M, N, dummy);
}

// ---------------------------------------------------------------------------- plans with a zero-stage map
// Step 1: one k-block list per (group, M tile, split-K slice) from the map (GemmArgs.klist).  One warp per list:
// 32 candidates at a time, ballot + prefix count compacts the set ones.
__global__ void __launch_bounds__(256) gemm_klist_kernel(const __grid_constant__ GemmArgs args, uint16_t* klist, int* counts,
                                                         int rows) {
  const int r = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (r >= rows) return;
  const int S = args.splitk, ks = r % S, mt = (r / S) % args.tiles_m, g = r / (S * args.tiles_m);
  const uint32_t* mrow = args.kmask[g] + (long long)mt * args.kmask_words;
  const int total_kb = args.nseg * args.kblocks;
  const int ncand = S > 1 ? (total_kb - ks + S - 1) / S : total_kb;
  uint16_t* out = klist + (long long)r * args.klist_pitch;
  int n = 0;
  for (int c0 = 0; c0 < ncand; c0 += 32) {
    const int c = c0 + lane;
    const int kb = S > 1 ? ks + c * S : c;
    bool set = false;
    if (c < ncand) {
      int kk = kb;
      while (kk >= args.kblocks) kk -= args.kblocks;
      set = (__ldg(mrow + (kk >> 5)) >> (kk & 31)) & 1u;
    }
    const uint32_t b = __ballot_sync(0xffffffffu, set);
    if (set) out[1 + n + __popc(b & ((1u << lane) - 1u))] = (uint16_t)kb;
    n += __popc(b);
  }
  if (lane == 0) {
    if (n == 0) { out[1] = (uint16_t)(S > 1 ? ks : 0); n = 1; }    // an all-zero stage: the accumulator must be written
    out[0] = (uint16_t)n;
    counts[r] = n;
  }
}
// Step 2: work-item order (GemmArgs.order): cost = list length, counting sort by descending cost, then serpentine
// dealing over the G persistent CTAs (round j hands out the next G items, every other round in reverse), so that each
// CTA receives a similar total.  One CTA; a few thousand items at most.
constexpr int kSchedMaxItems = 4096, kSchedMaxCost = 2048, kSchedThreads = 1024;
__global__ void __launch_bounds__(kSchedThreads) gemm_schedule_kernel(const __grid_constant__ GemmArgs args, const int* counts,
                                                                      int* order, int total, int G) {
  __shared__ int hist[kSchedMaxCost + 1];
  __shared__ unsigned short cost_of[kSchedMaxItems];
  __shared__ int sorted[kSchedMaxItems];
  for (int c = threadIdx.x; c <= kSchedMaxCost; c += kSchedThreads) hist[c] = 0;
  __syncthreads();
  for (int w = threadIdx.x; w < total; w += kSchedThreads) {
    const WorkItem it = decode_work(args, w, 0);
    int cost = it.g >= 0 ? counts[(it.g * args.tiles_m + it.mt) * args.splitk + it.ks] : it.kb1 - it.kb0;
    cost = cost < 1 ? 1 : (cost > kSchedMaxCost ? kSchedMaxCost : cost);
    cost_of[w] = (unsigned short)cost;
    atomicAdd(&hist[cost], 1);
  }
  __syncthreads();
  if (threadIdx.x == 0) {            // first position of every cost class, most expensive first
    int pos = 0;
    for (int c = kSchedMaxCost; c >= 1; --c) { const int n = hist[c]; hist[c] = pos; pos += n; }
  }
  __syncthreads();
  for (int w = threadIdx.x; w < total; w += kSchedThreads) sorted[atomicAdd(&hist[cost_of[w]], 1)] = w;
  __syncthreads();
  for (int p = threadIdx.x; p < total; p += kSchedThreads) {
    const int j = p / G, c = p - j * G;
    const int nr = min(G, total - j * G);
    order[j * G + ((j & 1) ? nr - 1 - c : c)] = sorted[p];
  }
}

void gemm_schedule(const GemmPlan& pl, cudaStream_t s) {
  if (!pl.klist) return;
  uint16_t* klist = static_cast<uint16_t*>(pl.klist.get());
  int* counts = reinterpret_cast<int*>(klist + (size_t)pl.klist_rows * pl.args.klist_pitch);
  const long long threads = (long long)pl.klist_rows * 32;
  gemm_klist_kernel<<<(int)((threads + 255) / 256), 256, 0, s>>>(pl.args, klist, counts, pl.klist_rows);
  if (pl.order)
    gemm_schedule_kernel<<<1, kSchedThreads, 0, s>>>(pl.args, counts, static_cast<int*>(pl.order.get()), (int)pl.items, pl.grid);
}

template <bool A_MN, bool B_MN, int EPI, int BN = 256, bool A_BITS = false, bool CG2 = false>
static cudaError_t set_smem() {
  return cudaFuncSetAttribute(gemm_tcgen05_kernel<A_MN, B_MN, EPI, BN, A_BITS, CG2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              (int)gemm_smem_bytes(EPI));
}
// (kind, epilogue, packed A) combinations that exist as CTA-pair kernels: the 40000-wide GEMMs of the step
static bool has_pair_variant(int kind, int epi, bool a_bits) {
  if (kind == GEMM_FWD) return a_bits ? (epi == EPI_STORE_F32 || epi == EPI_ACT) : (epi == EPI_BCE);
  if (kind == GEMM_WGRAD) return epi == EPI_STORE_F32;
  if (kind == GEMM_DGRAD) return !a_bits && epi == EPI_STORE_F32;
  return false;
}

std::string gemm_init() {
  if (!g_encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t err = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (err != cudaSuccess || fn == nullptr || qres != cudaDriverEntryPointSuccess)
      return std::string("cuTensorMapEncodeTiled not available: ") + cudaGetErrorString(err);
    g_encode = reinterpret_cast<EncodeTiledFn>(fn);
  }
  cudaError_t e = cudaSuccess;
  if (e == cudaSuccess) e = set_smem<false, true, EPI_STORE_F32>();
  if (e == cudaSuccess) e = set_smem<false, true, EPI_ACT>();
  if (e == cudaSuccess) e = set_smem<false, true, EPI_BCE>();
  if (e == cudaSuccess) e = set_smem<false, true, EPI_SIGMOID>();
  if (e == cudaSuccess) e = set_smem<false, false, EPI_STORE_F32>();
  if (e == cudaSuccess) e = set_smem<false, false, EPI_DACT>();
  if (e == cudaSuccess) e = set_smem<true, true, EPI_STORE_F32>();
  if (e == cudaSuccess) e = set_smem<true, true, EPI_SCATTER_F32>();
  // 128-wide tiles: the short-K layers (hidden fwd / dgrad, heads, the small split-K wgrads)
  if (e == cudaSuccess) e = set_smem<false, true, EPI_STORE_F32, 128>();
  if (e == cudaSuccess) e = set_smem<false, true, EPI_ACT, 128>();
  if (e == cudaSuccess) e = set_smem<false, false, EPI_STORE_F32, 128>();
  if (e == cudaSuccess) e = set_smem<false, false, EPI_DACT, 128>();
  if (e == cudaSuccess) e = set_smem<true, true, EPI_STORE_F32, 128>();
  // A as packed bits (the first encoder layer): forward (split-K partial sums or the fused activation), weight
  // gradient (plain or scattered to the owners' staging)
  if (e == cudaSuccess) e = set_smem<false, true, EPI_STORE_F32, 256, true>();
  if (e == cudaSuccess) e = set_smem<false, true, EPI_ACT, 256, true>();
  if (e == cudaSuccess) e = set_smem<true, true, EPI_STORE_F32, 256, true>();
  if (e == cudaSuccess) e = set_smem<true, true, EPI_SCATTER_F32, 256, true>();
  if (e == cudaSuccess) e = set_smem<true, true, EPI_ADAM, 256, true>();        // ... or consumed by the fused dual Adam
  // CTA pairs (tcgen05 cta_group::2)
  if (e == cudaSuccess) e = set_smem<false, true, EPI_BCE, 256, false, true>();
  if (e == cudaSuccess) e = set_smem<true, true, EPI_STORE_F32, 256, false, true>();
  if (e == cudaSuccess) e = set_smem<false, false, EPI_STORE_F32, 256, false, true>();
  if (e == cudaSuccess) e = set_smem<false, true, EPI_STORE_F32, 256, true, true>();
  if (e == cudaSuccess) e = set_smem<false, true, EPI_ACT, 256, true, true>();
  if (e == cudaSuccess) e = set_smem<true, true, EPI_STORE_F32, 256, true, true>();
  if (e == cudaSuccess) e = cudaFuncSetAttribute(chain_tcgen05_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kChainSmemBytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(chain_tcgen05_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kChainSmemBytes);
  if (e != cudaSuccess) return std::string("cudaFuncSetAttribute(smem) failed: ") + cudaGetErrorString(e);
  cudaFuncSetAttribute(gemm_schedule_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  cudaFuncSetAttribute(gemm_klist_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  cudaFuncSetAttribute(gemm_finish_kernel<EPI_ACT>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  cudaFuncSetAttribute(gemm_finish_kernel<EPI_DACT>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  kernels_init();
  return "";
}

// 2-D bf16 tensor map: `inner` contiguous elements per row, `outer` rows, row pitch ld elements, 128B swizzle
// (operands) or 64B swizzle (the 32-column output tiles of the BCE epilogue).
static std::string make_tmap(CUtensorMap* out, const void* ptr, long long inner, long long outer, long long ld,
                             int box_inner, int box_outer, CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  if (!g_encode) return "gemm_init() not called";
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0) return "tensor map: base not 16-byte aligned";
  if ((ld * 2) % 16 != 0) return "tensor map: row pitch not a multiple of 16 bytes";
  if (inner <= 0 || outer <= 0) return "tensor map: empty extent";
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[160];
    snprintf(buf, sizeof buf, "cuTensorMapEncodeTiled failed (%d) inner=%lld outer=%lld ld=%lld box=%dx%d", (int)r,
             inner, outer, ld, box_inner, box_outer);
    return buf;
  }
  return "";
}

// Output tile map over a row-major [rows, cols] matrix with pitch ld: box of 32 rows x 64 bytes (16 fp32 or 32 bf16
// columns), 64B swizzle -- the staging layout of TileOut in gemm_sm100.cuh.
static std::string make_tmap_out(CUtensorMap* out, void* ptr, bool f32, long long cols, long long rows, long long ld) {
  if (!g_encode) return "gemm_init() not called";
  const int eb = f32 ? 4 : 2;
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0) return "output map: base not 16-byte aligned";
  if ((ld * eb) % 16 != 0) return "output map: row pitch not a multiple of 16 bytes";
  if (cols <= 0 || rows <= 0) return "output map: empty extent";
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * eb};
  cuuint32_t box[2] = {f32 ? 16u : 32u, 32};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(out, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ptr, dims,
                        strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[160];
    snprintf(buf, sizeof buf, "cuTensorMapEncodeTiled(out) failed (%d) cols=%lld rows=%lld ld=%lld f32=%d", (int)r, cols,
             rows, ld, (int)f32);
    return buf;
  }
  return "";
}

// (Re)build the output maps of group g from args.epi[g].out / out_lo (null pointers are skipped: the caller patches
// them per call and comes back through gemm_set_output).
static std::string build_out_maps(GemmPlan& pl, int g) {
  if (pl.epi == EPI_SCATTER_F32) return "";
  GemmArgs& a = pl.args;
  const EpiArgs& e = a.epi[g];
  const bool f32 = pl.epi == EPI_STORE_F32 || pl.epi == EPI_SIGMOID;
  if (e.out) {
    std::string err = make_tmap_out(&a.tout[g][0], e.out, f32, a.N, a.M, e.ld_out);
    if (!err.empty()) return err;
  }
  if (!f32 && e.out_lo) {
    std::string err = make_tmap_out(&a.tout[g][1], e.out_lo, false, a.N, a.M, e.ld_out);
    if (!err.empty()) return err;
  }
  return "";
}

std::string gemm_set_output(GemmPlan* pl, int g, void* out, void* out_lo) {
  pl->args.epi[g].out = out;
  pl->args.epi[g].out_lo = out_lo;
  return build_out_maps(*pl, g);
}

// MN-major bf16 operand [rows = K, cols = MN, pitch ld] as a 3-D tensor {64 contiguous, K rows, MN/64 slabs}: one box
// {64, 64, slabs} fetches all the swizzle atoms of a pipeline stage with a single TMA instruction.  The last slab may
// reach past `cols` into the row padding (ld is a multiple of 64), which the consumers mask.
static std::string make_tmap_mn(CUtensorMap* out, const void* ptr, long long cols, long long rows, long long ld,
                                int box_k, int slabs) {
  if (!g_encode) return "gemm_init() not called";
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0) return "tensor map: base not 16-byte aligned";
  if (ld % 64 != 0) return "tensor map: MN-major operand needs a row pitch that is a multiple of 64 elements";
  if (cols <= 0 || rows <= 0 || cols > ld) return "tensor map: bad extent";
  cuuint64_t dims[3] = {64, (cuuint64_t)rows, (cuuint64_t)((cols + 63) / 64)};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, 128};
  cuuint32_t box[3] = {64, (cuuint32_t)box_k, (cuuint32_t)slabs};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[160];
    snprintf(buf, sizeof buf, "cuTensorMapEncodeTiled(3d) failed (%d) cols=%lld rows=%lld ld=%lld slabs=%d", (int)r, cols,
             rows, ld, slabs);
    return buf;
  }
  return "";
}

// Minimise waves * (k-blocks per item + fixed per-item overhead), see DESIGN.md "split-K".
int choose_splitk(int tiles, int total_kb, int num_sms, bool two_pass) {
  const double overhead = 8.0;   // epilogue + pipeline fill, in k-block units
  // a split with a non-linear epilogue costs two more launches (workspace zeroing + finishing pass), ~12 us ~ 27
  // k-block times (r01a launch list): never worth it for the 8-k-block decoder layers
  const double extra = two_pass ? 27.0 : 0.0;
  int best = 1;
  double best_cost = 1e30;
  int max_split = total_kb / 4;
  if (max_split > 64) max_split = 64;
  if (max_split < 1) max_split = 1;
  for (int s = 1; s <= max_split; ++s) {
    int per = (total_kb + s - 1) / s;
    if ((long long)(s - 1) * per >= total_kb) continue;   // would leave an empty split
    long long items = (long long)tiles * s;
    long long waves = (items + num_sms - 1) / num_sms;
    double cost = (double)waves * (per + overhead) * (s > 1 ? 1.02 : 1.0) + (s > 1 ? extra : 0.0);
    if (cost < best_cost - 1e-9) { best_cost = cost; best = s; }
  }
  return best;
}

std::string gemm_plan(const GemmProblem& p, int num_sms, GemmPlan* out) {
  GemmPlan& pl = *out;
  memset(&pl.args, 0, sizeof(pl.args));
  pl.kind = p.kind;
  pl.epi = p.epi;
  if (p.ngroups < 1 || p.ngroups > kMaxGroups || p.nseg < 1 || p.nseg > kMaxSegs) return "bad ngroups/nseg";
  const bool a_mn = p.kind == GEMM_WGRAD, b_mn = p.kind != GEMM_DGRAD;
  pl.a_bits = p.a_bits;
  if (p.a_bits) {
    if (p.kind == GEMM_DGRAD) return "a_bits: forward and wgrad kinds only";
    if (!(p.epi == EPI_STORE_F32 || (p.kind == GEMM_FWD && p.epi == EPI_ACT) || (p.kind == GEMM_WGRAD && p.epi == EPI_SCATTER_F32)))
      return "a_bits: no kernel variant for this epilogue";
    if (p.cluster_m > 1 || p.block_n == 128) return "a_bits: 256-wide tiles without clusters only";
    const long long tm = (p.M + kBlockM - 1) / kBlockM, kbl = (p.K + kBlockK - 1) / kBlockK;
    const long long need = a_mn ? 4 * tm : 2 * kbl;     // words a tile row / k-block row may touch
    for (int g = 0; g < p.ngroups; ++g) {
      const BitsA& ba = p.abits[g];
      if (!ba.base || (reinterpret_cast<uintptr_t>(ba.base) & 7) != 0) return "a_bits: base must be 8-byte aligned";
      if (ba.pitch % 2 != 0 || ba.pitch < need) return "a_bits: row pitch must be even and cover the padded extent";
      if (ba.rows < 0 || ba.rows > (a_mn ? p.K : p.M)) return "a_bits: more rows than the problem has";
    }
  }
  // Tile width.  128 for short-K problems whose 256-wide tiles would leave SMs without any tile (the decoder's hidden
  // layers at M = 4096: 64 tiles; every layer at the reference's batch of 100): twice the CTAs, half the exposed
  // epilogue (TileCfg in gemm_sm100.cuh; 10 -> 8 us per launch, profiles/r01m).  With one or two 256-wide tiles per SM
  // the narrower tile only re-reads A twice as often -- those layers are bound by the L2 -> SM operand traffic
  // (~690 cycles per k-block at either width, r01m trace) -- so they stay at 256.  Only the (kind, epilogue) pairs
  // instantiated at 128 qualify; clusters stay at 256.
  int bn = kBlockN;
  {
    const bool has128 = (p.epi == EPI_STORE_F32) || (p.kind == GEMM_FWD && p.epi == EPI_ACT) ||
                        (p.kind == GEMM_DGRAD && p.epi == EPI_DACT);
    const long long tiles256 = (long long)((p.M + kBlockM - 1) / kBlockM) * ((p.N + kBlockN - 1) / kBlockN) * p.ngroups;
    const int kb = (p.K + kBlockK - 1) / kBlockK * p.nseg;
    bool want = has128 && p.cluster_m <= 1 && p.N > 128 && tiles256 < (long long)num_sms && kb <= 32;
    if (p.block_n == 128) want = has128 && p.cluster_m <= 1;
    if (p.block_n == 256 || p.a_bits) want = false;
    if (want) bn = 128;
  }
  int cm_b = (b_mn && (p.cluster_m == 2 || p.cluster_m == 4)) ? p.cluster_m : 1;
  while (cm_b > 1 && ((p.M + kBlockM - 1) / kBlockM) % cm_b != 0) cm_b >>= 1;   // every CTA of a cluster gets a real M tile
  // CTA pairs: 256 x 256 tiles, each CTA loads half of the B tile's columns (an odd last M tile is paired with an empty one,
  // whose loads are clipped to zeros and whose rows the epilogue masks).  OPT-IN (MAGIC_CG2=1): the variants are
  // correct (selftest_gemm pair: every check passes on a B200) but measured about 2x SLOWER than the single-CTA kernels on
  // every shape of the step (dec-out BCE 0.334 vs 0.236 ms, first-layer wgrad 0.58 vs 0.27 ms, square 8192 651 vs 1448
  // TFLOP/s; step 3.45 vs 2.31 ms) -- with the MMAs removed the time stays, with the loads removed it drops to the
  // single-CTA time, i.e. the cta_group::2 loads signalling the leader's barrier are what is slow.  DESIGN.md section 4.
  static const bool cg2_enabled = getenv("MAGIC_CG2") && atoi(getenv("MAGIC_CG2")) != 0;
  const bool linear = p.epi == EPI_STORE_F32;   // SCATTER_F32 stores plain values to peer memory: never split
  const bool can_split = linear || (p.ws != nullptr && p.ngroups == 1 && (p.epi == EPI_ACT || p.epi == EPI_DACT));
  const int total_kb = (p.K + kBlockK - 1) / kBlockK * p.nseg;
  auto split_for = [&](int cmx) {               // split-K slices when `cmx` CTAs share a work item
    const int tiles_x = (((p.M + kBlockM - 1) / kBlockM + cmx - 1) / cmx) * ((p.N + bn - 1) / bn) * p.ngroups;
    int sp = 1;
    if (can_split) sp = p.force_splitk > 0 ? p.force_splitk : choose_splitk(tiles_x, total_kb, num_sms / cmx, !linear);
    int per = (total_kb + sp - 1) / sp;
    while (sp > 1 && (long long)(sp - 1) * per >= total_kb) { --sp; per = (total_kb + sp - 1) / sp; }
    return sp;
  };
  // the kernel that runs is the epilogue's own variant, or -- split along K with a finishing pass -- the fp32-store one
  const bool pair_direct = has_pair_variant(p.kind, p.epi, p.a_bits);
  const bool pair_two_pass = !linear && can_split && has_pair_variant(p.kind, EPI_STORE_F32, p.a_bits) && split_for(2) > 1;
  const bool pair = p.cta_pair && cg2_enabled && cm_b == 1 && bn == 256 && num_sms >= 2 &&
                    (pair_two_pass || (pair_direct && (linear || !can_split || split_for(2) == 1)));
  pl.cta_pair = pair;
  const int bdiv = pair ? 2 : cm_b;                 // B columns fetched per CTA = bn / bdiv
  for (int g = 0; g < p.ngroups; ++g)
    for (int s = 0; s < p.nseg; ++s) {
      const Mat& A = p.a[g][s];
      const Mat& B = p.b[g][s];
      std::string err;
      if (p.a_bits) {   // no tensor map: the expander warps read args.abits[g]
      } else if (!a_mn) {   // Mat[M,K]
        if (A.cols != p.K) return "A cols != K";
        err = make_tmap(&pl.args.ta[g][s], A.ptr, A.cols, A.rows, A.ld, kBlockK, kBlockM);
      } else {       // Mat[K,M]
        if (A.rows != p.K) return "A rows != K";
        err = make_tmap_mn(&pl.args.ta[g][s], A.ptr, A.cols, A.rows, A.ld, kBlockK, kBlockM / 64);
      }
      if (!err.empty()) return "A: " + err;
      if (!b_mn) {   // Mat[N,K]
        if (B.cols != p.K) return "B cols != K";
        err = make_tmap(&pl.args.tb[g][s], B.ptr, B.cols, B.rows, B.ld, kBlockK, pair ? bn / 2 : bn);
      } else {       // Mat[K,N]
        if (B.rows != p.K) return "B rows != K";
        err = make_tmap_mn(&pl.args.tb[g][s], B.ptr, B.cols, B.rows, B.ld, kBlockK, (bn / 64) / bdiv);
      }
      if (!err.empty()) return "B: " + err;
    }
  GemmArgs& a = pl.args;
  for (int g = 0; g < p.ngroups; ++g) a.abits[g] = p.abits[g];
  if (p.a_bits && p.kmask[0] != nullptr) {
    if (p.kmask_words * 32 < (p.K + kBlockK - 1) / kBlockK) return "kmask: rows shorter than the number of k-blocks";
    for (int g = 0; g < p.ngroups; ++g) {
      if (!p.kmask[g]) return "kmask: every group needs its map";
      a.kmask[g] = p.kmask[g];
    }
    a.kmask_words = p.kmask_words;
  }
  a.M = p.M;
  a.N = p.N;
  a.kblocks = (p.K + kBlockK - 1) / kBlockK;
  a.nseg = p.nseg;
  a.ngroups = p.ngroups;
  a.tiles_m = (p.M + kBlockM - 1) / kBlockM;
  a.tiles_n = (p.N + bn - 1) / bn;
  a.bn = bn;
  const int cm = pair ? 2 : cm_b;
  a.cluster_m = cm;
  a.tiles_m = (a.tiles_m + cm - 1) / cm;             // cluster rows from here on
  a.m_fastest = p.N > p.M ? 1 : 0;   // keep the larger operand's tiles adjacent in time so it is read from HBM once
  const int tiles = a.tiles_m * a.tiles_n * a.ngroups;
  const int split = split_for(cm);
  a.splitk = split;
  pl.two_pass = false;
  for (int g = 0; g < p.ngroups; ++g) a.epi[g] = p.e[g];
  if (split > 1) {
    if (linear) {
      for (int g = 0; g < p.ngroups; ++g) a.epi[g].atomic = 1;   // caller zeroes the destination
    } else {
      pl.two_pass = true;
      pl.ws = p.ws;
      pl.ws_ld = p.ws_ld;
      pl.finish_epi_kind = p.epi;
      pl.finish_epi[0] = p.e[0];
      EpiArgs st;
      memset(&st, 0, sizeof st);
      st.out = p.ws;
      st.ld_out = p.ws_ld;
      st.alpha = 1.f;
      st.atomic = 1;
      a.epi[0] = st;
      pl.epi = EPI_STORE_F32;
    }
  }
  for (int g = 0; g < p.ngroups; ++g) {
    std::string err = build_out_maps(pl, g);
    if (!err.empty()) return "out: " + err;
  }
  long long items = (long long)tiles * split;
  pl.flops = 2.0 * p.M * (double)p.N * p.K * p.nseg * p.ngroups;
  // ---- appended problems: descriptors into one device buffer, work items after the main problem's
  a.nextra = 0;
  a.xdev = nullptr;
  pl.xbuf.reset();
  if (!p.extra.empty()) {
    const bool x_scatter = p.epi == EPI_SCATTER_F32 && pl.epi == EPI_SCATTER_F32;      // every extra scatters like the main problem
    if (!((p.epi == EPI_STORE_F32 && pl.epi == EPI_STORE_F32) || x_scatter) || (cm != 1 && !pair))
      return "extra problems need EPI_STORE_F32 or EPI_SCATTER_F32 and no multicast cluster";
    if ((int)p.extra.size() > kMaxExtra) return "too many extra problems";
    std::vector<GemmExtraDev> host(p.extra.size());
    memset(host.data(), 0, sizeof(GemmExtraDev) * host.size());
    int item0 = 0;
    for (size_t i = 0; i < p.extra.size(); ++i) {
      const GemmExtra& x = p.extra[i];
      GemmExtraDev& d = host[i];
      for (int sgm = 0; sgm < p.nseg; ++sgm) {
        const Mat& A = x.a[sgm];
        const Mat& Bm = x.b[sgm];
        std::string err;
        if (!a_mn) {
          if (A.cols != x.K) return "extra A cols != K";
          err = make_tmap(&d.ta[sgm], A.ptr, A.cols, A.rows, A.ld, kBlockK, kBlockM);
        } else {
          if (A.rows != x.K) return "extra A rows != K";
          err = make_tmap_mn(&d.ta[sgm], A.ptr, A.cols, A.rows, A.ld, kBlockK, kBlockM / 64);
        }
        if (!err.empty()) return "extra A: " + err;
        if (!b_mn) {
          if (Bm.cols != x.K) return "extra B cols != K";
          err = make_tmap(&d.tb[sgm], Bm.ptr, Bm.cols, Bm.rows, Bm.ld, kBlockK, pair ? bn / 2 : bn);
        } else {
          if (Bm.rows != x.K) return "extra B rows != K";
          err = make_tmap_mn(&d.tb[sgm], Bm.ptr, Bm.cols, Bm.rows, Bm.ld, kBlockK, (bn / 64) / (pair ? 2 : 1));
        }
        if (!err.empty()) return "extra B: " + err;
      }
      d.epi = x.e;
      d.epi.atomic = 0;
      if (!x_scatter) {
        if ((reinterpret_cast<uintptr_t>(x.e.out) & 15) != 0 || (x.e.ld_out * 4) % 16 != 0) return "extra output not 16-byte aligned";
        std::string err = make_tmap_out(&d.tout[0], x.e.out, true, x.N, x.M, x.e.ld_out);
        if (!err.empty()) return "extra out: " + err;
      }
      const int tm = ((x.M + kBlockM - 1) / kBlockM + cm - 1) / cm, tn = (x.N + bn - 1) / bn;     // cluster rows
      a.x_item0[i] = item0;
      a.x_m[i] = x.M; a.x_n[i] = x.N; a.x_kb[i] = (x.K + kBlockK - 1) / kBlockK; a.x_tn[i] = tn;
      item0 += tm * tn;
      pl.flops += 2.0 * x.M * (double)x.N * x.K * p.nseg;
    }
    a.x_item0[p.extra.size()] = item0;
    a.nextra = (int)p.extra.size();
    void* dptr = nullptr;
    if (cudaMalloc(&dptr, sizeof(GemmExtraDev) * host.size()) != cudaSuccess) return "cudaMalloc(extra descriptors) failed";
    pl.xbuf = std::shared_ptr<void>(dptr, [](void* q) { cudaFree(q); });
    if (cudaMemcpy(dptr, host.data(), sizeof(GemmExtraDev) * host.size(), cudaMemcpyHostToDevice) != cudaSuccess)
      return "cudaMemcpy(extra descriptors) failed";
    a.xdev = static_cast<const GemmExtraDev*>(dptr);
    items += item0;
  }
  pl.items = items;
  pl.order.reset();
  pl.klist.reset();
  a.order = nullptr;
  a.klist = nullptr;
  if (a.kmask[0] != nullptr && cm == 1) {
    // k-block lists, dense (every candidate) until gemm_schedule has seen a map
    const int S = a.splitk, total_kb2 = a.nseg * a.kblocks;
    const int rows = a.ngroups * a.tiles_m * S;
    const int pitch = (int)((1 + (total_kb2 + S - 1) / S + 3) / 4 * 4);
    if (total_kb2 > 65535) return "kmask: too many k-blocks for 16-bit lists";
    std::vector<uint16_t> hl((size_t)rows * pitch, 0);
    std::vector<int> hc((size_t)rows, 0);
    for (int r = 0; r < rows; ++r) {
      const int ks = r % S;
      int n = 0;
      for (int kb = S > 1 ? ks : 0; kb < total_kb2; kb += S > 1 ? S : 1) hl[(size_t)r * pitch + 1 + n++] = (uint16_t)kb;
      hl[(size_t)r * pitch] = (uint16_t)n;
      hc[(size_t)r] = n;
    }
    const size_t lbytes = sizeof(uint16_t) * hl.size();     // a multiple of 8: the counts behind it stay aligned
    void* dptr = nullptr;
    if (cudaMalloc(&dptr, lbytes + sizeof(int) * hc.size()) != cudaSuccess) return "cudaMalloc(k-block lists) failed";
    pl.klist = std::shared_ptr<void>(dptr, [](void* q) { cudaFree(q); });
    if (cudaMemcpy(dptr, hl.data(), lbytes, cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMemcpy(static_cast<char*>(dptr) + lbytes, hc.data(), sizeof(int) * hc.size(), cudaMemcpyHostToDevice) != cudaSuccess)
      return "cudaMemcpy(k-block lists) failed";
    pl.klist_rows = rows;
    a.klist = static_cast<const uint16_t*>(dptr);
    a.klist_pitch = pitch;
  }
  if (a.klist != nullptr && items <= kSchedMaxItems) {       // identity until gemm_schedule sees a map
    std::vector<int> ident((size_t)items);
    for (long long i = 0; i < items; ++i) ident[(size_t)i] = (int)i;
    void* dptr = nullptr;
    if (cudaMalloc(&dptr, sizeof(int) * (size_t)items) != cudaSuccess) return "cudaMalloc(work order) failed";
    pl.order = std::shared_ptr<void>(dptr, [](void* q) { cudaFree(q); });
    if (cudaMemcpy(dptr, ident.data(), sizeof(int) * (size_t)items, cudaMemcpyHostToDevice) != cudaSuccess)
      return "cudaMemcpy(work order) failed";
    a.order = static_cast<const int*>(dptr);
  }
  const long long slots = num_sms / cm;              // refined at launch by the occupancy query when cm > 1
  pl.grid = (int)(items < slots ? items : slots) * cm;
  return "";
}

std::string gemm_fuse_adam(GemmPlan* plan, const AdamFuse& af) {
  GemmPlan& pl = *plan;
  const GemmArgs& a = pl.args;
  if (pl.kind != GEMM_WGRAD || pl.epi != EPI_STORE_F32 || !pl.a_bits || pl.cta_pair || pl.two_pass)
    return "fused Adam: needs the packed-operand wgrad kernel with the plain fp32 epilogue";
  if (a.ngroups != 2 || a.splitk != 1 || a.cluster_m != 1 || a.bn != 256 || a.klist != nullptr)
    return "fused Adam: needs two groups, whole-K work items, 256-wide tiles and no zero-stage lists";
  if (a.epi[0].atomic || a.epi[1].atomic) return "fused Adam: the gradient is accumulated (split-K)";
  if (a.N % 4 != 0 || af.ld % 4 != 0) return "fused Adam: the layer's width must be a multiple of 4";
  const long long pairs = (long long)a.tiles_m * a.tiles_n, grid = pl.grid;
  const long long main_items = 2 * pairs;
  if (pl.items < main_items || grid <= 0) return "fused Adam: inconsistent plan";
  // Pair rows: the CTAs walk w = cta, cta + grid, ...; rows 2i and 2i + 1 of that walk hold group 0 and group 1 of pair
  // i * grid + cta.  Only whole rows of pairs are laid out this way: the remaining pairs (fewer than `grid`) would
  // occupy two more rows with most CTAs idle, so their items follow one by one -- in the wave they share with the
  // appended problems -- and store their gradient; the caller's Adam launch covers their rows.
  const long long paired = pairs / grid * grid;
  const long long mt_fused = paired / a.tiles_n;                        // whole M tiles inside the paired region
  if (mt_fused <= 0) return "fused Adam: fewer tiles than CTAs";
  auto item = [&](long long pr, int g) -> int {                         // inverse of decode_work (main problem, splitk 1)
    const long long mt = pr / a.tiles_n, nt = pr % a.tiles_n;
    return (int)(a.m_fastest ? ((long long)g * a.tiles_n + nt) * a.tiles_m + mt : ((long long)g * a.tiles_m + mt) * a.tiles_n + nt);
  };
  std::vector<int> order((size_t)pl.items);
  for (long long w = 0; w < 2 * paired; ++w) {
    const long long row = w / grid, c = w % grid;
    order[(size_t)w] = item((row >> 1) * grid + c, (int)(row & 1));
  }
  long long w = 2 * paired;
  for (long long pr = paired; pr < pairs; ++pr)
    for (int g = 0; g < 2; ++g) order[(size_t)w++] = item(pr, g);
  for (; w < pl.items; ++w) order[(size_t)w] = (int)w;                  // appended problems
  void* dptr = nullptr;
  if (cudaMalloc(&dptr, sizeof(int) * order.size()) != cudaSuccess) return "cudaMalloc(fused Adam work order) failed";
  pl.order = std::shared_ptr<void>(dptr, [](void* q) { cudaFree(q); });
  if (cudaMemcpy(dptr, order.data(), sizeof(int) * order.size(), cudaMemcpyHostToDevice) != cudaSuccess)
    return "cudaMemcpy(fused Adam work order) failed";
  pl.args.order = static_cast<const int*>(dptr);
  pl.args.adam = af;
  pl.args.adam.rows_end = (int)(mt_fused * kBlockM);
  pl.adam_rows_end = mt_fused * kBlockM;
  pl.adam_fuse = true;
  return "";
}

template <bool A_MN, bool B_MN, int EPI, int BN = 256, bool A_BITS = false, bool CG2 = false>
static void launch_variant(const GemmPlan& pl, cudaStream_t s) {
  const int cm = pl.args.cluster_m;
  auto kernel = gemm_tcgen05_kernel<A_MN, B_MN, EPI, BN, A_BITS, CG2>;
  if (cm <= 1) {
    launch_chain(kernel, dim3(pl.grid), dim3(gemm_threads(EPI, A_BITS)), gemm_smem_bytes(EPI), s, pl.args);
    return;
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof cfg);
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cm;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;       // CTA pairs chain like every other step kernel
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.blockDim = dim3(gemm_threads(EPI, A_BITS));
  cfg.dynamicSmemBytes = gemm_smem_bytes(EPI);
  cfg.stream = s;
  cfg.attrs = attr;
  cfg.numAttrs = (CG2 && pdl_enabled()) ? 2 : 1;
  // A cluster lives inside one GPC, so fewer than num_sms / cm clusters may be co-resident; the persistent loop
  // strides by the number of clusters.
  static int max_clusters[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (max_clusters[cm] == 0) {
    cfg.gridDim = dim3(pl.grid);
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kernel, &cfg) != cudaSuccess || n <= 0) {
      (void)cudaGetLastError();
      n = pl.grid / cm;
    }
    max_clusters[cm] = n;
    if (getenv("MAGIC_VERBOSE")) fprintf(stderr, "[magic] cluster_m=%d%s: %d co-resident clusters\n", cm, CG2 ? " (CTA pairs)" : "", n);
  }
  long long ncl = pl.items < max_clusters[cm] ? pl.items : max_clusters[cm];
  cfg.gridDim = dim3((unsigned)(ncl * cm));
  cudaLaunchKernelEx(&cfg, kernel, pl.args);
}

std::string chain_plan(const std::vector<const GemmPlan*>& layers, const std::vector<Mat>& weights,
                       const std::vector<ChainOut>& outs, int num_sms, ChainPlan* plan) {
  ChainPlan& cp = *plan;
  cp.valid = false;
  memset(&cp.args, 0, sizeof cp.args);
  const int nl = (int)layers.size();
  if (nl < 2 || nl > kChainMaxLayers || outs.size() != layers.size() || weights.size() != layers.size()) return "chain: 2..4 layers";
  const int kind = layers[0]->kind;
  if (kind != GEMM_FWD && kind != GEMM_DGRAD) return "chain: forward or dgrad layers";
  cp.dgrad = kind == GEMM_DGRAD;
  cp.flops = 0;
  for (int l = 0; l < nl; ++l) {
    const GemmPlan& pl = *layers[l];
    const GemmArgs& a = pl.args;
    // a layer the planner split along K runs a linear store + a finishing pass: its real epilogue is the finishing one
    const int epi_kind = pl.two_pass ? pl.finish_epi_kind : pl.epi;
    const EpiArgs& e = pl.two_pass ? pl.finish_epi[0] : a.epi[0];
    if (pl.kind != kind || epi_kind != (cp.dgrad ? EPI_DACT : EPI_ACT)) return "chain: mixed kinds / epilogues";
    if (pl.a_bits || pl.cta_pair || a.ngroups != 1 || a.nseg != 1 || a.cluster_m != 1 || a.nextra != 0)
      return "chain: a layer is grouped, segmented (fp32 mode) or clustered";
    if (a.N != kChainN || a.kblocks > kChainKb) return "chain: layers must be 512 wide and at most 512 deep";
    if (a.M != layers[0]->args.M) return "chain: layers differ in M";
    if (l > 0 && a.kblocks != kChainKb) return "chain: inner layers must read the whole 512-wide slab";
    if (e.out_lo != nullptr || e.precise) return "chain: bf16 mode only";
    if (outs[l].ld != kChainN || outs[l].rows < a.M || static_cast<void*>(outs[l].ptr) != e.out) return "chain: output does not match the layer's plan";
    // weights as the layer's B operand, in 256-column boxes whatever tile width the layer's own plan chose
    const Mat& W = weights[l];
    std::string err = cp.dgrad ? make_tmap(&cp.args.tb[l], W.ptr, W.cols, W.rows, W.ld, kBlockK, 256)
                               : make_tmap_mn(&cp.args.tb[l], W.ptr, W.cols, W.rows, W.ld, kBlockK, 256 / 64);
    if (!err.empty()) return "chain weights: " + err;
    cp.args.epi[l] = e;
    cp.args.kb[l] = a.kblocks;
    err = make_tmap(&cp.args.tout[l], outs[l].ptr, kChainN, a.M, outs[l].ld, kBlockK, kBlockM);
    if (!err.empty()) return "chain out: " + err;
    cp.flops += pl.flops;
  }
  cp.args.ta0 = layers[0]->args.ta[0][0];                  // box {64 k, 128 m} at every tile width
  cp.args.nlayers = nl;
  cp.args.M = layers[0]->args.M;
  cp.args.n_slabs = (cp.args.M + kBlockM - 1) / kBlockM;
  // a handful of slabs would leave most SMs idle for the whole chain: the per-layer launches spread wider then
  if (cp.args.n_slabs < 16) return "chain: fewer than 16 slabs of 128 rows";
  cp.grid = cp.args.n_slabs < num_sms ? cp.args.n_slabs : num_sms;
  cp.valid = true;
  return "";
}

int chain_launch(const ChainPlan& cp, cudaStream_t s, std::string* err) {
  if (!cp.valid) { if (err) *err = "chain_launch: plan not valid"; return -1; }
  if (cp.dgrad) launch_chain(chain_tcgen05_kernel<true>, dim3(cp.grid), dim3(kChainThreads), kChainSmemBytes, s, cp.args);
  else launch_chain(chain_tcgen05_kernel<false>, dim3(cp.grid), dim3(kChainThreads), kChainSmemBytes, s, cp.args);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { if (err) *err = std::string("chain launch failed: ") + cudaGetErrorString(e); return -1; }
  return 1;
}

int gemm_launch(const GemmPlan& pl, cudaStream_t s, std::string* err) {
  int launches = 0;
  if (pl.two_pass) {
    zero_fill(pl.ws, sizeof(float) * pl.ws_ld * pl.args.M, s);
  }
  const int key = pl.kind * 8 + (pl.adam_fuse ? (int)EPI_ADAM : pl.epi) + (pl.args.bn == 128 ? 64 : 0) + (pl.a_bits ? 128 : 0) +
                  (pl.cta_pair ? 256 : 0);
  switch (key) {
    case 128 + GEMM_WGRAD * 8 + EPI_ADAM: launch_variant<true, true, EPI_ADAM, 256, true>(pl, s); break;
    case 256 + GEMM_FWD * 8 + EPI_BCE: launch_variant<false, true, EPI_BCE, 256, false, true>(pl, s); break;
    case 256 + GEMM_WGRAD * 8 + EPI_STORE_F32: launch_variant<true, true, EPI_STORE_F32, 256, false, true>(pl, s); break;
    case 256 + GEMM_DGRAD * 8 + EPI_STORE_F32: launch_variant<false, false, EPI_STORE_F32, 256, false, true>(pl, s); break;
    case 256 + 128 + GEMM_FWD * 8 + EPI_STORE_F32: launch_variant<false, true, EPI_STORE_F32, 256, true, true>(pl, s); break;
    case 256 + 128 + GEMM_FWD * 8 + EPI_ACT: launch_variant<false, true, EPI_ACT, 256, true, true>(pl, s); break;
    case 256 + 128 + GEMM_WGRAD * 8 + EPI_STORE_F32: launch_variant<true, true, EPI_STORE_F32, 256, true, true>(pl, s); break;
    case 128 + GEMM_FWD * 8 + EPI_STORE_F32: launch_variant<false, true, EPI_STORE_F32, 256, true>(pl, s); break;
    case 128 + GEMM_FWD * 8 + EPI_ACT: launch_variant<false, true, EPI_ACT, 256, true>(pl, s); break;
    case 128 + GEMM_WGRAD * 8 + EPI_STORE_F32: launch_variant<true, true, EPI_STORE_F32, 256, true>(pl, s); break;
    case 128 + GEMM_WGRAD * 8 + EPI_SCATTER_F32: launch_variant<true, true, EPI_SCATTER_F32, 256, true>(pl, s); break;
    case 64 + GEMM_FWD * 8 + EPI_STORE_F32: launch_variant<false, true, EPI_STORE_F32, 128>(pl, s); break;
    case 64 + GEMM_FWD * 8 + EPI_ACT: launch_variant<false, true, EPI_ACT, 128>(pl, s); break;
    case 64 + GEMM_DGRAD * 8 + EPI_STORE_F32: launch_variant<false, false, EPI_STORE_F32, 128>(pl, s); break;
    case 64 + GEMM_DGRAD * 8 + EPI_DACT: launch_variant<false, false, EPI_DACT, 128>(pl, s); break;
    case 64 + GEMM_WGRAD * 8 + EPI_STORE_F32: launch_variant<true, true, EPI_STORE_F32, 128>(pl, s); break;
    case GEMM_FWD * 8 + EPI_STORE_F32: launch_variant<false, true, EPI_STORE_F32>(pl, s); break;
    case GEMM_FWD * 8 + EPI_ACT: launch_variant<false, true, EPI_ACT>(pl, s); break;
    case GEMM_FWD * 8 + EPI_BCE: launch_variant<false, true, EPI_BCE>(pl, s); break;
    case GEMM_FWD * 8 + EPI_SIGMOID: launch_variant<false, true, EPI_SIGMOID>(pl, s); break;
    case GEMM_DGRAD * 8 + EPI_STORE_F32: launch_variant<false, false, EPI_STORE_F32>(pl, s); break;
    case GEMM_DGRAD * 8 + EPI_DACT: launch_variant<false, false, EPI_DACT>(pl, s); break;
    case GEMM_WGRAD * 8 + EPI_STORE_F32: launch_variant<true, true, EPI_STORE_F32>(pl, s); break;
    case GEMM_WGRAD * 8 + EPI_SCATTER_F32: launch_variant<true, true, EPI_SCATTER_F32>(pl, s); break;
    default:
      if (err) *err = "gemm_launch: unsupported (kind, epilogue) combination";
      return -1;
  }
  ++launches;
  if (pl.two_pass) {
    const int chunks = (pl.args.N + 31) / 32;
    const long long n = (long long)pl.args.M * chunks;
    const int bs = 256;
    const int gs = (int)((n + bs - 1) / bs);
    if (pl.finish_epi_kind == EPI_ACT)
      launch_chain(gemm_finish_kernel<EPI_ACT>, dim3(gs), dim3(bs), 0, s, pl.finish_epi[0], (const float*)pl.ws, pl.ws_ld,
                   pl.args.M, pl.args.N);
    else
      launch_chain(gemm_finish_kernel<EPI_DACT>, dim3(gs), dim3(bs), 0, s, pl.finish_epi[0], (const float*)pl.ws, pl.ws_ld,
                   pl.args.M, pl.args.N);
    ++launches;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    if (err) *err = std::string("gemm launch failed: ") + cudaGetErrorString(e);
    return -1;
  }
  return launches;
}

}  // namespace magic
