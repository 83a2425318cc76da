// Host-side planning and launch of the tcgen05 GEMM (gemm_sm100.cuh).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <memory>
#include <string>
#include <vector>
#include "gemm_sm100.cuh"
#include "chain_sm100.cuh"

namespace magic {

// A bf16 row-major matrix view in HBM: element (r, c) at ptr[r * ld + c].
struct Mat {
  const __nv_bfloat16* ptr = nullptr;
  long long rows = 0, cols = 0, ld = 0;
  Mat() {}
  Mat(const __nv_bfloat16* p, long long r, long long c, long long l) : ptr(p), rows(r), cols(c), ld(l) {}
};

enum GemmKind : int {
  GEMM_FWD = 0,    // D[M,N] = A[M,K] . B[K,N]        A: Mat[M,K]   B: Mat[K,N] (weights as stored)
  GEMM_DGRAD = 1,  // D[M,N] = A[M,K] . B[N,K]^T      A: Mat[M,K]   B: Mat[N,K]
  GEMM_WGRAD = 2,  // D[M,N] = A[K,M]^T . B[K,N]      A: Mat[K,M]   B: Mat[K,N]   (K = batch rows)
};

struct GemmPlan {
  GemmArgs args;
  int kind = 0, epi = 0, grid = 0;
  bool a_bits = false;   // the A_BITS kernel variant: the main problem's A operand is read as packed bits (args.abits)
  bool cta_pair = false; // the CG2 kernel variant (clusters of two CTAs, args.cluster_m == 2)
  long long items = 0;   // work items (per cluster when args.cluster_m > 1)
  // split-K with a non-linear epilogue: GEMM accumulates fp32 into ws (red.add), then finish applies epi
  bool two_pass = false;
  float* ws = nullptr;
  long long ws_ld = 0;
  EpiArgs finish_epi[kMaxGroups];
  int finish_epi_kind = 0;
  double flops = 0;
  std::shared_ptr<void> xbuf;   // device copy of the extras' descriptors (GemmExtraDev[nextra])
  std::shared_ptr<void> order;  // plans with a zero-stage map: the work-item permutation args.order points at (int[items])
  std::shared_ptr<void> klist;  // ... their k-block lists (args.klist, uint16[rows][klist_pitch]) followed by the list lengths (int[rows])
  int klist_rows = 0;
  bool adam_fuse = false;       // the EPI_ADAM kernel variant (gemm_fuse_adam): args.adam is live, args.order pairs the groups
  long long adam_rows_end = 0;  // rows [0, adam_rows_end) of the output are updated in the epilogue, the rest is stored
};

// A problem appended to another problem's launch (GemmExtraDev in gemm_sm100.cuh): same kind / epilogue / segment
// count, own shape, operands and output; never split along K.
struct GemmExtra {
  Mat a[kMaxSegs], b[kMaxSegs];
  EpiArgs e;
  int M = 0, N = 0, K = 0;
};

struct GemmProblem {
  int kind = GEMM_FWD;
  int epi = EPI_STORE_F32;
  int ngroups = 1, nseg = 1;
  Mat a[kMaxGroups][kMaxSegs], b[kMaxGroups][kMaxSegs];
  EpiArgs e[kMaxGroups];
  int M = 0, N = 0, K = 0;   // N: columns the epilogue writes (ACT: the output pitch)
  int force_splitk = 0;      // 0 = heuristic
  std::vector<GemmExtra> extra;   // appended problems (EPI_STORE_F32 only)
  int block_n = 0;           // tile width: 0 = heuristic, 128 or 256 forces it (self-test)
  int cluster_m = 1;         // CTAs per cluster along M sharing the B tile by TMA multicast (FWD / WGRAD kinds; 1, 2 or 4)
  bool cta_pair = false;     // ask for the CTA-pair kernel (tcgen05 cta_group::2, 256 x 256 tiles): used when the variant exists
  float* ws = nullptr;       // fp32 workspace [M, ws_ld] for two-pass split-K (ACT / DACT); may be null
  long long ws_ld = 0;
  // A as packed bits instead of a[g][*] (FWD: bits of A[M,K]; WGRAD: bits of A[K,M]; see BitsA in gemm_sm100.cuh).
  // Every precision segment reads the same bits (a binary operand is exact in bf16: its low part is zero), so with
  // nseg = 2 the segments are (A, B_hi), (A, B_lo).  Only b[g][*] is read from the problem; 256-wide tiles.
  bool a_bits = false;
  BitsA abits[kMaxGroups] = {};
  // optional zero-stage map of that operand (GemmArgs.kmask): kmask[g][tile_m * kmask_words + (kk >> 5)] bit kk & 31
  const uint32_t* kmask[kMaxGroups] = {};
  int kmask_words = 0;
};

// A chain of hidden layers run by ONE launch (chain_sm100.cuh).  Built from the per-layer plans it replaces: their weight
// tensor maps and epilogue arguments are reused, so the chain computes exactly what launching them one by one would.
struct ChainPlan {
  ChainArgs args;
  bool valid = false, dgrad = false;
  int grid = 0;
  double flops = 0;
};
struct ChainOut { __nv_bfloat16* ptr; long long rows, ld; };      // output activation of a layer: [rows, ld] bf16
// `layers` in execution order; `outs[l]` = where layer l writes.  Returns "" and sets plan->valid when the chain kernel
// applies (every layer: forward + EPI_ACT or dgrad + EPI_DACT, bf16, one group / segment, 512 wide, at most 8 k-blocks;
// 2..4 layers; at least 16 slabs of 128 rows); returns the reason otherwise (plan->valid = false: launch the layers one by one).
std::string chain_plan(const std::vector<const GemmPlan*>& layers, const std::vector<Mat>& weights /* each layer's B operand */,
                       const std::vector<ChainOut>& outs, int num_sms, ChainPlan* plan);
int chain_launch(const ChainPlan& plan, cudaStream_t stream, std::string* err);

// Returns empty string on success, else an error message.
std::string gemm_plan(const GemmProblem& p, int num_sms, GemmPlan* out);
// Enqueue (workspace zeroing, GEMM, finishing pass) on the stream.  Returns the number of kernels launched.
int gemm_launch(const GemmPlan& plan, cudaStream_t stream, std::string* err);
// Point group g's epilogue at another output matrix of the same shape and pitch (re-encodes its output tile maps).
std::string gemm_set_output(GemmPlan* plan, int g, void* out, void* out_lo);
// Turn a planned first-layer weight-gradient launch (wgrad kind, packed-bit A operand, two groups, EPI_STORE_F32, not
// split along K, no zero-stage lists) into its EPI_ADAM variant (AdamFuse in gemm_sm100.cuh): builds the work order that
// hands both groups of a tile to one CTA and sets plan->adam_rows_end.  af.rows_end / lr_t are filled in here / per launch.
// Returns the reason when the plan does not qualify (the plan is left as it was).
std::string gemm_fuse_adam(GemmPlan* plan, const AdamFuse& af);
// Plans with a zero-stage map: rebuild the work-item order from the map's current contents (one small kernel on the
// stream that just rebuilt the map).  No-op for other plans.
void gemm_schedule(const GemmPlan& plan, cudaStream_t stream);
std::string gemm_init();   // resolves cuTensorMapEncodeTiled, sets max dynamic smem on every variant
int choose_splitk(int tiles, int total_kb, int num_sms, bool two_pass = false);

}  // namespace magic
