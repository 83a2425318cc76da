// Persistent, warp-specialised tcgen05 GEMM for sm_100a.
//
//   D[M,N] = sum over segments s of  A_s[M,K] * B_s[K,N]     (bf16 in, fp32 accumulate in TMEM)
//
// One kernel serves every dense layer of the magic autoencoder step:
//   forward   y = x.W        A K-major (activations [M,K]),  B MN-major (weights [K,N] as the reference
//                             stores them, 8c/vae.py:93,347)
//   dgrad     dx = dy.W^T    A K-major, B K-major
//   wgrad     dW = x^T.dy    A MN-major, B MN-major (reduction over the batch)
// "Segments" concatenate along K: they carry the bf16x3 split of the fp32 mode (hi*hi + hi*lo + lo*hi)
// through one accumulator.  "Groups" batch independent problems of identical shape (the two optimisers'
// wgrads) into one launch.  Split-K spreads long reductions with few output tiles over all SMs; partial
// sums are combined with vector red.global.add.
//
// Roles (320 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + single-thread MMA issuer,
// warps 2..9 = epilogue (TMEM -> registers -> fused epilogue -> global).  Two TMEM accumulator stages
// (2 x 256 columns) let the epilogue of tile i overlap the MMAs of tile i+1.
#pragma once
#include <cuda.h>
#include "ptx.cuh"

namespace magic {

constexpr int kBlockM = 128;
constexpr int kBlockN = 256;   // default tile width; the kernel is also instantiated with BN = 128 (see TileCfg)
constexpr int kBlockK = 64;   // 64 bf16 = one 128-byte swizzle row
constexpr int kStages = 4;
constexpr int kUmmaK = 16;
constexpr int kMaxGroups = 2;
constexpr int kMaxSegs = 3;
// warp 0 TMA, warp 1 MMA, then the epilogue warps: 8 (two per TMEM lane quarter / scheduler) for the cheap
// epilogues, 16 for the sigmoid-BCE epilogue whose ~20 instructions per element need four warps per scheduler to
// hide MUFU and TMEM latencies.
__host__ __device__ constexpr int epi_warps(int epi) { return epi == 3 /*EPI_BCE*/ ? 16 : 8; }
// A_BITS kernels (packed first-layer operand) run FOUR epilogue warps, one per TMEM lane quarter: with their four
// expander warps that is 10 warps = 320 threads, the footprint of every other GEMM CTA, and a CTA of the batch generator /
// dataset feed (384 threads x 56 registers) fits on the same SM: 3 + 3 warps per sub-partition.  With eight (448 threads,
// 4 warps on two of the sub-partitions) it did not, and whatever part of the generator had not finished under the first
// half of the backward pass delayed the first-layer wgrad.  Their items are long (K = 8192 .. 13355 per tile), so an
// epilogue twice as slow still hides under the next item's main loop.
__host__ __device__ constexpr int epi_warps(int epi, bool a_bits) { return a_bits ? 4 : epi_warps(epi); }
__host__ __device__ constexpr int gemm_threads(int epi) { return 64 + 32 * epi_warps(epi); }
// A_BITS variants (the first encoder layer: its A operand is a binary image, kept in HBM as 1 bit per pixel): four more
// warps expand the packed bits of a pipeline stage into the bf16 swizzle layout the MMA reads -- see BitsA.
constexpr int kExpandWarps = 4;     // one per pipeline stage of the 256-wide tile configuration
__host__ __device__ constexpr int gemm_threads(int epi, bool a_bits) { return 64 + 32 * epi_warps(epi, a_bits) + (a_bits ? 32 * kExpandWarps : 0); }
constexpr uint32_t kStageBytesA = kBlockM * kBlockK * 2;   // 16 KiB
constexpr uint32_t kStageBytesB = kBlockN * kBlockK * 2;   // 32 KiB
constexpr uint32_t kStageBytes = kStageBytesA + kStageBytesB;
// Tile width BN = 256: 4 stages x 48 KiB, two 256-column accumulators.  BN = 128 (the short-K layers, where the
// epilogue of a tile outlasts its 8 k-blocks of MMA and the last one is exposed): 6 stages x 32 KiB -- the same
// 192 KiB -- and four 128-column accumulators, so a CTA's exposed final epilogue is half as long, the issuer may run
// three tiles ahead of the epilogue warps, and twice as many tiles fill the 148 SMs more evenly.
template <int BN>
struct TileCfg {
  static_assert(BN == 128 || BN == 256, "tile width");
  static constexpr int kStagesN = BN == 256 ? 4 : 6;
  static constexpr uint32_t kBytesB = BN * kBlockK * 2;
  static constexpr uint32_t kBytes = kStageBytesA + kBytesB;
  static constexpr int kAcc = 512 / BN;                    // accumulator stages in the 512 TMEM columns
  static_assert(kStagesN * kBytes == kStages * kStageBytes, "both tile widths use the same pipeline footprint");
};
constexpr uint32_t kAtomBytes = 64 * kBlockK * 2;          // one 64(MN) x 64(K) MN-major TMA box
constexpr uint32_t kTmemCols = 2 * kBlockN;                // 512: two accumulator stages
constexpr uint32_t kGemmSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
// EPI_SCATTER_F32 stages 8 rows x 128 B per epilogue warp so that every row segment leaves the SM as ONE 128-byte
// bulk copy (full NVLink packets) instead of eight 16-byte partial-sector stores from a strided warp
constexpr uint32_t kScatterStageBytes = 8 * 128;
// Every epilogue except EPI_SCATTER_F32 stages its 32-row chunks for TMA tile stores (see TileOut), 2 KiB per
// epilogue warp: 32 rows x 64 B, i.e. 32 bf16 columns or 16 fp32 columns per store.  16 KiB (32 KiB with the BCE
// epilogue's 16 warps) on top of the pipeline keeps the CTA at 210 KiB, so that a CTA of the batch generator
// (15 KiB, kernels.cu) still fits on the same SM and runs under the backward pass.
__host__ __device__ constexpr uint32_t tile_stage_bytes(int epi) { return 2048u * (uint32_t)epi_warps(epi); }
constexpr uint32_t kTileStageOffset = kStages * kStageBytes + 1024;   // after the barriers, 1024-byte aligned
__host__ __device__ constexpr uint32_t gemm_smem_bytes(int epi) {
  return kGemmSmemBytes + (epi == 5 /*EPI_SCATTER_F32*/ ? 8 * kScatterStageBytes : 1024 + tile_stage_bytes(epi));
}

enum EpiKind : int { EPI_STORE_F32 = 0, EPI_ACT = 1, EPI_DACT = 2, EPI_BCE = 3, EPI_SIGMOID = 4, EPI_SCATTER_F32 = 5,
                     EPI_ADAM = 6 };   // EPI_ADAM: kernel variant only (planned as EPI_STORE_F32, see AdamFuse)
constexpr int kMaxPeers = 8;
enum ActKind : int { ACT_NONE = 0, ACT_TANH = 1, ACT_ELU = 2 };

struct EpiArgs {
  void* out;                  // f32 (STORE_F32, SIGMOID) or bf16 (ACT, DACT, BCE)
  void* out_lo;               // bf16 low part for the split (fp32) mode, or nullptr
  long long ld_out;           // row pitch of out, elements
  const __nv_bfloat16* aux;   // DACT: forward activations y
  const uint32_t* xbits;      // BCE: packed target plane x (bit p & 31 of word p >> 5 of row r), row pitch xb_ld words
  int xb_ld;
  const __nv_bfloat16* aux_lo;
  long long ld_aux;
  int aux_split;              // DACT: aux row = r < aux_split ? r : r - aux_shift
  int aux_shift;
  float alpha;                // scale on the accumulator (STORE_F32/ACT/DACT), on the loss sum (BCE)
  int n_valid;                // ACT/DACT: cols < n_valid are data; ACT writes 1.0 at col n_valid (bias column)
  int act;                    // ActKind
  int atomic;                 // STORE_F32: red.add instead of store (split-K)
  int precise;                // 1: accurate tanhf/expm1f/logf;  0: MUFU approximations (bf16 mode)
  float* scalar;              // BCE: sum of the per-element loss (scaled by alpha) is added here
  // SCATTER_F32 (data-parallel wgrad fused with its reduce-scatter): the tile goes to the rank that owns its slice of
  // the flat parameter space, into that rank's staging buffer (peer memory over NVLink), slot `src_rank`:
  //   idx = flat_base + row * ld_out + col ;  granule q = idx >> gran_shift ;  owner o = q % world ;
  //   peer[o][src_rank * slice_len + ((q / world) << gran_shift) + (idx & (granule - 1))]
  // (granules are dealt round-robin so that every rank receives and updates the same share of every layer)
  float* peer[kMaxPeers];
  long long flat_base;
  int slice_len;
  int src_rank;
  int world;
  int gran_shift;
  int dbg;           // diagnostics (MAGIC_EPI_DBG): 1 = BCE epilogue without stores, 4 = no epilogue work at all (main loop only)
};

// "Extra" problems ride at the end of a launch's work list: same operand layouts, precision segments and epilogue kind
// as the main problem, but their own shapes, operands and outputs (the engine folds the small layers' weight
// gradients into the tail of the first layer's wgrad launch, whose last wave would otherwise leave half the SMs
// idle).  Their descriptors live in global memory so that ordinary launches do not carry them as parameters.
constexpr int kMaxExtra = 20;
struct alignas(128) GemmExtraDev {
  CUtensorMap ta[kMaxSegs];
  CUtensorMap tb[kMaxSegs];
  CUtensorMap tout[2];
  EpiArgs epi;
};

// EPI_ADAM (single GPU, first layer's wgrad): the weight gradient never leaves the SM.  The main problem has two groups
// (optimiser 1 = cost, optimiser 2 = metric: 8c/vae.py:314-321) and the work order (GemmArgs.order, built by
// gemm_fuse_adam) hands both groups of an output tile to the SAME CTA, group 0 first, so the same thread sees the same
// elements twice:   group 0:  m1, v1 <- ApplyAdam(g1) ; p <- p - d1        group 1:  m2, v2 <- ApplyAdam(g2) ; p <- p - d2 ;
// bf16 shadow <- p.  Same arithmetic, in the same order, as adam_dual_kernel (kernels.cu) on a stored gradient.  Tiles with
// m0 >= rows_end (the last, partly filled wave, whose pairs would cost a wave of their own) and the appended problems
// store their gradient as EPI_STORE_F32 does and are updated by the Adam launch that follows.
// Opt-in (magic_set_option "fuse_first_layer_adam"): on B200 the optimiser-state traffic slows the latency-bound main
// loop by as much as the separate Adam pass costs, so nothing is gained (profiles/r02q_*).
struct AdamFuse {
  float* p;                  // fp32 master weights of the layer, row pitch ld
  float* m[2]; float* v[2];  // the two optimisers' moments
  __nv_bfloat16* w_hi; __nv_bfloat16* w_lo;   // bf16 shadow (lo may be null)
  long long ld;
  int rows_end;              // output rows below this are updated in the epilogue (a multiple of 128)
  float lr_t[2];             // lr * sqrt(1-beta2^t) / (1-beta1^t), set per launch
  float beta1, beta2, eps;
  int dbg;                   // timing experiments (MAGIC_ADAM_DBG; results are wrong): 1 = no global accesses, 2 = no arithmetic, 4 = nothing
};

// The A operand as packed bits (A_BITS kernels): element (r, c) of a row-major {0,1} matrix is bit c & 31 of word
// base[r * pitch + (c >> 5)]; rows >= `rows` read as zero.  For the K-major kinds the matrix is A[M, K], for the wgrad
// kind it is A[K, M] (the stored orientation of the activations, as with the bf16 operands).  A 128 x 64 stage is
// 1 KiB of bits instead of 16 KiB of bf16 through L2; the expansion to bf16 happens on the way into shared memory.
struct BitsA {
  const uint32_t* base;
  long long pitch;     // words per row, even (rows are read as 8-byte pieces)
  int rows;
};

struct GemmArgs {
  CUtensorMap ta[kMaxGroups][kMaxSegs];
  CUtensorMap tb[kMaxGroups][kMaxSegs];
  EpiArgs epi[kMaxGroups];
  int M, N;          // rows < M and cols < N are written
  int kblocks;       // K blocks (of 64) per segment
  int nseg, ngroups, splitk;
  int tiles_m, tiles_n;
  int m_fastest;     // tile order: consecutive work items walk M first (B is the large operand) or N first
  int bn;            // tile width the plan was built for (128 or 256) = the kernel's BN
  int cluster_m;     // CTAs per cluster along M (1, 2 or 4): they share the N tile and receive B by TMA multicast
  const GemmExtraDev* xdev;   // [nextra] in global memory
  int nextra;
  int x_item0[kMaxExtra + 1]; // work-item offsets of the extras, relative to the end of the main problem's items
  int x_m[kMaxExtra], x_n[kMaxExtra], x_kb[kMaxExtra], x_tn[kMaxExtra];   // rows, cols, k-blocks per segment, N tiles
  long long* trace;  // diagnostics: [grid][64] clock64 stamps of the phases of each CTA (selftest "trace"), or nullptr
  CUtensorMap tout[kMaxGroups][2];   // output tiles (box 32 cols x 32 rows): [g][0] fp32 or bf16 hi plane, [g][1] lo plane
  BitsA abits[kMaxGroups];           // A_BITS kernels: the main problem's A operand (extras keep bf16 A through TMA)
  // Zero-stage map of the main problem's A operand (A_BITS kernels; null = none): bit kk of row (m0 / 128) of group g's
  // map says "k-block kk of this M tile holds at least one set pixel".  A stage whose 128 x 64 pixels are all zero adds
  // exact zeros to the accumulator, so it is neither fetched (B) nor expanded (A) nor multiplied.  The kernel does not
  // read the map: gemm_schedule() turns it into one list of k-blocks per (group, M tile, split-K slice) -- klist row
  // (g * tiles_m + mt) * splitk + ks: entry 0 = count (>= 1: an empty list is given the slice's first block, an
  // all-zero stage, so that the accumulator is written), entries 1.. = k-block indices over nseg * kblocks, ascending.
  // With lists the split-K slices interleave (slice s owns k-blocks s, s + splitk, ...), so that every slice sees the
  // same mix of the images' empty margins and busy centres.
  const uint32_t* kmask[kMaxGroups];
  int kmask_words;                   // words per map row
  const uint16_t* klist;             // null: every k-block of [kb0, kb1)
  int klist_pitch;                   // entries per row (a multiple of 4: rows are read 8 bytes at a time)
  // With a map the work items differ in length (an M tile in the images' empty margin is one forced k-block, one in
  // the centre runs all of them), so the static round-robin `w = cta, cta + grid, ...` would leave most CTAs waiting
  // for the unlucky ones.  order[w] (null = identity) is a permutation of the work items, rebuilt with every map
  // (gemm_schedule): items sorted by their number of set stages and dealt to the CTAs in serpentine order.
  const int* order;
  AdamFuse adam;        // EPI_ADAM kernels only
};

struct WorkItem {
  int g;             // group of the main problem, or -1 - e for extra problem e
  int m0, n0, kb0, kb1;
  int kstep;         // k-blocks kb0, kb0 + kstep, ... < kb1 (1 unless the split-K slices interleave)
  int mt, ks;        // main problem: M tile (cluster row) and split-K slice of the item
  int M, N, kblocks; // of the problem this item belongs to
};

// Work items are per CLUSTER: `tiles_m` counts cluster rows (cluster_m M-tiles each); CTA `crank` takes M-tile
// mt * cluster_m + crank of the item.  With cluster_m == 1 this is the plain per-CTA decode.
__device__ __forceinline__ WorkItem decode_work(const GemmArgs& a, int w, int crank = 0) {
  WorkItem it;
  const int main_work = a.tiles_m * a.tiles_n * a.ngroups * a.splitk;
  if (w >= main_work) {                 // an extra problem: whole-K items, N tiles fastest
    const int x = w - main_work;
    int e = 0;
    while (e + 1 < a.nextra && x >= a.x_item0[e + 1]) ++e;
    const int l = x - a.x_item0[e];
    it.g = -1 - e;
    it.m0 = ((l / a.x_tn[e]) * a.cluster_m + crank) * kBlockM;      // (x_item0 counts cluster rows, like tiles_m)
    it.n0 = (l % a.x_tn[e]) * a.bn;
    it.kb0 = 0;
    it.kb1 = a.nseg * a.x_kb[e];
    it.kstep = 1;
    it.mt = 0; it.ks = 0;
    it.M = a.x_m[e]; it.N = a.x_n[e]; it.kblocks = a.x_kb[e];
    return it;
  }
  it.M = a.M; it.N = a.N; it.kblocks = a.kblocks;
  int nt, mt, t;
  if (a.m_fastest) {
    mt = w % a.tiles_m; t = w / a.tiles_m;
    nt = t % a.tiles_n; t /= a.tiles_n;
  } else {
    nt = w % a.tiles_n; t = w / a.tiles_n;
    mt = t % a.tiles_m; t /= a.tiles_m;
  }
  mt = mt * a.cluster_m + crank;
  it.g = t % a.ngroups;
  int ks = t / a.ngroups;
  int total_kb = a.nseg * a.kblocks;
  int per = (total_kb + a.splitk - 1) / a.splitk;
  it.ks = ks;
  it.mt = mt;
  if (a.klist != nullptr && a.splitk > 1) {          // interleaved slices (see GemmArgs.kmask)
    it.kb0 = ks; it.kb1 = total_kb; it.kstep = a.splitk;
  } else {
    it.kb0 = ks * per;
    it.kb1 = min(total_kb, it.kb0 + per);
    it.kstep = 1;
  }
  it.m0 = mt * kBlockM;
  it.n0 = nt * a.bn;
  return it;
}

// The k-blocks one work item processes, in order: its klist row if the plan has lists, else every kstep-th block of
// [kb0, kb1).  Producer and expanders each walk the same sequence, so a skipped block consumes no pipeline stage;
// walking a list is one 8-byte load per four blocks and `skip(n)` is an index addition (the expanders look four
// blocks ahead).
struct KbWalk {
  const uint16_t* row;   // null: dense walk
  int i, n;              // list position / length
  int kb, kb1, kstep;
  uint2 c4;              // the four list entries that contain position i
  int c4_base;           // index of c4's first entry
  __device__ __forceinline__ void load() {
    if (i >= n) return;
    const int e = i + 1, base = e & ~3;                   // entry 0 is the count
    if (base != c4_base) { c4 = __ldg(reinterpret_cast<const uint2*>(row + base)); c4_base = base; }
    const uint32_t w = (e & 2) ? c4.y : c4.x;
    kb = (e & 1) ? (int)(w >> 16) : (int)(w & 0xFFFFu);
  }
  __device__ __forceinline__ void begin(const GemmArgs& a, const WorkItem& it) {
    kb = it.kb0; kb1 = it.kb1; kstep = it.kstep;
    row = nullptr;
    if (it.g >= 0 && a.klist != nullptr) {
      row = a.klist + ((long long)(it.g * a.tiles_m + it.mt) * a.splitk + it.ks) * a.klist_pitch;
      c4 = __ldg(reinterpret_cast<const uint2*>(row));
      c4_base = 0;
      n = (int)(c4.x & 0xFFFFu);
      i = 0;
      load();
    }
  }
  __device__ __forceinline__ bool done() const { return row ? i >= n : kb >= kb1; }
  __device__ __forceinline__ void skip(int k) {          // k blocks further
    if (row) { i += k; load(); }
    else kb += k * kstep;
  }
  __device__ __forceinline__ void step() { skip(1); }
  __device__ __forceinline__ int left() const { return row ? n - i : (kb1 - kb + kstep - 1) / kstep; }   // blocks still to go (>= 0)
};

__device__ __forceinline__ int work_item_at(const GemmArgs& a, int w) { return a.order ? __ldg(a.order + w) : w; }

// ---------------------------------------------------------------------------- epilogue math
// ACT / PRECISE are compile-time: a run-time branch inside the per-element loops would stop ptxas from
// interleaving the (independent) elements, and the epilogue warps have nobody else to hide latency behind.
template <int ACT, bool PRECISE>
__device__ __forceinline__ float act_fwd(float x) {
  if constexpr (ACT == ACT_TANH) {
    if constexpr (PRECISE) return tanhf(x);
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
  } else if constexpr (ACT == ACT_ELU) {
    if constexpr (PRECISE) return x > 0.f ? x : expm1f(x);
    return x > 0.f ? x : __expf(x) - 1.f;
  } else {
    return x;
  }
}
// derivative written in terms of the OUTPUT y: tanh' = 1 - y^2 ; elu' = 1 (y > 0) else y + 1
template <int ACT>
__device__ __forceinline__ float act_bwd_from_y(float y) {
  if constexpr (ACT == ACT_TANH) return 1.f - y * y;
  else if constexpr (ACT == ACT_ELU) return y > 0.f ? 1.f : y + 1.f;
  else return 1.f;
}
// softplus(b) and sigmoid(b) from t = exp(-|b|)
template <bool PRECISE>
__device__ __forceinline__ void softplus_sigmoid(float b, float& sp, float& sg) {
  if constexpr (PRECISE) {
    float t = expf(-fabsf(b));
    sp = fmaxf(b, 0.f) + log1pf(t);
    float r = 1.f / (1.f + t);
    sg = b >= 0.f ? r : t * r;
  } else {
    float t = __expf(-fabsf(b));
    // log1p(t), t in (0,1]: t*P(t) keeps relative accuracy for tiny t (confident pixels dominate a trained model)
    // degree-5 fit of log1p(t)/t on (0,1], max relative error 8.5e-6
    float p = fmaf(t, -0.023868885f, 0.10122161f);
    p = fmaf(t, p, -0.21004808f);
    p = fmaf(t, p, 0.32520646f);
    p = fmaf(t, p, -0.49936134f);
    p = fmaf(t, p, 0.9999916f);
    sp = fmaf(t, p, fmaxf(b, 0.f));
    float r = __fdividef(1.f, 1.f + t);
    sg = b >= 0.f ? r : t * r;
  }
}

// EPI_BCE for one 32-column chunk of one row.
// logits a -> loss += softplus(a) - x*a ; dA = sigmoid(a) - x      (x in {0,1}; 8c/vae.py:260-263)
// With s = a for x = 0 and s = -a for x = 1:  loss = softplus(s),  dA = sigmoid(s) carrying the sign of (1 - 2x).
// `xw` holds this row's 32 target bits (bit j <-> column col0 + j).  FULL: the whole chunk is inside the matrix, so
// no per-element column test is compiled in (it cost ~8 of 32 instructions per element, ncu r01f).
template <bool PRECISE, bool FULL>
__device__ __forceinline__ float bce_values(const EpiArgs& e, const uint32_t (&r)[32], int col0, int N, uint32_t xw,
                                            uint32_t (&hi)[16], uint32_t (&lo)[16]) {
  const int valid = FULL ? 32 : N - col0;
  if constexpr (PRECISE) {
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < 32; j += 2) {
      float d[2];
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const uint32_t sign = (xw << (31 - (j + q))) & 0x80000000u;
        const float sv = __uint_as_float(r[j + q] ^ sign);
        float sp, sg;
        softplus_sigmoid<true>(sv, sp, sg);
        const float dv = __uint_as_float(__float_as_uint(sg) ^ sign);
        const bool in = FULL || j + q < valid;
        d[q] = in ? dv : 0.f;
        acc += in ? sp : 0.f;
      }
      const uint32_t h2 = ptx::pack_bf16(d[0], d[1]);
      hi[j >> 1] = h2;
      lo[j >> 1] = ptx::pack_bf16(d[0] - ptx::bf16_lo(h2), d[1] - ptx::bf16_hi(h2));
    }
    return acc;
  } else {
    // Fast path, two columns per instruction (FFMA2/FADD2/FMUL2).  Per element:
    //   t = 2^(-|a| log2 e) ; rr = 1/(1+t) ; u = t rr = sigmoid(-|s|) ;  sigmoid(s) = 0.5 + sgn(s) (0.5 - u)
    //   softplus(s) = max(s, 0) + t P(t),  P = degree-3 minimax fit of log1p(t)/t on (0,1], max relative error 4.8e-4
    //   (the loss is a reported scalar; the gradient uses sigmoid only)
    const uint64_t k_nl2e = ptx::pk2(-1.4426950408889634f, -1.4426950408889634f), k_one = ptx::pk2(1.f, 1.f);
    const uint64_t k_mhalf = ptx::pk2(-0.5f, -0.5f), k_half = ptx::pk2(0.5f, 0.5f);
    const uint64_t c3 = ptx::pk2(-0.0762659232f, -0.0762659232f), c2 = ptx::pk2(0.256918443f, 0.256918443f);
    const uint64_t c1 = ptx::pk2(-0.487505179f, -0.487505179f), c0 = ptx::pk2(0.999668402f, 0.999668402f);
    uint64_t acc2 = ptx::pk2(0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 32; j += 2) {
      const uint32_t sg0 = (xw << (31 - j)) & 0x80000000u, sg1 = (xw << (30 - j)) & 0x80000000u;
      const uint32_t s0 = r[j] ^ sg0, s1 = r[j + 1] ^ sg1;                     // s = a or -a
      float d0, d1;
      {
        const uint64_t aa = ptx::pk2(__uint_as_float(r[j] & 0x7FFFFFFFu), __uint_as_float(r[j + 1] & 0x7FFFFFFFu));
        float e0, e1;
        ptx::unpk2(ptx::mul2(aa, k_nl2e), e0, e1);
        float t0, t1;
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t0) : "f"(e0));
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t1) : "f"(e1));
        const uint64_t t = ptx::pk2(t0, t1);
        float w0, w1;
        ptx::unpk2(ptx::add2(t, k_one), w0, w1);
        float r0, r1;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"(w0));
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r1) : "f"(w1));
        float h0, h1;                                                            // u - 0.5  (<= 0)
        ptx::unpk2(ptx::fma2(t, ptx::pk2(r0, r1), k_mhalf), h0, h1);
        // (0.5 - u) carrying the sign of s:  flip the sign of (u - 0.5) where s >= 0
        const float g0 = __uint_as_float(__float_as_uint(h0) ^ (~s0 & 0x80000000u));
        const float g1 = __uint_as_float(__float_as_uint(h1) ^ (~s1 & 0x80000000u));
        float q0, q1;                                                            // sigmoid(s)
        ptx::unpk2(ptx::add2(ptx::pk2(g0, g1), k_half), q0, q1);
        d0 = __uint_as_float(__float_as_uint(q0) ^ sg0);                         // sigmoid(a) - x
        d1 = __uint_as_float(__float_as_uint(q1) ^ sg1);
        uint64_t pp = ptx::fma2(t, c3, c2);
        pp = ptx::fma2(t, pp, c1);
        pp = ptx::fma2(t, pp, c0);
        const uint64_t relu = ptx::pk2(fmaxf(__uint_as_float(s0), 0.f), fmaxf(__uint_as_float(s1), 0.f));
        const uint64_t sp = ptx::fma2(t, pp, relu);
        if constexpr (FULL) {
          acc2 = ptx::add2(acc2, sp);
        } else {
          float p0, p1;
          ptx::unpk2(sp, p0, p1);
          acc2 = ptx::add2(acc2, ptx::pk2(j < valid ? p0 : 0.f, j + 1 < valid ? p1 : 0.f));
        }
      }
      if constexpr (!FULL) {
        d0 = j < valid ? d0 : 0.f;
        d1 = j + 1 < valid ? d1 : 0.f;
      }
      hi[j >> 1] = ptx::pack_bf16(d0, d1);
    }
    float a0, a1;
    ptx::unpk2(acc2, a0, a1);
    return a0 + a1;
  }
}

// ------------------------------------------------------------------ how a chunk leaves the SM
// In the GEMM kernel lane = accumulator row, so a direct warp-wide 16-byte store touches 32 different rows: 32
// half-written sectors per instruction, two request cycles each on the L1 -> crossbar port (58% busy in
// profiles/r01f) and a 2-3x slower epilogue on short-K shapes (profiles/r01g).  Instead every warp stages its
// 32-row chunk in shared memory in the TMA swizzle pattern (conflict-free for lane = row) and one lane issues a tile
// store -- or a tile reduce-add for split-K partial sums -- through a tensor map, which also clips at the matrix edge.
struct TileOut {
  const CUtensorMap* hi;   // fp32 map (box 16 x 32) for the fp32 epilogues, bf16 "hi" plane (box 32 x 32) otherwise
  const CUtensorMap* lo;   // bf16 "lo" plane (used when EpiArgs.out_lo is set)
  uint8_t* slot;           // this warp's 2 KiB staging area: 32 rows x 64 B, 64B swizzle
  int lane, row0;
  int reduce;              // fp32: add into the destination
};

// lane = row: its 64 bytes go to the slot (16-byte piece j of row l sits at piece j ^ ((l >> 1) & 3), the 64B swizzle,
// conflict-free), then one lane hands the 32 x 64 B tile to the TMA.
template <int OFF>
__device__ __forceinline__ void tile_emit_64(const TileOut& t, const CUtensorMap* map, const uint32_t* v, int col0,
                                             bool reduce) {
  // (the issuing lane is chosen by elect.sync -- the same lane every time for a full warp -- so that ptxas emits the
  //  UTMASTG / UTMAREDG directly, see ptx::elect_one; bulk groups belong to that thread)
  const bool issuer = ptx::elect_one();
  if (issuer) ptx::bulk_wait_read();        // earlier tile stores of this warp have finished reading the slot
  __syncwarp();
  uint8_t* rowp = t.slot + t.lane * 64;
  const int x = (t.lane >> 1) & 3;
#pragma unroll
  for (int j = 0; j < 4; ++j)
    ptx::st_shared_v4(rowp + ((j ^ x) << 4), v[OFF + 4 * j], v[OFF + 4 * j + 1], v[OFF + 4 * j + 2], v[OFF + 4 * j + 3]);
  ptx::fence_proxy_async_smem();
  __syncwarp();
  if (issuer) {
    if (reduce) ptx::tma_reduce_add_2d(map, t.slot, col0, t.row0);
    else ptx::tma_store_2d(map, t.slot, col0, t.row0);
    ptx::bulk_commit();
  }
}
__device__ __forceinline__ void tile_emit_f32(const TileOut& t, const uint32_t (&v)[32], int col0) {
  tile_emit_64<0>(t, t.hi, v, col0, t.reduce != 0);
  tile_emit_64<16>(t, t.hi, v, col0 + 16, t.reduce != 0);
}
__device__ __forceinline__ void tile_emit_bf16(const TileOut& t, const uint32_t (&hi)[16], const uint32_t (&lo)[16],
                                               bool with_lo, int col0) {
  tile_emit_64<0>(t, t.hi, hi, col0, false);
  if (with_lo) tile_emit_64<0>(t, t.lo, lo, col0, false);
}
// direct stores: the split-K finishing kernel (one thread per row chunk, consecutive threads on consecutive chunks)
__device__ __forceinline__ void direct_emit_bf16(const EpiArgs& e, const uint32_t (&hi)[16], const uint32_t (&lo)[16],
                                                 bool with_lo, int row, int col0) {
  uint4* o = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(e.out) + (long long)row * e.ld_out + col0);
#pragma unroll
  for (int q = 0; q < 4; ++q) o[q] = make_uint4(hi[4 * q], hi[4 * q + 1], hi[4 * q + 2], hi[4 * q + 3]);
  if (with_lo) {
    uint4* ol = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(e.out_lo) + (long long)row * e.ld_out + col0);
#pragma unroll
    for (int q = 0; q < 4; ++q) ol[q] = make_uint4(lo[4 * q], lo[4 * q + 1], lo[4 * q + 2], lo[4 * q + 3]);
  }
}

// One 32-column chunk of one accumulator row.  TILE: called by all 32 lanes of an epilogue warp of the GEMM kernel
// (rows >= M and columns >= N are clipped by the tensor map; `row_ok` guards reads and the loss).  !TILE: one thread,
// direct stores, row < M guaranteed by the caller.
template <int EPI, int ACT, bool PRECISE, bool TILE>
__device__ __forceinline__ void epilogue_impl(const EpiArgs& e, const TileOut& t, const uint32_t (&r)[32], int row,
                                              bool row_ok, int col0, int N, float& loss_acc, uint32_t xw) {
  if constexpr (EPI == EPI_STORE_F32) {
    const float alpha = e.alpha;
    if constexpr (TILE) {
      uint32_t v[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(__uint_as_float(r[j]) * alpha);
      tile_emit_f32(t, v, col0);
    } else {
      float* out = reinterpret_cast<float*>(e.out) + (long long)row * e.ld_out + col0;
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < N) {
          const float v = __uint_as_float(r[j]) * alpha;
          if (e.atomic) atomicAdd(out + j, v);
          else out[j] = v;
        }
    }
  } else if constexpr (EPI == EPI_SCATTER_F32) {
    // handled by scatter_chunk() in the kernel (its own staging area and peer pointers)
  } else if constexpr (EPI == EPI_SIGMOID) {
    uint32_t v[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      float sp, sg;
      softplus_sigmoid<true>(__uint_as_float(r[j]), sp, sg);
      v[j] = __float_as_uint(sg);
    }
    if constexpr (TILE) {
      tile_emit_f32(t, v, col0);
    } else {
      float* out = reinterpret_cast<float*>(e.out) + (long long)row * e.ld_out + col0;
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < N) out[j] = __uint_as_float(v[j]);
    }
  } else if constexpr (EPI == EPI_ACT) {
    uint32_t hi[16], lo[16];
#pragma unroll
    for (int j = 0; j < 32; j += 2) {
      float v[2];
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        int c = col0 + j + q;
        float x = __uint_as_float(r[j + q]) * e.alpha;
        v[q] = c < e.n_valid ? act_fwd<ACT, PRECISE>(x) : (c == e.n_valid ? 1.f : 0.f);
      }
      uint32_t h = ptx::pack_bf16(v[0], v[1]);
      hi[j >> 1] = h;
      lo[j >> 1] = ptx::pack_bf16(v[0] - ptx::bf16_lo(h), v[1] - ptx::bf16_hi(h));
    }
    if constexpr (TILE) tile_emit_bf16(t, hi, lo, e.out_lo != nullptr, col0);
    else direct_emit_bf16(e, hi, lo, e.out_lo != nullptr, row, col0);
  } else if constexpr (EPI == EPI_DACT) {
    uint32_t y[16], yl[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) { y[q] = 0u; yl[q] = 0u; }
    if (row_ok) {
      int ar = row < e.aux_split ? row : row - e.aux_shift;
      const uint4* ap = reinterpret_cast<const uint4*>(e.aux + (long long)ar * e.ld_aux + col0);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 tt = ap[q];
        y[4 * q] = tt.x; y[4 * q + 1] = tt.y; y[4 * q + 2] = tt.z; y[4 * q + 3] = tt.w;
      }
      if (e.aux_lo) {
        const uint4* alp = reinterpret_cast<const uint4*>(e.aux_lo + (long long)ar * e.ld_aux + col0);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint4 tt = alp[q];
          yl[4 * q] = tt.x; yl[4 * q + 1] = tt.y; yl[4 * q + 2] = tt.z; yl[4 * q + 3] = tt.w;
        }
      }
    }
    uint32_t hi[16], lo[16];
#pragma unroll
    for (int j = 0; j < 32; j += 2) {
      float y0 = ptx::bf16_lo(y[j >> 1]) + ptx::bf16_lo(yl[j >> 1]);
      float y1 = ptx::bf16_hi(y[j >> 1]) + ptx::bf16_hi(yl[j >> 1]);
      float v0 = (col0 + j < e.n_valid) ? __uint_as_float(r[j]) * e.alpha * act_bwd_from_y<ACT>(y0) : 0.f;
      float v1 = (col0 + j + 1 < e.n_valid) ? __uint_as_float(r[j + 1]) * e.alpha * act_bwd_from_y<ACT>(y1) : 0.f;
      uint32_t h = ptx::pack_bf16(v0, v1);
      hi[j >> 1] = h;
      lo[j >> 1] = ptx::pack_bf16(v0 - ptx::bf16_lo(h), v1 - ptx::bf16_hi(h));
    }
    if constexpr (TILE) tile_emit_bf16(t, hi, lo, e.out_lo != nullptr, col0);
    else direct_emit_bf16(e, hi, lo, e.out_lo != nullptr, row, col0);
  } else if constexpr (EPI == EPI_BCE) {
    uint32_t hi[16], lo[16];
    float acc;
    if (col0 + 32 <= N) acc = bce_values<PRECISE, true>(e, r, col0, N, xw, hi, lo);
    else acc = bce_values<PRECISE, false>(e, r, col0, N, xw, hi, lo);
    if (row_ok) loss_acc += acc;
    if (e.dbg & 1) return;     // diagnostic: no stores
    if constexpr (TILE) tile_emit_bf16(t, hi, lo, PRECISE && e.out_lo != nullptr, col0);
    else direct_emit_bf16(e, hi, lo, PRECISE && e.out_lo != nullptr, row, col0);
  }
}

// EPI_SCATTER_F32: this thread's 32 consecutive fp32 of row `row` go to the owning rank's staging slot (see EpiArgs).
// Four passes of 8 lanes: registers -> 128-byte shared-memory slot -> one cp.async.bulk of the row segment.
__device__ __forceinline__ void scatter_chunk(const EpiArgs& e, const uint32_t (&r)[32], int row, int col0, int M, int N,
                                              uint8_t* stage /* this warp's kScatterStageBytes */, int lane) {
  const int valid = min(32, N - col0);                      // floats of this chunk inside the matrix
  const bool active = row < M && valid > 0;
  const long long idx = e.flat_base + (long long)row * e.ld_out + col0;
  const int q = active ? (int)(idx >> e.gran_shift) : 0;
  const int o = q % e.world;
  float* dst = e.peer[o] + ((long long)e.src_rank * e.slice_len + ((long long)(q / e.world) << e.gran_shift) +
                            (idx & ((1LL << e.gran_shift) - 1)));
  const float alpha = e.alpha;
  if ((valid & 3) != 0) {                                    // ragged tail that is not 16-byte granular: plain stores
    if (active)
      for (int j = 0; j < valid; ++j) dst[j] = __uint_as_float(r[j]) * alpha;
    return;
  }
#pragma unroll
  for (int pass = 0; pass < 4; ++pass) {
    if ((lane >> 3) == pass) {
      float4* slot = reinterpret_cast<float4*>(stage + (lane & 7) * 128);
#pragma unroll
      for (int j = 0; j < 32; j += 4)
        slot[j >> 2] = make_float4(__uint_as_float(r[j]) * alpha, __uint_as_float(r[j + 1]) * alpha,
                                   __uint_as_float(r[j + 2]) * alpha, __uint_as_float(r[j + 3]) * alpha);
      ptx::fence_proxy_async_smem();
      if (active) ptx::bulk_store(dst, slot, (uint32_t)valid * 4u);
      ptx::bulk_commit();
      ptx::bulk_wait_read();                                  // the slot may be overwritten by the next pass / chunk
    }
    __syncwarp();
  }
}

// EPI_ADAM: one 32 x 32 chunk of the accumulator (lane = row in `r`), 16 columns at a time.  The half chunk is turned
// round in the warp's staging slot (written lane = row in the tile stores' 64B swizzle, read back as 8 rows x 4 pieces of
// 16 bytes) so that every global access of the update is whole 32-byte sectors: four lanes cover the 64 contiguous bytes
// a row owns in each of m, v, p (and the 32 bytes of the bf16 shadow).  Straight from the accumulator layout every lane
// would walk its own 128-byte line 16 bytes at a time -- 32 lines per instruction, partial-sector stores: the launch
// took 1.08 ms in the step that way, 0.82 ms this way, 0.52 ms without the update (why: AdamFuse above).
__device__ __forceinline__ float adam_fused_one(float g, float& m, float& v, float lr_t, float b1, float b2, float eps) {
  m = b1 * m + (1.f - b1) * g;                      // (adam_one in kernels.cu, term by term)
  v = b2 * v + (1.f - b2) * g * g;
  return lr_t * m / (sqrtf(v) + eps);
}
// The update of one staged half chunk (32 rows x 16 columns in `slot`).  Not inlined, arguments by value: inlined into
// the eight places of the unrolled epilogue loop the IEEE division and square root grow the kernel from 4.7 k to 13.3 k
// instructions (6.2 k this way) for no gain -- the epilogue is nowhere near its issue limit.
static __device__ __noinline__ void adam_rows(float* p, float* mg, float* vg, __nv_bfloat16* w_hi, __nv_bfloat16* w_lo,
                                       long long ld, float lr, float b1, float b2, float eps, float alpha, int last_group,
                                       const uint8_t* slot, int lane, int row0, int col0, int M, int N, int dbg) {
  const int c4 = lane & 3, rsub = lane >> 2;
  const int col = col0 + 4 * c4;
  if (col >= N) return;                             // N % 4 == 0 (gemm_fuse_adam)
#pragma unroll 1
  for (int it = 0; it < 4; it += 2) {
    float4 gv[2], m4[2], v4[2], p4[2];
    long long idx[2];
    bool ok[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int lr_ = (it + h) * 8 + rsub;          // row inside the chunk
      const int row = row0 + lr_;
      ok[h] = row < M;
      idx[h] = (long long)row * ld + col;
      gv[h] = *reinterpret_cast<const float4*>(slot + lr_ * 64 + ((c4 ^ ((lr_ >> 1) & 3)) << 4));
      m4[h] = v4[h] = p4[h] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ok[h] && !(dbg & 1)) {
        // streaming (evict-first) accesses: 0.9 GB of optimiser state passes through the L2 during this launch and
        // should not push out what the main loop lives on (the B operand, 16 MB re-read by every M tile)
        m4[h] = __ldcs(reinterpret_cast<const float4*>(mg + idx[h]));
        v4[h] = __ldcs(reinterpret_cast<const float4*>(vg + idx[h]));
        p4[h] = __ldcs(reinterpret_cast<const float4*>(p + idx[h]));
      }
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if (!ok[h]) continue;
      if (!(dbg & 2)) {
        p4[h].x -= adam_fused_one(gv[h].x * alpha, m4[h].x, v4[h].x, lr, b1, b2, eps);
        p4[h].y -= adam_fused_one(gv[h].y * alpha, m4[h].y, v4[h].y, lr, b1, b2, eps);
        p4[h].z -= adam_fused_one(gv[h].z * alpha, m4[h].z, v4[h].z, lr, b1, b2, eps);
        p4[h].w -= adam_fused_one(gv[h].w * alpha, m4[h].w, v4[h].w, lr, b1, b2, eps);
      }
      if (dbg & 1) { if (p4[h].x == 123.f) p[idx[h]] = 0.f; continue; }
      __stcs(reinterpret_cast<float4*>(mg + idx[h]), m4[h]);
      __stcs(reinterpret_cast<float4*>(vg + idx[h]), v4[h]);
      __stcs(reinterpret_cast<float4*>(p + idx[h]), p4[h]);
      if (last_group) {
        const uint32_t h0 = ptx::pack_bf16(p4[h].x, p4[h].y), h1 = ptx::pack_bf16(p4[h].z, p4[h].w);
        __stcs(reinterpret_cast<uint2*>(w_hi + idx[h]), make_uint2(h0, h1));
        if (w_lo) {
          const uint32_t l0 = ptx::pack_bf16(p4[h].x - ptx::bf16_lo(h0), p4[h].y - ptx::bf16_hi(h0));
          const uint32_t l1 = ptx::pack_bf16(p4[h].z - ptx::bf16_lo(h1), p4[h].w - ptx::bf16_hi(h1));
          __stcs(reinterpret_cast<uint2*>(w_lo + idx[h]), make_uint2(l0, l1));
        }
      }
    }
  }
}
template <int OFF>
__device__ __forceinline__ void adam_half(const AdamFuse& af, int g, float alpha, const uint32_t (&r)[32], uint8_t* slot,
                                          int lane, int row0, int col0, int M, int N) {
  if (af.dbg & 4) return;
  if (ptx::elect_one()) ptx::bulk_wait_read();      // a tile store of an earlier (stored) item may still read the slot (same lane: tile_emit_64)
  __syncwarp();
  {
    uint8_t* rowp = slot + lane * 64;
    const int x = (lane >> 1) & 3;
#pragma unroll
    for (int j = 0; j < 4; ++j)
      ptx::st_shared_v4(rowp + ((j ^ x) << 4), r[OFF + 4 * j], r[OFF + 4 * j + 1], r[OFF + 4 * j + 2], r[OFF + 4 * j + 3]);
  }
  __syncwarp();
  adam_rows(af.p, af.m[g], af.v[g], af.w_hi, af.w_lo, af.ld, af.lr_t[g], af.beta1, af.beta2, af.eps, alpha, g == 1, slot,
            lane, row0, col0, M, N, af.dbg);
  __syncwarp();                                      // the slot is rewritten by the next half chunk
}
__device__ __forceinline__ void adam_chunk(const AdamFuse& af, int g, float alpha, const uint32_t (&r)[32], uint8_t* slot,
                                           int lane, int row0, int col0, int M, int N) {
  adam_half<0>(af, g, alpha, r, slot, lane, row0, col0, M, N);
  adam_half<16>(af, g, alpha, r, slot, lane, row0, col0 + 16, M, N);
}

// Run-time (act, precise) -> compile-time variant, decided once per 32-column chunk (warp-uniform).
template <int EPI, bool TILE>
__device__ __forceinline__ void epilogue_chunk(const EpiArgs& e, const TileOut& t, const uint32_t (&r)[32], int row,
                                               int col0, int M, int N, float& loss_acc, uint32_t xw = 0u) {
  const bool row_ok = row < M;
  if constexpr (!TILE) {
    if (!row_ok) return;
  }
  if constexpr (EPI == EPI_ACT || EPI == EPI_DACT) {
    if (e.act == ACT_TANH) {
      if (e.precise) epilogue_impl<EPI, ACT_TANH, true, TILE>(e, t, r, row, row_ok, col0, N, loss_acc, xw);
      else epilogue_impl<EPI, ACT_TANH, false, TILE>(e, t, r, row, row_ok, col0, N, loss_acc, xw);
    } else {
      if (e.precise) epilogue_impl<EPI, ACT_ELU, true, TILE>(e, t, r, row, row_ok, col0, N, loss_acc, xw);
      else epilogue_impl<EPI, ACT_ELU, false, TILE>(e, t, r, row, row_ok, col0, N, loss_acc, xw);
    }
  } else if constexpr (EPI == EPI_BCE) {
    if (e.precise) epilogue_impl<EPI, ACT_NONE, true, TILE>(e, t, r, row, row_ok, col0, N, loss_acc, xw);
    else epilogue_impl<EPI, ACT_NONE, false, TILE>(e, t, r, row, row_ok, col0, N, loss_acc, xw);
  } else {
    epilogue_impl<EPI, ACT_NONE, true, TILE>(e, t, r, row, row_ok, col0, N, loss_acc, xw);
  }
}

static_assert(sizeof(GemmArgs) <= 4096, "kernel parameter space (kept under the classic 4 KiB)");

// 16 packed pixels -> 16 bf16 (0.0 / 1.0) in 8 words.  u carries bit i at position i and again at position i + 15, so
// (u >> 2j) & 0x00010001 holds elements 2j and 2j + 1 in bit 0 of the two halves; times 0x3F80 makes each a bf16 1.0.
__device__ __forceinline__ void expand16(uint32_t v16, uint32_t (&w)[8]) {
  const uint32_t u = v16 | (v16 << 15);
#pragma unroll
  for (int j = 0; j < 8; ++j) w[j] = ((u >> (2 * j)) & 0x00010001u) * 0x3F80u;
}

// ---------------------------------------------------------------------------- the kernel
// CG2: the CTA-pair variant (tcgen05 cta_group::2, ptx.cuh): launched as clusters of two, args.cluster_m == 2; a work
// item is a 256 x BN tile, CTA `crank` of the pair owns its rows [128 crank, 128 crank + 128) -- A tile, accumulator,
// epilogue -- and loads columns [BN/2 crank, BN/2 crank + BN/2) of the B tile; the leader (crank 0) issues the MMAs.
// Six pipeline stages of 16 + 16 KiB instead of four of 16 + 32.
template <bool A_MN, bool B_MN, int EPI, int BN, bool A_BITS = false, bool CG2 = false>
// 112 registers: three of this kernel's warps (3 x 32 x 112) and three warps of the batch generator (3 x 32 x 56) fit
// the 16K registers of one SM sub-partition, so the generator's CTA stays co-resident with every GEMM variant; it is
// the 576-thread BCE variant has five warps on a sub-partition, hence 96.
__global__ void __maxnreg__(epi_warps(EPI, A_BITS) == 16 ? 96 : 112) gemm_tcgen05_kernel(const __grid_constant__ GemmArgs args) {
#define MAGIC_TR(slot)                                                                  \
  do {                                                                                  \
    if (args.trace) args.trace[(size_t)blockIdx.x * 64 + (slot)] = clock64();           \
  } while (0)
  if (threadIdx.x == 0) {
    MAGIC_TR(0);
    if (args.trace) {
      unsigned long long gt;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
      args.trace[(size_t)blockIdx.x * 64 + 11] = (long long)gt;
    }
  }
  static_assert(!CG2 || BN == 256, "CTA pairs run 256 x 256 tiles");
  using Cfg = TileCfg<CG2 ? BN / 2 : BN>;  // per-CTA share of a stage: a pair member holds half of the B tile's columns
  constexpr int kNS = Cfg::kStagesN;       // pipeline stages
  constexpr int kAcc = 512 / BN;           // accumulator stages
  extern __shared__ uint8_t smem_raw[];
  // 128-byte swizzle atoms need 1024-byte aligned stage bases
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kNS * kStageBytesA;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);   // = kNS * Cfg::kBytes
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kNS;
  uint64_t* tfull_bar = bars + 2 * kNS;
  uint64_t* tempty_bar = bars + 2 * kNS + kAcc;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * kNS + 2 * kAcc);   // (2*6 + 2*4) * 8 + 4 <= 192 bytes
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total_work = args.tiles_m * args.tiles_n * args.ngroups * args.splitk +
                         (args.nextra > 0 ? args.x_item0[args.nextra] : 0);
  const int CM = args.cluster_m;
  const int crank = CM > 1 ? (int)ptx::cluster_ctarank() : 0;
  const int work0 = (int)blockIdx.x / CM, work_stride = (int)gridDim.x / CM;   // this cluster's first item / number of clusters
  const uint16_t cmask = (uint16_t)((1u << CM) - 1u);

  if (warp == 0 && lane == 0) {
    for (int g = 0; g < args.ngroups; ++g)
      for (int s = 0; s < args.nseg; ++s) {
        ptx::tma_prefetch_desc(&args.ta[g][s]);
        ptx::tma_prefetch_desc(&args.tb[g][s]);
      }
    for (int i = 0; i < kNS; ++i) {
      // arrivals on a stage's "full" barrier: the producer's expect_tx (+ the stage's expander warp); in a CTA pair the
      // leader's barrier also takes the other CTA's producer (and expander) -- only the leader's is ever waited on
      ptx::mbar_init(&full_bar[i], (A_BITS ? 2 : 1) * (CG2 ? 2 : 1));
      ptx::mbar_init(&empty_bar[i], CG2 ? 1 : CM);   // one arrival per committing CTA (multicast commit); a pair: the leader's
    }
    for (int i = 0; i < kAcc; ++i) {
      ptx::mbar_init(&tfull_bar[i], 1);
      ptx::mbar_init(&tempty_bar[i], CG2 ? 2 * epi_warps(EPI, A_BITS) : epi_warps(EPI, A_BITS) * 32);   // pair: one arrival per epilogue warp of both CTAs
    }
    ptx::fence_barrier_init();
  } else if (warp == 1) {
    if constexpr (CG2) { ptx::tmem_alloc_pair(tmem_ptr, kTmemCols); ptx::tmem_relinquish_pair(); }
    else { ptx::tmem_alloc(tmem_ptr, kTmemCols); ptx::tmem_relinquish(); }
  }
  ptx::tc_fence_before();
  if (CM > 1) ptx::cluster_sync();       // barriers of every CTA are initialised before anybody multicasts into them
  else __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  if (threadIdx.x == 0) MAGIC_TR(1);
  ptx::grid_dep_wait();     // everything above overlapped the previous kernel's tail; its outputs are visible from here
  if (threadIdx.x == 0) MAGIC_TR(2);

  // pair: the leader's barriers as seen from either CTA (its own address space for the leader itself)
  [[maybe_unused]] const bool leader = !CG2 || crank == 0;
  auto leader_addr = [&](const void* local) { return CG2 ? ptx::mapa_u32(ptx::smem_u32(local), 0u) : ptx::smem_u32(local); };
  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (one thread; in a pair: one per CTA)
    // This thread's per-k-block instruction chain paces the whole pipeline (one warp, dependent scalar code: ~6 cycles per
    // instruction; ncu r02l: with ~150 instructions per k-block -- a division by kblocks, the list walker, flags -- the
    // tensor pipe was 58 % busy), so everything that does not change from one k-block to the next is hoisted to the work
    // item: tensor maps, byte counts, the segment / k offset (stepped, not divided).  The thread is chosen with elect.sync
    // (ptx::elect_one): inside, ptxas issues UTMALDG / UTCHMMA / UTCBAR directly; under `lane == 0` it wrapped every one of
    // them in an ELECT ... BRA.U.ANY loop over the active threads (85 -> 45 instructions per k-block in the MMA issuer).
    if (ptx::elect_one()) {
      uint32_t stage = 0, phase = 0;
      bool first_item = true;
      for (int w = work0; w < total_work; w += work_stride) {
        const WorkItem it = decode_work(args, work_item_at(args, w), crank);
        const bool main_p = it.g >= 0;
        const CUtensorMap* ta0 = main_p ? &args.ta[it.g][0] : &args.xdev[-1 - it.g].ta[0];     // [segment]
        const CUtensorMap* tb0 = main_p ? &args.tb[it.g][0] : &args.xdev[-1 - it.g].tb[0];
        const bool a_by_tma = !A_BITS || !main_p;   // A_BITS: the main problem's A stage is written by the expander warps
        const uint32_t tx = a_by_tma ? Cfg::kBytes : Cfg::kBytesB;
        const bool no_loads = CG2 && (args.epi[0].dbg & 16) != 0;     // timing experiment (pairs): barriers only
        // one k-block: wait for the stage, arm its barrier, issue the loads
        auto issue = [&](const CUtensorMap* ta, const CUtensorMap* tb, int k0) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem_a + stage * kStageBytesA;
          uint8_t* sb = smem_b + stage * Cfg::kBytesB;
          if constexpr (CG2) {
            // both CTAs' bytes are counted on the leader's barrier: it expects the pair's total, the other CTA only arrives
            const uint32_t fb = leader_addr(&full_bar[stage]);
            if (leader) ptx::mbar_expect_tx(&full_bar[stage], no_loads ? 0u : 2u * tx);
            else ptx::mbar_arrive_cluster(fb);
            if (!no_loads) {
              if (a_by_tma) {
                if constexpr (!A_MN) ptx::tma_load_2d_pair(ta, fb, sa, k0, it.m0);                 // box {64 k, 128 m}
                else ptx::tma_load_3d_pair(ta, fb, sa, 0, k0, it.m0 >> 6);                         // box {64 m, 64 k, 2 slabs}
              }
              const int nh = it.n0 + crank * (BN / 2);                                             // this CTA's half of the B tile
              if constexpr (!B_MN) ptx::tma_load_2d_pair(tb, fb, sb, k0, nh);                      // box {64 k, BN/2 n}
              else ptx::tma_load_3d_pair(tb, fb, sb, 0, k0, nh >> 6);                              // box {64 n, 64 k, BN/128 slabs}
            }
          } else {
            ptx::mbar_expect_tx(&full_bar[stage], tx);
            if (a_by_tma) {
              if constexpr (!A_MN) ptx::tma_load_2d(ta, &full_bar[stage], sa, k0, it.m0);          // box {64 k, 128 m}
              else ptx::tma_load_3d(ta, &full_bar[stage], sa, 0, k0, it.m0 >> 6);                  // box {64 m, 64 k, 2 slabs}
            }
            if constexpr (!B_MN) {
              ptx::tma_load_2d(tb, &full_bar[stage], sb, k0, it.n0);                               // box {64 k, BN n}
            } else if (CM == 1) {
              ptx::tma_load_3d(tb, &full_bar[stage], sb, 0, k0, it.n0 >> 6);                       // box {64 n, 64 k, BN/64 slabs}
            } else {
              // the cluster shares this B tile: each CTA fetches 4/CM slabs and multicasts them to all
              const int per = (BN / 64) / CM;
              ptx::tma_load_3d_multicast(tb, &full_bar[stage], sb + crank * per * kAtomBytes, 0, k0,
                                         (it.n0 >> 6) + crank * per, cmask);
            }
          }
          if (++stage == kNS) { stage = 0; phase ^= 1; }
        };
        if (first_item) MAGIC_TR(3);
        first_item = false;
        if (!(main_p && args.klist != nullptr)) {
          // every k-block of [kb0, kb1): (segment, k offset) stepped along
          int seg = it.kb0 / it.kblocks;
          int k0 = (it.kb0 - seg * it.kblocks) * kBlockK;
          const int kend = it.kblocks * kBlockK;
          for (int kb = it.kb0; kb < it.kb1; ++kb) {
            issue(ta0 + seg, tb0 + seg, k0);
            k0 += kBlockK;
            if (k0 >= kend) { k0 = 0; ++seg; }
          }
        } else {
          // the item's k-block list (zero-stage skipping)
          KbWalk walk;
          walk.begin(args, it);
          while (!walk.done()) {
            const int kb = walk.kb;
            walk.step();
            const int seg = kb / it.kblocks;
            issue(ta0 + seg, tb0 + seg, (kb - seg * it.kblocks) * kBlockK);
          }
        }
      }
      // the other CTA of a pair issues no MMAs: its trigger for the next kernel of the stream comes when its loads are out
      if (CG2 && !leader) ptx::grid_dep_launch();
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (one thread; in a pair: the leader's)
    if (leader && ptx::elect_one()) {
      constexpr uint32_t idesc = ptx::make_idesc_bf16(CG2 ? 2 * kBlockM : kBlockM, BN, A_MN, B_MN);
      // K-major SW128: 8-row groups 1024 B apart, K advance = 32 B inside the swizzle row.
      // MN-major SW128: 64-element MN atoms kBlockK*128 B apart (LBO), 8-k-row groups 1024 B apart (SBO),
      //                 K advance = 16 rows * 128 B.
      constexpr uint32_t a_lbo = A_MN ? kAtomBytes : 0, b_lbo = B_MN ? kAtomBytes : 0;
      constexpr uint32_t a_kadv = A_MN ? kUmmaK * 128 : kUmmaK * 2, b_kadv = B_MN ? kUmmaK * 128 : kUmmaK * 2;
      // Descriptors: everything but the start address is constant; the k advance adds to the 14-bit start-address
      // field (address >> 4), which cannot carry out of it (shared memory ends below 256 KiB).
      constexpr uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
      constexpr uint32_t a_lo_c = ((a_lbo >> 4) & 0x3FFFu) << 16, b_lo_c = ((b_lbo >> 4) & 0x3FFFu) << 16;
      // (the mask drops the window bits a shared-memory address carries above its 18-bit offset in a cluster launch)
      const uint32_t a_addr0 = (ptx::smem_u32(smem_a) >> 4) & 0x3FFFu, b_addr0 = (ptx::smem_u32(smem_b) >> 4) & 0x3FFFu;
      const int n_items = work0 < total_work ? (total_work - work0 + work_stride - 1) / work_stride : 0;
      uint32_t stage = 0, phase = 0, iter = 0;
      // Like the producer, this thread's per-k-block instruction chain paces the pipeline: the loop body is the barrier
      // wait, two descriptor words, the four MMAs and the commit.  How many k-blocks a work item has is known up front
      // (its k range, or the length of its list) -- no per-stage flag from the producer.
      for (; (int)iter < n_items; ++iter) {
        const WorkItem it = decode_work(args, work_item_at(args, work0 + (int)iter * work_stride), crank);
        int n_kb;                                             // >= 1 (planner, gemm_klist_kernel)
        if (it.g >= 0 && args.klist != nullptr)
          n_kb = (int)__ldg(args.klist + ((long long)(it.g * args.tiles_m + it.mt) * args.splitk + it.ks) * args.klist_pitch);
        else
          n_kb = (it.kb1 - it.kb0 + it.kstep - 1) / it.kstep;
        const uint32_t acc = iter % kAcc, acc_phase = (iter / kAcc) & 1;
        ptx::mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        ptx::tc_fence_after();
        if (iter < 8) MAGIC_TR(16 + iter * 5 + 0);
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int i = 0; i < n_kb; ++i) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          const uint32_t a_lo = (a_addr0 + stage * (kStageBytesA >> 4)) | a_lo_c;
          const uint32_t b_lo = (b_addr0 + stage * (Cfg::kBytesB >> 4)) | b_lo_c;
          const bool no_mma = CG2 && (args.epi[0].dbg & 32) != 0;   // timing experiment (pairs): barriers and loads only
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k) {
            const uint64_t da = (static_cast<uint64_t>(desc_hi) << 32) | (a_lo + k * (a_kadv >> 4));
            const uint64_t db = (static_cast<uint64_t>(desc_hi) << 32) | (b_lo + k * (b_kadv >> 4));
            if constexpr (CG2) { if (!no_mma) ptx::umma_bf16_pair(d_tmem, da, db, idesc, (i > 0 || k > 0) ? 1u : 0u); }
            else ptx::umma_bf16(d_tmem, da, db, idesc, (i > 0 || k > 0) ? 1u : 0u);
          }
          if constexpr (CG2) ptx::umma_commit_pair(&empty_bar[stage], (uint16_t)3);   // frees the stage in both CTAs
          else if (CM > 1) ptx::umma_commit_multicast(&empty_bar[stage], cmask);
          else ptx::umma_commit(&empty_bar[stage]);   // frees the smem stage when these MMAs retire
          if (++stage == kNS) { stage = 0; phase ^= 1; }
        }
        if constexpr (CG2) ptx::umma_commit_pair(&tfull_bar[acc], (uint16_t)3);   // both CTAs' epilogues
        else ptx::umma_commit(&tfull_bar[acc]);   // accumulator complete -> epilogue
        if (iter == 0) MAGIC_TR(5);
        if (iter < 8) MAGIC_TR(16 + iter * 5 + 1);
      }
      MAGIC_TR(6);
      if (args.trace) args.trace[(size_t)blockIdx.x * 64 + 12] = iter;
      // All of this CTA's MMAs are issued: from here only its last epilogue remains, so the next kernel of the step
      // may become resident on SMs as they free and run its prologue (launch.cuh).  Triggering earlier parks the
      // following kernels' CTAs on the SMs for the whole duration and starves the generator's side stream.
      ptx::grid_dep_launch();
    }
  } else if (warp < 2 + epi_warps(EPI, A_BITS)) {
    // ------------------------------------------------------------ epilogue warps
    // TMEM lane quarter = warp % 4 (hardware rule); the two warps of a quarter split the 32-column chunks
    // even/odd, so every scheduler has two epilogue warps to overlap latencies with.
    const int q = warp & 3;
    constexpr int kGroups = epi_warps(EPI, A_BITS) / 4;  // warps per lane quarter; chunk c belongs to group c % kGroups
    constexpr int kMaxOwn = (BN / 32 + kGroups - 1) / kGroups;
    const int grp = (warp - 2) >> 2;
    uint32_t iter = 0;
    float loss_acc = 0.f;
    // BCE: a row's packed target bits for the chunks this warp owns.  They are fetched ONE TILE AHEAD: loaded right
    // before the accumulator wait, their ~1.3 us of DRAM latency sat on the critical path of every tile (r01i trace).
    uint32_t xw[kMaxOwn], xw_next[kMaxOwn];
    auto load_xw = [&](int w_item, uint32_t (&dst)[kMaxOwn]) {
      const WorkItem nit = decode_work(args, work_item_at(args, w_item), crank);
      const EpiArgs& ne = args.epi[nit.g];
      const int nrow = nit.m0 + q * 32 + lane;
      const int nn = (min(BN, args.N - nit.n0) + 31) >> 5;
#pragma unroll
      for (int k = 0; k < kMaxOwn; ++k) {
        const int c = grp + k * kGroups;
        dst[k] = (nrow < args.M && c < nn) ? __ldg(ne.xbits + (long long)nrow * ne.xb_ld + ((nit.n0 >> 5) + c)) : 0u;
      }
    };
    if constexpr (EPI == EPI_BCE) {
      if (work0 < total_work) load_xw(work0, xw_next);
    }
    for (int w = work0; w < total_work; w += work_stride, ++iter) {
      const WorkItem it = decode_work(args, work_item_at(args, w), crank);
      const EpiArgs& e = it.g >= 0 ? args.epi[it.g] : args.xdev[-1 - it.g].epi;
      const uint32_t acc = iter % kAcc, acc_phase = (iter / kAcc) & 1;
      const int row = it.m0 + q * 32 + lane;
      const int ncols = min(BN, it.N - it.n0);
      const int nchunks = (ncols + 31) >> 5;
      if constexpr (EPI == EPI_BCE) {
#pragma unroll
        for (int k = 0; k < kMaxOwn; ++k) xw[k] = xw_next[k];
        if (w + work_stride < total_work) load_xw(w + work_stride, xw_next);
      }
      ptx::mbar_wait(&tfull_bar[acc], acc_phase);
      ptx::tc_fence_after();
      if (warp == 2 && lane == 0 && w == work0) MAGIC_TR(7);
      if (warp == 2 && lane == 0 && iter < 8) MAGIC_TR(16 + iter * 5 + 2);
      bool arrived = false;
      // hand the accumulator stage back to the MMA issuer: per thread on this CTA's barrier, or -- in a pair -- one
      // arrival per warp on the leader's, which waits for the epilogue warps of both CTAs
      auto release_acc = [&](uint32_t a_) {
        ptx::tc_fence_before();
        if constexpr (CG2) {
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive_cluster(leader_addr(&tempty_bar[a_]));
        } else {
          ptx::mbar_arrive(&tempty_bar[a_]);
        }
      };
      // Two chunks are read per TMEM wait: a tcgen05.ld queues behind the MMAs already issued for the next tile, so
      // every separate load-wait round trip costs ~1-2 us during which the accumulator stays locked (r01i: the BCE
      // kernel ran 0.043 ms slower than its main loop with an epilogue that only read TMEM).  With 16 epilogue
      // warps the two chunks are all a warp owns, so the accumulator is handed back before any math starts.
      auto process = [&](const uint32_t (&r)[32], int k) {
        const int c = grp + k * kGroups;
        uint32_t xwk = 0u;
        if constexpr (EPI == EPI_BCE) xwk = xw[k];
        if constexpr (EPI == EPI_SCATTER_F32) {
          scatter_chunk(e, r, row, it.n0 + c * 32, it.M, it.N,
                        smem + kStages * kStageBytes + 256 + (warp - 2) * kScatterStageBytes, lane);
        } else if (EPI == EPI_ADAM && it.g >= 0 && it.m0 < args.adam.rows_end) {
          adam_chunk(args.adam, it.g, e.alpha, r, smem + kTileStageOffset + (warp - 2) * 2048, lane, it.m0 + q * 32,
                     it.n0 + c * 32, it.M, it.N);
        } else {
          TileOut t;
          t.hi = it.g >= 0 ? &args.tout[it.g][0] : &args.xdev[-1 - it.g].tout[0];
          t.lo = it.g >= 0 ? &args.tout[it.g][1] : &args.xdev[-1 - it.g].tout[1];
          t.slot = smem + kTileStageOffset + (warp - 2) * 2048;
          t.lane = lane;
          t.row0 = it.m0 + q * 32;
          t.reduce = e.atomic;
          epilogue_chunk<EPI == EPI_ADAM ? EPI_STORE_F32 : EPI, true>(e, t, r, row, it.n0 + c * 32, it.M, it.N, loss_acc, xwk);
        }
      };
      constexpr int kBatch = epi_warps(EPI, A_BITS) == 16 ? 2 : 1;   // 8-warp epilogues own 4 chunks: one at a time (registers)
#pragma unroll
      for (int k = 0; k < kMaxOwn; k += kBatch) {
        const int c0 = grp + k * kGroups, c1 = grp + (k + 1) * kGroups;
        if (c0 >= nchunks || (e.dbg & 4)) break;   // dbg 4: mainloop-only timing, the accumulator is never read
        const bool two = kBatch == 2 && (k + 1 < kMaxOwn) && c1 < nchunks;
        uint32_t ra[32], rb[32];
        const uint32_t tbase = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
        ptx::tmem_ld_32x32(tbase + c0 * 32, ra);
        if (two) ptx::tmem_ld_32x32(tbase + c1 * 32, rb);
        ptx::tmem_ld_wait();
        if (grp + (k + kBatch) * kGroups >= nchunks || k + kBatch >= kMaxOwn) {   // this warp's last TMEM read of the accumulator
          release_acc(acc);
          arrived = true;
          if (warp == 2 && lane == 0 && iter < 8) MAGIC_TR(16 + iter * 5 + 3);
        }
        process(ra, k);
        if (two) process(rb, k + 1);
      }
      if (!arrived) release_acc(acc);
      if (warp == 2 && lane == 0 && iter < 8) MAGIC_TR(16 + iter * 5 + 4);
    }
    if (warp == 2 && lane == 0) MAGIC_TR(8);
    ptx::bulk_wait_all();   // every tile store / row segment this thread issued has left before the CTA exits
    if (warp == 2 && lane == 0) MAGIC_TR(9);
    if constexpr (EPI == EPI_BCE) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) loss_acc += __shfl_xor_sync(0xffffffffu, loss_acc, o);
      // one atomic per CTA (not per warp: 2368 same-address atomics serialise in L2 at the kernel's tail): the warp
      // sums meet in the staging slots, which are idle after bulk_wait_all()
      if (lane == 0) *reinterpret_cast<float*>(smem + kTileStageOffset + (warp - 2) * 2048) = loss_acc;
    }
  } else if constexpr (A_BITS) {
    // ------------------------------------------------------------ bit expanders (A_BITS kernels only)
    // The A stage is 128 swizzle rows of 128 bytes: accumulator row r of the tile for the K-major kinds (64
    // consecutive k), k-row r & 63 of slab r >> 6 for the MN-major kind (64 consecutive m).  Either way row r is 64
    // consecutive bits of one row of the packed matrix and lands at byte r * 128 of the stage, 16-byte piece c at
    // position c ^ (r & 7) -- exactly what the TMA's 128-byte swizzle would have written from a bf16 matrix.
    // Warp e owns pipeline stage e (k-blocks e, e + 4, ... of this CTA's sequence; lane l writes rows l, l + 32,
    // l + 64, l + 96), so four stages are being expanded at once, each with its own wait -> store -> fence -> arrive
    // round trip (one group walking the stages in turn was bound by that latency: 920 cycles per k-block).  The bits of
    // a warp's next k-block are fetched while it expands the current one -- four k-block times ahead of their use.
    // (a second expander warp per stage -- measured with eight epilogue warps: 576 threads -- did not fit the 16 K registers of an SM sub-partition at
    // this kernel's 112 registers, and one warp per stage already keeps up with the MMAs: profiles/r01p)
    static_assert(kNS >= kExpandWarps, "expander warp e serves the k-blocks e, e + 4, ... of the CTA's sequence");
    const int ew = warp - (2 + epi_warps(EPI, A_BITS));
    struct Cursor { KbWalk walk; int w, m0, g, kblocks; bool valid; };
    auto open = [&](Cursor& c) {
      c.valid = false;
      while (c.w < total_work) {
        const WorkItem it = decode_work(args, work_item_at(args, c.w), crank);
        c.walk.begin(args, it);
        if (!c.walk.done()) { c.m0 = it.m0; c.g = it.g; c.kblocks = it.kblocks; c.valid = true; return; }
        c.w += work_stride;
      }
    };
    auto advance = [&](Cursor& c, int n) {       // n k-blocks further along the CTA's sequence of (work item, k-block)
      while (c.valid && n > 0) {
        const int left = c.walk.left();
        if (n < left) { c.walk.skip(n); return; }
        n -= left;
        c.w += work_stride;
        open(c);
      }
    };
    auto fetch = [&](const Cursor& c, uint2 (&v)[4]) {
#pragma unroll
      for (int q = 0; q < 4; ++q) v[q] = make_uint2(0u, 0u);
      if (c.valid && c.g >= 0) {
        const BitsA& ba = args.abits[c.g];
        int kk = c.walk.kb;
        while (kk >= c.kblocks) kk -= c.kblocks;          // k-block inside its precision segment: every segment reads the same A
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int r = lane + 32 * q;
          int row, word;
          if constexpr (!A_MN) { row = c.m0 + r; word = kk * (kBlockK / 32); }
          else { row = kk * kBlockK + (r & 63); word = (c.m0 >> 5) + 2 * (r >> 6); }
          if (row < ba.rows) v[q] = __ldg(reinterpret_cast<const uint2*>(ba.base + (long long)row * ba.pitch + word));
        }
      }
    };
    Cursor cur, pre;
    cur.w = work0; open(cur);
    advance(cur, ew);
    pre = cur;
    advance(pre, 4);
    uint2 now[4], nxt[4];
    fetch(cur, now);
    const int x = lane & 7;
    uint32_t pos = (uint32_t)ew;                          // position in the CTA's k-block sequence: stage pos % kNS
    while (cur.valid) {
      fetch(pre, nxt);
      const uint32_t st = pos % kNS, phase = (pos / kNS) & 1u;
      uint8_t* const stage_base = smem_a + st * kStageBytesA + lane * 128;
      ptx::mbar_wait(&empty_bar[st], phase ^ 1);          // the MMAs that read this stage have retired
      if (cur.g >= 0 && !(args.epi[0].dbg & 8)) {     // dbg 8: timing experiment, the stage is left as it is
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint8_t* dst = stage_base + q * (32 * 128);
          const uint32_t half[4] = {now[q].x & 0xFFFFu, now[q].x >> 16, now[q].y & 0xFFFFu, now[q].y >> 16};
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            uint32_t w[8];
            expand16(half[h], w);
            ptx::st_shared_v4(dst + (((2 * h) ^ x) << 4), w[0], w[1], w[2], w[3]);
            ptx::st_shared_v4(dst + (((2 * h + 1) ^ x) << 4), w[4], w[5], w[6], w[7]);
          }
        }
        ptx::fence_proxy_async_smem();                    // generic-proxy writes -> visible to the MMA's async-proxy reads
      }
      __syncwarp();
      if (lane == 0) {
        if constexpr (CG2) ptx::mbar_arrive_cluster(leader_addr(&full_bar[st]));
        else ptx::mbar_arrive(&full_bar[st]);
      }
      pos += kExpandWarps;
#pragma unroll
      for (int q = 0; q < 4; ++q) now[q] = nxt[q];
      cur = pre;
      advance(pre, 4);
    }
  }

  ptx::tc_fence_before();
  if (CM > 1) ptx::cluster_sync();       // no CTA leaves while a peer may still multicast into it or arrive on its barriers
  else __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    if constexpr (CG2) ptx::tmem_dealloc_pair(tmem_base, kTmemCols);
    else ptx::tmem_dealloc(tmem_base, kTmemCols);
  }
  if constexpr (EPI == EPI_BCE) {
    if (threadIdx.x == 64 && work0 < total_work) {
      float sum = 0.f;
#pragma unroll
      for (int w = 0; w < epi_warps(EPI, A_BITS); ++w) sum += *reinterpret_cast<const float*>(smem + kTileStageOffset + w * 2048);
      atomicAdd(args.epi[0].scalar, sum * args.epi[0].alpha);
    }
  }
  if (threadIdx.x == 0) MAGIC_TR(10);
#undef MAGIC_TR
}

}  // namespace magic
