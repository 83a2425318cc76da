// Non-GEMM kernels of the step: latent layer (reparameterisation, KL, deformation, metric loss and the four
// upstream latent gradients), fused dual Adam, input layout conversion, synthetic pair generator.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace magic {

enum MetricKind : int { METRIC_COSINE = 0, METRIC_SQDIFF = 1 };

// Row order of every [3B, *] activation buffer: [rot ; lock ; key]  (so both optimisers' wgrads read a
// contiguous 2B-row range: cost -> rows [0,2B), metric -> rows [B,3B)).
// Row order of every [4B, *] backward buffer: [rot|cost ; lock|cost ; lock|metric ; key|metric].
struct LatentArgs {
  int B;                 // local batch (pairs)
  int L;                 // latent dims
  float inv_bg;          // 1 / global batch
  float def_weight;      // 10 (8c/vae.py:287) or 100 (11a/vae.py:293)
  int metric;            // MetricKind
  int reciprocal;        // 9a/vae.py:301
  const float* heads;    // [3B, ld_heads]: mu at cols [0,L), log sigma^2 at cols [L,2L)
  long long ld_heads;
  const float* eps;      // [3, B, L] in API order (lock, rot, key), or nullptr -> internal RNG
  uint64_t seed, step;   // internal RNG counters (seed, step, rank folded into seed by the host)
  float* z;              // [3B, L] sampled latents (row order rot, lock, key)
  __nv_bfloat16* zdec;   // [B, ld_zdec] decoder input: z_lock, 1.0 at col L (bias column), zero pad
  __nv_bfloat16* zdec_lo;
  long long ld_zdec;
  const float* labels;   // [B]
  float* colsq;          // [2, L]  sum_i z_lock^2, sum_i z_key^2   (cosine; all-reduced across ranks)
  float* colS;           // [L]     sum_i g_i a^_ij b^_ij           (cosine backward; all-reduced)
  float* g;              // [B]     d training_loss / d dist_i (pre-reciprocal chain applied)
  float* distance;       // [B]     f(dist_i) as returned by get_predictions
  float* sums;           // [8]: 0 rec (written by the BCE epilogue), 1 kl, 2 def, 3 training  -- already / Bg
  const float* dz_dec;   // [B, ld_dz] decoder dgrad into z_lock, unscaled (sum over pixels of (sigma-x).W)
  long long ld_dz;
  __nv_bfloat16* dheads; // [4B, ld_dheads]: d mu at cols [0,L), d log sigma^2 at [L,2L)
  __nv_bfloat16* dheads_lo;
  long long ld_dheads;
  // bias gradients of the two heads, summed here in fp32 (they are column sums of dheads in which the deformation terms
  // of the rot and lock streams cancel exactly; summing the bf16-rounded rows instead loses the remainder):
  float* g1_bias;        // [2L] bias row of the heads' gradient, optimiser 1: d mu | d log sigma^2 (zero before the launch)
  float* g2_bias;        // [2L] optimiser 2
};
constexpr int kLatBwdSlab = 32;   // rows per CTA of latent_backward

void latent_forward(const LatentArgs& a, cudaStream_t s);     // z, zdec, kl, def, colsq (cosine) ; sqdiff: also metric
void latent_metric(const LatentArgs& a, cudaStream_t s);      // distance, training loss, g, colS (needs final colsq)
void latent_backward(const LatentArgs& a, cudaStream_t s);    // dheads for the four backward streams, head bias gradients
void zero_fill3(float* p0, int n0, float* p1, int n1, float* p2, int n2, cudaStream_t s);   // three float ranges, one launch
// cosine_distance(a, b) of 8c/vae.py:438-452 on two device arrays [B, L]; colsq_ws: 2L floats of scratch; out[B]
void cosine_distance_rows(const float* a, const float* b, int B, int L, float* colsq_ws, float* out, cudaStream_t s);
// The same on a static device array instead of caller-supplied scratch (no allocation anywhere): L <= kCosineStaticMaxL.
// Returns nullptr on success, else a message.  Calls on one device share the array: overlapping calls race.
constexpr int kCosineStaticMaxL = 4096;
const char* cosine_distance_rows_static(const float* a, const float* b, int B, int L, float* out, cudaStream_t s);

struct AdamArgs {
  float* p;                       // fp32 master parameters, flat
  const float* g1; const float* g2;
  float* m1; float* v1; float* m2; float* v2;
  __nv_bfloat16* w_hi; __nv_bfloat16* w_lo;   // bf16 shadow (lo may be null)
  long long begin;                // first element of the range to update (multiple of 4); the range ends at n_total
  long long n_both;               // leading elements updated by both optimisers (encoder)
  long long n_total;              // all elements (encoder + used decoder)
  float lr_t1, lr_t2;             // lr * sqrt(1-beta2^t) / (1-beta1^t)
  float beta1, beta2, eps;
  int thin;                       // > 0: cap the grid at this many CTAs (grid-stride), for a launch that runs beside GEMMs
};
void adam_dual(const AdamArgs& a, cudaStream_t s);
// Data parallel: sharded reduce + dual Adam + shadow all-gather over peer memory.
constexpr int kMaxRanks = 8;
struct AdamShardArgs {
  long long n_total;              // flat parameter space [0, n_total)
  long long n_both;               // elements [0, n_both) are trained by both optimisers
  int world, rank, slice_len;     // slice_len = owned elements per rank (granules dealt round-robin, see EpiArgs)
  int gran_shift;
  const float* stage1; const float* stage2;   // local staging [world][slice_len]: every rank's partial g1 / g2 of this slice
  float *p, *m1, *v1, *m2, *v2;   // flat buffers (only the owned range is touched)
  float *g1, *g2;                 // reduced gradients, written at the owned indices (diagnostics / parity)
  __nv_bfloat16* w_hi[kMaxRanks]; // shadow weights of every rank (peer pointers; [rank] is local)
  __nv_bfloat16* w_lo[kMaxRanks]; // null in the bf16 mode
  float lr_t1, lr_t2, beta1, beta2, eps;
  int apply;                      // 0: reduce only (magic_step_grads_only)
};
void adam_dual_sharded(const AdamShardArgs& a, cudaStream_t s);
struct PushArgs {
  const float* src;               // local flat gradient buffer
  float* peer[kMaxRanks];         // staging base of every rank
  long long begin, end;           // flat range, multiples of 4
  int slice_len, rank, world, gran_shift;
};
void push_range(const PushArgs& a, cudaStream_t s);
void mask_unowned(float* buf, long long n, int world, int rank, int gran_shift, cudaStream_t s);
void shadow_refresh(const float* p, __nv_bfloat16* hi, __nv_bfloat16* lo, long long n, cudaStream_t s);

// X [B, 3*D] fp32 channel-interleaved (8c/vae.py:166, ch0 lock, ch1 rot, ch2 key) -> planes bf16 [3B, ld]
// rows [rot; lock; key], 1.0 at col D, zeros after.
// *not_binary (device) is set to 1 when a value is neither 0 nor 1 (uint8 feeds: nor 255): the first layer reads the images as bits
void deinterleave_f32(const float* x, int B, int D, __nv_bfloat16* planes, long long ld, int* not_binary, cudaStream_t s);
void deinterleave_u8(const uint8_t* x, int B, int D, __nv_bfloat16* planes, long long ld, int* not_binary, cudaStream_t s);
void planar_u8(const uint8_t* lock, const uint8_t* rot, const uint8_t* key, int B, int D, __nv_bfloat16* planes,
               long long ld, int* not_binary, cudaStream_t s);
void planes_to_interleaved_f32(const __nv_bfloat16* planes, long long ld, int B, int D, float* x, cudaStream_t s);

// On-device synthetic shape pairs (replaces 8c/overlap_input.py): random filled shapes on a
// side x side canvas, nearest-neighbour rotated copy (8c/utils.py:329-340), integer overlap label.
// Output: packed rows xb3[3B, xb_ld] in the order [rot; lock; key] (pixel p = bit p & 31 of word p >> 5, bias column D
// set): the A operand of the first layer's GEMMs and, rows [B, 2B), the target the sigmoid-BCE epilogue reads
// (8c/vae.py:260-263).  `planes` may be null; otherwise the same rows are also written as bf16 [3B, ld].
void generate_pairs(uint64_t seed, uint64_t step, int B, int side, __nv_bfloat16* planes, long long ld, uint32_t* xb3,
                    int xb_ld, float* labels, int* rot_cs_sn, int max_ctas, cudaStream_t s);
void pairs_u8(const uint8_t* lock, const uint8_t* key, int B, int side, uint64_t seed, uint64_t step,
              __nv_bfloat16* planes, long long ld, int* rot_cs_sn, cudaStream_t s);
// bit-packed pairs: pixel p = bit (p & 31) of word p >> 5 of the pair's row (wpr words per row); fills xb3 like the
// generator, and `planes` when it is not null
void pairs_bits(const uint32_t* lock, const uint32_t* key, int wpr, int B, int side, uint64_t seed, uint64_t step,
                __nv_bfloat16* planes, long long ld, uint32_t* xb3, int xb_ld, int* rot_cs_sn, int max_ctas, cudaStream_t s);
// Small all-reduce (sum) over peer memory, for the handful of floats the data-parallel step exchanges (batch-axis
// statistics of the cosine metric, the loss scalars, its two barriers): every rank stores its n values into slot `rank`
// of every peer's exchange area, raises a flag there, waits for all flags in its own area and sums the slots in rank
// order (same result on every rank, bit for bit).  One CTA; ~5 us against ~25 us for an NCCL call of this size, and the
// step has four of them on its critical path.  Two areas alternate (epoch parity) so that a rank one call ahead does not
// overwrite what a slower rank is still reading.  A peer that never arrives traps after kXchgTimeoutNs instead of hanging.
constexpr int kXchgMaxFloats = 8192;
constexpr unsigned long long kXchgTimeoutNs = 10000000000ull;
// layout of one rank's exchange area (floats): [2 parities][world][kXchgMaxFloats] data, then [2][kMaxRanks] flags (u32)
inline size_t xchg_area_floats(int world) { return (size_t)2 * world * kXchgMaxFloats + 2 * 8; }
struct XchgArgs {
  float* peer[8];                 // every rank's exchange area (own one included), as mapped here
  float* buf;                     // local values in, sums out
  int n, world, rank;
  unsigned epoch;                 // 1, 2, 3, ... : the same sequence on every rank
};
void peer_allreduce_small(const XchgArgs& a, cudaStream_t s);
// Gather of dataset rows that live in PINNED HOST memory (mapped into the device's address space): row idx[b] of
// lock_all / key_all ([n_rows, wpr] words, wpr % 4 == 0) -> row b of out_lock / out_key in HBM.  The GPU reads the rows over
// PCIe itself; a thin grid (one 128-thread CTA on some of the SMs, <= 32 registers) that fits beside every GEMM CTA.
void gather_rows_host(const uint32_t* lock_all, const uint32_t* key_all, const int32_t* idx, int B, int wpr, long long n_rows,
                      uint32_t* out_lock, uint32_t* out_key, cudaStream_t s);
void pack_planes_bits(const __nv_bfloat16* planes, long long ld, int rows, int D, uint32_t* xb, int xb_ld, cudaStream_t s);
void unpack_bits_planes(const uint32_t* xb, int xb_ld, int rows, __nv_bfloat16* planes, long long ld, cudaStream_t s);

// Zero-stage maps of packed rows for the first layer's GEMMs (GemmArgs.kmask): which 128 x 64 operand stages hold a
// set pixel.  fwd: mask[ceil(rows / 128)][mwords] over k-blocks of 64 pixels; wgrad: mask[tiles_m][mwords] over
// k-blocks of 64 of the `krows` rows starting at xb, M tiles of 128 pixels.  mask size a multiple of 16 bytes.
void build_kmask_fwd(const uint32_t* xb, int xb_ld, int rows, int kblocks, uint32_t* mask, int mwords, cudaStream_t s);
void build_kmask_wgrad(const uint32_t* xb, int xb_ld, int krows, int tiles_m, uint32_t* mask, int mwords, cudaStream_t s);

void zero_fill(void* ptr, size_t bytes, cudaStream_t s);   // kernel-based memset(0): ptr 16-byte aligned, bytes % 4 == 0
void kernels_init();                                        // one-time function attributes (shared-memory carveout)

}  // namespace magic
