// Non-GEMM kernels of the step (see kernels.h).  All are HBM/latency-bound elementwise or reduction kernels.
#include "kernels.h"
#include "ptx.cuh"
#include "launch.cuh"
#include <math.h>
#include <cstring>

namespace magic {

// ------------------------------------------------------------------ counter-based RNG
__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__device__ __forceinline__ uint64_t rng_key(uint64_t seed, uint64_t step, uint64_t a, uint64_t b) {
  return splitmix64(splitmix64(splitmix64(seed ^ 0xA5A5A5A5ull) + step) + a * 0x100000001B3ull + b);
}
__device__ __forceinline__ float u01(uint32_t bits) { return (bits >> 8) * (1.0f / 16777216.0f) + (0.5f / 16777216.0f); }
// one N(0,1) draw from a 64-bit key (Box-Muller)
__device__ __forceinline__ float normal_from_key(uint64_t k) {
  float u1 = u01((uint32_t)k), u2 = u01((uint32_t)(k >> 32));
  return sqrtf(-2.f * logf(u1)) * cospif(2.f * u2);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void store_split(__nv_bfloat16* hi, __nv_bfloat16* lo, long long idx, float v) {
  __nv_bfloat16 h = __float2bfloat16(v);
  hi[idx] = h;
  if (lo) lo[idx] = __float2bfloat16(v - __bfloat162float(h));
}

// ------------------------------------------------------------------ zero fill
// A kernel instead of cudaMemsetAsync so that every launch of the step shares one shared-memory carveout (below).
__global__ void __launch_bounds__(256) zero_kernel(uint4* p, long long n16, uint32_t* tail, int ntail) {
  ptx::grid_dep_wait();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n16) p[i] = make_uint4(0u, 0u, 0u, 0u);
  if (i < ntail) tail[i] = 0u;
}
// three ranges in one launch (the step's statistics and the two bias rows latent_backward adds into)
__global__ void __launch_bounds__(256) zero3_kernel(float* p0, int n0, float* p1, int n1, float* p2, int n2) {
  ptx::grid_dep_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n0) p0[i] = 0.f;
  if (i < n1) p1[i] = 0.f;
  if (i < n2) p2[i] = 0.f;
}
void zero_fill3(float* p0, int n0, float* p1, int n1, float* p2, int n2, cudaStream_t s) {
  const int n = n0 > n1 ? (n0 > n2 ? n0 : n2) : (n1 > n2 ? n1 : n2);
  if (n > 0) launch_chain(zero3_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, s, p0, n0, p1, n1, p2, n2);
}
void zero_fill(void* ptr, size_t bytes, cudaStream_t s) {   // ptr 16-byte aligned, bytes a multiple of 4
  const long long n16 = (long long)(bytes / 16);
  const int ntail = (int)((bytes % 16) / 4);
  const long long n = n16 > ntail ? n16 : ntail;
  if (n == 0) return;
  launch_chain(zero_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, s, reinterpret_cast<uint4*>(ptr), n16,
                                                     reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(ptr) + n16 * 16), ntail);
}

// ------------------------------------------------------------------ latent forward
// (8c/vae.py:358-371, 275-288)  Thread <-> latent column, CTA <-> slab of kLatSlab pairs: loads are coalesced along
// the latent axis and the batch-axis sums of squares the cosine metric needs stay in registers until the end.
constexpr int kLatWarps = 8;
constexpr int kLatRowsPerWarp = 2;
constexpr int kLatSlab = 8;    // 512 CTAs at B = 4096

__global__ void __launch_bounds__(256) latent_forward_kernel(LatentArgs a) {
  ptx::grid_dep_wait();
  const int L = a.L, B = a.B;
  const int j = blockIdx.x * 256 + threadIdx.x;
  const int i0 = blockIdx.y * kLatSlab, i1 = min(B, i0 + kLatSlab);
  const bool cosine = a.metric == METRIC_COSINE;
  float kl = 0.f, def = 0.f, sq_l = 0.f, sq_k = 0.f;
  if (j < L) {
    for (int i = i0; i < i1; ++i) {
      const float* h_rot = a.heads + (long long)i * a.ld_heads;
      const float* h_lock = a.heads + (long long)(B + i) * a.ld_heads;
      const float* h_key = a.heads + (long long)(2 * B + i) * a.ld_heads;
      float e_lock, e_rot, e_key;
      if (a.eps) {
        e_lock = a.eps[((long long)0 * B + i) * L + j];
        e_rot = a.eps[((long long)1 * B + i) * L + j];
        e_key = a.eps[((long long)2 * B + i) * L + j];
      } else {
        const uint64_t c = (uint64_t)i * L + j;
        e_lock = normal_from_key(rng_key(a.seed, a.step, 0, c));
        e_rot = normal_from_key(rng_key(a.seed, a.step, 1, c));
        e_key = normal_from_key(rng_key(a.seed, a.step, 2, c));
      }
      const float mu = h_lock[j], lv = h_lock[L + j];
      const float ex = expf(lv);
      const float zl = mu + sqrtf(ex) * e_lock;
      const float zr = h_rot[j] + sqrtf(expf(h_rot[L + j])) * e_rot;
      const float zk = h_key[j] + sqrtf(expf(h_key[L + j])) * e_key;
      a.z[(long long)i * L + j] = zr;
      a.z[(long long)(B + i) * L + j] = zl;
      a.z[(long long)(2 * B + i) * L + j] = zk;
      store_split(a.zdec, a.zdec_lo, (long long)i * a.ld_zdec + j, zl);
      if (j == 0) store_split(a.zdec, a.zdec_lo, (long long)i * a.ld_zdec + L, 1.0f);   // bias column
      kl += -0.5f * (1.f + lv - mu * mu - ex);
      const float d = zl - zr;
      def += d * d;
      sq_l += zl * zl;
      sq_k += zk * zk;
    }
    if (cosine) {
      atomicAdd(&a.colsq[j], sq_l);
      atomicAdd(&a.colsq[L + j], sq_k);
    }
  }
  // one atomic pair per CTA: with one per warp, 8192 same-address atomics serialised in L2 (~30 us of a 40 us kernel)
  __shared__ float red[2][8];
  kl = warp_sum(kl);
  def = warp_sum(def);
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = kl; red[1][threadIdx.x >> 5] = def; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float k2 = 0.f, d2 = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) { k2 += red[0][w]; d2 += red[1][w]; }
    if (k2 != 0.f || d2 != 0.f) {
      atomicAdd(&a.sums[1], k2 * a.inv_bg);
      atomicAdd(&a.sums[2], d2 * a.def_weight * a.inv_bg);
    }
  }
}

// ------------------------------------------------------------------ metric loss (8c/vae.py:300-307, 438-452)
__global__ void __launch_bounds__(kLatWarps * 32) latent_metric_kernel(LatentArgs a) {
  ptx::grid_dep_wait();
  extern __shared__ float sm[];   // cosine: [L] column sums S_j | [L] 1/||a_j|| | [L] 1/||b_j||
  const int L = a.L, B = a.B;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool cosine = a.metric == METRIC_COSINE;
  float* const s_ra = sm + L;
  float* const s_rb = sm + 2 * L;
  if (cosine) {
    for (int j = threadIdx.x; j < L; j += blockDim.x) {
      sm[j] = 0.f;
      s_ra[j] = rsqrtf(fmaxf(a.colsq[j], 1e-12f));
      s_rb[j] = rsqrtf(fmaxf(a.colsq[L + j], 1e-12f));
    }
    __syncthreads();
  }
  float tl = 0.f;
  for (int rr = 0; rr < kLatRowsPerWarp; ++rr) {
    const int i = (blockIdx.x * kLatWarps + warp) * kLatRowsPerWarp + rr;
    if (i >= B) break;
    const float* zl = a.z + (long long)(B + i) * L;
    const float* zk = a.z + (long long)(2 * B + i) * L;
    float dist = 0.f;
    for (int j = lane; j < L; j += 32) {
      if (cosine) {
        dist += (zl[j] * s_ra[j]) * (zk[j] * s_rb[j]);
      } else {
        const float d = zl[j] - zk[j];
        dist += d * d;
      }
    }
    dist = warp_sum(dist);
    const float f = a.reciprocal ? 1.f / dist : dist;
    const float err = f - a.labels[i];
    const float dfd = a.reciprocal ? -f * f : 1.f;       // d(1/d)/dd = -1/d^2
    const float gi = 2.f * err * a.inv_bg * dfd;
    if (lane == 0) {
      a.distance[i] = f;
      a.g[i] = gi;
      tl += err * err;
    }
    if (cosine) {
      for (int j = lane; j < L; j += 32) atomicAdd(&sm[j], gi * (zl[j] * s_ra[j]) * (zk[j] * s_rb[j]));
    }
  }
  __shared__ float red_tl[kLatWarps];
  if (lane == 0) red_tl[warp] = tl;
  __syncthreads();                     // (also orders the sm[] atomics before the column pass below)
  if (threadIdx.x == 0) {
    float t2 = 0.f;
#pragma unroll
    for (int w = 0; w < kLatWarps; ++w) t2 += red_tl[w];
    if (t2 != 0.f) atomicAdd(&a.sums[3], t2 * a.inv_bg);
  }
  if (cosine) {
    for (int j = threadIdx.x; j < L; j += blockDim.x) atomicAdd(&a.colS[j], sm[j]);
  }
}

// ------------------------------------------------------------------ cosine_distance(a, b) as a free function
// (8c/vae.py:438-452: tf.nn.l2_normalize(x, 0) normalises over axis 0 = the BATCH axis, then the products are summed
// over the latent axis.)  Two small kernels: column sums of squares of a and b, then one warp per row.
__global__ void __launch_bounds__(256) cosine_colsq_kernel(const float* __restrict__ a, const float* __restrict__ b, int B, int L,
                                                           float* colsq /* [2L], zeroed */) {
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j >= L) return;
  float sa = 0.f, sb = 0.f;
  for (int i = blockIdx.y; i < B; i += gridDim.y) {
    const float x = a[(long long)i * L + j], y = b[(long long)i * L + j];
    sa += x * x;
    sb += y * y;
  }
  atomicAdd(&colsq[j], sa);
  atomicAdd(&colsq[L + j], sb);
}
__global__ void __launch_bounds__(256) cosine_rows_kernel(const float* __restrict__ a, const float* __restrict__ b, int B, int L,
                                                          const float* __restrict__ colsq, float* __restrict__ out) {
  const int i = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (i >= B) return;
  float d = 0.f;
  for (int j = lane; j < L; j += 32)
    d += (a[(long long)i * L + j] * rsqrtf(fmaxf(colsq[j], 1e-12f))) * (b[(long long)i * L + j] * rsqrtf(fmaxf(colsq[L + j], 1e-12f)));
  d = warp_sum(d);
  if (lane == 0) out[i] = d;
}
void cosine_distance_rows(const float* a, const float* b, int B, int L, float* colsq_ws, float* out, cudaStream_t s) {
  cudaMemsetAsync(colsq_ws, 0, sizeof(float) * 2 * (size_t)L, s);
  const int slabs = B < 64 ? 1 : 64;
  cosine_colsq_kernel<<<dim3((L + 255) / 256, slabs), 256, 0, s>>>(a, b, B, L, colsq_ws);
  cosine_rows_kernel<<<(B + 7) / 8, 256, 0, s>>>(a, b, B, L, colsq_ws, out);
}

// ------------------------------------------------------------------ cosine_distance without any allocation
// (No counterpart in the reference: added for callers that cannot call cudaMalloc --
// code under CUDA graph capture, tight serving loops, harnesses that forbid it.)  The same two kernels as
// cosine_distance_rows(); their 2L floats of scratch (column sums of squares) live in a file-scope device array, 32 KiB
// that every GPU which loads this library holds until the process exits.  In host code the array's NAME is only a
// host-side stand-in: the device address is looked up once per device with cudaGetSymbolAddress (that device current)
// and that pointer is what the kernels receive.  All calls on a device share the array and it keeps the previous call's
// contents: every call clears the part it uses on its stream first; calls that overlap on the GPU (two streams) race.
static __device__ __align__(256) float g_cosine_scratch[2 * kCosineStaticMaxL];
static float* g_cosine_scratch_ptr[64] = {};       // per device ordinal, filled on first use
const char* cosine_distance_rows_static(const float* a, const float* b, int B, int L, float* out, cudaStream_t s) {
  if (L < 1 || L > kCosineStaticMaxL) return "L exceeds the static scratch (maximum 4096 latent columns)";
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return "no current CUDA device (ordinal 0..63)";
  float* ws = g_cosine_scratch_ptr[dev];
  if (ws == nullptr) {                             // set-up, once per device: the symbol's address ON this device
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_cosine_scratch);
    if (e != cudaSuccess || p == nullptr) { (void)cudaGetLastError(); return "cudaGetSymbolAddress(scratch) failed"; }
    ws = g_cosine_scratch_ptr[dev] = static_cast<float*>(p);
  }
  if (cudaMemsetAsync(ws, 0, sizeof(float) * 2 * (size_t)L, s) != cudaSuccess) { (void)cudaGetLastError(); return "clearing the scratch failed"; }
  const int slabs = B < 64 ? 1 : 64;
  cosine_colsq_kernel<<<dim3((L + 255) / 256, slabs), 256, 0, s>>>(a, b, B, L, ws);
  cosine_rows_kernel<<<(B + 7) / 8, 256, 0, s>>>(a, b, B, L, ws, out);
  return cudaGetLastError() == cudaSuccess ? nullptr : "kernel launch failed";
}

// ------------------------------------------------------------------ latent backward
// dheads rows: [rot|cost ; lock|cost ; lock|metric ; key|metric], cols [d mu | d log sigma^2].
// d z / d mu = 1 ;  d z / d lv = 0.5 * sqrt(exp(lv)) * eps = 0.5 * (z - mu).
// CTA = 64 latent columns x 4 row groups; a slab of kLatBwdSlab rows per CTA, so each thread walks kLatBwdSlab / 4 rows.
__global__ void __launch_bounds__(256) latent_backward_kernel(LatentArgs a) {
  ptx::grid_dep_wait();
  const int L = a.L, B = a.B;
  const int tc = threadIdx.x & 63, rg = threadIdx.x >> 6;
  const int j = blockIdx.x * 64 + tc;
  const int i0 = blockIdx.y * kLatBwdSlab, i1 = min(B, i0 + kLatBwdSlab);
  float bsum[4] = {0.f, 0.f, 0.f, 0.f};                         // g1: d mu, d lv ; g2: d mu, d lv
  if (j < L) {
    float sa = 0.f, sb = 0.f, ra = 0.f, rb = 0.f, S = 0.f;
    if (a.metric == METRIC_COSINE) {
      sa = a.colsq[j]; sb = a.colsq[L + j];
      ra = rsqrtf(fmaxf(sa, 1e-12f)); rb = rsqrtf(fmaxf(sb, 1e-12f));
      S = a.colS[j];
    }
    const long long ld = a.ld_dheads;
    for (int i = i0 + rg; i < i1; i += 4) {
      const float* h_rot = a.heads + (long long)i * a.ld_heads;
      const float* h_lock = a.heads + (long long)(B + i) * a.ld_heads;
      const float* h_key = a.heads + (long long)(2 * B + i) * a.ld_heads;
      const float zr = a.z[(long long)i * L + j], zl = a.z[(long long)(B + i) * L + j], zk = a.z[(long long)(2 * B + i) * L + j];
      const float mu = h_lock[j], lv = h_lock[L + j];
      // --- optimiser 1: cost
      const float ddef = 2.f * a.def_weight * (zl - zr) * a.inv_bg;
      const float dz_l0 = a.dz_dec[(long long)i * a.ld_dz + j] * a.inv_bg;       // through the decoder
      const float dz_l1 = dz_l0 + ddef;
      const float dmu_l1 = dz_l1 + mu * a.inv_bg;
      const float dlv_l1 = dz_l1 * 0.5f * (zl - mu) + 0.5f * (expf(lv) - 1.f) * a.inv_bg;
      const float dz_r1 = -ddef;
      const float dlv_r1 = dz_r1 * 0.5f * (zr - h_rot[j]);
      // --- optimiser 2: metric
      float da, db;
      const float gi = a.g[i];
      if (a.metric == METRIC_COSINE) {
        const float ah = zl * ra, bh = zk * rb;
        // through x * rsqrt(max(sum x^2, eps)): the normalisation term vanishes when the clamp is active
        da = (gi * bh - (sa > 1e-12f ? ah * S : 0.f)) * ra;
        db = (gi * ah - (sb > 1e-12f ? bh * S : 0.f)) * rb;
      } else {
        da = gi * 2.f * (zl - zk);
        db = -da;
      }
      const float dlv_l2 = da * 0.5f * (zl - mu);
      const float dlv_k2 = db * 0.5f * (zk - h_key[j]);
      store_split(a.dheads, a.dheads_lo, (long long)i * ld + j, dz_r1);
      store_split(a.dheads, a.dheads_lo, (long long)i * ld + L + j, dlv_r1);
      store_split(a.dheads, a.dheads_lo, (long long)(B + i) * ld + j, dmu_l1);
      store_split(a.dheads, a.dheads_lo, (long long)(B + i) * ld + L + j, dlv_l1);
      store_split(a.dheads, a.dheads_lo, (long long)(2 * B + i) * ld + j, da);
      store_split(a.dheads, a.dheads_lo, (long long)(2 * B + i) * ld + L + j, dlv_l2);
      store_split(a.dheads, a.dheads_lo, (long long)(3 * B + i) * ld + j, db);
      store_split(a.dheads, a.dheads_lo, (long long)(3 * B + i) * ld + L + j, dlv_k2);
      bsum[0] += dz_l0 + mu * a.inv_bg;             // = dz_r1 + dmu_l1 with the deformation terms cancelled exactly
      bsum[1] += dlv_r1 + dlv_l1;
      bsum[2] += da + db;
      bsum[3] += dlv_l2 + dlv_k2;
    }
  }
  // the heads' bias gradients: the four row groups meet in shared memory, one atomic per (column, sum) and CTA into
  // the bias rows of g1 / g2 (cleared with the step's statistics, Engine::run_forward)
  __shared__ float red[4][4][64];
#pragma unroll
  for (int q = 0; q < 4; ++q) red[rg][q][tc] = bsum[q];
  __syncthreads();
  if (rg == 0 && j < L) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float t = red[0][q][tc] + red[1][q][tc] + red[2][q][tc] + red[3][q][tc];
      float* dst = (q < 2 ? a.g1_bias : a.g2_bias) + ((q & 1) ? L + j : j);
      if (t != 0.f) atomicAdd(dst, t);
    }
  }
}

static int lat_blocks(int B) { return (B + kLatWarps * kLatRowsPerWarp - 1) / (kLatWarps * kLatRowsPerWarp); }

void latent_forward(const LatentArgs& a, cudaStream_t s) {
  launch_chain(latent_forward_kernel, dim3((a.L + 255) / 256, (a.B + kLatSlab - 1) / kLatSlab), dim3(256), 0, s, a);
}
void latent_metric(const LatentArgs& a, cudaStream_t s) {
  launch_chain(latent_metric_kernel, dim3(lat_blocks(a.B)), dim3(kLatWarps * 32), sizeof(float) * 3 * a.L, s, a);
}
void latent_backward(const LatentArgs& a, cudaStream_t s) {
  launch_chain(latent_backward_kernel, dim3((a.L + 63) / 64, (a.B + kLatBwdSlab - 1) / kLatBwdSlab), dim3(256), 0, s, a);
}

// ------------------------------------------------------------------ fused dual Adam (8c/vae.py:314-321)
// TF1 ApplyAdam:  m <- b1 m + (1-b1) g ; v <- b2 v + (1-b2) g^2 ; p <- p - lr_t m / (sqrt(v) + eps),
// lr_t = lr sqrt(1-b2^t)/(1-b1^t) (host).  Both optimisers' updates are applied in ONE pass over p
// (W <- W - d1 - d2) and the bf16 shadow weights the GEMMs read are re-emitted in the same pass.
__device__ __forceinline__ float adam_one(float g, float& m, float& v, float lr_t, float b1, float b2, float eps) {
  m = b1 * m + (1.f - b1) * g;
  v = b2 * v + (1.f - b2) * g * g;
  return lr_t * m / (sqrtf(v) + eps);
}

__global__ void __launch_bounds__(256) adam_dual_kernel(AdamArgs a) {
  ptx::grid_dep_wait();
  // grid-stride: the full-width launch covers the range in one pass; the side-stream launch (a.thin) uses two CTAs
  // per SM so that it leaves room for the GEMM CTAs it runs beside
  const long long stride = (long long)gridDim.x * blockDim.x * 4;
  for (long long i4 = a.begin + ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4; i4 < a.n_total; i4 += stride) {
    float4 p = *reinterpret_cast<const float4*>(a.p + i4);
    float pv[4] = {p.x, p.y, p.z, p.w};
    {
      const float4 g = *reinterpret_cast<const float4*>(a.g1 + i4);
      float4 m = *reinterpret_cast<const float4*>(a.m1 + i4), v = *reinterpret_cast<const float4*>(a.v1 + i4);
      pv[0] -= adam_one(g.x, m.x, v.x, a.lr_t1, a.beta1, a.beta2, a.eps);
      pv[1] -= adam_one(g.y, m.y, v.y, a.lr_t1, a.beta1, a.beta2, a.eps);
      pv[2] -= adam_one(g.z, m.z, v.z, a.lr_t1, a.beta1, a.beta2, a.eps);
      pv[3] -= adam_one(g.w, m.w, v.w, a.lr_t1, a.beta1, a.beta2, a.eps);
      *reinterpret_cast<float4*>(a.m1 + i4) = m;
      *reinterpret_cast<float4*>(a.v1 + i4) = v;
    }
    if (i4 < a.n_both) {
      const float4 g = *reinterpret_cast<const float4*>(a.g2 + i4);
      float4 m = *reinterpret_cast<const float4*>(a.m2 + i4), v = *reinterpret_cast<const float4*>(a.v2 + i4);
      pv[0] -= adam_one(g.x, m.x, v.x, a.lr_t2, a.beta1, a.beta2, a.eps);
      pv[1] -= adam_one(g.y, m.y, v.y, a.lr_t2, a.beta1, a.beta2, a.eps);
      pv[2] -= adam_one(g.z, m.z, v.z, a.lr_t2, a.beta1, a.beta2, a.eps);
      pv[3] -= adam_one(g.w, m.w, v.w, a.lr_t2, a.beta1, a.beta2, a.eps);
      *reinterpret_cast<float4*>(a.m2 + i4) = m;
      *reinterpret_cast<float4*>(a.v2 + i4) = v;
    }
    *reinterpret_cast<float4*>(a.p + i4) = make_float4(pv[0], pv[1], pv[2], pv[3]);
    const uint32_t h0 = ptx::pack_bf16(pv[0], pv[1]), h1 = ptx::pack_bf16(pv[2], pv[3]);
    *reinterpret_cast<uint2*>(a.w_hi + i4) = make_uint2(h0, h1);
    if (a.w_lo) {
      const uint32_t l0 = ptx::pack_bf16(pv[0] - ptx::bf16_lo(h0), pv[1] - ptx::bf16_hi(h0));
      const uint32_t l1 = ptx::pack_bf16(pv[2] - ptx::bf16_lo(h1), pv[3] - ptx::bf16_hi(h1));
      *reinterpret_cast<uint2*>(a.w_lo + i4) = make_uint2(l0, l1);
    }
  }
}

static long long adam_blocks(const AdamArgs& a) {
  const long long n4 = (a.n_total - a.begin + 3) / 4;
  if (n4 <= 0) return 0;
  long long blocks = (n4 + 255) / 256;
  if (a.thin > 0 && blocks > a.thin) blocks = a.thin;
  return blocks;
}
// ------------------------------------------------------------------ small all-reduce over peer memory
__global__ void __launch_bounds__(256) peer_allreduce_small_kernel(XchgArgs a) {
  ptx::grid_dep_wait();
  const int par = (int)(a.epoch & 1u);
  const size_t slot = ((size_t)par * a.world + a.rank) * kXchgMaxFloats;
  for (int i = threadIdx.x; i < a.n; i += blockDim.x) {
    const float v = a.buf[i];
    for (int r = 0; r < a.world; ++r) a.peer[r][slot + i] = v;
  }
  __threadfence_system();
  __syncthreads();
  if ((int)threadIdx.x < a.world) {
    unsigned* flag = reinterpret_cast<unsigned*>(a.peer[threadIdx.x] + (size_t)2 * a.world * kXchgMaxFloats) + par * 8 + a.rank;
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(a.epoch) : "memory");
    // ... and wait for rank threadIdx.x's flag in the own area
    const unsigned* mine = reinterpret_cast<const unsigned*>(a.peer[a.rank] + (size_t)2 * a.world * kXchgMaxFloats) + par * 8 + threadIdx.x;
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
      unsigned v;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
      if (v == a.epoch) break;
      unsigned long long t1;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > kXchgTimeoutNs) __trap();              // a peer is gone: fail the context rather than hang the box
    }
  }
  __syncthreads();
  const float* own = a.peer[a.rank] + (size_t)par * a.world * kXchgMaxFloats;
  for (int i = threadIdx.x; i < a.n; i += blockDim.x) {
    float sum = 0.f;
    for (int r = 0; r < a.world; ++r) sum += __ldcv(own + (size_t)r * kXchgMaxFloats + i);
    a.buf[i] = sum;
  }
}
void peer_allreduce_small(const XchgArgs& a, cudaStream_t s) {
  launch_chain(peer_allreduce_small_kernel, dim3(1), dim3(256), 0, s, a);
}

// ------------------------------------------------------------------ dataset rows from pinned host memory
// 32 registers, 4 warps: one warp per SM sub-partition next to the fullest GEMM CTA of the step (BCE: 576 threads x 96
// registers leave 1024 registers on two of the four sub-partitions).  With 39 registers the gather's CTAs and the GEMMs'
// took turns on their SMs: 2.31 instead of 2.0 ms per end-to-end step.
constexpr int kGatherCtas = 64, kGatherThreads = 128, kGatherUnroll = 2;
__global__ void __maxnreg__(32) gather_rows_host_kernel(const uint4* __restrict__ lock_all, const uint4* __restrict__ key_all,
                                                                          const int32_t* __restrict__ idx, int B, int row_u4,
                                                                          long long n_rows, uint4* __restrict__ out_lock,
                                                                          uint4* __restrict__ out_key) {
  const long long per_plane = (long long)B * row_u4, total = 2 * per_plane;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < total; i0 += stride * kGatherUnroll) {
    uint4 v[kGatherUnroll];
#pragma unroll
    for (int u = 0; u < kGatherUnroll; ++u) {                 // all loads of the round in flight before the first store
      const long long i = i0 + u * stride;
      v[u] = make_uint4(0u, 0u, 0u, 0u);
      if (i < total) {
        const bool key = i >= per_plane;
        const long long j = key ? i - per_plane : i;
        const int b = (int)(j / row_u4), c = (int)(j - (long long)b * row_u4);
        const long long r = idx[b];
        if (r >= 0 && r < n_rows) v[u] = __ldcs((key ? key_all : lock_all) + r * row_u4 + c);
      }
    }
#pragma unroll
    for (int u = 0; u < kGatherUnroll; ++u) {
      const long long i = i0 + u * stride;
      if (i < total) {
        if (i >= per_plane) out_key[i - per_plane] = v[u];
        else out_lock[i] = v[u];
      }
    }
  }
}
void gather_rows_host(const uint32_t* lock_all, const uint32_t* key_all, const int32_t* idx, int B, int wpr, long long n_rows,
                      uint32_t* out_lock, uint32_t* out_key, cudaStream_t s) {
  static const int ctas = getenv("MAGIC_GATHER_CTAS") ? atoi(getenv("MAGIC_GATHER_CTAS")) : kGatherCtas;     // tuning knob
  gather_rows_host_kernel<<<ctas, kGatherThreads, 0, s>>>(reinterpret_cast<const uint4*>(lock_all),
                                                                  reinterpret_cast<const uint4*>(key_all), idx, B, wpr / 4, n_rows,
                                                                  reinterpret_cast<uint4*>(out_lock), reinterpret_cast<uint4*>(out_key));
}

void adam_dual(const AdamArgs& a, cudaStream_t s) {
  const long long blocks = adam_blocks(a);
  if (blocks > 0) launch_chain(adam_dual_kernel, dim3((unsigned)blocks), dim3(256), 0, s, a);
}

// ------------------------------------------------------------------ data parallel: sharded reduce + dual Adam
// Each rank owns one contiguous slice of the flat parameter space.  The wgrad GEMMs of all ranks have stored their
// partial gradients into this rank's staging buffer stage[src][slice] (EPI_SCATTER_F32 / push_range below).  This
// kernel finishes the reduce-scatter (sum over src), applies both TF1 Adams on the owned slice only (optimiser
// state traffic / world) and all-gathers the result by storing the new bf16 shadow weights into every rank's
// shadow buffer over NVLink.
__global__ void __launch_bounds__(256) adam_dual_sharded_kernel(AdamShardArgs a) {
  ptx::grid_dep_wait();
  const long long loc = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;     // position in the owned slice
  if (loc >= a.slice_len) return;
  // owned granules are dealt round-robin over the ranks: local granule loc >> gs is flat granule (that * world + rank)
  const long long i4 = ((((loc >> a.gran_shift) * a.world) + a.rank) << a.gran_shift) + (loc & ((1LL << a.gran_shift) - 1));
  if (i4 >= a.n_total) return;
  float4 g1 = make_float4(0.f, 0.f, 0.f, 0.f), g2 = g1;
  const bool both = i4 < a.n_both;
  for (int s = 0; s < a.world; ++s) {
    const float4 t = *reinterpret_cast<const float4*>(a.stage1 + (long long)s * a.slice_len + loc);
    g1.x += t.x; g1.y += t.y; g1.z += t.z; g1.w += t.w;
  }
  if (both)
    for (int s = 0; s < a.world; ++s) {
      const float4 t = *reinterpret_cast<const float4*>(a.stage2 + (long long)s * a.slice_len + loc);
      g2.x += t.x; g2.y += t.y; g2.z += t.z; g2.w += t.w;
    }
  *reinterpret_cast<float4*>(a.g1 + i4) = g1;
  if (both) *reinterpret_cast<float4*>(a.g2 + i4) = g2;
  if (!a.apply) return;
  float4 p = *reinterpret_cast<const float4*>(a.p + i4);
  float pv[4] = {p.x, p.y, p.z, p.w};
  {
    float4 m = *reinterpret_cast<const float4*>(a.m1 + i4), v = *reinterpret_cast<const float4*>(a.v1 + i4);
    pv[0] -= adam_one(g1.x, m.x, v.x, a.lr_t1, a.beta1, a.beta2, a.eps);
    pv[1] -= adam_one(g1.y, m.y, v.y, a.lr_t1, a.beta1, a.beta2, a.eps);
    pv[2] -= adam_one(g1.z, m.z, v.z, a.lr_t1, a.beta1, a.beta2, a.eps);
    pv[3] -= adam_one(g1.w, m.w, v.w, a.lr_t1, a.beta1, a.beta2, a.eps);
    *reinterpret_cast<float4*>(a.m1 + i4) = m;
    *reinterpret_cast<float4*>(a.v1 + i4) = v;
  }
  if (both) {
    float4 m = *reinterpret_cast<const float4*>(a.m2 + i4), v = *reinterpret_cast<const float4*>(a.v2 + i4);
    pv[0] -= adam_one(g2.x, m.x, v.x, a.lr_t2, a.beta1, a.beta2, a.eps);
    pv[1] -= adam_one(g2.y, m.y, v.y, a.lr_t2, a.beta1, a.beta2, a.eps);
    pv[2] -= adam_one(g2.z, m.z, v.z, a.lr_t2, a.beta1, a.beta2, a.eps);
    pv[3] -= adam_one(g2.w, m.w, v.w, a.lr_t2, a.beta1, a.beta2, a.eps);
    *reinterpret_cast<float4*>(a.m2 + i4) = m;
    *reinterpret_cast<float4*>(a.v2 + i4) = v;
  }
  *reinterpret_cast<float4*>(a.p + i4) = make_float4(pv[0], pv[1], pv[2], pv[3]);
  const uint32_t h0 = ptx::pack_bf16(pv[0], pv[1]), h1 = ptx::pack_bf16(pv[2], pv[3]);
  const uint2 hv = make_uint2(h0, h1);
  uint2 lv = make_uint2(0u, 0u);
  const bool split = a.w_lo[0] != nullptr;
  if (split) {
    lv.x = ptx::pack_bf16(pv[0] - ptx::bf16_lo(h0), pv[1] - ptx::bf16_hi(h0));
    lv.y = ptx::pack_bf16(pv[2] - ptx::bf16_lo(h1), pv[3] - ptx::bf16_hi(h1));
  }
  for (int r = 0; r < a.world; ++r) {           // r == rank is the local copy; the others are peer stores
    *reinterpret_cast<uint2*>(a.w_hi[r] + i4) = hv;
    if (split) *reinterpret_cast<uint2*>(a.w_lo[r] + i4) = lv;
  }
}
void adam_dual_sharded(const AdamShardArgs& a, cudaStream_t s) {
  const long long n4 = ((long long)a.slice_len + 3) / 4;
  if (n4 <= 0) return;
  launch_chain(adam_dual_sharded_kernel, dim3((unsigned)((n4 + 255) / 256)), dim3(256), 0, s, a);
}

// local partial gradients of the small layers (accumulated with split-K red.add) -> the owners' staging slots
__global__ void __launch_bounds__(256) push_range_kernel(PushArgs a) {
  ptx::grid_dep_wait();
  const long long i4 = a.begin + ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i4 >= a.end) return;
  const long long q = i4 >> a.gran_shift;
  const int o = (int)(q % a.world);
  float* dst = a.peer[o] + ((long long)a.rank * a.slice_len + ((q / a.world) << a.gran_shift) + (i4 & ((1LL << a.gran_shift) - 1)));
  *reinterpret_cast<float4*>(dst) = *reinterpret_cast<const float4*>(a.src + i4);
}

// zero every granule this rank does not own: a following all-reduce(sum) then assembles the full buffer on every rank
// (host accessors of sharded state: magic_get_param / get_slot / get_grad)
__global__ void __launch_bounds__(256) mask_unowned_kernel(float* buf, long long n, int world, int rank, int gran_shift) {
  const long long i4 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i4 >= n) return;
  if ((int)((i4 >> gran_shift) % world) != rank) *reinterpret_cast<float4*>(buf + i4) = make_float4(0.f, 0.f, 0.f, 0.f);
}
void mask_unowned(float* buf, long long n, int world, int rank, int gran_shift, cudaStream_t s) {
  const long long n4 = (n + 3) / 4;
  mask_unowned_kernel<<<(int)((n4 + 255) / 256), 256, 0, s>>>(buf, n, world, rank, gran_shift);
}
void push_range(const PushArgs& a, cudaStream_t s) {
  const long long n4 = (a.end - a.begin + 3) / 4;
  if (n4 <= 0) return;
  launch_chain(push_range_kernel, dim3((unsigned)((n4 + 255) / 256)), dim3(256), 0, s, a);
}

__global__ void __launch_bounds__(256) shadow_refresh_kernel(const float* __restrict__ p, __nv_bfloat16* hi,
                                                             __nv_bfloat16* lo, long long n) {
  const long long i4 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i4 >= n) return;
  const float4 v = *reinterpret_cast<const float4*>(p + i4);
  const uint32_t h0 = ptx::pack_bf16(v.x, v.y), h1 = ptx::pack_bf16(v.z, v.w);
  *reinterpret_cast<uint2*>(hi + i4) = make_uint2(h0, h1);
  if (lo) {
    const uint32_t l0 = ptx::pack_bf16(v.x - ptx::bf16_lo(h0), v.y - ptx::bf16_hi(h0));
    const uint32_t l1 = ptx::pack_bf16(v.z - ptx::bf16_lo(h1), v.w - ptx::bf16_hi(h1));
    *reinterpret_cast<uint2*>(lo + i4) = make_uint2(l0, l1);
  }
}
void shadow_refresh(const float* p, __nv_bfloat16* hi, __nv_bfloat16* lo, long long n, cudaStream_t s) {
  const long long n4 = (n + 3) / 4;
  shadow_refresh_kernel<<<(int)((n4 + 255) / 256), 256, 0, s>>>(p, hi, lo, n);
}

// ------------------------------------------------------------------ input layout (8c/vae.py:166-179)
// Each thread converts 8 pixels of one pair: 24 interleaved floats -> 8 bf16 into each of the three planes.
__global__ void __launch_bounds__(256) deinterleave_kernel(const float* __restrict__ x, int B, int D,
                                                           __nv_bfloat16* planes, long long ld, int* not_binary) {
  const int per_row = (int)(ld / 8);
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)B * per_row) return;
  const int i = (int)(idx / per_row);
  const int p0 = (int)(idx % per_row) * 8;
  uint32_t lock[4], rot[4], key[4];
  if (p0 + 8 <= D) {
    const float4* src = reinterpret_cast<const float4*>(x + ((long long)i * D + p0) * 3);
    float f[24];
#pragma unroll
    for (int q = 0; q < 6; ++q) {
      const float4 v = src[q];
      f[4 * q] = v.x; f[4 * q + 1] = v.y; f[4 * q + 2] = v.z; f[4 * q + 3] = v.w;
    }
    bool bad = false;
#pragma unroll
    for (int q = 0; q < 24; ++q) bad |= !(f[q] == 0.f || f[q] == 1.f);
    if (bad) *not_binary = 1;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      lock[q] = ptx::pack_bf16(f[6 * q], f[6 * q + 3]);
      rot[q] = ptx::pack_bf16(f[6 * q + 1], f[6 * q + 4]);
      key[q] = ptx::pack_bf16(f[6 * q + 2], f[6 * q + 5]);
    }
  } else {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float v[2][3];
      for (int h = 0; h < 2; ++h) {
        const int p = p0 + 2 * q + h;
        for (int c = 0; c < 3; ++c)
          v[h][c] = p < D ? x[((long long)i * D + p) * 3 + c] : (p == D ? 1.f : 0.f);   // bias column, then zero pad
        for (int c = 0; c < 3; ++c)
          if (!(v[h][c] == 0.f || v[h][c] == 1.f)) *not_binary = 1;
      }
      lock[q] = ptx::pack_bf16(v[0][0], v[1][0]);
      rot[q] = ptx::pack_bf16(v[0][1], v[1][1]);
      key[q] = ptx::pack_bf16(v[0][2], v[1][2]);
    }
  }
  *reinterpret_cast<uint4*>(planes + (long long)i * ld + p0) = make_uint4(rot[0], rot[1], rot[2], rot[3]);
  *reinterpret_cast<uint4*>(planes + (long long)(B + i) * ld + p0) = make_uint4(lock[0], lock[1], lock[2], lock[3]);
  *reinterpret_cast<uint4*>(planes + (long long)(2 * B + i) * ld + p0) = make_uint4(key[0], key[1], key[2], key[3]);
}
void deinterleave_f32(const float* x, int B, int D, __nv_bfloat16* planes, long long ld, int* not_binary, cudaStream_t s) {
  const long long n = (long long)B * (ld / 8);
  deinterleave_kernel<<<(int)((n + 255) / 256), 256, 0, s>>>(x, B, D, planes, ld, not_binary);
}

// X [B, 3*D] uint8 channel-interleaved ({0,1} or {0,255}; any other byte raises *not_binary) -> bf16 planes
// (the compact host feed: 1/4 of the fp32 bytes over PCIe).  Each thread converts 8 pixels (24 bytes).
__global__ void __launch_bounds__(256) deinterleave_u8_kernel(const uint8_t* __restrict__ x, int B, int D,
                                                              __nv_bfloat16* planes, long long ld, int* not_binary) {
  const int per_row = (int)(ld / 8);
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)B * per_row) return;
  const int i = (int)(idx / per_row);
  const int p0 = (int)(idx % per_row) * 8;
  uint32_t lock[4] = {0, 0, 0, 0}, rot[4] = {0, 0, 0, 0}, key[4] = {0, 0, 0, 0};
  if (p0 + 8 <= D) {
    const uint2* src = reinterpret_cast<const uint2*>(x + ((long long)i * D + p0) * 3);   // 24-byte groups: 8-byte aligned
    uint8_t b[24];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      const uint2 v = src[q];
      *reinterpret_cast<uint2*>(b + 8 * q) = v;
    }
    bool bad = false;
#pragma unroll
    for (int q = 0; q < 24; ++q) bad |= b[q] > 1 && b[q] != 255;
    if (bad) *not_binary = 1;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const uint32_t one = (e & 1) ? 0x3F800000u : 0x00003F80u;
      if (b[3 * e]) lock[e >> 1] |= one;
      if (b[3 * e + 1]) rot[e >> 1] |= one;
      if (b[3 * e + 2]) key[e >> 1] |= one;
    }
  } else if (p0 == D) {
    lock[0] = rot[0] = key[0] = 0x00003F80u;   // bias column (1.0), then zero pad
  }
  *reinterpret_cast<uint4*>(planes + (long long)i * ld + p0) = make_uint4(rot[0], rot[1], rot[2], rot[3]);
  *reinterpret_cast<uint4*>(planes + (long long)(B + i) * ld + p0) = make_uint4(lock[0], lock[1], lock[2], lock[3]);
  *reinterpret_cast<uint4*>(planes + (long long)(2 * B + i) * ld + p0) = make_uint4(key[0], key[1], key[2], key[3]);
}
void deinterleave_u8(const uint8_t* x, int B, int D, __nv_bfloat16* planes, long long ld, int* not_binary, cudaStream_t s) {
  const long long n = (long long)B * (ld / 8);
  deinterleave_u8_kernel<<<(int)((n + 255) / 256), 256, 0, s>>>(x, B, D, planes, ld, not_binary);
}

__global__ void __launch_bounds__(256) planar_u8_kernel(const uint8_t* __restrict__ lock, const uint8_t* __restrict__ rot,
                                                        const uint8_t* __restrict__ key, int B, int D,
                                                        __nv_bfloat16* planes, long long ld, int* not_binary) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)3 * B * ld) return;
  const int c = (int)(idx % ld);
  const long long r = idx / ld;
  const int plane = (int)(r / B), i = (int)(r % B);
  const uint8_t* src = plane == 0 ? rot : (plane == 1 ? lock : key);
  const uint8_t b = c < D ? src[(long long)i * D + c] : (uint8_t)(c == D ? 1 : 0);
  if (b > 1 && b != 255) *not_binary = 1;
  planes[idx] = __float2bfloat16(b ? 1.f : 0.f);
}
void planar_u8(const uint8_t* lock, const uint8_t* rot, const uint8_t* key, int B, int D, __nv_bfloat16* planes,
               long long ld, int* not_binary, cudaStream_t s) {
  const long long n = (long long)3 * B * ld;
  planar_u8_kernel<<<(int)((n + 255) / 256), 256, 0, s>>>(lock, rot, key, B, D, planes, ld, not_binary);
}

__global__ void __launch_bounds__(256) planes_to_interleaved_kernel(const __nv_bfloat16* __restrict__ planes, long long ld,
                                                                    int B, int D, float* x) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)B * D * 3) return;
  const int c = (int)(idx % 3);
  const long long t = idx / 3;
  const int p = (int)(t % D), i = (int)(t / D);
  const int plane = c == 0 ? 1 : (c == 1 ? 0 : 2);   // ch0 lock, ch1 rot, ch2 key  <-  rows [rot; lock; key]
  x[idx] = __bfloat162float(planes[((long long)plane * B + i) * ld + p]);
}
void planes_to_interleaved_f32(const __nv_bfloat16* planes, long long ld, int B, int D, float* x, cudaStream_t s) {
  const long long n = (long long)B * D * 3;
  planes_to_interleaved_kernel<<<(int)((n + 255) / 256), 256, 0, s>>>(planes, ld, B, D, x);
}

// ------------------------------------------------------------------ synthetic pair generator
// One CTA per pair, four phases, all on 1-bit-per-pixel bitmaps in shared memory (pixel p = y*side + x is bit
// p & 31 of word p >> 5):
//   1. scanline rasterisation of lock and key.  A shape is the union of 4 convex primitives -- two ellipses (the
//      body) and two thin rotated boxes (the limbs) -- scattered around a frame centre that is itself uniform over
//      the middle of the canvas, so every (primitive, row) is one x-interval found in closed form.  The constants
//      are fitted to the reference's data (overlap_micro.zip, 200 images of 200 x 200): filled area 710..7213 px,
//      mean 3402 (8.5 %); bounding box ~70..115 px; centroids anywhere in 35..160; 83 % of the canvas set in at
//      least one image.  tests/test_generator.py compares area, bounding-box and centroid quantiles, the occupancy
//      profile and the first layer's operand-stage density (0.94 / 0.96 on real batches) with the committed fixture.
//   2. rot = nearest-neighbour rotation of lock about ((W-1)/2,(H-1)/2) by a random angle, zero fill (the
//      arithmetic of tf.contrib.image.rotate, 8c/utils.py:337-338; third-party, parity unpinned).  The inverse
//      map is evaluated in Q14 fixed point on doubled coordinates so that the CPU check in tests/ reproduces it
//      bit for bit:  sx = (CS*(2x-W+1) - SN*(2y-W+1) + (W<<14)) >> 15,  sy = (SN*(2x-W+1) + CS*(2y-W+1) + (W<<14)) >> 15;
//   3. label = the "maximum possible overlap area between the two shapes" (README.md:19-27; the reference's dataset
//      generator is not in the repository), searched over rotations x translations of the key:
//        overlap(r, t) = popcount(lock AND T(key)),  T: lock pixel (x, y) reads key pixel
//          (cKx, cKy) +- ((cs X - sn Y + 8192) >> 14, (sn X + cs Y + 8192) >> 14),  X = x - cLx - tx, Y = y - cLy - ty
//      with cL / cK the rounded pixel centroids of lock and key ((2 sum + area) / (2 area)), (cs, sn) the Q14 constants
//      of r * 45 degrees for r = 0..3 and the minus sign (a point reflection through cK) giving r + 4.
//      The SEARCH runs on every 4th pixel in x and y (1/16 of the lock): first the 8 rotations with the centroids
//      aligned (t = 0), then the first best rotation at (+-s, 0) and (0, +-s), s = (3 W + 50) / 100 pixels; the label is
//      the EXACT overlap (all pixels) of the first best of those five candidates.  On the reference's 100 real pairs
//      this reaches 91 % of the stored OVERLAP_AREAS label on average (median 95 %; tests/test_shape_stats.py);
//      rotations about the canvas centre alone, the previous definition, reached 66 %.
//   4. the three bitmaps go out as they are: packed rows [rot; lock; key] of xb3 with the bias column D set -- the
//      first layer's GEMMs read them as their A operand (BitsA, gemm_sm100.cuh) and the BCE epilogue reads the lock
//      rows as its target.  With `planes` != nullptr (MAGIC_BITS_A=0) they are also expanded to bf16 operand planes
//      (1.0 in the bias column, zero pad) with 16-byte coalesced stores.
constexpr int kPrims = 4;
constexpr int kLabelRots = 8;
constexpr int kGenThreads = 384;
constexpr int kMaxBitWords = 2048;                       // canvas up to 256 x 256
// Q14 (cos, sin) of r * 45 degrees, r = 0..3; rotation r + 4 is DEFINED as the point reflection of rotation r's
// source pixel (through the key's centroid in the label search).
__host__ __device__ constexpr int rot_cs(int r) { return r == 0 ? 16384 : r == 1 ? 11585 : r == 2 ? 0 : -11585; }
__host__ __device__ constexpr int rot_sn(int r) { return r == 0 ? 0 : r == 1 ? 11585 : r == 2 ? 16384 : 11585; }

struct Prim { float cx, cy, a, b, cs, sn; int box; };

__device__ void make_prims(Prim* out, uint64_t key, float side) {
  const float c = 0.5f * (side - 1.f);
  const uint64_t rf = splitmix64(key + 63);
  const float fx = c + (u01((uint32_t)rf) - 0.5f) * 0.45f * side;           // frame centre: 55..145 of 200
  const float fy = c + (u01((uint32_t)(rf >> 32)) - 0.5f) * 0.45f * side;
  const float thick = 0.45f + 0.95f * u01((uint32_t)splitmix64(key + 62));   // per-shape thickness: spreads the area like the data's (1376 / 3378 / 5387 px at 10 / 50 / 90 %)
  for (int k = 0; k < kPrims; ++k) {
    const uint64_t r0 = splitmix64(key + 3 * k), r1 = splitmix64(key + 3 * k + 1), r2 = splitmix64(key + 3 * k + 2);
    Prim& q = out[k];
    q.cx = fx + (u01((uint32_t)r0) - 0.5f) * 0.30f * side;                   // +-30 around the frame centre
    q.cy = fy + (u01((uint32_t)(r0 >> 32)) - 0.5f) * 0.30f * side;
    q.box = k >= 2 ? 1 : 0;
    if (!q.box) {                                                             // body: ellipse, half-axes (5..36) x thickness
      q.a = (0.025f + 0.155f * u01((uint32_t)r1)) * side * thick;
      q.b = (0.025f + 0.155f * u01((uint32_t)(r1 >> 32))) * side * thick;
    } else {                                                                  // limb: half-length 22..52, half-width (3..9) x thickness
      q.a = (0.11f + 0.15f * u01((uint32_t)r1)) * side;
      q.b = (0.015f + 0.03f * u01((uint32_t)(r1 >> 32))) * side * thick;
    }
    sincospif(2.f * u01((uint32_t)r2), &q.sn, &q.cs);
  }
}

// x-interval [lo, hi] (relative to cx) of primitive q on the row at vertical offset dy; false if empty
__device__ __forceinline__ bool prim_row_interval(const Prim& q, float dy, float& lo, float& hi) {
  if (!q.box) {
    const float ia = 1.f / (q.a * q.a), ib = 1.f / (q.b * q.b);
    const float A = q.cs * q.cs * ia + q.sn * q.sn * ib;
    const float Bh = dy * q.cs * q.sn * (ia - ib);
    const float C = dy * dy * (q.sn * q.sn * ia + q.cs * q.cs * ib) - 1.f;
    const float disc = Bh * Bh - A * C;
    if (disc < 0.f) return false;
    const float sq = sqrtf(disc), inv = 1.f / A;
    lo = (-Bh - sq) * inv;
    hi = (-Bh + sq) * inv;
    return true;
  }
  lo = -1e9f; hi = 1e9f;
  // |cs dx + sn dy| <= a   and   |-sn dx + cs dy| <= b
  const float al[2] = {q.cs, -q.sn}, be[2] = {q.sn * dy, q.cs * dy}, ga[2] = {q.a, q.b};
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    if (fabsf(al[c]) < 1e-6f) {
      if (fabsf(be[c]) > ga[c]) return false;
    } else {
      const float i = 1.f / al[c];
      const float x0 = (-ga[c] - be[c]) * i, x1 = (ga[c] - be[c]) * i;
      lo = fmaxf(lo, fminf(x0, x1));
      hi = fminf(hi, fmaxf(x0, x1));
    }
  }
  return lo <= hi;
}

// OR the bit range [p0, p1] (inclusive, linear pixel indices) into a shared-memory bitmap
__device__ __forceinline__ void or_bit_range(uint32_t* bm, int p0, int p1) {
  const int w0 = p0 >> 5, w1 = p1 >> 5;
  const uint32_t m0 = 0xFFFFFFFFu << (p0 & 31), m1 = 0xFFFFFFFFu >> (31 - (p1 & 31));
  if (w0 == w1) { atomicOr(&bm[w0], m0 & m1); return; }
  atomicOr(&bm[w0], m0);
  for (int w = w0 + 1; w < w1; ++w) atomicOr(&bm[w], 0xFFFFFFFFu);
  atomicOr(&bm[w1], m1);
}

// linear index of the source pixel of (x, y) under the Q14 inverse rotation (cs, sn) about the canvas centre;
// -1 outside the canvas
__device__ __forceinline__ int rotated_index(int x, int y, int side, int cs, int sn) {
  const int X2 = 2 * x - side + 1, Y2 = 2 * y - side + 1;
  const int sx = (cs * X2 - sn * Y2 + (side << 14)) >> 15;
  const int sy = (sn * X2 + cs * Y2 + (side << 14)) >> 15;
  if ((unsigned)sx >= (unsigned)side || (unsigned)sy >= (unsigned)side) return -1;
  return sy * side + sx;
}
__device__ __forceinline__ uint32_t bit_at(const uint32_t* bm, int sp) {
  return sp < 0 ? 0u : (bm[sp >> 5] >> (sp & 31)) & 1u;
}

// The 32 pixels p0 .. p0 + 31 of the rotated image as one word, by ONE thread: the two Q15 source coordinates of
// rotated_index() are linear in x, so along a row they advance by (2 cs, 2 sn) per pixel -- two adds, two shifts, the
// bounds test and one shared-memory read per pixel, bit for bit the same map.  (One warp per word with a ballot spent
// ~100 warp instructions on 32 pixels, most of them warp-uniform bookkeeping: 44 % of the generator's instructions,
// profiles/r03q_generate_ncu_full_summary.csv.)  (x0, y0) = position of pixel p0; a word may wrap onto following rows.
__device__ __forceinline__ uint32_t rotated_word(const uint32_t* bm, int side, int D, int cs, int sn, int p0, int x0, int y0) {
  const int k0 = side << 14;
  int x = x0;
  int Y2 = 2 * y0 - side + 1;
  int X2 = 2 * x0 - side + 1;
  int ax = cs * X2 - sn * Y2 + k0, ay = sn * X2 + cs * Y2 + k0;
  const int n = min(32, D - p0);
  uint32_t word = 0u;
  for (int j = 0; j < n; ++j) {
    const int sx = ax >> 15, sy = ay >> 15;
    if ((unsigned)sx < (unsigned)side && (unsigned)sy < (unsigned)side) {
      const int sp = sy * side + sx;
      word |= ((bm[sp >> 5] >> (sp & 31)) & 1u) << j;
    }
    ax += 2 * cs; ay += 2 * sn;
    if (++x == side) {                                  // next row starts at x = 0
      x = 0; Y2 += 2; X2 = 1 - side;
      ax = cs * X2 - sn * Y2 + k0; ay = sn * X2 + cs * Y2 + k0;
    }
  }
  return word;
}

// nibble -> four bf16 values (two packed words): bit k of the nibble -> 1.0 (0x3F80) in element k
__device__ __forceinline__ uint2 expand_nibble(uint32_t n) {
  uint2 r;
  r.x = ((n & 1u) ? 0x00003F80u : 0u) | ((n & 2u) ? 0x3F800000u : 0u);
  r.y = ((n & 4u) ? 0x00003F80u : 0u) | ((n & 8u) ? 0x3F800000u : 0u);
  return r;
}
__device__ __forceinline__ uint4 expand_byte(uint32_t b) {
  const uint2 lo = expand_nibble(b & 15u), hi = expand_nibble((b >> 4) & 15u);
  return make_uint4(lo.x, lo.y, hi.x, hi.y);
}

__device__ uint4 g_expand_lut[256];           // byte of 8 pixels -> 8 bf16 (0.0 / 1.0), filled by kernels_init()
__global__ void fill_expand_lut_kernel() { g_expand_lut[threadIdx.x] = expand_byte(threadIdx.x); }

__global__ void __launch_bounds__(kGenThreads) generate_pairs_kernel(uint64_t seed, uint64_t step, int B, int side,
                                                                     uint32_t inv_side /* ceil(2^32 / side) */,
                                                                     __nv_bfloat16* planes, long long ld,
                                                                     uint32_t* xb3, int xb_ld, float* labels,
                                                                     int* rot_cs_sn) {
  const int D = side * side;
  const int nwords = (D + 31) >> 5;
  // dynamic shared memory, sized by the canvas so that a generator CTA fits on an SM next to a resident GEMM CTA:
  // [lock | key | rot bitmaps, (nwords + 8) words each] = 15 KiB at 200 x 200; the byte -> 8 bf16 table (4 KiB) is
  // read through L1 from global memory (g_expand_lut) to stay under what a 210 KiB GEMM CTA leaves free
  extern __shared__ uint4 gen_smem[];
  const uint4* __restrict__ lut = g_expand_lut;
  const int bm_pitch = (nwords + 8 + 3) & ~3;
  uint32_t* const bm_base = reinterpret_cast<uint32_t*>(gen_smem);
  uint32_t* const bm[3] = {bm_base, bm_base + bm_pitch, bm_base + 2 * bm_pitch};   // 0 lock, 1 key, 2 rot
  __shared__ Prim prims[2][kPrims];
  __shared__ int rot_q14[2];
  __shared__ int counts[kLabelRots + 5];      // label search: 8 rotations, 4 shifts of the best one, exact overlap of the winner
  __shared__ int rot_box[4];                   // x0, x1, y0, y1: every pixel of the rotated lock lies inside
  __shared__ int cent[6];                      // lock: area, sum x, sum y ; key: the same
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // persistent: the grid is at most one CTA per SM, so the generator shares every SM with the step's GEMM CTA
  // instead of displacing it
  for (int i = blockIdx.x; i < B; i += gridDim.x) {
  for (int w = threadIdx.x; w < nwords + 8; w += kGenThreads) bm[0][w] = bm[1][w] = bm[2][w] = 0u;
  if (threadIdx.x == 0) {
    make_prims(prims[0], rng_key(seed, step, 11, (uint64_t)i * 64), (float)side);
    float sn, cs;
    sincospif(2.f * u01((uint32_t)rng_key(seed, step, 13, i)), &sn, &cs);   // angle ~ U[0, 2 pi)
    rot_q14[0] = __float2int_rn(cs * 16384.f);
    rot_q14[1] = __float2int_rn(sn * 16384.f);
    if (rot_cs_sn) { rot_cs_sn[2 * i] = rot_q14[0]; rot_cs_sn[2 * i + 1] = rot_q14[1]; }
    // bounding box of the rotated lock: destination = R(-theta) source about the canvas centre, so a primitive centred
    // at c + p lands around c + (cs px + sn py, -sn px + cs py); nearest-neighbour resampling moves a pixel by < 1.5
    const float c = 0.5f * (side - 1.f);
    float bx0 = 1e9f, bx1 = -1e9f, by0 = 1e9f, by1 = -1e9f;
    for (int k = 0; k < kPrims; ++k) {
      const Prim& q = prims[0][k];
      const float ext = (q.box ? sqrtf(q.a * q.a + q.b * q.b) : fmaxf(q.a, q.b)) + 2.f;
      const float px = q.cx - c, py = q.cy - c;
      const float dx = c + (cs * px + sn * py), dy = c + (-sn * px + cs * py);
      bx0 = fminf(bx0, dx - ext); bx1 = fmaxf(bx1, dx + ext);
      by0 = fminf(by0, dy - ext); by1 = fmaxf(by1, dy + ext);
    }
    rot_box[0] = max(0, (int)floorf(bx0)); rot_box[1] = min(side - 1, (int)ceilf(bx1));
    rot_box[2] = max(0, (int)floorf(by0)); rot_box[3] = min(side - 1, (int)ceilf(by1));
  }
  if (threadIdx.x == 32) make_prims(prims[1], rng_key(seed, step, 12, (uint64_t)i * 64), (float)side);
  if (threadIdx.x < kLabelRots + 5) counts[threadIdx.x] = 0;
  if (threadIdx.x >= 64 && threadIdx.x < 70) cent[threadIdx.x - 64] = 0;
  __syncthreads();

  // ---- 1. scanline rasterisation: one work item per (shape, row)
  for (int it = threadIdx.x; it < 2 * side; it += kGenThreads) {
    const int s = it >= side ? 1 : 0, y = it - s * side;
#pragma unroll
    for (int k = 0; k < kPrims; ++k) {
      const Prim& q = prims[s][k];
      float lo, hi;
      if (!prim_row_interval(q, (float)y - q.cy, lo, hi)) continue;
      const int x0 = max(0, (int)ceilf(q.cx + lo)), x1 = min(side - 1, (int)floorf(q.cx + hi));
      if (x0 <= x1) or_bit_range(bm_base + s * bm_pitch, y * side + x0, y * side + x1);
    }
  }
  __syncthreads();

  // ---- 2. rotated lock: one thread per 32-pixel word (rotated_word).  Only words that reach into the bounding box of
  //         the rotated lock can receive pixels.
  {
    const int cs = rot_q14[0], sn = rot_q14[1];
    const int xlo = rot_box[0], xhi = rot_box[1], ylo = rot_box[2], yhi = rot_box[3];
    const int w_begin = (ylo * side) >> 5, w_end = min(nwords, ((yhi * side + side - 1) >> 5) + 1);
    for (int w = w_begin + (int)threadIdx.x; w < w_end; w += kGenThreads) {
      const int p0 = w * 32;
      const int y0 = (int)__umulhi((uint32_t)p0, inv_side), x0 = p0 - y0 * side;
      // a word lies on one row, or on the end of one and the start of the next (side >= 32; smaller canvases: every word)
      if (side >= 32 && (x0 + 31 < side ? (x0 > xhi || x0 + 31 < xlo) : (x0 > xhi && x0 + 31 - side < xlo))) continue;
      bm[2][w] = rotated_word(bm[0], side, D, cs, sn, p0, x0, y0);
    }
  }
  // ---- 3. label (see the header comment).  3a: area and coordinate sums of lock and key, one thread per word:
  //         sum_j j*bit_j of a word is a weighted popcount (five masks), a word straddles at most two rows.
  {
    int acc[6] = {0, 0, 0, 0, 0, 0};                       // lock: area, sum x, sum y ; key: the same
    for (int w = threadIdx.x; w < nwords; w += kGenThreads) {
      const int p0 = w * 32;
      const int y0 = (int)__umulhi((uint32_t)p0, inv_side), x0 = p0 - y0 * side;
      const int nfirst = min(32, side - x0);               // bits [0, nfirst) lie on row y0, the rest on row y0 + 1
      const uint32_t first = nfirst >= 32 ? 0xFFFFFFFFu : ((1u << nfirst) - 1u);
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        const uint32_t v = bm[s][w];
        if (v == 0u) continue;
        const int n = __popc(v), nhi = __popc(v & ~first);
        const int sj = __popc(v & 0xAAAAAAAAu) + 2 * __popc(v & 0xCCCCCCCCu) + 4 * __popc(v & 0xF0F0F0F0u) +
                       8 * __popc(v & 0xFF00FF00u) + 16 * __popc(v & 0xFFFF0000u);
        acc[3 * s] += n;
        acc[3 * s + 1] += n * x0 + sj - nhi * side;
        acc[3 * s + 2] += n * y0 + nhi;
      }
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const int v = __reduce_add_sync(0xffffffffu, acc[k]);
      if (lane == 0 && v) atomicAdd(&cent[k], v);
    }
  }
  __syncthreads();
  const int areaL = max(cent[0], 1), areaK = max(cent[3], 1);
  const int cLx = (2 * cent[1] + areaL) / (2 * areaL), cLy = (2 * cent[2] + areaL) / (2 * areaL);
  const int cKx = (2 * cent[4] + areaK) / (2 * areaK), cKy = (2 * cent[5] + areaK) / (2 * areaK);
  const int zbit = nwords << 5;                            // a bit of the zero padding behind every bitmap
  // 3b: search, on every 4th pixel of the lock in x and y: the 8 rotations with the centroids aligned.  Lane <-> one
  //     sample; its count is masked by its lock bit; out-of-canvas sources read the zero padding (branch-free).
  const int nsx = (side + 3) >> 2, nsamp = nsx * nsx;
  {
    int cnt[kLabelRots];
#pragma unroll
    for (int r = 0; r < kLabelRots; ++r) cnt[r] = 0;
    for (int q0 = warp * 32; q0 < nsamp; q0 += kGenThreads) {
      const int q = q0 + lane, qy = q / nsx, x = 4 * (q - qy * nsx), y = 4 * qy;
      const int p = y * side + x;
      const uint32_t lb = q < nsamp ? (bm[0][p >> 5] >> (p & 31)) & 1u : 0u;
      if (__ballot_sync(0xffffffffu, lb != 0u) == 0u) continue;
      const int X = x - cLx, Y = y - cLy;
#pragma unroll
      for (int r = 0; r < kLabelRots / 2; ++r) {
        const int da = (rot_cs(r) * X - rot_sn(r) * Y + 8192) >> 14, db = (rot_sn(r) * X + rot_cs(r) * Y + 8192) >> 14;
        const int ax = cKx + da, ay = cKy + db, bx = cKx - da, by = cKy - db;
        const int spa = ((unsigned)ax < (unsigned)side && (unsigned)ay < (unsigned)side) ? ay * side + ax : zbit;
        const int spb = ((unsigned)bx < (unsigned)side && (unsigned)by < (unsigned)side) ? by * side + bx : zbit;
        cnt[r] += (int)((bm[1][spa >> 5] >> (spa & 31)) & lb);
        cnt[r + 4] += (int)((bm[1][spb >> 5] >> (spb & 31)) & lb);
      }
    }
#pragma unroll
    for (int r = 0; r < kLabelRots; ++r) {
      const int v = __reduce_add_sync(0xffffffffu, cnt[r]);
      if (lane == 0 && v) atomicAdd(&counts[r], v);
    }
  }
  __syncthreads();
  // 3c: the first best rotation again, shifted by +-s along x and y (same samples)
  int rb = 0;
#pragma unroll
  for (int r = 1; r < kLabelRots; ++r) rb = counts[r] > counts[rb] ? r : rb;
  const int rr = rb & 3;
  const int bcs = rr == 0 ? rot_cs(0) : rr == 1 ? rot_cs(1) : rr == 2 ? rot_cs(2) : rot_cs(3);
  const int bsn = rr == 0 ? rot_sn(0) : rr == 1 ? rot_sn(1) : rr == 2 ? rot_sn(2) : rot_sn(3);
  const int sgn = rb >= 4 ? -1 : 1;
  const int sh = (3 * side + 50) / 100;
  {
    int cnt[4] = {0, 0, 0, 0};
    for (int q0 = warp * 32; q0 < nsamp; q0 += kGenThreads) {
      const int q = q0 + lane, qy = q / nsx, x = 4 * (q - qy * nsx), y = 4 * qy;
      const int p = y * side + x;
      const uint32_t lb = q < nsamp ? (bm[0][p >> 5] >> (p & 31)) & 1u : 0u;
      if (__ballot_sync(0xffffffffu, lb != 0u) == 0u) continue;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int tx = c == 0 ? sh : c == 1 ? -sh : 0, ty = c == 2 ? sh : c == 3 ? -sh : 0;
        const int X = x - cLx - tx, Y = y - cLy - ty;
        const int ax = cKx + sgn * ((bcs * X - bsn * Y + 8192) >> 14), ay = cKy + sgn * ((bsn * X + bcs * Y + 8192) >> 14);
        const int sp = ((unsigned)ax < (unsigned)side && (unsigned)ay < (unsigned)side) ? ay * side + ax : zbit;
        cnt[c] += (int)((bm[1][sp >> 5] >> (sp & 31)) & lb);
      }
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int v = __reduce_add_sync(0xffffffffu, cnt[c]);
      if (lane == 0 && v) atomicAdd(&counts[kLabelRots + c], v);
    }
  }
  __syncthreads();
  // 3d: the exact overlap (every lock pixel) of the first best of the five candidates (t = 0 first)
  {
    int tb = -1, best = counts[rb];
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if (counts[kLabelRots + c] > best) { best = counts[kLabelRots + c]; tb = c; }
    const int tx = tb == 0 ? sh : tb == 1 ? -sh : 0, ty = tb == 2 ? sh : tb == 3 ? -sh : 0;
    // one thread per lock word (the map is linear in x: u, v advance by (bcs, bsn) per pixel; same integers as 3b / 3c)
    int cnt = 0;
    for (int w = threadIdx.x; w < nwords; w += kGenThreads) {
      const uint32_t lw = bm[0][w];
      if (lw == 0u) continue;
      const int p0 = w * 32;
      const int y0 = (int)__umulhi((uint32_t)p0, inv_side);
      int x = p0 - y0 * side, Y = y0 - cLy - ty;
      int X = x - cLx - tx;
      int u = bcs * X - bsn * Y + 8192, v = bsn * X + bcs * Y + 8192;
      const int last = 31 - __clz(lw);                     // nothing to count behind the last set pixel
      for (int j = 0; j <= last; ++j) {
        const int ax = cKx + sgn * (u >> 14), ay = cKy + sgn * (v >> 14);
        const int sp = ((unsigned)ax < (unsigned)side && (unsigned)ay < (unsigned)side) ? ay * side + ax : zbit;
        cnt += (int)((bm[1][sp >> 5] >> (sp & 31)) & (lw >> j) & 1u);
        u += bcs; v += bsn;
        if (++x == side) {                                 // the word continues on the next row
          x = 0; ++Y; X = -cLx - tx;
          u = bcs * X - bsn * Y + 8192; v = bsn * X + bcs * Y + 8192;
        }
      }
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if (lane == 0 && cnt) atomicAdd(&counts[kLabelRots + 4], cnt);
  }
  __syncthreads();
  if (threadIdx.x == 0) labels[i] = (float)counts[kLabelRots + 4];

  // ---- 4. packed rows [rot; lock; key] with the bias bit (D % 32 == 0: the bitmaps end on a word boundary)
  for (int w = threadIdx.x; w < xb_ld; w += kGenThreads) {
    const uint32_t bias = w == (D >> 5) ? 1u << (D & 31) : 0u;
    const bool in = w < nwords;
    xb3[(long long)i * xb_ld + w] = (in ? bm[2][w] : 0u) | bias;
    xb3[(long long)(B + i) * xb_ld + w] = (in ? bm[0][w] : 0u) | bias;
    xb3[(long long)(2 * B + i) * xb_ld + w] = (in ? bm[1][w] : 0u) | bias;
  }
  // ---- 4b. (optional) bf16 operand planes: 8 pixels (one bitmap byte) -> one 16-byte store per plane
  if (planes) {
  const int chunks = (int)(ld / 8);
  const uint8_t* bl = reinterpret_cast<const uint8_t*>(bm[0]);
  const uint8_t* bk = reinterpret_cast<const uint8_t*>(bm[1]);
  const uint8_t* br = reinterpret_cast<const uint8_t*>(bm[2]);
  uint4* prot = reinterpret_cast<uint4*>(planes + (long long)i * ld);
  uint4* plock = reinterpret_cast<uint4*>(planes + (long long)(B + i) * ld);
  uint4* pkey = reinterpret_cast<uint4*>(planes + (long long)(2 * B + i) * ld);
  const int data_chunks = D / 8;               // D % 8 == 0 (host checks)
  for (int ch = threadIdx.x; ch < chunks; ch += kGenThreads) {
    uint4 vl, vr, vk;
    if (ch < data_chunks) {
      vl = __ldg(lut + bl[ch]); vr = __ldg(lut + br[ch]); vk = __ldg(lut + bk[ch]);
    } else {
      vl = vr = vk = make_uint4(ch == data_chunks ? 0x00003F80u : 0u, 0u, 0u, 0u);   // bias column (1.0), then zero pad
    }
    prot[ch] = vr; plock[ch] = vl; pkey[ch] = vk;
  }
  }
  __syncthreads();                             // the bitmaps are reused by the next pair
  }
}

void generate_pairs(uint64_t seed, uint64_t step, int B, int side, __nv_bfloat16* planes, long long ld, uint32_t* xb3,
                    int xb_ld, float* labels, int* rot_cs_sn, int max_ctas, cudaStream_t s) {
  const uint32_t inv_side = (uint32_t)((0x100000000ull + (uint64_t)side - 1) / (uint64_t)side);
  // same (maximum) shared-memory carveout as the GEMM kernels: an SM only changes its carveout when it is idle, so a
  // different preference here would make every overlapped GEMM CTA wait for the generator to drain
  const int nwords = (side * side + 31) / 32;
  const size_t smem = 3 * (size_t)((nwords + 8 + 3) & ~3) * sizeof(uint32_t);
  generate_pairs_kernel<<<B < max_ctas ? B : max_ctas, kGenThreads, smem, s>>>(seed, step, B, side, inv_side, planes, ld, xb3, xb_ld, labels,
                                                     rot_cs_sn);
}

// ------------------------------------------------------------------ N2: dataset pairs -> operand planes
// (8c/overlap_input.py:181-195: lock, random_rotation(lock), key concatenated; 8c/utils.py:329-340: the angle is
// uniform in [0, 2 pi), nearest-neighbour, zero fill.)  lock / key are uint8 [B, D] (any non-zero byte is a set
// pixel); the rotated plane is sampled on the fly with the generator's Q14 integer rotation, so it is reproducible
// bit for bit from the exported (cos, sin).  One CTA per (pair, 1024-column chunk).
__global__ void __launch_bounds__(256) pairs_u8_kernel(const uint8_t* __restrict__ lock, const uint8_t* __restrict__ key,
                                                       int B, int side, uint64_t seed, uint64_t step,
                                                       __nv_bfloat16* planes, long long ld, int* rot_cs_sn) {
  __shared__ int q14[2];
  const int i = blockIdx.y, D = side * side;
  if (threadIdx.x == 0) {
    float sn, cs;
    sincospif(2.f * u01((uint32_t)rng_key(seed, step, 13, i)), &sn, &cs);   // same draw as generate_pairs_kernel
    q14[0] = __float2int_rn(cs * 16384.f);
    q14[1] = __float2int_rn(sn * 16384.f);
    if (rot_cs_sn && blockIdx.x == 0) { rot_cs_sn[2 * i] = q14[0]; rot_cs_sn[2 * i + 1] = q14[1]; }
  }
  __syncthreads();
  const int cs = q14[0], sn = q14[1];
  const uint8_t* L = lock + (long long)i * D;
  const uint8_t* K = key + (long long)i * D;
  __nv_bfloat16* prot = planes + (long long)i * ld;
  __nv_bfloat16* plock = planes + (long long)(B + i) * ld;
  __nv_bfloat16* pkey = planes + (long long)(2 * B + i) * ld;
  const __nv_bfloat16 one = __float2bfloat16(1.f), zero = __float2bfloat16(0.f);
  for (int c = blockIdx.x * 1024 + threadIdx.x; c < min((long long)(blockIdx.x + 1) * 1024, ld); c += 256) {
    __nv_bfloat16 vl = zero, vk = zero, vr = zero;
    if (c < D) {
      vl = L[c] ? one : zero;
      vk = K[c] ? one : zero;
      const int y = c / side, x = c - y * side;
      const int sp = rotated_index(x, y, side, cs, sn);
      vr = (sp >= 0 && L[sp]) ? one : zero;
    } else if (c == D) {
      vl = vk = vr = one;                                  // bias column
    }
    prot[c] = vr; plock[c] = vl; pkey[c] = vk;
  }
}
void pairs_u8(const uint8_t* lock, const uint8_t* key, int B, int side, uint64_t seed, uint64_t step,
              __nv_bfloat16* planes, long long ld, int* rot_cs_sn, cudaStream_t s) {
  pairs_u8_kernel<<<dim3((unsigned)((ld + 1023) / 1024), B), 256, 0, s>>>(lock, key, B, side, seed, step, planes, ld, rot_cs_sn);
}

// The same pipeline fed with bit-packed planes: lock / key are [B, wpr] 32-bit words, pixel p = bit (p & 31) of word
// p >> 5 (numpy packbits(bitorder='little')): 10 KB per pair instead of 80 KB as uint8 or 480 KB as the float32
// array the reference feeds.  Persistent CTAs, one pair at a time: the lock's bitmap goes to shared memory, lock and
// key words pass straight through to their packed rows of xb3 (tail bits masked, bias column set), and each warp
// assembles 32 rotated pixels into one word with a ballot -- the generator's phase 2; only words that reach into the
// bounding box of the rotated lock are visited (the lock's own box is measured from the bitmap while it is loaded and
// its corners are rotated).  Output is the packed rows only (the GEMMs' A operand and the BCE
// target).
// Same footprint as the generator (384 threads, at most 56 registers, one CTA per SM): three of its warps fit beside
// the GEMM's warps in a sub-partition's 16 K registers, so it runs UNDER the step instead of making GEMM CTAs wait for
// a persistent grid to drain (256 threads x 2 CTAs per SM did exactly that: 0.82 ms per batch, e2e 2.41 -> 2.58 ms).
constexpr int kPairsBitsThreads = kGenThreads;
__global__ void __maxnreg__(56) pairs_bits_kernel(const uint32_t* __restrict__ lock,
                                                                       const uint32_t* __restrict__ key, int wpr, int B,
                                                                       int side, uint32_t inv_side, uint64_t seed,
                                                                       uint64_t step, uint32_t* xb3, int xb_ld,
                                                                       int* rot_cs_sn) {
  extern __shared__ uint32_t bm[];              // the lock's bitmap: nwords words (5 KB at 200 x 200)
  __shared__ int q14[2];
  __shared__ int box[4];                        // the lock's bounding box: min x, max x, min y, max y
  __shared__ int rbox[4];                       // ... of the rotated lock: x0, x1, y0, y1
  const int D = side * side;
  const int nwords = (D + 31) >> 5;
  const uint32_t tail = (D & 31) ? ((1u << (D & 31)) - 1u) : 0xffffffffu;     // bits past D in the last word are ignored
  const int bias_w = D >> 5;
  const uint32_t bias_b = 1u << (D & 31);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = blockIdx.x; i < B; i += gridDim.x) {
    const uint32_t* L = lock + (long long)i * wpr;
    const uint32_t* K = key + (long long)i * wpr;
    uint32_t* orot = xb3 + (long long)i * xb_ld;
    uint32_t* olock = xb3 + (long long)(B + i) * xb_ld;
    uint32_t* okey = xb3 + (long long)(2 * B + i) * xb_ld;
    if (threadIdx.x == 0) {
      float sn, cs;
      sincospif(2.f * u01((uint32_t)rng_key(seed, step, 13, i)), &sn, &cs);   // same draw as generate_pairs_kernel
      q14[0] = __float2int_rn(cs * 16384.f);
      q14[1] = __float2int_rn(sn * 16384.f);
      if (rot_cs_sn) { rot_cs_sn[2 * i] = q14[0]; rot_cs_sn[2 * i + 1] = q14[1]; }
      box[0] = side; box[1] = -1; box[2] = side; box[3] = -1;
    }
    __syncthreads();
    int bx0 = side, bx1 = -1, by0 = side, by1 = -1;
    auto far = [&](int p) {
      const int y = (int)__umulhi((uint32_t)p, inv_side), x = p - y * side;
      bx0 = min(bx0, x); bx1 = max(bx1, x); by0 = min(by0, y); by1 = max(by1, y);
    };
    for (int w = threadIdx.x; w < xb_ld; w += kPairsBitsThreads) {
      const uint32_t m = w == nwords - 1 ? tail : 0xffffffffu;
      const uint32_t bias = w == bias_w ? bias_b : 0u;
      const uint32_t vl = w < nwords ? __ldg(L + w) & m : 0u;
      const uint32_t vk = w < nwords ? __ldg(K + w) & m : 0u;
      if (w < nwords) bm[w] = vl;
      olock[w] = vl | bias;
      okey[w] = vk | bias;
      orot[w] = bias;                           // the words inside the rotated lock's box are overwritten below
      if (vl) {
        const int p0 = w * 32;
        const int y0 = (int)__umulhi((uint32_t)p0, inv_side);
        if (p0 - y0 * side + 31 < side) {       // one canvas row: its extreme set pixels are the first and the last one
          far(p0 + __ffs(vl) - 1);
          far(p0 + 31 - __clz(vl));
        } else {                                // the word wraps onto the next row(s): look at every set pixel
          for (uint32_t v = vl; v; v &= v - 1) far(p0 + __ffs(v) - 1);
        }
      }
    }
    if (bx1 >= 0) { atomicMin(&box[0], bx0); atomicMax(&box[1], bx1); atomicMin(&box[2], by0); atomicMax(&box[3], by1); }
    __syncthreads();
    const int cs = q14[0], sn = q14[1];
    if (threadIdx.x == 0) {
      // destination = R(-theta) source about the canvas centre: rotate the four corners of the lock's box;
      // nearest-neighbour rounding moves a sample by less than 1.5 pixels
      const float c = 0.5f * (side - 1.f), fc = cs * (1.f / 16384.f), fs = sn * (1.f / 16384.f);
      float x0 = 1e9f, x1 = -1e9f, y0 = 1e9f, y1 = -1e9f;
      for (int k = 0; k < 4; ++k) {
        const float px = (float)box[k & 1] - c, py = (float)box[2 + (k >> 1)] - c;
        const float dx = c + (fc * px + fs * py), dy = c + (-fs * px + fc * py);
        x0 = fminf(x0, dx); x1 = fmaxf(x1, dx); y0 = fminf(y0, dy); y1 = fmaxf(y1, dy);
      }
      const bool empty = box[1] < 0;
      rbox[0] = empty ? side : max(0, (int)floorf(x0 - 2.f)); rbox[1] = empty ? -1 : min(side - 1, (int)ceilf(x1 + 2.f));
      rbox[2] = empty ? side : max(0, (int)floorf(y0 - 2.f)); rbox[3] = empty ? -1 : min(side - 1, (int)ceilf(y1 + 2.f));
    }
    __syncthreads();
    const int xlo = rbox[0], xhi = rbox[1], lo = rbox[2], hi = rbox[3];
    const int w_begin = (max(lo, 0) * side) >> 5, w_end = hi < 0 ? 0 : min(nwords, ((hi * side + side - 1) >> 5) + 1);
    for (int w = w_begin + (int)threadIdx.x; w < w_end; w += kPairsBitsThreads) {     // one thread per word (rotated_word)
      const int p0 = w * 32;
      const int y0 = (int)__umulhi((uint32_t)p0, inv_side);
      const int x0 = p0 - y0 * side;
      // a word lies on one row, or on the end of one and the start of the next (side >= 32; smaller canvases: every word)
      if (side >= 32 && (x0 + 31 < side ? (x0 > xhi || x0 + 31 < xlo) : (x0 > xhi && x0 + 31 - side < xlo))) continue;
      orot[w] = rotated_word(bm, side, D, cs, sn, p0, x0, y0) | (w == bias_w ? bias_b : 0u);
    }
    __syncthreads();                            // the bitmap is reused by the next pair
  }
}
void pairs_bits(const uint32_t* lock, const uint32_t* key, int wpr, int B, int side, uint64_t seed, uint64_t step,
                __nv_bfloat16* planes, long long ld, uint32_t* xb3, int xb_ld, int* rot_cs_sn, int max_ctas, cudaStream_t s) {
  const uint32_t inv_side = (uint32_t)((0x100000000ull + (uint64_t)side - 1) / (uint64_t)side);
  pairs_bits_kernel<<<B < max_ctas ? B : max_ctas, kPairsBitsThreads, sizeof(uint32_t) * (size_t)((side * side + 31) / 32), s>>>(lock, key, wpr, B, side, inv_side, seed, step, xb3,
                                                                              xb_ld, rot_cs_sn);
  if (planes) unpack_bits_planes(xb3, xb_ld, 3 * B, planes, ld, s);     // MAGIC_BITS_A=0: the bf16 planes as well
}

// bf16 {0,1} operand planes [rows, ld] -> packed rows xb[rows, xb_ld] with the bias column D set, for the loaders
// that produce planes (array feeds).  One warp per 32-pixel word.
__global__ void __launch_bounds__(256) pack_planes_bits_kernel(const __nv_bfloat16* __restrict__ planes, long long ld, int rows,
                                                               int D, uint32_t* xb, int xb_ld) {
  const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (gw >= (long long)rows * xb_ld) return;
  const int i = (int)(gw / xb_ld), w = (int)(gw % xb_ld);
  const int p = w * 32 + lane;
  const bool bit = p < D ? __bfloat162float(planes[(long long)i * ld + p]) != 0.f : p == D;
  const uint32_t word = __ballot_sync(0xffffffffu, bit);
  if (lane == 0) xb[gw] = word;
}
void pack_planes_bits(const __nv_bfloat16* planes, long long ld, int rows, int D, uint32_t* xb, int xb_ld, cudaStream_t s) {
  const long long threads = (long long)rows * xb_ld * 32;
  pack_planes_bits_kernel<<<(int)((threads + 255) / 256), 256, 0, s>>>(planes, ld, rows, D, xb, xb_ld);
}
// packed rows -> bf16 operand planes (export of a batch that was only ever held as bits): 8 pixels per thread
__global__ void __launch_bounds__(256) unpack_bits_planes_kernel(const uint32_t* __restrict__ xb, int xb_ld, int rows,
                                                                 __nv_bfloat16* planes, long long ld) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int chunks = (int)(ld >> 3);
  if (idx >= (long long)rows * chunks) return;
  const int i = (int)(idx / chunks), ch = (int)(idx % chunks);
  const uint32_t b = reinterpret_cast<const uint8_t*>(xb + (long long)i * xb_ld)[ch];
  reinterpret_cast<uint4*>(planes + (long long)i * ld)[ch] = expand_byte(b);
}
void unpack_bits_planes(const uint32_t* xb, int xb_ld, int rows, __nv_bfloat16* planes, long long ld, cudaStream_t s) {
  const long long n = (long long)rows * (ld >> 3);
  unpack_bits_planes_kernel<<<(int)((n + 255) / 256), 256, 0, s>>>(xb, xb_ld, rows, planes, ld);
}


// ------------------------------------------------------------------ zero-stage maps of the packed input rows
// (GemmArgs.kmask, gemm_sm100.cuh.)  Forward GEMM: M tile t = packed rows [128 t, 128 t + 128), k-block kk = pixels
// [64 kk, 64 kk + 64) = words 2 kk, 2 kk + 1; lane <-> k-block, so a warp reads 256 contiguous bytes per row.
__global__ void __launch_bounds__(256) kmask_fwd_kernel(const uint32_t* __restrict__ xb, int xb_ld, int rows, int kblocks,
                                                        uint32_t* mask, int mwords, int tiles) {
  const int gw = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (gw >= tiles * mwords) return;
  const int t = gw / mwords, j = gw % mwords;
  const int kk = 32 * j + lane;
  uint32_t acc = 0u;
  if (kk < kblocks) {
    const int r1 = min(rows, t * 128 + 128);
#pragma unroll 8
    for (int r = t * 128; r < r1; ++r) {
      const uint2 v = __ldg(reinterpret_cast<const uint2*>(xb + (long long)r * xb_ld + 2 * kk));
      acc |= v.x | v.y;
    }
  }
  const uint32_t word = __ballot_sync(0xffffffffu, acc != 0u);
  if (lane == 0) mask[(long long)t * mwords + j] = word;
}
// Weight gradient (A = X^T): M tile tm = pixels [128 tm, 128 tm + 128) = words 4 tm .. 4 tm + 3, k-block kb = rows
// [64 kb, 64 kb + 64) of the group's row range.  lane <-> M tile (512 contiguous bytes per row), one warp per
// (32 tiles, k-block); every lane ORs its bit into its own tile's map word (the map is cleared first).
__global__ void __launch_bounds__(256) kmask_wg_kernel(const uint32_t* __restrict__ xb, int xb_ld, int krows, int tiles_m,
                                                       uint32_t* mask, int mwords, int kblocks) {
  const int gw = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  const int chunks = (tiles_m + 31) >> 5;
  if (gw >= chunks * kblocks) return;
  const int kb = gw / chunks, tm = (gw % chunks) * 32 + lane;
  if (tm >= tiles_m) return;
  uint32_t acc = 0u;
  const int r1 = min(krows, kb * 64 + 64);
#pragma unroll 8
  for (int r = kb * 64; r < r1; ++r) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(xb + (long long)r * xb_ld + 4 * tm));
    acc |= v.x | v.y | v.z | v.w;
  }
  if (acc) atomicOr(&mask[(long long)tm * mwords + (kb >> 5)], 1u << (kb & 31));
}
void build_kmask_fwd(const uint32_t* xb, int xb_ld, int rows, int kblocks, uint32_t* mask, int mwords, cudaStream_t s) {
  const int tiles = (rows + 127) / 128;
  const long long threads = (long long)tiles * mwords * 32;
  kmask_fwd_kernel<<<(int)((threads + 255) / 256), 256, 0, s>>>(xb, xb_ld, rows, kblocks, mask, mwords, tiles);
}
void build_kmask_wgrad(const uint32_t* xb, int xb_ld, int krows, int tiles_m, uint32_t* mask, int mwords, cudaStream_t s) {
  const int kblocks = (krows + 63) / 64;
  zero_fill(mask, sizeof(uint32_t) * (size_t)tiles_m * mwords, s);
  const long long threads = (long long)((tiles_m + 31) / 32) * kblocks * 32;
  kmask_wg_kernel<<<(int)((threads + 255) / 256), 256, 0, s>>>(xb, xb_ld, krows, tiles_m, mask, mwords, kblocks);
}

// An SM changes its shared-memory carveout only when it is idle.  The GEMM kernels need the maximum carveout and the
// generator shares SMs with them from its own stream, so every kernel of the library asks for the same carveout:
// no launch ever waits for co-resident CTAs of the other stream to drain.
#define MAGIC_MAX_CARVEOUT(k) \
  cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared)
void kernels_init() {
  MAGIC_MAX_CARVEOUT(zero_kernel);
  MAGIC_MAX_CARVEOUT(zero3_kernel);
  MAGIC_MAX_CARVEOUT(latent_forward_kernel);
  MAGIC_MAX_CARVEOUT(latent_metric_kernel);
  MAGIC_MAX_CARVEOUT(latent_backward_kernel);
  MAGIC_MAX_CARVEOUT(peer_allreduce_small_kernel);
  MAGIC_MAX_CARVEOUT(gather_rows_host_kernel);   // (it lives beside GEMM CTAs for most of a step: same carveout or the SM serialises them)
  MAGIC_MAX_CARVEOUT(adam_dual_kernel);
  MAGIC_MAX_CARVEOUT(adam_dual_sharded_kernel);
  MAGIC_MAX_CARVEOUT(push_range_kernel);
  MAGIC_MAX_CARVEOUT(mask_unowned_kernel);
  MAGIC_MAX_CARVEOUT(shadow_refresh_kernel);
  MAGIC_MAX_CARVEOUT(deinterleave_kernel);
  MAGIC_MAX_CARVEOUT(deinterleave_u8_kernel);
  MAGIC_MAX_CARVEOUT(planar_u8_kernel);
  MAGIC_MAX_CARVEOUT(pairs_u8_kernel);
  MAGIC_MAX_CARVEOUT(pairs_bits_kernel);
  MAGIC_MAX_CARVEOUT(planes_to_interleaved_kernel);
  MAGIC_MAX_CARVEOUT(generate_pairs_kernel);
  MAGIC_MAX_CARVEOUT(pack_planes_bits_kernel);
  MAGIC_MAX_CARVEOUT(unpack_bits_planes_kernel);
  MAGIC_MAX_CARVEOUT(kmask_fwd_kernel);
  MAGIC_MAX_CARVEOUT(kmask_wg_kernel);
  fill_expand_lut_kernel<<<1, 256>>>();
  cudaDeviceSynchronize();      // init time only: the generator reads the table from its own stream
  cudaGetLastError();
}

}  // namespace magic
