// Non-GEMM kernels of the step (see kernels.h).  All are HBM/latency-bound elementwise or reduction kernels.
#include "kernels.h"
#include "ptx.cuh"
#include <math.h>

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

// ------------------------------------------------------------------ latent forward
// One warp per pair; lanes stride over latent columns.  (8c/vae.py:358-371, 275-288)
constexpr int kLatWarps = 8;
constexpr int kLatRowsPerWarp = 4;

__global__ void __launch_bounds__(kLatWarps * 32) latent_forward_kernel(LatentArgs a) {
  extern __shared__ float sm[];   // [2, L] column sums of squares
  const int L = a.L, B = a.B;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool cosine = a.metric == METRIC_COSINE;
  if (cosine) {
    for (int j = threadIdx.x; j < 2 * L; j += blockDim.x) sm[j] = 0.f;
    __syncthreads();
  }
  float kl = 0.f, def = 0.f;
  for (int rr = 0; rr < kLatRowsPerWarp; ++rr) {
    const int i = (blockIdx.x * kLatWarps + warp) * kLatRowsPerWarp + rr;
    if (i >= B) break;
    const float* h_rot = a.heads + (long long)i * a.ld_heads;
    const float* h_lock = a.heads + (long long)(B + i) * a.ld_heads;
    const float* h_key = a.heads + (long long)(2 * B + i) * a.ld_heads;
    for (int j = lane; j < L; j += 32) {
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
      kl += -0.5f * (1.f + lv - mu * mu - ex);
      const float d = zl - zr;
      def += d * d;
      if (cosine) {
        atomicAdd(&sm[j], zl * zl);
        atomicAdd(&sm[L + j], zk * zk);
      }
    }
    if (lane == 0) store_split(a.zdec, a.zdec_lo, (long long)i * a.ld_zdec + L, 1.0f);   // bias column
  }
  kl = warp_sum(kl);
  def = warp_sum(def);
  if (lane == 0) {
    atomicAdd(&a.sums[1], kl * a.inv_bg);
    atomicAdd(&a.sums[2], def * a.def_weight * a.inv_bg);
  }
  if (cosine) {
    __syncthreads();
    for (int j = threadIdx.x; j < 2 * L; j += blockDim.x) atomicAdd(&a.colsq[j], sm[j]);
  }
}

// ------------------------------------------------------------------ metric loss (8c/vae.py:300-307, 438-452)
__global__ void __launch_bounds__(kLatWarps * 32) latent_metric_kernel(LatentArgs a) {
  extern __shared__ float sm[];   // [L] column sums S_j (cosine)
  const int L = a.L, B = a.B;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool cosine = a.metric == METRIC_COSINE;
  if (cosine) {
    for (int j = threadIdx.x; j < L; j += blockDim.x) sm[j] = 0.f;
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
        const float ra = rsqrtf(fmaxf(a.colsq[j], 1e-12f)), rb = rsqrtf(fmaxf(a.colsq[L + j], 1e-12f));
        dist += (zl[j] * ra) * (zk[j] * rb);
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
      for (int j = lane; j < L; j += 32) {
        const float ra = rsqrtf(fmaxf(a.colsq[j], 1e-12f)), rb = rsqrtf(fmaxf(a.colsq[L + j], 1e-12f));
        atomicAdd(&sm[j], gi * (zl[j] * ra) * (zk[j] * rb));
      }
    }
  }
  if (lane == 0 && tl != 0.f) atomicAdd(&a.sums[3], tl * a.inv_bg);
  if (cosine) {
    __syncthreads();
    for (int j = threadIdx.x; j < L; j += blockDim.x) atomicAdd(&a.colS[j], sm[j]);
  }
}

// ------------------------------------------------------------------ latent backward
// dheads rows: [rot|cost ; lock|cost ; lock|metric ; key|metric], cols [d mu | d log sigma^2].
// d z / d mu = 1 ;  d z / d lv = 0.5 * sqrt(exp(lv)) * eps = 0.5 * (z - mu).
__global__ void __launch_bounds__(256) latent_backward_kernel(LatentArgs a) {
  const int L = a.L, B = a.B;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)B * L) return;
  const int i = (int)(idx / L), j = (int)(idx % L);
  const float* h_rot = a.heads + (long long)i * a.ld_heads;
  const float* h_lock = a.heads + (long long)(B + i) * a.ld_heads;
  const float* h_key = a.heads + (long long)(2 * B + i) * a.ld_heads;
  const float zr = a.z[(long long)i * L + j], zl = a.z[(long long)(B + i) * L + j], zk = a.z[(long long)(2 * B + i) * L + j];
  const float mu = h_lock[j], lv = h_lock[L + j];
  // --- optimiser 1: cost
  const float ddef = 2.f * a.def_weight * (zl - zr) * a.inv_bg;
  const float dz_l1 = a.dz_dec[(long long)i * a.ld_dz + j] * a.inv_bg + ddef;
  const float dmu_l1 = dz_l1 + mu * a.inv_bg;
  const float dlv_l1 = dz_l1 * 0.5f * (zl - mu) + 0.5f * (expf(lv) - 1.f) * a.inv_bg;
  const float dz_r1 = -ddef;
  const float dlv_r1 = dz_r1 * 0.5f * (zr - h_rot[j]);
  // --- optimiser 2: metric
  float da, db;
  const float gi = a.g[i];
  if (a.metric == METRIC_COSINE) {
    const float sa = a.colsq[j], sb = a.colsq[L + j];
    const float ra = rsqrtf(fmaxf(sa, 1e-12f)), rb = rsqrtf(fmaxf(sb, 1e-12f));
    const float ah = zl * ra, bh = zk * rb;
    const float S = a.colS[j];
    // through x * rsqrt(max(sum x^2, eps)): the normalisation term vanishes when the clamp is active
    da = (gi * bh - (sa > 1e-12f ? ah * S : 0.f)) * ra;
    db = (gi * ah - (sb > 1e-12f ? bh * S : 0.f)) * rb;
  } else {
    da = gi * 2.f * (zl - zk);
    db = -da;
  }
  const float dlv_l2 = da * 0.5f * (zl - mu);
  const float dlv_k2 = db * 0.5f * (zk - h_key[j]);
  const long long ld = a.ld_dheads;
  store_split(a.dheads, a.dheads_lo, (long long)i * ld + j, dz_r1);
  store_split(a.dheads, a.dheads_lo, (long long)i * ld + L + j, dlv_r1);
  store_split(a.dheads, a.dheads_lo, (long long)(B + i) * ld + j, dmu_l1);
  store_split(a.dheads, a.dheads_lo, (long long)(B + i) * ld + L + j, dlv_l1);
  store_split(a.dheads, a.dheads_lo, (long long)(2 * B + i) * ld + j, da);
  store_split(a.dheads, a.dheads_lo, (long long)(2 * B + i) * ld + L + j, dlv_l2);
  store_split(a.dheads, a.dheads_lo, (long long)(3 * B + i) * ld + j, db);
  store_split(a.dheads, a.dheads_lo, (long long)(3 * B + i) * ld + L + j, dlv_k2);
}

static int lat_blocks(int B) { return (B + kLatWarps * kLatRowsPerWarp - 1) / (kLatWarps * kLatRowsPerWarp); }

void latent_forward(const LatentArgs& a, cudaStream_t s) {
  latent_forward_kernel<<<lat_blocks(a.B), kLatWarps * 32, sizeof(float) * 2 * a.L, s>>>(a);
}
void latent_metric(const LatentArgs& a, cudaStream_t s) {
  latent_metric_kernel<<<lat_blocks(a.B), kLatWarps * 32, sizeof(float) * a.L, s>>>(a);
}
void latent_backward(const LatentArgs& a, cudaStream_t s) {
  const long long n = (long long)a.B * a.L;
  latent_backward_kernel<<<(int)((n + 255) / 256), 256, 0, s>>>(a);
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
  const long long i4 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i4 >= a.n_total) return;
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

void adam_dual(const AdamArgs& a, cudaStream_t s) {
  const long long n4 = (a.n_total + 3) / 4;
  adam_dual_kernel<<<(int)((n4 + 255) / 256), 256, 0, s>>>(a);
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
                                                           __nv_bfloat16* planes, long long ld) {
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
void deinterleave_f32(const float* x, int B, int D, __nv_bfloat16* planes, long long ld, cudaStream_t s) {
  const long long n = (long long)B * (ld / 8);
  deinterleave_kernel<<<(int)((n + 255) / 256), 256, 0, s>>>(x, B, D, planes, ld);
}

__global__ void __launch_bounds__(256) planar_u8_kernel(const uint8_t* __restrict__ lock, const uint8_t* __restrict__ rot,
                                                        const uint8_t* __restrict__ key, int B, int D,
                                                        __nv_bfloat16* planes, long long ld) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)3 * B * ld) return;
  const int c = (int)(idx % ld);
  const long long r = idx / ld;
  const int plane = (int)(r / B), i = (int)(r % B);
  const uint8_t* src = plane == 0 ? rot : (plane == 1 ? lock : key);
  const float v = c < D ? (float)src[(long long)i * D + c] : (c == D ? 1.f : 0.f);
  planes[idx] = __float2bfloat16(v);
}
void planar_u8(const uint8_t* lock, const uint8_t* rot, const uint8_t* key, int B, int D, __nv_bfloat16* planes,
               long long ld, cudaStream_t s) {
  const long long n = (long long)3 * B * ld;
  planar_u8_kernel<<<(int)((n + 255) / 256), 256, 0, s>>>(lock, rot, key, B, D, planes, ld);
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
// A shape is the union of 3 analytic primitives (ellipse or rotated box) near the canvas centre; mean filled
// area ~8.5 % of the canvas like the reference's data (overlap_micro.zip: 710..7213 px of 40000).
// rot  = nearest-neighbour rotation of lock about ((W-1)/2,(H-1)/2), zero fill  (tf.contrib.image.rotate,
//        8c/utils.py:337-338; third-party arithmetic, parity unpinned)
// label = max over 8 rotations of the key about the centre of |lock AND rot(key)|  (README.md:19-27: the
//        "maximum overlap area"; the reference's dataset generator is not in the repository).
constexpr int kPrims = 3;
constexpr int kLabelRots = 8;
// A primitive as an affine map to its unit frame: u = ax*x + ay*y + a0 ; v = bx*x + by*y + b0 ;
// ellipse: u^2 + v^2 <= 1 ; box: max(|u|,|v|) <= 1.
struct Prim { float ax, ay, a0, bx, by, b0; int box; };
struct Shape {
  Prim p[kPrims];
  float xmin, xmax, ymin, ymax;   // pixel bounding box of the union
};
__device__ __forceinline__ bool prim_hit(const Prim& q, float x, float y) {
  const float u = fmaf(q.ax, x, fmaf(q.ay, y, q.a0)), v = fmaf(q.bx, x, fmaf(q.by, y, q.b0));
  return q.box ? fmaxf(fabsf(u), fabsf(v)) <= 1.f : fmaf(u, u, v * v) <= 1.f;
}
__device__ __forceinline__ bool inside(const Shape& s, float x, float y) {
  if (x < s.xmin || x > s.xmax || y < s.ymin || y > s.ymax) return false;
  bool in = false;
#pragma unroll
  for (int k = 0; k < kPrims; ++k) in |= prim_hit(s.p[k], x, y);
  return in;
}
__device__ void make_shape(Shape& s, uint64_t key, float side) {
  s.xmin = s.ymin = 1e9f;
  s.xmax = s.ymax = -1e9f;
  const float c = 0.5f * (side - 1.f);
  for (int k = 0; k < kPrims; ++k) {
    const uint64_t r0 = splitmix64(key + 3 * k), r1 = splitmix64(key + 3 * k + 1), r2 = splitmix64(key + 3 * k + 2);
    const float cx = c + (u01((uint32_t)r0) - 0.5f) * 0.24f * side;
    const float cy = c + (u01((uint32_t)(r0 >> 32)) - 0.5f) * 0.24f * side;
    const float a = (0.05f + 0.11f * u01((uint32_t)r1)) * side;
    const float b = (0.05f + 0.11f * u01((uint32_t)(r1 >> 32))) * side;
    float sn, cs;
    sincospif(2.f * u01((uint32_t)r2), &sn, &cs);
    Prim& q = s.p[k];
    q.box = (int)((r2 >> 40) & 1);
    q.ax = cs / a; q.ay = sn / a; q.a0 = -(cx * cs + cy * sn) / a;
    q.bx = -sn / b; q.by = cs / b; q.b0 = (cx * sn - cy * cs) / b;
    const float rad = q.box ? sqrtf(a * a + b * b) : fmaxf(a, b);
    s.xmin = fminf(s.xmin, cx - rad); s.xmax = fmaxf(s.xmax, cx + rad);
    s.ymin = fminf(s.ymin, cy - rad); s.ymax = fmaxf(s.ymax, cy + rad);
  }
}

__global__ void __launch_bounds__(256, 3) generate_pairs_kernel(uint64_t seed, uint64_t step, int B, int side,
                                                                __nv_bfloat16* planes, long long ld, float* labels) {
  const int i = blockIdx.x;
  const int D = side * side;
  __shared__ Shape lock_s, key_s;
  __shared__ Prim key_rot[kLabelRots][kPrims];   // key primitives composed with the label rotations
  __shared__ float rot_cs, rot_sn;
  __shared__ int counts[kLabelRots];
  const float c = 0.5f * (side - 1.f);
  if (threadIdx.x == 0) {
    make_shape(lock_s, rng_key(seed, step, 11, (uint64_t)i * 64), (float)side);
    float sn, cs;
    sincospif(2.f * u01((uint32_t)rng_key(seed, step, 13, i)), &sn, &cs);   // angle ~ U[0, 2 pi)
    rot_cs = cs;
    rot_sn = sn;
  }
  if (threadIdx.x == 32) make_shape(key_s, rng_key(seed, step, 12, (uint64_t)i * 64), (float)side);
  if (threadIdx.x < kLabelRots) counts[threadIdx.x] = 0;
  __syncthreads();
  if (threadIdx.x < kLabelRots * kPrims) {
    // test point = R_r (x - c) + c ; fold R_r and the translation into the primitive's affine map
    const int r = threadIdx.x / kPrims, k = threadIdx.x % kPrims;
    float rs, rc;
    sincospif(2.f * r / kLabelRots, &rs, &rc);
    const Prim q = key_s.p[k];
    Prim o;
    o.box = q.box;
    const float tx = c - rc * c + rs * c, ty = c - rs * c - rc * c;   // kx = rc x - rs y + tx ; ky = rs x + rc y + ty
    o.ax = q.ax * rc + q.ay * rs; o.ay = -q.ax * rs + q.ay * rc; o.a0 = q.ax * tx + q.ay * ty + q.a0;
    o.bx = q.bx * rc + q.by * rs; o.by = -q.bx * rs + q.by * rc; o.b0 = q.bx * tx + q.by * ty + q.b0;
    key_rot[r][k] = o;
  }
  __syncthreads();
  int cnt[kLabelRots];
#pragma unroll
  for (int r = 0; r < kLabelRots; ++r) cnt[r] = 0;
  const float rcs = rot_cs, rsn = rot_sn;
  const float smax = side - 1.f;
  const int chunks = (int)(ld / 8);
  const int chunks_per_row = side / 8;     // host guarantees side % 8 == 0
  for (int ch = threadIdx.x; ch < chunks; ch += blockDim.x) {
    uint32_t wl[4] = {0, 0, 0, 0}, wr[4] = {0, 0, 0, 0}, wk[4] = {0, 0, 0, 0};
    const int p0 = ch * 8;
    if (p0 < D) {
      const int yi = ch / chunks_per_row, x0 = (ch - yi * chunks_per_row) * 8;
      const float y = (float)yi;
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float x = (float)(x0 + e);
        const bool l = inside(lock_s, x, y);
        const bool k = inside(key_s, x, y);
        // inverse map of the rotation, nearest source pixel, zero outside the canvas
        const float sx = rintf(fmaf(rcs, x - c, fmaf(-rsn, y - c, c))), sy = rintf(fmaf(rsn, x - c, fmaf(rcs, y - c, c)));
        const bool rr = sx >= 0.f && sy >= 0.f && sx <= smax && sy <= smax && inside(lock_s, sx, sy);
        const uint32_t one = (e & 1) ? 0x3F800000u : 0x00003F80u;   // bf16 1.0 in the high / low half
        if (l) wl[e >> 1] |= one;
        if (rr) wr[e >> 1] |= one;
        if (k) wk[e >> 1] |= one;
        if (l) {
#pragma unroll
          for (int r = 0; r < kLabelRots; ++r) {
            bool in = false;
#pragma unroll
            for (int q = 0; q < kPrims; ++q) in |= prim_hit(key_rot[r][q], x, y);
            cnt[r] += in ? 1 : 0;
          }
        }
      }
    } else if (p0 == D) {
      wl[0] = wr[0] = wk[0] = 0x00003F80u;   // bias column (1.0) then zero pad
    }
    *reinterpret_cast<uint4*>(planes + (long long)i * ld + p0) = make_uint4(wr[0], wr[1], wr[2], wr[3]);
    *reinterpret_cast<uint4*>(planes + (long long)(B + i) * ld + p0) = make_uint4(wl[0], wl[1], wl[2], wl[3]);
    *reinterpret_cast<uint4*>(planes + (long long)(2 * B + i) * ld + p0) = make_uint4(wk[0], wk[1], wk[2], wk[3]);
  }
#pragma unroll
  for (int r = 0; r < kLabelRots; ++r) {
    int v = cnt[r];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(&counts[r], v);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int best = 0;
    for (int r = 0; r < kLabelRots; ++r) best = max(best, counts[r]);
    labels[i] = (float)best;
  }
}

void generate_pairs(uint64_t seed, uint64_t step, int B, int side, __nv_bfloat16* planes, long long ld, float* labels,
                    cudaStream_t s) {
  generate_pairs_kernel<<<B, 256, 0, s>>>(seed, step, B, side, planes, ld, labels);
}


}  // namespace magic
