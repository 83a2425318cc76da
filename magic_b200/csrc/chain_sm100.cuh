// One persistent kernel for a CHAIN of hidden layers (8c/vae.py:347-355 forward; their dgrads backward).
//
// The encoder's hidden layers are 512-wide (500 + bias column, padded): per 128-row slab a layer is 128 x 512 x 512,
// ~8 k cycles of MMA -- as separate launches each costs 26-38 us for 3-6 us of tensor work (CTA start-up, pipeline fill,
// one exposed epilogue, the activations' round trip through L2).  Here a CTA keeps its 128 x 512 activation slab in
// shared memory, in exactly the K-major 128-byte-swizzle layout the MMA reads (8 k-blocks of 16 KiB), across all layers:
//
//   slab <- TMA(first layer's input)                                  [one load per k-block]
//   for every layer:  D[128 x 512] (TMEM, all 512 columns) = slab . W_l    W_l streamed in 256 n x 64 k stages (3 x 32 KiB)
//                     epilogue: act / act' -> bf16 -> written back INTO the slab (all MMAs of the layer have retired)
//                     the slab's 8 k-block atoms leave for HBM as 8 TMA tile stores (the backward pass needs every layer)
//
// Same MMA atoms in the same k order and the same epilogue arithmetic as gemm_tcgen05_kernel with 256-wide tiles and no
// split-K, so the results are bit-identical to the per-layer launches (tests/test_gpu_parity.py::test_hidden_layer_chains).
// bf16 mode only (the fp32 mode's hi + lo slabs would not fit); layers must be 512 wide and at most 512 deep.
//
// Roles (576 threads): warp 0 = weight producer (one thread), warp 1 = TMEM allocator + one thread that loads the slab,
// issues the MMAs and the output stores, warps 2..17 = epilogue (four per TMEM lane quarter, four 32-column chunks each).
#pragma once
#include "gemm_sm100.cuh"

namespace magic {

constexpr int kChainMaxLayers = 4;
constexpr int kChainN = 512;                 // width (row pitch) of every activation of a chain
constexpr int kChainKb = kChainN / kBlockK;  // k-blocks of a resident slab
constexpr int kChainStages = 3;              // weight stages
constexpr int kChainEpiWarps = 16;
constexpr int kChainThreads = 64 + 32 * kChainEpiWarps;
constexpr uint32_t kChainAtomBytes = kBlockM * kBlockK * 2;              // one k-block of the slab: 16 KiB
constexpr uint32_t kChainSlabBytes = kChainKb * kChainAtomBytes;          // 128 KiB
constexpr uint32_t kChainStageBytes = 256 * kBlockK * 2;                  // 256 n x 64 k: 32 KiB
constexpr uint32_t kChainSmemBytes = kChainSlabBytes + kChainStages * kChainStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;

struct ChainArgs {
  CUtensorMap ta0;                          // first layer's input [M, K0] bf16, box {64 k, 128 m}, 128B swizzle
  CUtensorMap tb[kChainMaxLayers];          // weights of layer l: forward [K, N] MN-major box {64 n, 64 k, 4}; dgrad [N, K] box {64 k, 256 n}
  CUtensorMap tout[kChainMaxLayers];        // output of layer l [M, 512] bf16, box {64, 128}, 128B swizzle (= the slab's atoms)
  EpiArgs epi[kChainMaxLayers];             // EPI_ACT (forward) / EPI_DACT (dgrad) arguments of layer l
  int kb[kChainMaxLayers];                  // k-blocks of layer l (layer 0: of its input; later layers: of the slab)
  int nlayers, M, n_slabs;
};
static_assert(sizeof(ChainArgs) <= 4096, "kernel parameter space");

// One 32-column chunk of one row: accumulator -> activation (forward) or gradient x act'(y) (dgrad), packed bf16.
template <bool DGRAD, int ACT>
__device__ __forceinline__ void chain_values(const EpiArgs& e, const uint32_t (&r)[32], const uint32_t (&y)[16], int col0,
                                             uint32_t (&hi)[16]) {
#pragma unroll
  for (int j = 0; j < 32; j += 2) {
    float v[2];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int c = col0 + j + q;
      const float x = __uint_as_float(r[j + q]) * e.alpha;
      if constexpr (!DGRAD) {
        v[q] = c < e.n_valid ? act_fwd<ACT, false>(x) : (c == e.n_valid ? 1.f : 0.f);
      } else {
        const float yy = q == 0 ? ptx::bf16_lo(y[j >> 1]) : ptx::bf16_hi(y[j >> 1]);
        v[q] = c < e.n_valid ? x * act_bwd_from_y<ACT>(yy) : 0.f;
      }
    }
    hi[j >> 1] = ptx::pack_bf16(v[0], v[1]);
  }
}

template <bool DGRAD>
__global__ void __maxnreg__(96) chain_tcgen05_kernel(const __grid_constant__ ChainArgs args) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* slab = smem;                                        // 8 k-block atoms of 128 rows x 128 B
  uint8_t* smem_b = smem + kChainSlabBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_b + kChainStages * kChainStageBytes);
  uint64_t* full_bar = bars;                                   // [3] weight stage landed
  uint64_t* empty_bar = bars + kChainStages;                   // [3] weight stage consumed
  uint64_t* a_full = bars + 2 * kChainStages;                  // first layer's input landed in the slab
  uint64_t* tfull = a_full + 1;                                // accumulator of a layer complete
  uint64_t* slab_ready = a_full + 2;                           // epilogue has rewritten the slab (and is done with TMEM)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(a_full + 3);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nl = args.nlayers;
  if (warp == 0 && lane == 0) {
    ptx::tma_prefetch_desc(&args.ta0);
    for (int l = 0; l < nl; ++l) { ptx::tma_prefetch_desc(&args.tb[l]); ptx::tma_prefetch_desc(&args.tout[l]); }
    for (int i = 0; i < kChainStages; ++i) { ptx::mbar_init(&full_bar[i], 1); ptx::mbar_init(&empty_bar[i], 1); }
    ptx::mbar_init(a_full, 1);
    ptx::mbar_init(tfull, 1);
    ptx::mbar_init(slab_ready, kChainEpiWarps);
    ptx::fence_barrier_init();
  } else if (warp == 1) {
    ptx::tmem_alloc(tmem_ptr, 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  ptx::grid_dep_wait();

  if (warp == 0) {
    // ------------------------------------------------------------ weight producer: every slab streams every layer again (L2-resident)
    if (ptx::elect_one()) {                                    // (elect.sync: direct UTMALDG / UTCHMMA, see gemm_sm100.cuh)
      uint32_t stage = 0, phase = 0;
      for (int w = blockIdx.x; w < args.n_slabs; w += gridDim.x)
        for (int l = 0; l < nl; ++l)
          for (int h = 0; h < 2; ++h)
            for (int kb = 0; kb < args.kb[l]; ++kb) {
              ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
              ptx::mbar_expect_tx(&full_bar[stage], kChainStageBytes);
              uint8_t* sb = smem_b + stage * kChainStageBytes;
              if constexpr (DGRAD) ptx::tma_load_2d(&args.tb[l], &full_bar[stage], sb, kb * kBlockK, h * 256);          // box {64 k, 256 n}
              else ptx::tma_load_3d(&args.tb[l], &full_bar[stage], sb, 0, kb * kBlockK, (h * 256) >> 6);                // box {64 n, 64 k, 4 slabs}
              if (++stage == kChainStages) { stage = 0; phase ^= 1; }
            }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ slab loads, MMAs, output stores (one thread)
    if (ptx::elect_one()) {
      constexpr uint32_t idesc = ptx::make_idesc_bf16(kBlockM, 256, false, !DGRAD);
      constexpr uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
      constexpr uint32_t b_lbo = DGRAD ? 0u : kAtomBytes;
      constexpr uint32_t b_lo_c = ((b_lbo >> 4) & 0x3FFFu) << 16;
      constexpr uint32_t a_kadv = kUmmaK * 2, b_kadv = DGRAD ? kUmmaK * 2 : kUmmaK * 128;
      const uint32_t a_addr0 = (ptx::smem_u32(slab) >> 4) & 0x3FFFu, b_addr0 = (ptx::smem_u32(smem_b) >> 4) & 0x3FFFu;
      uint32_t stage = 0, phase = 0, n = 0, it = 0;            // n: (slab, layer) counter = phase of tfull / slab_ready
      for (int w = blockIdx.x; w < args.n_slabs; w += gridDim.x, ++it) {
        const int m0 = w * kBlockM;
        if (it > 0) ptx::bulk_wait_read();                     // the previous slab's last output stores have read the slab
        ptx::mbar_expect_tx(a_full, (uint32_t)args.kb[0] * kChainAtomBytes);
        for (int kb = 0; kb < args.kb[0]; ++kb) ptx::tma_load_2d(&args.ta0, a_full, slab + kb * kChainAtomBytes, kb * kBlockK, m0);
        ptx::mbar_wait(a_full, it & 1);
        for (int l = 0; l < nl; ++l) {
          if (l > 0) {                                         // layer l-1's output is in the slab: send it to HBM, then read it as A
            ptx::mbar_wait(slab_ready, (n - 1) & 1);
            ptx::tc_fence_after();
            for (int kb = 0; kb < kChainKb; ++kb) ptx::tma_store_2d(&args.tout[l - 1], slab + kb * kChainAtomBytes, kb * kBlockK, m0);
            ptx::bulk_commit();
          }
          for (int h = 0; h < 2; ++h)
            for (int kb = 0; kb < args.kb[l]; ++kb) {
              ptx::mbar_wait(&full_bar[stage], phase);
              ptx::tc_fence_after();
              const uint32_t a_lo = a_addr0 + kb * (kChainAtomBytes >> 4);
              const uint32_t b_lo = (b_addr0 + stage * (kChainStageBytes >> 4)) | b_lo_c;
#pragma unroll
              for (int k = 0; k < kBlockK / kUmmaK; ++k) {
                const uint64_t da = (static_cast<uint64_t>(desc_hi) << 32) | (a_lo + k * (a_kadv >> 4));
                const uint64_t db = (static_cast<uint64_t>(desc_hi) << 32) | (b_lo + k * (b_kadv >> 4));
                ptx::umma_bf16(tmem_base + h * 256, da, db, idesc, (kb > 0 || k > 0) ? 1u : 0u);
              }
              ptx::umma_commit(&empty_bar[stage]);
              if (++stage == kChainStages) { stage = 0; phase ^= 1; }
            }
          if (l > 0) ptx::bulk_wait_read();                    // the stores above have read the slab before the epilogue rewrites it
          ptx::umma_commit(tfull);
          ++n;
        }
        ptx::mbar_wait(slab_ready, (n - 1) & 1);               // the last layer's output
        ptx::tc_fence_after();
        for (int kb = 0; kb < kChainKb; ++kb) ptx::tma_store_2d(&args.tout[nl - 1], slab + kb * kChainAtomBytes, kb * kBlockK, m0);
        ptx::bulk_commit();
      }
      ptx::grid_dep_launch();
      ptx::bulk_wait_all();
    }
  } else {
    // ------------------------------------------------------------ epilogue warps
    const int q = warp & 3, grp = (warp - 2) >> 2;              // TMEM lane quarter, chunk group: chunks grp, grp + 4, ...
    uint32_t n = 0;
    for (int w = blockIdx.x; w < args.n_slabs; w += gridDim.x) {
      const int row = w * kBlockM + q * 32 + lane;
      const int r_in = q * 32 + lane;
      uint8_t* const row_base = slab + r_in * 128;
      const int x = r_in & 7;
      for (int l = 0; l < nl; ++l, ++n) {
        const EpiArgs& e = args.epi[l];
        ptx::mbar_wait(tfull, n & 1);
        ptx::tc_fence_after();
#pragma unroll 1
        for (int k = 0; k < kChainN / 32 / 4; ++k) {
          const int c = grp + 4 * k, col0 = c * 32;
          uint32_t r[32], y[16], hi[16];
          ptx::tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + col0, r);
          if constexpr (DGRAD) {
#pragma unroll
            for (int i = 0; i < 16; ++i) y[i] = 0u;
            if (row < args.M) {
              const int ar = row < e.aux_split ? row : row - e.aux_shift;
              const uint4* ap = reinterpret_cast<const uint4*>(e.aux + (long long)ar * e.ld_aux + col0);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint4 t = ap[i];
                y[4 * i] = t.x; y[4 * i + 1] = t.y; y[4 * i + 2] = t.z; y[4 * i + 3] = t.w;
              }
            }
          }
          ptx::tmem_ld_wait();
          if (e.act == ACT_TANH) chain_values<DGRAD, ACT_TANH>(e, r, y, col0, hi);
          else chain_values<DGRAD, ACT_ELU>(e, r, y, col0, hi);
          // 64 bytes of this row into k-block col0 / 64 of the slab: 16-byte piece p of the swizzle row sits at p ^ (row & 7)
          uint8_t* dst = row_base + (col0 >> 6) * kChainAtomBytes;
          const int p0 = (col0 & 63) >> 3;
#pragma unroll
          for (int i = 0; i < 4; ++i)
            ptx::st_shared_v4(dst + (((p0 + i) ^ x) << 4), hi[4 * i], hi[4 * i + 1], hi[4 * i + 2], hi[4 * i + 3]);
        }
        ptx::fence_proxy_async_smem();                          // generic-proxy writes -> the MMA's and the TMA store's reads
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(slab_ready);
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace magic
