// The training engine: parameter layout, activations, GEMM plans and the step (see engine.h).
#include "engine.h"
#include "launch.cuh"
#include "ptx.cuh"
#include <cmath>
#include <cstdio>
#include <cstring>
#include <dlfcn.h>

namespace magic {

#define MAGIC_CUDA(x)                                                                                    \
  do {                                                                                                   \
    cudaError_t e_ = (x);                                                                                \
    if (e_ != cudaSuccess) return std::string(#x) + ": " + cudaGetErrorString(e_);                       \
  } while (0)
#define MAGIC_TRY(x)                 \
  do {                               \
    std::string s_ = (x);            \
    if (!s_.empty()) return s_;      \
  } while (0)

static long long round_up(long long v, long long m) { return (v + m - 1) / m * m; }

// ------------------------------------------------------------------ small device helpers
__global__ void fill_kernel(float* p, long long n, float v) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// Xavier-uniform window init (8c/utils.py:346-353): U(-b, b), b = sqrt(6 / (fan_in + fan_out)).
__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__global__ void xavier_kernel(float* w, long long ld, int nrows, int col0, int ncols, float bound, uint64_t key) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)nrows * ncols) return;
  const int r = (int)(i / ncols), c = (int)(i % ncols);
  const uint32_t bits = (uint32_t)(mix64(key + (uint64_t)i) >> 32);
  const float u = (bits >> 8) * (1.0f / 16777216.0f);
  w[(long long)r * ld + col0 + c] = (2.f * u - 1.f) * bound;
}

// losses5 = [cost, training, rec, kl, def] from sums = [rec, kl, def, training]
__global__ void finalize_losses_kernel(const float* sums, float* out) {
  ptx::grid_dep_wait();
  if (threadIdx.x == 0) {
    out[0] = sums[0] + sums[1] + sums[2];
    out[1] = sums[3];
    out[2] = sums[0];
    out[3] = sums[1];
    out[4] = sums[2];
  }
}

// user z [B,L] fp32 -> decoder input bf16 [B, ld] with the bias column (generate(), 8c/vae.py:416-428)
__global__ void z_to_zdec_kernel(const float* z, int B, int L, bf16* hi, bf16* lo, long long ld) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)B * (L + 1)) return;
  const int i = (int)(idx / (L + 1)), j = (int)(idx % (L + 1));
  const float v = j < L ? z[(long long)i * L + j] : 1.f;
  const bf16 h = __float2bfloat16(v);
  hi[(long long)i * ld + j] = h;
  if (lo) lo[(long long)i * ld + j] = __float2bfloat16(v - __bfloat162float(h));
}

// ------------------------------------------------------------------ NCCL through dlopen (optional at run time)
typedef int (*nccl_get_uid_t)(void*);
struct NcclUid { char b[128]; };
typedef int (*nccl_init_rank2_t)(void**, int, NcclUid, int);
typedef int (*nccl_allreduce_t)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*nccl_allgather_t)(const void*, void*, size_t, int, void*, cudaStream_t);
typedef int (*nccl_destroy_t)(void*);
typedef const char* (*nccl_errstr_t)(int);
struct NcclApi {
  void* lib = nullptr;
  nccl_get_uid_t get_uid = nullptr;
  nccl_init_rank2_t init_rank = nullptr;
  nccl_allreduce_t allreduce = nullptr;
  nccl_allgather_t allgather = nullptr;
  nccl_destroy_t destroy = nullptr;
  nccl_errstr_t errstr = nullptr;
};
static NcclApi g_nccl;
static std::string load_nccl() {
  if (g_nccl.lib) return "";
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (g_nccl.lib) break;
  }
  if (!g_nccl.lib) return "NCCL not found (dlopen libnccl.so.2 failed): data parallelism unavailable";
  g_nccl.get_uid = (nccl_get_uid_t)dlsym(g_nccl.lib, "ncclGetUniqueId");
  g_nccl.init_rank = (nccl_init_rank2_t)dlsym(g_nccl.lib, "ncclCommInitRank");
  g_nccl.allreduce = (nccl_allreduce_t)dlsym(g_nccl.lib, "ncclAllReduce");
  g_nccl.allgather = (nccl_allgather_t)dlsym(g_nccl.lib, "ncclAllGather");
  g_nccl.destroy = (nccl_destroy_t)dlsym(g_nccl.lib, "ncclCommDestroy");
  g_nccl.errstr = (nccl_errstr_t)dlsym(g_nccl.lib, "ncclGetErrorString");
  if (!g_nccl.get_uid || !g_nccl.init_rank || !g_nccl.allreduce) return "NCCL symbols missing";
  return "";
}
std::string nccl_unique_id(void* out128) {
  MAGIC_TRY(load_nccl());
  int r = g_nccl.get_uid(out128);
  if (r != 0) return std::string("ncclGetUniqueId: ") + (g_nccl.errstr ? g_nccl.errstr(r) : "error");
  return "";
}

// ------------------------------------------------------------------ Engine: allocation / layout
Engine::~Engine() {
  // the process-lifetime handle frees with the context; explicit frees keep repeated create/destroy clean
  if (gen_stream_) cudaStreamSynchronize(gen_stream_);
  cudaStreamSynchronize(stream);
  void* ptrs[] = {p_, g1_, g2_, m1_, v1_, m2_, v2_, w_hi_, w_lo_, in_[0].X.hi, in_[1].X.hi, heads_, z_, zdec_.hi, zdec_.lo, D1_.hi, D1_.lo,
                  D2_.hi, D2_.lo, dA_.hi, dA_.lo, dD2_.hi, dD2_.lo, dD1_.hi, dD1_.lo, dz_dec_, dheads_.hi, dheads_.lo, ws_,
                  statbuf_, g_row_, dist_, losses_dev_, stage_x_, stage_labels_, stage_eps_,
                  in_[0].xb3, in_[1].xb3, in_[0].kmask, in_[1].kmask, in_[0].labels, in_[1].labels, in_[0].rot, in_[1].rot, stage_, barrier_buf_, bits_stage_, not_binary_, gather_idx_};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  for (auto& a : H_) { if (a.hi) cudaFree(a.hi); if (a.lo) cudaFree(a.lo); }
  for (auto& a : dH_) { if (a.hi) cudaFree(a.hi); if (a.lo) cudaFree(a.lo); }
  for (int r = 0; r < kMaxRanks; ++r) {
    if (r == cfg.rank) continue;
    if (peer_stage_[r]) cudaIpcCloseMemHandle(peer_stage_[r]);
    if (peer_whi_[r]) cudaIpcCloseMemHandle(peer_whi_[r]);
    if (peer_wlo_[r]) cudaIpcCloseMemHandle(peer_wlo_[r]);
  }
  if (comm_ && g_nccl.destroy) g_nccl.destroy(comm_);
  if (comm_stream_) cudaStreamDestroy(comm_stream_);
  if (gen_stream_) cudaStreamDestroy(gen_stream_);
  if (aux_stream_) { cudaStreamSynchronize(aux_stream_); cudaStreamDestroy(aux_stream_); }
  if (ev_aux_fork_) cudaEventDestroy(ev_aux_fork_);
  if (read_stream_) { cudaStreamSynchronize(read_stream_); cudaStreamDestroy(read_stream_); }
  if (ev_losses_) cudaEventDestroy(ev_losses_);
  if (h_losses_) cudaFreeHost(h_losses_);
  if (ev_aux_join_) cudaEventDestroy(ev_aux_join_);
  if (ev_gate_) cudaEventDestroy(ev_gate_);
  for (auto& b : in_) { if (b.ready) cudaEventDestroy(b.ready); if (b.consumed) cudaEventDestroy(b.consumed); }
  for (auto& e : ev_bwd_) if (e) cudaEventDestroy(e);
  if (ev_comm_) cudaEventDestroy(ev_comm_);
}

std::string Engine::alloc(void** p, size_t bytes) {
  if (bytes == 0) bytes = 16;
  MAGIC_CUDA(cudaMalloc(p, bytes));
  MAGIC_CUDA(cudaMemsetAsync(*p, 0, bytes, stream));
  stats.hbm_bytes += (int64_t)bytes;
  return "";
}

std::string Engine::alloc_act(Act* a, long long rows, int width_plus_bias) {
  a->rows = rows;
  a->ld = round_up(width_plus_bias, 64);
  MAGIC_TRY(alloc((void**)&a->hi, sizeof(bf16) * rows * a->ld));
  if (split) MAGIC_TRY(alloc((void**)&a->lo, sizeof(bf16) * rows * a->ld));
  return "";
}

std::string Engine::init(const magic_config& c, cudaStream_t s) {
  cfg = c;
  stream = s;
  if (c.D <= 0 || c.D % 8 != 0) return "config: D must be a positive multiple of 8";
  if (c.n_enc < 1 || c.n_enc > MAGIC_MAX_ENC_LAYERS) return "config: n_enc out of range";   // 8c/vae.py:31
  if (c.L < 1) return "config: L must be >= 1";
  if (c.act != MAGIC_ACT_TANH && c.act != MAGIC_ACT_ELU) return "config: unknown activation";
  if (c.metric != MAGIC_METRIC_COSINE && c.metric != MAGIC_METRIC_SQDIFF)
    return "Invalid distance metric: " + std::to_string(c.metric);                           // 8c/vae.py:305
  if (c.batch_local < 1 || c.batch_global < c.batch_local) return "config: bad batch sizes";
  if (c.world < 1 || c.world > kMaxRanks || c.rank < 0 || c.rank >= c.world) return "config: bad rank/world (world <= 8)";
  int dev = 0;
  MAGIC_CUDA(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  MAGIC_CUDA(cudaGetDeviceProperties(&prop, dev));
  if (prop.major != 10) return std::string("this library is built for sm_100a (B200); found ") + prop.name;
  num_sms = prop.multiProcessorCount;
  MAGIC_TRY(gemm_init());
  cudaFuncSetAttribute(finalize_losses_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  cudaFuncSetAttribute(z_to_zdec_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);

  E = c.n_enc; B = c.batch_local; D = c.D; L = c.L;
  split = c.precision == MAGIC_PRECISION_FP32;
  inv_bg = 1.0f / (float)c.batch_global;

  // ---- layers and the flat parameter layout: [enc0 .. enc(E-1) | heads] [dec h1 | dec h2 | dec out]
  lin.clear();
  long long off = 0;
  auto add = [&](int Kv, int Nv) {
    Linear l;
    l.Kv = Kv; l.Nv = Nv; l.ld = round_up(Nv, 64); l.off = off;
    off += l.size();
    lin.push_back(l);
    return (int)lin.size() - 1;
  };
  add(D, c.enc[0]);
  for (int i = 1; i < E; ++i) add(c.enc[i - 1], c.enc[i]);
  iHead = add(c.enc[E - 1], 2 * L);
  n_both = off;
  iD1 = add(L, c.dec[0]);
  iD2 = add(c.dec[0], c.dec[1]);
  iOut = add(c.dec[1], D);
  n_total = off;

  // ---- reference variables as windows (creation order of 8c/vae.py:89-147, dead tensors omitted)
  views.clear();
  auto view = [&](const std::string& name, int layer, bool bias, int col0, int ncols, int opts) {
    ParamView v;
    v.name = name; v.layer = layer; v.is_bias = bias; v.col0 = col0; v.ncols = ncols; v.optimisers = opts;
    v.row0 = bias ? lin[layer].Kv : 0;
    v.nrows = bias ? 1 : lin[layer].Kv;
    views.push_back(v);
  };
  char nm[64];
  for (int i = 0; i < E; ++i) { snprintf(nm, sizeof nm, "encoder_h%d_weights", i); view(nm, i, false, 0, lin[i].Nv, 3); }
  view("encoder_out_mean_weights", iHead, false, 0, L, 3);
  view("encoder_out_log_sigma_weights", iHead, false, L, L, 3);
  for (int i = 0; i < E; ++i) { snprintf(nm, sizeof nm, "encoder_b%d_biases", i); view(nm, i, true, 0, lin[i].Nv, 3); }
  view("encoder_out_mean_biases", iHead, true, 0, L, 3);
  view("encoder_out_log_sigma_biases", iHead, true, L, L, 3);
  view("decoder_h1_weights", iD1, false, 0, lin[iD1].Nv, 1);
  view("decoder_h2_weights", iD2, false, 0, lin[iD2].Nv, 1);
  view("decoder_out_mean_weights", iOut, false, 0, lin[iOut].Nv, 1);
  view("decoder_b1_biases", iD1, true, 0, lin[iD1].Nv, 1);
  view("decoder_b2_biases", iD2, true, 0, lin[iD2].Nv, 1);
  view("decoder_out_mean_biases", iOut, true, 0, lin[iOut].Nv, 1);
  stats.n_params1 = stats.n_params2 = 0;
  for (auto& v : views) {
    const long long n = (long long)v.nrows * v.ncols;
    stats.n_params1 += n;
    if (v.optimisers & 2) stats.n_params2 += n;
  }
  {
    // MAC = 11 enc + 3 dec - 4 D H0, enc = sum of encoder layer MACs incl. both heads
    double enc = (double)D * c.enc[0], dec = (double)L * c.dec[0] + (double)c.dec[0] * c.dec[1] + (double)c.dec[1] * D;
    for (int i = 1; i < E; ++i) enc += (double)c.enc[i - 1] * c.enc[i];
    enc += 2.0 * c.enc[E - 1] * L;
    stats.model_flops_per_pair = 2.0 * (11.0 * enc + 3.0 * dec - 4.0 * (double)D * c.enc[0]);
  }

  // ---- parameter-shaped buffers
  // data parallel: the flat space is cut into `world` equal slices (owner r = idx / slice_len); buffers are sized
  // world * slice_len so that slices can be all-gathered in place
  slice_len_ = (int)(round_up(n_total, (long long)kDpGranule * c.world) / c.world);
  alloc_len_ = c.world > 1 ? (long long)c.world * slice_len_ : n_total;
  const long long alloc2 = c.world > 1 ? alloc_len_ : n_both;
  MAGIC_TRY(alloc((void**)&p_, sizeof(float) * alloc_len_));
  MAGIC_TRY(alloc((void**)&g1_, sizeof(float) * alloc_len_));
  MAGIC_TRY(alloc((void**)&m1_, sizeof(float) * alloc_len_));
  MAGIC_TRY(alloc((void**)&v1_, sizeof(float) * alloc_len_));
  MAGIC_TRY(alloc((void**)&g2_, sizeof(float) * alloc2));
  MAGIC_TRY(alloc((void**)&m2_, sizeof(float) * alloc2));
  MAGIC_TRY(alloc((void**)&v2_, sizeof(float) * alloc2));
  MAGIC_TRY(alloc((void**)&w_hi_, sizeof(bf16) * alloc_len_));
  if (split) MAGIC_TRY(alloc((void**)&w_lo_, sizeof(bf16) * alloc_len_));
  if (c.world > 1) {
    xchg_off_ = (size_t)2 * alloc_len_;
    MAGIC_TRY(alloc((void**)&stage_, sizeof(float) * (2 * alloc_len_ + xchg_area_floats(c.world))));
    MAGIC_TRY(alloc((void**)&barrier_buf_, sizeof(float) * 8));
  }
  for (size_t vi = 0; vi < views.size(); ++vi) {
    const ParamView& v = views[vi];
    if (v.is_bias) continue;   // biases start at zero (8c/vae.py:119)
    const Linear& l = lin[v.layer];
    const float bound = sqrtf(6.0f / (float)(l.Kv + v.ncols));
    const long long n = (long long)v.nrows * v.ncols;
    xavier_kernel<<<(int)((n + 255) / 256), 256, 0, stream>>>(p_ + l.off, l.ld, v.nrows, v.col0, v.ncols, bound,
                                                               c.seed * 0x9E3779B97F4A7C15ull + vi * 0x1000000000ull);
  }
  shadow_dirty_ = true;

  // ---- activations
  // two input buffers: the generator fills one on its own stream while the step consumes the other
  // packed rows: wide enough for the 64-column k-blocks of the forward GEMM and the 128-row tiles of the weight
  // gradient (both reach past column D, the bias column, into zero padding)
  xb_ld_ = (int)round_up(4 * ((D + 1 + 127) / 128 + 1), 8);   // (+ 1 tile: a CTA pair's empty partner tile reads zeros there)
  if (const char* ba = getenv("MAGIC_BITS_A")) bits_a_ = atoi(ba) != 0;
  if (const char* sz = getenv("MAGIC_SKIP_ZERO")) skip_zero_ = atoi(sz) != 0;
  km_fwd_tiles_ = (3 * B + 127) / 128;
  km_fwd_kb_ = (D + 1 + 63) / 64;
  km_fwd_words_ = (km_fwd_kb_ + 31) / 32;
  km_wg_tiles_ = (D + 1 + 127) / 128;
  km_wg_words_ = ((2 * B + 63) / 64 + 31) / 32;
  km_wg_off_[0] = (size_t)round_up((long long)km_fwd_tiles_ * km_fwd_words_, 4);
  km_wg_off_[1] = km_wg_off_[0] + (size_t)round_up((long long)km_wg_tiles_ * km_wg_words_, 4);
  km_total_ = km_wg_off_[1] + (size_t)round_up((long long)km_wg_tiles_ * km_wg_words_, 4);
  for (auto& b : in_) {
    MAGIC_TRY(alloc_act(&b.X, 3LL * B, D + 1));
    if (b.X.lo) { cudaFree(b.X.lo); stats.hbm_bytes -= (int64_t)(sizeof(bf16) * b.X.rows * b.X.ld); b.X.lo = nullptr; }  // pixels are exact in bf16
    MAGIC_TRY(alloc((void**)&b.xb3, sizeof(uint32_t) * 3 * (size_t)B * xb_ld_));
    b.xbits = b.xb3 + (size_t)B * xb_ld_;
    MAGIC_TRY(alloc((void**)&b.kmask, sizeof(uint32_t) * km_total_));
    MAGIC_TRY(alloc((void**)&b.rot, sizeof(int) * 2 * (size_t)B));
    MAGIC_TRY(alloc((void**)&b.labels, sizeof(float) * B));
    MAGIC_CUDA(cudaEventCreateWithFlags(&b.ready, cudaEventDisableTiming));
    MAGIC_CUDA(cudaEventCreateWithFlags(&b.consumed, cudaEventDisableTiming));
  }
  MAGIC_TRY(alloc((void**)&not_binary_, 16));
  MAGIC_CUDA(cudaStreamCreateWithFlags(&gen_stream_, cudaStreamNonBlocking));
  MAGIC_CUDA(cudaStreamCreateWithFlags(&aux_stream_, cudaStreamNonBlocking));
  MAGIC_CUDA(cudaEventCreateWithFlags(&ev_aux_fork_, cudaEventDisableTiming));
  MAGIC_CUDA(cudaStreamCreateWithFlags(&read_stream_, cudaStreamNonBlocking));
  MAGIC_CUDA(cudaEventCreateWithFlags(&ev_losses_, cudaEventDisableTiming));
  MAGIC_CUDA(cudaMallocHost((void**)&h_losses_, sizeof(float) * 8));
  MAGIC_CUDA(cudaEventCreateWithFlags(&ev_aux_join_, cudaEventDisableTiming));
  if (const char* ea = getenv("MAGIC_EARLY_ADAM")) early_adam_ok_ = atoi(ea) != 0;
  MAGIC_CUDA(cudaEventCreateWithFlags(&ev_gate_, cudaEventDisableTiming));
  if (const char* g = getenv("MAGIC_GEN_GATE")) gen_gate_ = atoi(g);        // tuning knob, default 1
  H_.resize(E);
  dH_.resize(E);
  for (int i = 0; i < E; ++i) {
    MAGIC_TRY(alloc_act(&H_[i], 3LL * B, c.enc[i] + 1));
    MAGIC_TRY(alloc_act(&dH_[i], 4LL * B, c.enc[i] + 1));
  }
  ld_heads_ = round_up(2 * L, 64);
  MAGIC_TRY(alloc((void**)&heads_, sizeof(float) * 3 * B * ld_heads_));
  MAGIC_TRY(alloc((void**)&z_, sizeof(float) * 3 * B * L));
  MAGIC_TRY(alloc_act(&zdec_, B, L + 1));
  MAGIC_TRY(alloc_act(&D1_, B, c.dec[0] + 1));
  MAGIC_TRY(alloc_act(&D2_, B, c.dec[1] + 1));
  MAGIC_TRY(alloc_act(&dA_, B, D));
  MAGIC_TRY(alloc_act(&dD2_, B, c.dec[1] + 1));
  MAGIC_TRY(alloc_act(&dD1_, B, c.dec[0] + 1));
  ld_dz_ = round_up(L, 64);
  MAGIC_TRY(alloc((void**)&dz_dec_, sizeof(float) * B * ld_dz_));
  MAGIC_TRY(alloc_act(&dheads_, 4LL * B, 2 * L));
  ws_ld_ = 64;
  for (int i = 0; i < E; ++i) ws_ld_ = std::max(ws_ld_, H_[i].ld);
  ws_ld_ = std::max(ws_ld_, std::max(D1_.ld, D2_.ld));
  MAGIC_TRY(alloc((void**)&ws_, sizeof(float) * 4 * B * ws_ld_));
  MAGIC_TRY(alloc((void**)&statbuf_, sizeof(float) * (8 + 3 * L)));
  MAGIC_TRY(alloc((void**)&g_row_, sizeof(float) * B));
  MAGIC_TRY(alloc((void**)&dist_, sizeof(float) * B));
  MAGIC_TRY(alloc((void**)&losses_dev_, sizeof(float) * 8));
  for (auto& e : ev_bwd_) MAGIC_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  MAGIC_CUDA(cudaEventCreateWithFlags(&ev_comm_, cudaEventDisableTiming));
  MAGIC_TRY(build_plans());
  MAGIC_CUDA(cudaStreamSynchronize(stream));
  return "";
}

// ------------------------------------------------------------------ plans
Mat Engine::wmat(int layer, bool lo) const {
  const Linear& l = lin[layer];
  return Mat((lo ? w_lo_ : w_hi_) + l.off, l.rows(), l.Nv, l.ld);
}
Mat Engine::amat(const Act& a, bool lo, long long row0, long long rows, long long cols) {
  return Mat((lo ? a.lo : a.hi) + row0 * a.ld, rows, cols, a.ld);
}
// bf16 mode: one segment.  fp32 mode: a*b ~ a_hi*b_hi + a_hi*b_lo + a_lo*b_hi (a_lo*b_lo ~ 2^-18 dropped);
// an exact A (binary pixels) needs only the first two.
int Engine::add_segments(GemmProblem& p, int g, Mat a_hi, Mat a_lo, Mat b_hi, Mat b_lo, bool a_exact) {
  p.a[g][0] = a_hi; p.b[g][0] = b_hi;
  int n = 1;
  if (split) {
    p.a[g][1] = a_hi; p.b[g][1] = b_lo; n = 2;
    if (!a_exact) { p.a[g][2] = a_lo; p.b[g][2] = b_hi; n = 3; }
  }
  p.nseg = n;
  return n;
}

std::string Engine::build_plans() {
  const magic_config& c = cfg;
  const int prec = split ? 1 : 0;
  double flops = 0;
  auto finish = [&](GemmProblem& p, GemmPlan* pl, const char* what) -> std::string {
    std::string err = gemm_plan(p, num_sms, pl);
    if (!err.empty()) return std::string(what) + ": " + err;
    flops += pl->flops;
    return "";
  };
  auto act_epi = [&](const Act& out, int n_valid) {
    EpiArgs e;
    memset(&e, 0, sizeof e);
    e.out = out.hi; e.out_lo = out.lo; e.ld_out = out.ld; e.alpha = 1.f; e.n_valid = n_valid; e.act = c.act;
    e.precise = prec; e.aux_split = 1 << 30;
    return e;
  };
  auto f32_epi = [&](float* out, long long ld, float alpha) {
    EpiArgs e;
    memset(&e, 0, sizeof e);
    e.out = out; e.ld_out = ld; e.alpha = alpha; e.aux_split = 1 << 30;
    return e;
  };
  auto dact_epi = [&](const Act& out, const Act& y, int n_valid, int aux_split, int aux_shift) {
    EpiArgs e = act_epi(out, n_valid);
    e.aux = y.hi; e.aux_lo = y.lo; e.ld_aux = y.ld; e.aux_split = aux_split; e.aux_shift = aux_shift;
    return e;
  };

  // ---------------- forward: encoder (three streams stacked, M = 3B), heads
  fwd_enc_.assign(E + 1, GemmPlan());
  for (int i = 0; i < E; ++i) {
    for (int alt = 0; alt < (i == 0 ? 2 : 1); ++alt) {
      GemmProblem p;
      p.kind = GEMM_FWD; p.epi = EPI_ACT; p.M = 3 * B; p.K = lin[i].Kv + 1; p.N = (int)H_[i].ld;
      const Act& in = i == 0 ? in_[alt].X : H_[i - 1];
      add_segments(p, 0, amat(in, false, 0, 3LL * B, p.K), amat(in, true, 0, 3LL * B, p.K), wmat(i, false), wmat(i, true),
                   i == 0);
      p.e[0] = act_epi(H_[i], lin[i].Nv);
      p.ws = ws_; p.ws_ld = H_[i].ld;
      p.cta_pair = i == 0;       // the 40000-deep first layer: CTA pairs (tcgen05 cta_group::2) where the variant exists
      if (i == 0 && bits_a_) {   // the images stay packed in HBM: 1 KiB instead of 16 KiB per 128 x 64 stage
        p.a_bits = true;
        p.abits[0].base = in_[alt].xb3; p.abits[0].pitch = xb_ld_; p.abits[0].rows = 3 * B;
        p.kmask[0] = in_[alt].kmask; p.kmask_words = km_fwd_words_;
      }
      MAGIC_TRY(finish(p, alt == 0 ? &fwd_enc_[i] : &fwd_enc0_alt_, "fwd enc"));
      if (alt == 1) flops -= fwd_enc0_alt_.flops;
    }
  }
  {
    GemmProblem p;
    p.kind = GEMM_FWD; p.epi = EPI_STORE_F32; p.M = 3 * B; p.K = lin[iHead].Kv + 1; p.N = 2 * L;
    const Act& in = H_[E - 1];
    add_segments(p, 0, amat(in, false, 0, 3LL * B, p.K), amat(in, true, 0, 3LL * B, p.K), wmat(iHead, false),
                 wmat(iHead, true), false);
    p.e[0] = f32_epi(heads_, ld_heads_, 1.f);
    p.force_splitk = 1;
    MAGIC_TRY(finish(p, &fwd_enc_[E], "fwd heads"));
  }
  // ---------------- forward: decoder (lock stream only, 8c/vae.py:221-226)
  auto dec_fwd = [&](GemmPlan* pl, const Act& in, int layer, const Act& out) -> std::string {
    GemmProblem p;
    p.kind = GEMM_FWD; p.epi = EPI_ACT; p.M = B; p.K = lin[layer].Kv + 1; p.N = (int)out.ld;
    add_segments(p, 0, amat(in, false, 0, B, p.K), amat(in, true, 0, B, p.K), wmat(layer, false), wmat(layer, true), false);
    p.e[0] = act_epi(out, lin[layer].Nv);
    p.ws = ws_; p.ws_ld = out.ld;
    return finish(p, pl, "fwd dec");
  };
  MAGIC_TRY(dec_fwd(&pl_d1_, zdec_, iD1, D1_));
  MAGIC_TRY(dec_fwd(&pl_d2_, D1_, iD2, D2_));
  for (int v = 0; v < 3; ++v) {   // 0: BCE against input buffer 0; 1: sigmoid (inference); 2: BCE against input buffer 1
    GemmProblem p;
    p.kind = GEMM_FWD; p.epi = v != 1 ? EPI_BCE : EPI_SIGMOID; p.M = B; p.K = lin[iOut].Kv + 1; p.N = D;
    add_segments(p, 0, amat(D2_, false, 0, B, p.K), amat(D2_, true, 0, B, p.K), wmat(iOut, false), wmat(iOut, true), false);
    EpiArgs e;
    memset(&e, 0, sizeof e);
    if (v != 1) {
      e.out = dA_.hi; e.out_lo = dA_.lo; e.ld_out = dA_.ld;
      e.xbits = in_[v / 2].xbits; e.xb_ld = xb_ld_;             // packed lock plane (the reconstruction target)
      e.alpha = inv_bg; e.scalar = statbuf_ + 0; e.precise = prec;
      if (const char* d = getenv("MAGIC_EPI_DBG")) e.dbg = atoi(d);   // timing experiments only: results are wrong
    } else {
      e.out = nullptr; e.ld_out = D;                             // output pointer patched per call
    }
    e.aux_split = 1 << 30;
    p.e[0] = e;
    p.cta_pair = true;
    // MAGIC_CLUSTER_OUT=2|4: CTAs along M form a cluster and receive the weight tile by one multicast load.  Measured
    // on B200 (profiles/r01g): no gain at cluster sizes <= 4 (L2 already merges the concurrent unicast requests) and
    // only 33 clusters of 4 are co-resident (132 of 148 SMs), so the default stays 1.
    {
      const char* cm = getenv("MAGIC_CLUSTER_OUT");
      p.cluster_m = cm ? atoi(cm) : 1;
    }
    GemmPlan* dst = v == 0 ? &pl_out_bce_ : (v == 1 ? &pl_out_sig_ : &pl_out_bce_alt_);
    MAGIC_TRY(finish(p, dst, "fwd dec out"));
    if (v != 0) flops -= dst->flops;
  }

  // ---------------- backward of `cost` through the decoder
  zero_.clear();
  std::vector<GemmProblem>& small_wg = small_wg_;
  small_wg_.clear(); small_wg_layer_.clear();
  GemmProblem enc0_wg[2];
  auto wgrad1 = [&](GemmPlan* pl, const Act& x, int xcols, const Act& dy, int layer) -> std::string {
    GemmProblem p;
    p.kind = GEMM_WGRAD; p.epi = EPI_STORE_F32; p.M = xcols; p.N = lin[layer].Nv; p.K = B;
    add_segments(p, 0, amat(x, false, 0, B, xcols), amat(x, true, 0, B, xcols), amat(dy, false, 0, B, p.N),
                 amat(dy, true, 0, B, p.N), false);
    p.e[0] = f32_epi(g1_ + lin[layer].off, lin[layer].ld, inv_bg);
    p.cta_pair = layer == iOut;
    MAGIC_TRY(finish(p, pl, "wgrad dec"));
    if (layer != iOut) { small_wg.push_back(p); small_wg_layer_.push_back(layer); }
    if (pl->args.splitk > 1) zero_.push_back({g1_ + lin[layer].off, sizeof(float) * (size_t)lin[layer].size()});
    return "";
  };
  auto dgrad1 = [&](GemmPlan* pl, const Act& dy, int layer, const Act& y, const Act& out) -> std::string {
    GemmProblem p;
    p.kind = GEMM_DGRAD; p.epi = EPI_DACT; p.M = B; p.K = lin[layer].Nv; p.N = (int)out.ld;
    add_segments(p, 0, amat(dy, false, 0, B, p.K), amat(dy, true, 0, B, p.K), wmat(layer, false), wmat(layer, true), false);
    p.e[0] = dact_epi(out, y, lin[layer].Kv, 1 << 30, 0);
    p.ws = ws_; p.ws_ld = out.ld;
    p.cta_pair = layer == iOut;
    return finish(p, pl, "dgrad dec");
  };
  MAGIC_TRY(wgrad1(&pl_wg_out_, D2_, c.dec[1] + 1, dA_, iOut));
  MAGIC_TRY(dgrad1(&pl_dg_out_, dA_, iOut, D2_, dD2_));
  MAGIC_TRY(wgrad1(&pl_wg_h2_, D1_, c.dec[0] + 1, dD2_, iD2));
  MAGIC_TRY(dgrad1(&pl_dg_h2_, dD2_, iD2, D1_, dD1_));
  MAGIC_TRY(wgrad1(&pl_wg_h1_, zdec_, L + 1, dD1_, iD1));
  {
    GemmProblem p;   // d z_lock (unscaled), fp32
    p.kind = GEMM_DGRAD; p.epi = EPI_STORE_F32; p.M = B; p.K = lin[iD1].Nv; p.N = L;
    add_segments(p, 0, amat(dD1_, false, 0, B, p.K), amat(dD1_, true, 0, B, p.K), wmat(iD1, false), wmat(iD1, true), false);
    p.e[0] = f32_epi(dz_dec_, ld_dz_, 1.f);
    p.force_splitk = 1;
    MAGIC_TRY(finish(p, &pl_dg_h1_, "dgrad dec h1"));
  }

  // ---------------- backward through the encoder: four streams [rot|cost ; lock|cost ; lock|metric ; key|metric]
  auto wgrad2 = [&](GemmPlan* pl, const Act& x, int xcols, bool x_exact, const Act& dy, int layer,
                    const uint32_t* xb = nullptr) -> std::string {
    GemmProblem p;
    p.kind = GEMM_WGRAD; p.epi = EPI_STORE_F32; p.ngroups = 2; p.M = xcols; p.N = lin[layer].Nv; p.K = 2 * B;
    if (xb) {   // packed rows [rot; lock; key]: optimiser 1 reads rows [0, 2B), optimiser 2 rows [B, 3B)
      p.a_bits = true;
      p.abits[0].base = xb; p.abits[0].pitch = xb_ld_; p.abits[0].rows = 2 * B;
      p.abits[1].base = xb + (size_t)B * xb_ld_; p.abits[1].pitch = xb_ld_; p.abits[1].rows = 2 * B;
      const uint32_t* km = xb == in_[0].xb3 ? in_[0].kmask : in_[1].kmask;
      p.kmask[0] = km + km_wg_off_[0]; p.kmask[1] = km + km_wg_off_[1]; p.kmask_words = km_wg_words_;
    }
    // optimiser 1 (cost): x rows [rot; lock] = [0, 2B), dy rows [0, 2B).  optimiser 2 (metric): x rows [lock; key] =
    // [B, 3B), dy rows [2B, 4B).
    add_segments(p, 0, amat(x, false, 0, 2LL * B, xcols), amat(x, true, 0, 2LL * B, xcols), amat(dy, false, 0, 2LL * B, p.N),
                 amat(dy, true, 0, 2LL * B, p.N), x_exact);
    add_segments(p, 1, amat(x, false, B, 2LL * B, xcols), amat(x, true, B, 2LL * B, xcols),
                 amat(dy, false, 2LL * B, 2LL * B, p.N), amat(dy, true, 2LL * B, 2LL * B, p.N), x_exact);
    p.e[0] = f32_epi(g1_ + lin[layer].off, lin[layer].ld, 1.f);
    p.e[1] = f32_epi(g2_ + lin[layer].off, lin[layer].ld, 1.f);
    p.cta_pair = layer == 0;
    MAGIC_TRY(finish(p, pl, "wgrad enc"));
    if (layer != 0) { small_wg.push_back(p); small_wg_layer_.push_back(layer); }
    else enc0_wg[pl == &pl_wg_enc0_alt_ ? 1 : 0] = p;
    if (pl->args.splitk > 1) {
      zero_.push_back({g1_ + lin[layer].off, sizeof(float) * (size_t)lin[layer].size()});
      zero_.push_back({g2_ + lin[layer].off, sizeof(float) * (size_t)lin[layer].size()});
    }
    return "";
  };
  auto dgrad2 = [&](GemmPlan* pl, const Act& dy, int layer, const Act& y, const Act& out) -> std::string {
    GemmProblem p;
    p.kind = GEMM_DGRAD; p.epi = EPI_DACT; p.M = 4 * B; p.K = lin[layer].Nv; p.N = (int)out.ld;
    add_segments(p, 0, amat(dy, false, 0, 4LL * B, p.K), amat(dy, true, 0, 4LL * B, p.K), wmat(layer, false),
                 wmat(layer, true), false);
    p.e[0] = dact_epi(out, y, lin[layer].Kv, 2 * B, B);   // y rows: r < 2B -> r ; else r - B
    p.ws = ws_; p.ws_ld = out.ld;
    return finish(p, pl, "dgrad enc");
  };
  // (without the ones column: the heads' bias gradients are summed in fp32 by latent_backward, see LatentArgs.g1_bias)
  MAGIC_TRY(wgrad2(&pl_wg_head_, H_[E - 1], c.enc[E - 1], false, dheads_, iHead));
  MAGIC_TRY(dgrad2(&pl_dg_head_, dheads_, iHead, H_[E - 1], dH_[E - 1]));
  pl_wg_enc_.assign(E, GemmPlan());
  pl_dg_enc_.assign(E, GemmPlan());
  for (int i = E - 1; i >= 1; --i) {
    MAGIC_TRY(wgrad2(&pl_wg_enc_[i], H_[i - 1], c.enc[i - 1] + 1, false, dH_[i], i));
    MAGIC_TRY(dgrad2(&pl_dg_enc_[i], dH_[i], i, H_[i - 1], dH_[i - 1]));
  }
  MAGIC_TRY(wgrad2(&pl_wg_enc_[0], in_[0].X, D + 1, true, dH_[0], 0, bits_a_ ? in_[0].xb3 : nullptr));
  {
    const size_t nz = zero_.size();
    MAGIC_TRY(wgrad2(&pl_wg_enc0_alt_, in_[1].X, D + 1, true, dH_[0], 0, bits_a_ ? in_[1].xb3 : nullptr));
    zero_.resize(nz);                        // same gradient windows as the first variant
    flops -= pl_wg_enc0_alt_.flops;
  }
  // Single GPU: the small layers' weight gradients (76 whole-K tiles at the 8d shape) ride in the last, partly empty
  // wave of the first-layer launch instead of six launches of their own; no split-K, so no gradient windows to clear.
  fuse_wg_ = false;
  {
    const char* fw = getenv("MAGIC_FUSE_WGRAD");
    bool want = c.world == 1 && !(fw && atoi(fw) == 0);
    size_t nx = 0;
    for (auto& sp : small_wg) { nx += sp.ngroups; want = want && sp.nseg == enc0_wg[0].nseg; }
    if (want && nx <= (size_t)kMaxExtra) {
      for (int v = 0; v < 2; ++v) {
        GemmProblem p = enc0_wg[v];
        for (auto& sp : small_wg)
          for (int g = 0; g < sp.ngroups; ++g) {
            GemmExtra x;
            for (int sg = 0; sg < sp.nseg; ++sg) { x.a[sg] = sp.a[g][sg]; x.b[sg] = sp.b[g][sg]; }
            x.e = sp.e[g]; x.M = sp.M; x.N = sp.N; x.K = sp.K;
            p.extra.push_back(x);
          }
        MAGIC_TRY(finish(p, &pl_wg_enc0_fused_[v], "wgrad enc0 + small layers"));
        flops -= pl_wg_enc0_fused_[v].flops;
      }
      fuse_wg_ = true;
    }
  }
  // ---------------- the first layer's dual Adam inside that launch's epilogue
  fuse_adam_ok_ = fuse_adam_plans_ok_ = false;
  {
    const char* fa = getenv("MAGIC_FUSE_ADAM");
    if (c.world == 1 && bits_a_ && !(fa && atoi(fa) < 0)) {          // (MAGIC_FUSE_ADAM=-1: do not even build the plans)
      AdamFuse af;
      memset(&af, 0, sizeof af);
      const long long o = lin[0].off;
      af.p = p_ + o; af.m[0] = m1_ + o; af.v[0] = v1_ + o; af.m[1] = m2_ + o; af.v[1] = v2_ + o;
      af.w_hi = w_hi_ + o; af.w_lo = w_lo_ ? w_lo_ + o : nullptr;
      af.ld = lin[0].ld; af.beta1 = c.beta1; af.beta2 = c.beta2; af.eps = c.eps;
      if (const char* d = getenv("MAGIC_ADAM_DBG")) af.dbg = atoi(d);   // timing experiments only: results are wrong
      std::string why;
      for (int v = 0; v < 2 && why.empty(); ++v) {
        GemmProblem p = enc0_wg[v];
        if (fuse_wg_)
          for (auto& sp : small_wg)
            for (int g = 0; g < sp.ngroups; ++g) {
              GemmExtra x;
              for (int sg = 0; sg < sp.nseg; ++sg) { x.a[sg] = sp.a[g][sg]; x.b[sg] = sp.b[g][sg]; }
              x.e = sp.e[g]; x.M = sp.M; x.N = sp.N; x.K = sp.K;
              p.extra.push_back(x);
            }
        for (auto& k : p.kmask) k = nullptr;             // dense: no k-block lists (the zero-stage option keeps the plain launch)
        p.kmask_words = 0;
        why = finish(p, &pl_wg_enc0_adam_[v], "wgrad enc0 + Adam");
        flops -= pl_wg_enc0_adam_[v].flops;
        if (why.empty()) why = gemm_fuse_adam(&pl_wg_enc0_adam_[v], af);
      }
      fuse_adam_plans_ok_ = why.empty();
      // OFF unless asked for (MAGIC_FUSE_ADAM=1 or magic_set_option): correct, but no faster.  The main loop of this
      // launch lives on L2 hits for its B operand and is latency-bound (4 stages in flight); with 1.2 GB of optimiser
      // state streaming through the same L2 / HBM queues its tensor pipe drops from 77 % to 53 % busy, and the launch
      // grows by what the separate Adam pass costs (0.47 -> 0.66 ms alone, 0.52 -> 0.82 ms in the step;
      // profiles/r02q_*).  DESIGN.md section 4, "measured and rejected".
      fuse_adam_ok_ = fuse_adam_plans_ok_ && fa && atoi(fa) != 0;
      if (getenv("MAGIC_VERBOSE"))
        fprintf(stderr, "[magic] first-layer Adam in the wgrad epilogue: %s\n",
                fuse_adam_ok_ ? ("rows < " + std::to_string(pl_wg_enc0_adam_[0].adam_rows_end)).c_str() : why.c_str());
    }
  }
  // ---------------- hidden-layer chains (MAGIC_CHAIN=0: every layer its own launch)
  chain_fwd_.valid = chain_dg_.valid = false;
  chain_dg_has_head_ = false;
  {
    // MAGIC_CHAIN: 0 = every layer its own launch, 1 = forward chain only (default), 2 = forward and dgrad chains.  The
    // dgrad chain sits where the next batch's generator / dataset feed runs on its own stream: its 230 KB of shared memory
    // cannot share an SM with one of their CTAs, and the step with a new batch every iteration loses 5 % with it
    // (measured: 1.838 M pairs/s at level 2, 1.943 M at level 1, 1.940 M at level 0).
    const char* ch = getenv("MAGIC_CHAIN");
    const int chain_level = ch ? atoi(ch) : 1;
    if (!split && chain_level > 0) {
      std::vector<const GemmPlan*> ls;
      std::vector<Mat> ws;
      std::vector<ChainOut> outs;
      for (int i = 1; i < E; ++i) { ls.push_back(&fwd_enc_[i]); ws.push_back(wmat(i, false)); outs.push_back({H_[i].hi, H_[i].rows, H_[i].ld}); }
      std::string why = chain_plan(ls, ws, outs, num_sms, &chain_fwd_);
      if (getenv("MAGIC_VERBOSE")) fprintf(stderr, "[magic] forward chain: %s\n", chain_fwd_.valid ? "on" : why.c_str());
      ls.clear(); ws.clear(); outs.clear();
      chain_dg_has_head_ = pl_dg_head_.args.kblocks <= kChainKb && E <= kChainMaxLayers;
      if (chain_dg_has_head_) { ls.push_back(&pl_dg_head_); ws.push_back(wmat(iHead, false)); outs.push_back({dH_[E - 1].hi, dH_[E - 1].rows, dH_[E - 1].ld}); }
      for (int i = E - 1; i >= 1; --i) { ls.push_back(&pl_dg_enc_[i]); ws.push_back(wmat(i, false)); outs.push_back({dH_[i - 1].hi, dH_[i - 1].rows, dH_[i - 1].ld}); }
      why = chain_level >= 2 ? chain_plan(ls, ws, outs, num_sms, &chain_dg_) : std::string("off (MAGIC_CHAIN < 2)");
      if (chain_level >= 2 && !chain_dg_.valid && chain_dg_has_head_) {         // the heads' dgrad does not fit the chain: try the hidden layers alone
        chain_dg_has_head_ = false;
        ls.erase(ls.begin()); ws.erase(ws.begin()); outs.erase(outs.begin());
        why = chain_plan(ls, ws, outs, num_sms, &chain_dg_);
      }
      if (!chain_dg_.valid) chain_dg_has_head_ = false;
      if (getenv("MAGIC_VERBOSE")) fprintf(stderr, "[magic] dgrad chain: %s%s\n", chain_dg_.valid ? "on" : why.c_str(),
                                           chain_dg_.valid && chain_dg_has_head_ ? " (with the heads' dgrad)" : "");
    }
  }
  stats.gemm_flops_per_step = flops;
  return "";
}

// ------------------------------------------------------------------ inputs
// Every writer targets the buffer that is NOT loaded and then makes it the loaded one; `ready` orders the generator's
// stream before the step's stream, `consumed` orders the step's last read before the generator's next write.
// A staged host feed becomes a batch: rotation + layout kernel on the input stream.  From a training step it is launched
// behind the step's gate (the end of its forward pass), like the generator: it then shares the GPU with the backward
// GEMMs instead of landing on the next forward pass -- whose first-layer GEMM is latency-bound and does not tolerate
// a neighbour (launched at feed time: 2.30 ms per end-to-end step against 2.02 ms with the batch resident).
void Engine::flush_expand(InBuf& b, bool behind_gate) {
  if (!b.expand_pending) return;
  b.expand_pending = false;
  const int side = (int)llround(sqrt((double)D));
  if (behind_gate) cudaStreamWaitEvent(gen_stream_, ev_gate_, 0);
  prof_begin("input.pairs_bits", 0, gen_stream_);
  pairs_bits(bits_stage_, bits_stage_ + (size_t)B * b.exp_wpr, b.exp_wpr, B, side,
             b.exp_seed + 0x9E3779B97F4A7C15ull * (uint64_t)cfg.rank, b.exp_step, bits_a_ ? nullptr : b.X.hi, b.X.ld, b.xb3,
             xb_ld_, b.rot, num_sms, gen_stream_);
  build_kmasks(b, gen_stream_);
  prof_end(gen_stream_);
  cudaEventRecord(b.ready, gen_stream_);
  b.ready_pending = true;
}
void Engine::wait_ready(InBuf& b) {
  flush_expand(b, false);
  if (b.ready_pending) { cudaStreamWaitEvent(stream, b.ready, 0); b.ready_pending = false; }
}
void Engine::mark_consumed(InBuf& b) {
  cudaEventRecord(b.consumed, stream);
  b.consumed_pending = true;
}
// The engine keeps images as 1 bit per pixel (first-layer operand, BCE target).  The reference's pipeline only ever
// produces {0,1} (PNG {0,255} / 255, nearest-neighbour rotation: 8c/overlap_input.py:117, 8c/utils.py:337-338), but its
// placeholder would accept any float: an array feed with another value must fail loudly, not be thresholded.
std::string Engine::check_binary(const char* who) {
  int flag = 0;
  MAGIC_CUDA(cudaMemcpyAsync(&flag, not_binary_, sizeof(int), cudaMemcpyDeviceToHost, stream));
  MAGIC_CUDA(cudaStreamSynchronize(stream));
  if (flag) {
    MAGIC_CUDA(cudaMemsetAsync(not_binary_, 0, sizeof(int), stream));
    in_[cur_ ^ 1].filled = false;
    return std::string(who) + ": X holds values other than 0 and 1 (uint8: 0, 1, 255); the B200 backend trains on binary "
           "images, as the reference's input pipeline produces them (8c/overlap_input.py:117)";
  }
  return "";
}
std::string Engine::load_interleaved(const float* d_x) {
  if (!d_x) return "load_interleaved: null input";
  InBuf& b = in_[cur_ ^ 1];
  wait_ready(b);
  prof_begin("input.deinterleave");
  deinterleave_f32(d_x, B, D, b.X.hi, b.X.ld, not_binary_, stream);
  pack_planes_bits(b.X.hi, b.X.ld, 3 * B, D, b.xb3, xb_ld_, stream);
  build_kmasks(b, stream);
  b.planes_valid = true;
  prof_end();
  MAGIC_TRY(check_binary("load_interleaved"));
  MAGIC_CUDA(cudaGetLastError());
  mark_consumed(b);
  b.has_labels = false;
  b.has_rot = false;
  b.filled = true;
  cur_ ^= 1;
  return "";
}
std::string Engine::load_interleaved_u8(const uint8_t* d_x) {
  if (!d_x) return "load_interleaved_u8: null input";
  InBuf& b = in_[cur_ ^ 1];
  wait_ready(b);
  prof_begin("input.deinterleave");
  deinterleave_u8(d_x, B, D, b.X.hi, b.X.ld, not_binary_, stream);
  pack_planes_bits(b.X.hi, b.X.ld, 3 * B, D, b.xb3, xb_ld_, stream);
  build_kmasks(b, stream);
  b.planes_valid = true;
  prof_end();
  MAGIC_TRY(check_binary("load_interleaved_u8"));
  MAGIC_CUDA(cudaGetLastError());
  mark_consumed(b);
  b.has_labels = false;
  b.has_rot = false;
  b.filled = true;
  cur_ ^= 1;
  return "";
}
std::string Engine::load_planar_u8(const uint8_t* l, const uint8_t* r, const uint8_t* k) {
  if (!l || !r || !k) return "load_planar_u8: null input";
  InBuf& b = in_[cur_ ^ 1];
  wait_ready(b);
  planar_u8(l, r, k, B, D, b.X.hi, b.X.ld, not_binary_, stream);
  pack_planes_bits(b.X.hi, b.X.ld, 3 * B, D, b.xb3, xb_ld_, stream);
  build_kmasks(b, stream);
  b.planes_valid = true;
  MAGIC_TRY(check_binary("load_planar_u8"));
  MAGIC_CUDA(cudaGetLastError());
  mark_consumed(b);
  b.has_labels = false;
  b.has_rot = false;
  b.filled = true;
  cur_ ^= 1;
  return "";
}
std::string Engine::load_pairs_u8(const uint8_t* d_lock, const uint8_t* d_key, uint64_t seed, uint64_t step_idx) {
  if (!d_lock || !d_key) return "load_pairs_u8: null input";
  const int side = (int)llround(sqrt((double)D));
  if (side * side != D) return "load_pairs_u8: D must be a square canvas";
  InBuf& b = in_[cur_ ^ 1];
  wait_ready(b);
  pairs_u8(d_lock, d_key, B, side, seed + 0x9E3779B97F4A7C15ull * (uint64_t)cfg.rank, step_idx, b.X.hi, b.X.ld, b.rot, stream);
  pack_planes_bits(b.X.hi, b.X.ld, 3 * B, D, b.xb3, xb_ld_, stream);
  build_kmasks(b, stream);
  b.planes_valid = true;
  MAGIC_CUDA(cudaGetLastError());
  mark_consumed(b);
  b.has_labels = false;
  b.has_rot = true;
  b.filled = true;
  cur_ ^= 1;
  return "";
}
std::string Engine::load_pairs_bits(const uint32_t* d_lock, const uint32_t* d_key, int wpr, uint64_t seed, uint64_t step_idx) {
  if (!d_lock || !d_key) return "load_pairs_bits: null input";
  const int side = (int)llround(sqrt((double)D));
  if (side * side != D || side > 256) return "load_pairs_bits: D must be a square canvas with side <= 256";
  if (wpr < (D + 31) / 32) return "load_pairs_bits: words_per_row is smaller than ceil(D / 32)";
  InBuf& b = in_[cur_ ^ 1];
  wait_ready(b);
  pairs_bits(d_lock, d_key, wpr, B, side, seed + 0x9E3779B97F4A7C15ull * (uint64_t)cfg.rank, step_idx,
             bits_a_ ? nullptr : b.X.hi, b.X.ld, b.xb3, xb_ld_, b.rot, num_sms, stream);
  build_kmasks(b, stream);
  b.planes_valid = !bits_a_;
  MAGIC_CUDA(cudaGetLastError());
  mark_consumed(b);
  b.has_labels = false;
  b.has_rot = true;
  b.filled = true;
  cur_ ^= 1;
  return "";
}
// Host-buffer variant for a training loop: copies and expands on the input stream (the generator's), so the batch
// of step n+1 crosses PCIe and is unpacked while step n runs; with activate = 0 the batch stays staged in the
// not-loaded buffer until swap_batch makes it the loaded one.
std::string Engine::feed_pairs_bits_host(const uint32_t* h_lock, const uint32_t* h_key, const float* h_labels, int wpr,
                                         uint64_t seed, uint64_t step_idx, bool activate) {
  if (!h_lock || !h_key) return "feed_pairs_bits_host: null input";
  const int side = (int)llround(sqrt((double)D));
  if (side * side != D || side > 256) return "feed_pairs_bits_host: D must be a square canvas with side <= 256";
  if (wpr < (D + 31) / 32) return "feed_pairs_bits_host: words_per_row is smaller than ceil(D / 32)";
  const size_t plane = sizeof(uint32_t) * (size_t)B * wpr;
  if (bits_stage_bytes_ < 2 * plane) {
    flush_expand(in_[0], false); flush_expand(in_[1], false);     // a staged batch still waiting in the old staging area
    MAGIC_CUDA(cudaStreamSynchronize(gen_stream_));
    if (bits_stage_) cudaFree(bits_stage_);
    bits_stage_ = nullptr; bits_stage_bytes_ = 0;
    MAGIC_CUDA(cudaMalloc((void**)&bits_stage_, 2 * plane));
    bits_stage_bytes_ = 2 * plane;
  }
  InBuf& b = in_[cur_ ^ 1];
  flush_expand(in_[cur_], false);       // one staging area: a feed still waiting in the loaded buffer is expanded before it is overwritten
  b.expand_pending = false;             // ... and one waiting in the target buffer is replaced by this one
  if (b.ready_pending) { MAGIC_CUDA(cudaStreamWaitEvent(gen_stream_, b.ready, 0)); b.ready_pending = false; }
  if (b.consumed_pending) { MAGIC_CUDA(cudaStreamWaitEvent(gen_stream_, b.consumed, 0)); b.consumed_pending = false; }
  uint32_t* d_lock = bits_stage_;
  uint32_t* d_key = bits_stage_ + (size_t)B * wpr;
  MAGIC_CUDA(cudaMemcpyAsync(d_lock, h_lock, plane, cudaMemcpyHostToDevice, gen_stream_));
  MAGIC_CUDA(cudaMemcpyAsync(d_key, h_key, plane, cudaMemcpyHostToDevice, gen_stream_));
  if (h_labels) MAGIC_CUDA(cudaMemcpyAsync(b.labels, h_labels, sizeof(float) * B, cudaMemcpyHostToDevice, gen_stream_));
  b.expand_pending = true;              // (a staged batch that was never made the loaded one is simply replaced)
  b.exp_seed = seed; b.exp_step = step_idx; b.exp_wpr = wpr;
  b.planes_valid = !bits_a_;
  // activate: the caller trains on this batch next -> expand now.  Staged: the copies cross PCIe now, the kernel
  // follows from the next step (behind its forward pass) or from whoever touches the buffer first (wait_ready).
  if (activate) flush_expand(b, false);
  MAGIC_CUDA(cudaGetLastError());
  b.has_labels = h_labels != nullptr;
  b.has_rot = true;
  b.filled = true;
  if (activate) cur_ ^= 1;
  return "";
}
// The same with the GATHER on the device: the caller keeps the whole packed dataset in pinned host memory and passes the
// rows of this batch as indices; a thin kernel on the input stream pulls the rows over PCIe into the staging area (41 MB per
// 4096 pairs, the same bytes the copy engine moved before) and the host does no per-step work beyond 16 KB of indices and
// labels.  With the gather on the host, eight ranks on one 32-core box were bound by their index_select threads (5.2 ms
// per end-to-end step against 2.3 ms of GPU time, profiles/r02z_bench_n8.json).
std::string Engine::feed_pairs_bits_gather(const uint32_t* h_lock_all, const uint32_t* h_key_all, int wpr, long long n_rows,
                                           const int32_t* h_idx, const float* h_labels, uint64_t seed, uint64_t step_idx,
                                           bool activate) {
  if (!h_lock_all || !h_key_all || !h_idx) return "feed_pairs_bits_gather: null input";
  const int side = (int)llround(sqrt((double)D));
  if (side * side != D || side > 256) return "feed_pairs_bits_gather: D must be a square canvas with side <= 256";
  if (wpr < (D + 31) / 32 || wpr % 4 != 0) return "feed_pairs_bits_gather: words_per_row must cover D pixels and be a multiple of 4";
  if (n_rows <= 0) return "feed_pairs_bits_gather: empty dataset";
  void *d_lock_all = nullptr, *d_key_all = nullptr;
  if (cudaHostGetDevicePointer(&d_lock_all, const_cast<uint32_t*>(h_lock_all), 0) != cudaSuccess ||
      cudaHostGetDevicePointer(&d_key_all, const_cast<uint32_t*>(h_key_all), 0) != cudaSuccess) {
    (void)cudaGetLastError();
    return "feed_pairs_bits_gather: the dataset arrays must be pinned host memory (cudaHostAlloc / cudaHostRegister / torch pin_memory)";
  }
  const size_t plane = sizeof(uint32_t) * (size_t)B * wpr;
  if (bits_stage_bytes_ < 2 * plane) {
    flush_expand(in_[0], false); flush_expand(in_[1], false);     // a staged batch still waiting in the old staging area
    MAGIC_CUDA(cudaStreamSynchronize(gen_stream_));
    if (bits_stage_) cudaFree(bits_stage_);
    bits_stage_ = nullptr; bits_stage_bytes_ = 0;
    MAGIC_CUDA(cudaMalloc((void**)&bits_stage_, 2 * plane));
    bits_stage_bytes_ = 2 * plane;
  }
  if (!gather_idx_) MAGIC_TRY(alloc((void**)&gather_idx_, sizeof(int32_t) * B));
  InBuf& b = in_[cur_ ^ 1];
  flush_expand(in_[cur_], false);
  b.expand_pending = false;
  if (b.ready_pending) { MAGIC_CUDA(cudaStreamWaitEvent(gen_stream_, b.ready, 0)); b.ready_pending = false; }
  if (b.consumed_pending) { MAGIC_CUDA(cudaStreamWaitEvent(gen_stream_, b.consumed, 0)); b.consumed_pending = false; }
  MAGIC_CUDA(cudaMemcpyAsync(gather_idx_, h_idx, sizeof(int32_t) * B, cudaMemcpyHostToDevice, gen_stream_));
  if (h_labels) MAGIC_CUDA(cudaMemcpyAsync(b.labels, h_labels, sizeof(float) * B, cudaMemcpyHostToDevice, gen_stream_));
  prof_begin("input.gather_rows", 0, gen_stream_);
  gather_rows_host(static_cast<const uint32_t*>(d_lock_all), static_cast<const uint32_t*>(d_key_all), gather_idx_, B, wpr, n_rows,
                   bits_stage_, bits_stage_ + (size_t)B * wpr, gen_stream_);
  prof_end(gen_stream_);
  b.expand_pending = true;
  b.exp_seed = seed; b.exp_step = step_idx; b.exp_wpr = wpr;
  b.planes_valid = !bits_a_;
  if (activate) flush_expand(b, false);
  MAGIC_CUDA(cudaGetLastError());
  b.has_labels = h_labels != nullptr;
  b.has_rot = true;
  b.filled = true;
  if (activate) cur_ ^= 1;
  return "";
}
std::string Engine::generate(uint64_t seed, uint64_t step_idx) {
  const int side = (int)llround(sqrt((double)D));
  if (side * side != D || side % 8 != 0 || side > 256)
    return "generate: D must be a square canvas with side % 8 == 0 and side <= 256";
  InBuf& b = in_[cur_ ^ 1];
  b.expand_pending = false;             // a staged host feed in this buffer is replaced
  // the generator runs on its own stream so that it overlaps the (tensor-bound) step still in flight on `stream`
  if (b.consumed_pending) { MAGIC_CUDA(cudaStreamWaitEvent(gen_stream_, b.consumed, 0)); b.consumed_pending = false; }
  if (gate_pending_ && gen_gate_) { MAGIC_CUDA(cudaStreamWaitEvent(gen_stream_, ev_gate_, 0)); }
  gate_pending_ = false;
  // fold the rank into the seed so every rank draws different pairs
  prof_begin("input.generate", 0, gen_stream_);
  generate_pairs(seed + 0x9E3779B97F4A7C15ull * (uint64_t)cfg.rank, step_idx, B, side, bits_a_ ? nullptr : b.X.hi, b.X.ld,
                 b.xb3, xb_ld_, b.labels, b.rot, num_sms, gen_stream_);
  build_kmasks(b, gen_stream_);
  b.planes_valid = !bits_a_;
  prof_end(gen_stream_);
  MAGIC_CUDA(cudaGetLastError());
  MAGIC_CUDA(cudaEventRecord(b.ready, gen_stream_));
  b.ready_pending = true;
  b.has_labels = true;
  b.filled = true;
  cur_ ^= 1;
  return "";
}
// Make the OTHER resident input buffer the loaded one again (it keeps the batch of the load/generate call before the
// last one).  Lets a caller alternate between two batches that stay in HBM -- the reference re-feeds a held batch the
// same way (8c/main.py:66-80 evaluates on a batch it dequeued earlier).
std::string Engine::swap_batch() {
  if (!in_[cur_ ^ 1].filled) return "swap_batch: the other input buffer has never been loaded";
  cur_ ^= 1;
  return "";
}
std::string Engine::export_rotation(int32_t* d_cs_sn) {
  if (!d_cs_sn) return "export_rotation: null output";
  InBuf& b = in_[cur_];
  if (!b.has_labels && !b.has_rot) return "export_rotation: the loaded batch was not rotated on the device";
  wait_ready(b);
  MAGIC_CUDA(cudaMemcpyAsync(d_cs_sn, b.rot, sizeof(int) * 2 * (size_t)B, cudaMemcpyDeviceToDevice, stream));
  mark_consumed(b);
  return "";
}
std::string Engine::export_latents(float* d_z3) {
  if (!d_z3) return "export_latents: null output";
  const size_t n = sizeof(float) * (size_t)B * L;          // z_ rows are [rot; lock; key]; the API order is lock, rot, key
  MAGIC_CUDA(cudaMemcpyAsync(d_z3, z_ + (size_t)B * L, n, cudaMemcpyDeviceToDevice, stream));
  MAGIC_CUDA(cudaMemcpyAsync(d_z3 + (size_t)B * L, z_, n, cudaMemcpyDeviceToDevice, stream));
  MAGIC_CUDA(cudaMemcpyAsync(d_z3 + 2 * (size_t)B * L, z_ + 2 * (size_t)B * L, n, cudaMemcpyDeviceToDevice, stream));
  return "";
}
std::string Engine::export_batch(float* d_x, float* d_labels) {
  InBuf& b = in_[cur_];
  wait_ready(b);
  if (d_x) {
    if (!b.planes_valid) {     // the batch only exists as packed rows: expand it once
      unpack_bits_planes(b.xb3, xb_ld_, 3 * B, b.X.hi, b.X.ld, stream);
      b.planes_valid = true;
    }
    planes_to_interleaved_f32(b.X.hi, b.X.ld, B, D, d_x, stream);
  }
  if (d_labels) MAGIC_CUDA(cudaMemcpyAsync(d_labels, b.labels, sizeof(float) * B, cudaMemcpyDeviceToDevice, stream));
  MAGIC_CUDA(cudaGetLastError());
  mark_consumed(b);
  return "";
}

// Which 128 x 64 stages of the packed rows hold a set pixel, for the forward GEMM and for the two weight-gradient
// groups (rows [0, 2B) and [B, 3B)).  Runs behind whatever wrote the packed rows, on the same stream.
void Engine::build_kmasks(InBuf& b, cudaStream_t s, bool force) {
  if (!bits_a_) return;
  // Off by default: on the reference's images 94-97 % of the stages hold a set pixel (tests/test_shape_stats.py), so the
  // maps, lists and work order of every batch cost more than the skipped stages save; sparse data sets switch it on.
  if (!skip_zero_ && !force) { b.kmask_valid = false; return; }
  b.kmask_valid = true;
  prof_begin("input.kmask", 0, s);
  build_kmask_fwd(b.xb3, xb_ld_, 3 * B, km_fwd_kb_, b.kmask, km_fwd_words_, s);
  for (int g = 0; g < 2; ++g)
    build_kmask_wgrad(b.xb3 + (size_t)g * B * xb_ld_, xb_ld_, 2 * B, km_wg_tiles_, b.kmask + km_wg_off_[g], km_wg_words_, s);
  // the plans that read this buffer get their work items re-dealt by the new maps (GemmArgs.order)
  const int alt = &b == &in_[1] ? 1 : 0;
  gemm_schedule(alt ? fwd_enc0_alt_ : fwd_enc_[0], s);
  gemm_schedule(dp_fused_ ? pl_wg_enc0_sc_[alt] : fuse_wg_ ? pl_wg_enc0_fused_[alt] : (alt ? pl_wg_enc0_alt_ : pl_wg_enc_[0]), s);
  prof_end(s);
}
std::string Engine::set_option(const char* name, int value) {
  if (!name) return "set_option: null name";
  if (std::string(name) == "skip_zero_stages") { skip_zero_ = value != 0; return ""; }
  if (std::string(name) == "fuse_first_layer_adam") {
    if (value != 0 && !fuse_adam_plans_ok_) return "set_option: fuse_first_layer_adam is not available for this handle (data parallel, or bf16 planes as the first layer's operand)";
    fuse_adam_ok_ = value != 0;
    return "";
  }
  return std::string("set_option: unknown option ") + name;
}
std::string Engine::read_losses(float* h_losses5) {
  if (!h_losses5) return "read_losses: null output";
  if (!losses_src_) return "read_losses: no step with a loss buffer has been enqueued on this handle";
  MAGIC_CUDA(cudaStreamWaitEvent(read_stream_, ev_losses_, 0));
  MAGIC_CUDA(cudaMemcpyAsync(h_losses_, losses_src_, sizeof(float) * 5, cudaMemcpyDeviceToHost, read_stream_));
  MAGIC_CUDA(cudaStreamSynchronize(read_stream_));
  memcpy(h_losses5, h_losses_, sizeof(float) * 5);
  return "";
}
std::string Engine::input_density(double* out2) {
  if (!out2) return "input_density: null output";
  out2[0] = out2[1] = 1.0;
  if (!bits_a_) return "";
  InBuf& b = in_[cur_];
  if (!b.filled) return "input_density: no batch loaded";
  wait_ready(b);
  if (!b.kmask_valid) build_kmasks(b, stream, true);      // maps are only kept up to date while skip_zero_stages is on
  std::vector<uint32_t> h(km_total_);
  MAGIC_CUDA(cudaMemcpyAsync(h.data(), b.kmask, sizeof(uint32_t) * km_total_, cudaMemcpyDeviceToHost, stream));
  MAGIC_CUDA(cudaStreamSynchronize(stream));
  long long set = 0;
  for (size_t i = 0; i < (size_t)km_fwd_tiles_ * km_fwd_words_; ++i) set += __builtin_popcount(h[i]);
  out2[0] = (double)set / ((double)km_fwd_tiles_ * km_fwd_kb_);
  set = 0;
  for (int g = 0; g < 2; ++g)
    for (size_t i = 0; i < (size_t)km_wg_tiles_ * km_wg_words_; ++i) set += __builtin_popcount(h[km_wg_off_[g] + i]);
  out2[1] = (double)set / (2.0 * km_wg_tiles_ * ((2 * B + 63) / 64));
  return "";
}
// The first layer's plans carry their zero-stage maps; with the option off they run without (every stage processed).
std::string Engine::launch_maybe_dense(GemmPlan& pl, int* launches, const char* tag) {
  if ((skip_zero_ && in_[cur_].kmask_valid) || pl.args.klist == nullptr) return launch(pl, launches, tag);
  GemmPlan dense = pl;
  for (auto& k : dense.args.kmask) k = nullptr;
  dense.args.klist = nullptr;
  dense.args.order = nullptr;
  return launch(dense, launches, tag);
}

// ------------------------------------------------------------------ step
void Engine::refresh_shadow() {
  if (!shadow_dirty_) return;
  shadow_refresh(p_, w_hi_, w_lo_, n_total, stream);
  shadow_dirty_ = false;
}

void Engine::profile_enable(bool on) { profile_ = on; }
void Engine::prof_begin(const char* tag, double flops, cudaStream_t on) {
  if (!profile_) return;
  if (!on) on = stream;
  ProfRec r;
  r.tag = tag; r.flops = flops;
  for (cudaEvent_t* e : {&r.a, &r.b}) {
    if (!prof_pool_.empty()) { *e = prof_pool_.back(); prof_pool_.pop_back(); }
    else cudaEventCreate(e);
  }
  cudaEventRecord(r.a, on);
  prof_open_.push_back((int)prof_pending_.size());     // begin / end pairs nest (a loader times itself and its map kernels)
  prof_pending_.push_back(r);
}
void Engine::prof_end(cudaStream_t on) {
  if (!profile_ || prof_open_.empty()) return;
  cudaEventRecord(prof_pending_[prof_open_.back()].b, on ? on : stream);
  prof_open_.pop_back();
}
std::string Engine::profile_report() {
  cudaStreamSynchronize(gen_stream_);
  cudaStreamSynchronize(stream);
  for (auto& r : prof_pending_) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.a, r.b) != cudaSuccess) { cudaGetLastError(); ms = 0.f; }   // a pair left open by an error path
    ProfAcc* acc = nullptr;
    for (auto& a : prof_acc_) if (a.tag == r.tag) { acc = &a; break; }
    if (!acc) { prof_acc_.push_back(ProfAcc()); acc = &prof_acc_.back(); acc->tag = r.tag; }
    acc->ms += ms; acc->n += 1; acc->flops += r.flops;
    prof_pool_.push_back(r.a); prof_pool_.push_back(r.b);
  }
  prof_pending_.clear();
  prof_open_.clear();
  std::string out;
  char buf[256];
  for (auto& a : prof_acc_) {
    snprintf(buf, sizeof buf, "%s %.6f %lld %.6e\n", a.tag.c_str(), a.ms, a.n, a.flops);
    out += buf;
  }
  prof_acc_.clear();
  return out;
}

std::string Engine::launch(GemmPlan& pl, int* launches, const char* tag) {
  std::string err;
  prof_begin(tag, pl.flops);
  const int n = gemm_launch(pl, stream, &err);
  prof_end();
  if (n < 0) return err;
  *launches += n;
  return "";
}

std::string Engine::launch_chain_plan(ChainPlan& cp, int* launches, const char* tag) {
  std::string err;
  prof_begin(tag, cp.flops);
  const int n = chain_launch(cp, stream, &err);
  prof_end();
  if (n < 0) return err;
  *launches += n;
  return "";
}

LatentArgs Engine::latent_args(const float* d_eps3, const float* d_labels) {
  LatentArgs a;
  memset(&a, 0, sizeof a);
  a.B = B; a.L = L; a.inv_bg = inv_bg; a.def_weight = cfg.def_weight;
  a.metric = cfg.metric == MAGIC_METRIC_COSINE ? METRIC_COSINE : METRIC_SQDIFF;
  a.reciprocal = cfg.reciprocal;
  a.heads = heads_; a.ld_heads = ld_heads_;
  a.eps = d_eps3;
  a.seed = cfg.seed + 0x51ED270B9F2D4C11ull * (uint64_t)(cfg.rank + 1);
  a.step = (uint64_t)t;
  a.z = z_;
  a.zdec = zdec_.hi; a.zdec_lo = zdec_.lo; a.ld_zdec = zdec_.ld;
  a.labels = d_labels;
  a.sums = statbuf_; a.colsq = statbuf_ + 8; a.colS = statbuf_ + 8 + 2 * L;
  a.g = g_row_; a.distance = dist_;
  a.dz_dec = dz_dec_; a.ld_dz = ld_dz_;
  a.dheads = dheads_.hi; a.dheads_lo = dheads_.lo; a.ld_dheads = dheads_.ld;
  const long long bias_row = lin[iHead].off + (long long)lin[iHead].Kv * lin[iHead].ld;      // row Kv of the heads' matrix
  a.g1_bias = g1_ + bias_row; a.g2_bias = g2_ + bias_row;
  return a;
}

std::string Engine::allreduce(float* buf, size_t n, cudaStream_t s) {
  if (cfg.world == 1) return "";
  if (!comm_) return "world > 1 but magic_comm_init was not called";
  if (xchg_ok_ && n <= (size_t)kXchgMaxFloats) {       // the step's small exchanges: over peer memory (kernels.h)
    XchgArgs a;
    memset(&a, 0, sizeof a);
    for (int r = 0; r < cfg.world; ++r) a.peer[r] = peer_stage_[r] + xchg_off_;
    a.buf = buf; a.n = (int)n; a.world = cfg.world; a.rank = cfg.rank; a.epoch = ++xchg_epoch_;
    peer_allreduce_small(a, s);
    return "";
  }
  const int r = g_nccl.allreduce(buf, buf, n, /*ncclFloat32*/ 7, /*ncclSum*/ 0, comm_, s);
  if (r != 0) return std::string("ncclAllReduce: ") + (g_nccl.errstr ? g_nccl.errstr(r) : "error");
  return "";
}

std::string Engine::run_forward(const float* d_eps3, bool train, int* launches) {
  refresh_shadow();
  {   // the step's statistics and the heads' bias-gradient rows (latent_backward adds into them)
    const long long bias_row = lin[iHead].off + (long long)lin[iHead].Kv * lin[iHead].ld;
    zero_fill3(statbuf_, 8 + 3 * L, g1_ + bias_row, train ? 2 * L : 0, g2_ + bias_row, train ? 2 * L : 0, stream);
  }
  InBuf& ib = in_[cur_];
  wait_ready(ib);
  for (int i = 0; i <= E; ++i) {
    if (chain_fwd_.valid && i >= 1 && i < E) {            // hidden layers 1 .. E-1: one launch
      if (i == 1) MAGIC_TRY(launch_chain_plan(chain_fwd_, launches, "gemm.fwd_enc_hidden"));
      continue;
    }
    MAGIC_TRY(launch_maybe_dense(i == 0 && cur_ == 1 ? fwd_enc0_alt_ : fwd_enc_[i], launches,
                                 i == 0 ? "gemm.fwd_enc0" : (i == E ? "gemm.fwd_heads" : "gemm.fwd_enc_hidden")));
  }
  const float* labels = have_labels_ext_ ? labels_ext_ : ib.labels;
  LatentArgs la = latent_args(d_eps3, labels);
  prof_begin("latent.forward"); latent_forward(la, stream); prof_end();
  const bool cosine = cfg.metric == MAGIC_METRIC_COSINE;
  if (cosine) MAGIC_TRY(allreduce(statbuf_ + 8, 2 * (size_t)L, stream));        // batch-axis norms over the GLOBAL batch
  prof_begin("latent.metric"); latent_metric(la, stream); prof_end();
  if (cosine && train) MAGIC_TRY(allreduce(statbuf_ + 8 + 2 * L, (size_t)L, stream));
  *launches += 2;
  MAGIC_TRY(launch(pl_d1_, launches, "gemm.fwd_dec_hidden"));
  MAGIC_TRY(launch(pl_d2_, launches, "gemm.fwd_dec_hidden"));
  if (train) {
    MAGIC_TRY(launch(cur_ == 1 ? pl_out_bce_alt_ : pl_out_bce_, launches, "gemm.fwd_dec_out_bce"));
    // gate for the next batch's generator: let it share the GPU with the backward GEMMs (tensor-bound, HBM half
    // idle) rather than with the HBM-bound Adam pass of the previous step
    if (gen_gate_ == 1) { MAGIC_CUDA(cudaEventRecord(ev_gate_, stream)); gate_pending_ = true; }
    flush_expand(in_[cur_ ^ 1], gen_gate_ == 1);      // a staged host feed: its kernel goes behind the same gate
  } else {
    mark_consumed(ib);
  }
  MAGIC_CUDA(cudaGetLastError());
  return "";
}

std::string Engine::run_backward(int* launches) {
  const bool fw = fuse_wg_ || (dp_fused_ && dp_small_fused_);   // small-layer wgrads folded into the first-layer launch
  if (!fw) {
    prof_begin("memset.grad_windows");
    for (auto& z : zero_) zero_fill(z.ptr, z.bytes, stream);
    prof_end();
  }
  const bool dp = cfg.world > 1 && !dp_fused_;     // NCCL all-reduce buckets (baseline path)
  MAGIC_TRY(launch(dp_fused_ ? pl_wg_out_sc_ : pl_wg_out_, launches, "gemm.wgrad_dec_out"));
  MAGIC_TRY(launch(pl_dg_out_, launches, "gemm.dgrad_dec_out"));
  if (!fw) MAGIC_TRY(launch(pl_wg_h2_, launches, "gemm.wgrad_small"));
  MAGIC_TRY(launch(pl_dg_h2_, launches, "gemm.dgrad_small"));
  if (!fw) MAGIC_TRY(launch(pl_wg_h1_, launches, "gemm.wgrad_small"));
  if (dp) {   // bucket 0: all decoder gradients are final -> reduce them while the encoder backward runs
    MAGIC_CUDA(cudaEventRecord(ev_bwd_[0], stream));
    MAGIC_CUDA(cudaStreamWaitEvent(comm_stream_, ev_bwd_[0], 0));
    MAGIC_TRY(allreduce(g1_ + n_both, (size_t)(n_total - n_both), comm_stream_));
  }
  MAGIC_TRY(launch(pl_dg_h1_, launches, "gemm.dgrad_small"));
  if (early_adam_) {
    // The decoder's gradients are final and nothing in this step reads its weights any more: update them now, on
    // the side stream, while the (tensor-bound) encoder backward keeps the main stream busy.
    MAGIC_CUDA(cudaEventRecord(ev_aux_fork_, stream));
    MAGIC_CUDA(cudaStreamWaitEvent(aux_stream_, ev_aux_fork_, 0));
    AdamArgs a = adam_;
    a.begin = early_begin_;
    a.thin = 2 * num_sms;
    prof_begin("adam.decoder", 0, aux_stream_); adam_dual(a, aux_stream_); prof_end(aux_stream_);
    MAGIC_CUDA(cudaEventRecord(ev_aux_join_, aux_stream_));
    *launches += 1;
  }
  LatentArgs la = latent_args(nullptr, in_[cur_].labels);
  prof_begin("latent.backward"); latent_backward(la, stream); prof_end();
  *launches += 1;
  if (!fw) MAGIC_TRY(launch(pl_wg_head_, launches, "gemm.wgrad_small"));
  if (!(chain_dg_.valid && chain_dg_has_head_)) MAGIC_TRY(launch(pl_dg_head_, launches, "gemm.dgrad_small"));
  if (chain_dg_.valid) {                                    // dgrads through (the heads and) hidden layers E-1 .. 1: one launch
    MAGIC_TRY(launch_chain_plan(chain_dg_, launches, "gemm.dgrad_small"));
    if (!fw)
      for (int i = E - 1; i >= 1; --i) MAGIC_TRY(launch(pl_wg_enc_[i], launches, "gemm.wgrad_small"));
  } else {
    for (int i = E - 1; i >= 1; --i) {
      if (!fw) MAGIC_TRY(launch(pl_wg_enc_[i], launches, "gemm.wgrad_small"));
      MAGIC_TRY(launch(pl_dg_enc_[i], launches, "gemm.dgrad_small"));
    }
  }
  if (dp) {   // bucket 1: hidden encoder layers + heads, both optimisers
    MAGIC_CUDA(cudaEventRecord(ev_bwd_[1], stream));
    MAGIC_CUDA(cudaStreamWaitEvent(comm_stream_, ev_bwd_[1], 0));
    const long long o1 = lin[0].size();
    MAGIC_TRY(allreduce(g1_ + o1, (size_t)(n_both - o1), comm_stream_));
    MAGIC_TRY(allreduce(g2_ + o1, (size_t)(n_both - o1), comm_stream_));
  }
  if (dp_fused_) {
    // the small layers accumulated locally (split-K red.add): hand their partial sums to the owners' staging slots.
    // Ranges: g1 over [end of layer 0, start of decoder out), g2 over [end of layer 0, n_both).
    PushArgs pa;
    memset(&pa, 0, sizeof pa);
    pa.slice_len = slice_len_; pa.rank = cfg.rank; pa.world = cfg.world; pa.gran_shift = kDpGranShift;
    prof_begin("dp.push_small");
    // (with the small layers in the scatter launch only the heads' bias rows -- summed by latent_backward -- are local)
    const long long bias_row = lin[iHead].off + (long long)lin[iHead].Kv * lin[iHead].ld;
    pa.src = g1_; pa.begin = dp_small_fused_ ? bias_row : lin[0].size(); pa.end = dp_small_fused_ ? bias_row + lin[iHead].ld : lin[iOut].off;
    for (int r = 0; r < cfg.world; ++r) pa.peer[r] = peer_stage_[r];
    push_range(pa, stream);
    pa.src = g2_; pa.begin = dp_small_fused_ ? bias_row : lin[0].size(); pa.end = dp_small_fused_ ? bias_row + lin[iHead].ld : n_both;
    for (int r = 0; r < cfg.world; ++r) pa.peer[r] = peer_stage_[r] + alloc_len_;
    push_range(pa, stream);
    prof_end();
    *launches += 2;
  }
  if (gen_gate_ == 2) { MAGIC_CUDA(cudaEventRecord(ev_gate_, stream)); gate_pending_ = true; }
  if (fuse_adam_) {
    GemmPlan& pl = pl_wg_enc0_adam_[cur_];
    pl.args.adam.lr_t[0] = adam_.lr_t1; pl.args.adam.lr_t[1] = adam_.lr_t2;
    MAGIC_TRY(launch(pl, launches, "gemm.wgrad_enc0"));
  } else {
    MAGIC_TRY(launch_maybe_dense(dp_fused_ ? pl_wg_enc0_sc_[cur_] : fw ? pl_wg_enc0_fused_[cur_] : (cur_ == 1 ? pl_wg_enc0_alt_ : pl_wg_enc_[0]),
                                 launches, "gemm.wgrad_enc0"));
  }
  mark_consumed(in_[cur_]);                 // last read of the input buffer in the step
  if (dp) {   // bucket 2: the two 40000-wide first-layer gradients
    MAGIC_CUDA(cudaEventRecord(ev_bwd_[2], stream));
    MAGIC_CUDA(cudaStreamWaitEvent(comm_stream_, ev_bwd_[2], 0));
    MAGIC_TRY(allreduce(g1_, (size_t)lin[0].size(), comm_stream_));
    MAGIC_TRY(allreduce(g2_, (size_t)lin[0].size(), comm_stream_));
    MAGIC_CUDA(cudaEventRecord(ev_comm_, comm_stream_));
    MAGIC_CUDA(cudaStreamWaitEvent(stream, ev_comm_, 0));
  }
  MAGIC_CUDA(cudaGetLastError());
  return "";
}

std::string Engine::step(const float* d_labels, const float* d_eps3, float* d_losses5, float* d_distance, bool apply) {
  if (!d_labels && !in_[cur_].has_labels) return "step: no labels (pass d_labels or call magic_generate_batch first)";
  if (d_labels) { labels_ext_ = d_labels; have_labels_ext_ = true; } else { have_labels_ext_ = false; }
  int launches = 0;
  const double b1 = cfg.beta1, b2 = cfg.beta2;
  const double corr = sqrt(1.0 - pow(b2, (double)(t + 1))) / (1.0 - pow(b1, (double)(t + 1)));
  memset(&adam_, 0, sizeof adam_);
  adam_.p = p_; adam_.g1 = g1_; adam_.g2 = g2_; adam_.m1 = m1_; adam_.v1 = v1_; adam_.m2 = m2_; adam_.v2 = v2_;
  adam_.w_hi = w_hi_; adam_.w_lo = w_lo_;
  adam_.n_both = n_both; adam_.n_total = n_total;
  adam_.lr_t1 = (float)(cfg.lr1 * corr);
  adam_.lr_t2 = (float)(cfg.lr2 * corr);
  adam_.beta1 = cfg.beta1; adam_.beta2 = cfg.beta2; adam_.eps = cfg.eps;
  early_adam_ = apply && cfg.world == 1 && early_adam_ok_;
  // with the small wgrads folded into the last launch only the decoder's output layer (98% of the decoder) is final early
  early_begin_ = fuse_wg_ ? lin[iOut].off : n_both;
  if (early_begin_ % 4) early_adam_ = false;
  // the first layer's update rides in its wgrad launch (the zero-stage option keeps the plain launch and its lists)
  fuse_adam_ = apply && cfg.world == 1 && fuse_adam_ok_ && !(skip_zero_ && in_[cur_].kmask_valid);
  std::string err = run_forward(d_eps3, true, &launches);
  // One rank: the loss scalars are complete once the forward pass is (BCE epilogue, latent_forward, latent_metric), so
  // they are published here and read_losses() need not wait for the backward pass.  Data parallel: after the all-reduce
  // at the end of the step (which the fused path also uses as its barrier).
  const bool early_losses = cfg.world == 1 && d_losses5 != nullptr;
  auto publish_losses = [&]() {
    launch_chain(finalize_losses_kernel, dim3(1), dim3(32), 0, stream, (const float*)statbuf_, d_losses5);
    ++launches;
    cudaEventRecord(ev_losses_, stream);
    losses_src_ = d_losses5;
  };
  if (err.empty() && early_losses) publish_losses();
  if (err.empty()) err = run_backward(&launches);
  have_labels_ext_ = false;
  if (!err.empty()) return err;
  MAGIC_TRY(allreduce(statbuf_, 4, stream));   // loss scalars: local sums / Bg add up to the global means
  if (d_losses5 && !early_losses) publish_losses();
  if (d_distance) MAGIC_CUDA(cudaMemcpyAsync(d_distance, dist_, sizeof(float) * B, cudaMemcpyDeviceToDevice, stream));
  if (dp_fused_) {
    // The loss all-reduce above is also the barrier that says: every rank's partial gradients have landed in the
    // owners' staging buffers (kernel completion makes peer stores visible; NCCL orders the ranks).
    AdamShardArgs a;
    memset(&a, 0, sizeof a);
    a.n_total = n_total; a.gran_shift = kDpGranShift;
    a.n_both = n_both; a.world = cfg.world; a.rank = cfg.rank; a.slice_len = slice_len_;
    a.stage1 = stage_; a.stage2 = stage_ + alloc_len_;
    a.p = p_; a.m1 = m1_; a.v1 = v1_; a.m2 = m2_; a.v2 = v2_; a.g1 = g1_; a.g2 = g2_;
    for (int r = 0; r < cfg.world; ++r) { a.w_hi[r] = peer_whi_[r]; a.w_lo[r] = peer_wlo_[r]; }
    a.lr_t1 = (float)(cfg.lr1 * corr); a.lr_t2 = (float)(cfg.lr2 * corr);
    a.beta1 = cfg.beta1; a.beta2 = cfg.beta2; a.eps = cfg.eps;
    a.apply = apply ? 1 : 0;
    prof_begin("adam.dual_sharded"); adam_dual_sharded(a, stream); prof_end();
    ++launches;
    // second barrier: all owners have pushed the new shadow weights (and are done reading their staging) before
    // any rank starts the next forward pass / the next step's gradient scatter
    MAGIC_TRY(allreduce(barrier_buf_, 1, stream));
    if (apply) { ++t; stats.steps = t; master_stale_ = true; }
    grads_sharded_ = true;
  } else if (apply) {
    ++t;
    AdamArgs a = adam_;
    if (early_adam_) a.n_total = early_begin_;   // the rest of the decoder range is already in flight on aux_stream_
    if (fuse_adam_) a.begin = lin[0].off + pl_wg_enc0_adam_[cur_].adam_rows_end * lin[0].ld;   // rows below: done in the wgrad epilogue
    prof_begin("adam.dual");
    adam_dual(a, stream);
    prof_end();
    ++launches;
    if (early_adam_) MAGIC_CUDA(cudaStreamWaitEvent(stream, ev_aux_join_, 0));
    stats.steps = t;
  }
  MAGIC_CUDA(cudaGetLastError());
  stats.launches_last_step = launches;
  return "";
}

std::string Engine::partial_fit_host(const void* h_x, bool x_is_u8, const float* h_labels, const float* h_eps3,
                                     float* h_losses5, float* h_distance) {
  if (!h_x || !h_labels || !h_losses5) return "partial_fit_host: null argument";
  const size_t nx = (size_t)B * 3 * D;
  const size_t xbytes = x_is_u8 ? nx : nx * sizeof(float);
  if (stage_x_bytes_ < xbytes) {
    if (stage_x_) { MAGIC_CUDA(cudaStreamSynchronize(stream)); cudaFree(stage_x_); stage_x_ = nullptr; }
    MAGIC_TRY(alloc((void**)&stage_x_, xbytes));
    stage_x_bytes_ = xbytes;
  }
  if (!stage_labels_) MAGIC_TRY(alloc((void**)&stage_labels_, sizeof(float) * B));
  if (h_eps3 && !stage_eps_) MAGIC_TRY(alloc((void**)&stage_eps_, sizeof(float) * 3 * B * L));
  MAGIC_CUDA(cudaMemcpyAsync(stage_x_, h_x, xbytes, cudaMemcpyHostToDevice, stream));
  MAGIC_CUDA(cudaMemcpyAsync(stage_labels_, h_labels, sizeof(float) * B, cudaMemcpyHostToDevice, stream));
  if (h_eps3) MAGIC_CUDA(cudaMemcpyAsync(stage_eps_, h_eps3, sizeof(float) * 3 * B * L, cudaMemcpyHostToDevice, stream));
  if (x_is_u8) MAGIC_TRY(load_interleaved_u8(static_cast<const uint8_t*>(stage_x_)));
  else MAGIC_TRY(load_interleaved(static_cast<const float*>(stage_x_)));
  MAGIC_TRY(step(stage_labels_, h_eps3 ? stage_eps_ : nullptr, losses_dev_, nullptr, true));
  MAGIC_CUDA(cudaMemcpyAsync(h_losses5, losses_dev_, sizeof(float) * 5, cudaMemcpyDeviceToHost, stream));
  if (h_distance) MAGIC_CUDA(cudaMemcpyAsync(h_distance, dist_, sizeof(float) * B, cudaMemcpyDeviceToHost, stream));
  MAGIC_CUDA(cudaStreamSynchronize(stream));
  return "";
}

// ------------------------------------------------------------------ inference (8c/vae.py:407-434)
std::string Engine::forward(int what, const float* d_eps3, float* d_out) {
  if (!d_out) return "forward: null output";
  int launches = 0;
  have_labels_ext_ = false;
  MAGIC_TRY(run_forward(d_eps3, false, &launches));
  if (what == MAGIC_FWD_DISTANCE) {
    MAGIC_CUDA(cudaMemcpyAsync(d_out, dist_, sizeof(float) * B, cudaMemcpyDeviceToDevice, stream));
  } else if (what == MAGIC_FWD_ZMEAN) {   // z_mean of the lock stream: heads rows [B, 2B), cols [0, L)
    MAGIC_CUDA(cudaMemcpy2DAsync(d_out, sizeof(float) * L, heads_ + (long long)B * ld_heads_, sizeof(float) * ld_heads_,
                                 sizeof(float) * L, B, cudaMemcpyDeviceToDevice, stream));
  } else if (what == MAGIC_FWD_RECON) {
    MAGIC_TRY(gemm_set_output(&pl_out_sig_, 0, d_out, nullptr));   // d_out must be 16-byte aligned
    MAGIC_TRY(launch(pl_out_sig_, &launches, "gemm.fwd_dec_out_sigmoid"));
  } else {
    return "forward: unknown output selector";
  }
  return "";
}

std::string Engine::decode(const float* d_z, float* d_out) {
  if (!d_z || !d_out) return "decode: null argument";
  refresh_shadow();
  const long long n = (long long)B * (L + 1);
  z_to_zdec_kernel<<<(int)((n + 255) / 256), 256, 0, stream>>>(d_z, B, L, zdec_.hi, zdec_.lo, zdec_.ld);
  int launches = 0;
  MAGIC_TRY(launch(pl_d1_, &launches, "gemm.fwd_dec_hidden"));
  MAGIC_TRY(launch(pl_d2_, &launches, "gemm.fwd_dec_hidden"));
  MAGIC_TRY(gemm_set_output(&pl_out_sig_, 0, d_out, nullptr));     // d_out must be 16-byte aligned
  MAGIC_TRY(launch(pl_out_sig_, &launches, "gemm.fwd_dec_out_sigmoid"));
  MAGIC_CUDA(cudaGetLastError());
  return "";
}

// ------------------------------------------------------------------ data parallel
std::string Engine::comm_init(const void* uid) {
  MAGIC_TRY(load_nccl());
  NcclUid id;
  memcpy(&id, uid, sizeof id);
  const int r = g_nccl.init_rank(&comm_, cfg.world, id, cfg.rank);
  if (r != 0) return std::string("ncclCommInitRank: ") + (g_nccl.errstr ? g_nccl.errstr(r) : "error");
  nccl_ = &g_nccl;
  MAGIC_CUDA(cudaStreamCreateWithFlags(&comm_stream_, cudaStreamNonBlocking));
  return "";
}

// ------------------------------------------------------------------ fused data parallelism over peer memory
struct PeerBlob {                       // what one rank tells the others (kPeerBlobBytes)
  cudaIpcMemHandle_t stage, whi, wlo;
  int has_lo, device;
};
static_assert(sizeof(PeerBlob) <= Engine::kPeerBlobBytes, "peer blob too large");

std::string Engine::comm_export_peer(void* h_blob) {
  if (cfg.world < 2) return "comm_export_peer: world == 1";
  PeerBlob b;
  memset(&b, 0, sizeof b);
  MAGIC_CUDA(cudaIpcGetMemHandle(&b.stage, stage_));
  MAGIC_CUDA(cudaIpcGetMemHandle(&b.whi, w_hi_));
  if (w_lo_) { MAGIC_CUDA(cudaIpcGetMemHandle(&b.wlo, w_lo_)); b.has_lo = 1; }
  MAGIC_CUDA(cudaGetDevice(&b.device));
  memset(h_blob, 0, kPeerBlobBytes);
  memcpy(h_blob, &b, sizeof b);
  return "";
}

std::string Engine::comm_import_peers(const void* h_blobs) {
  if (cfg.world < 2) return "comm_import_peers: world == 1";
  if (!comm_) return "comm_import_peers: call magic_comm_init first";
  const char* mode = getenv("MAGIC_DP_MODE");
  if (mode && std::string(mode) == "allreduce") return "";       // keep the NCCL all-reduce baseline path
  for (int r = 0; r < cfg.world; ++r) {
    if (r == cfg.rank) {
      peer_stage_[r] = stage_; peer_whi_[r] = w_hi_; peer_wlo_[r] = w_lo_;
      continue;
    }
    PeerBlob b;
    memcpy(&b, static_cast<const char*>(h_blobs) + (size_t)r * kPeerBlobBytes, sizeof b);
    void* ptr = nullptr;
    MAGIC_CUDA(cudaIpcOpenMemHandle(&ptr, b.stage, cudaIpcMemLazyEnablePeerAccess));
    peer_stage_[r] = static_cast<float*>(ptr);
    MAGIC_CUDA(cudaIpcOpenMemHandle(&ptr, b.whi, cudaIpcMemLazyEnablePeerAccess));
    peer_whi_[r] = static_cast<bf16*>(ptr);
    if (b.has_lo) {
      MAGIC_CUDA(cudaIpcOpenMemHandle(&ptr, b.wlo, cudaIpcMemLazyEnablePeerAccess));
      peer_wlo_[r] = static_cast<bf16*>(ptr);
    }
  }
  MAGIC_TRY(build_dp_plans());
  dp_fused_ = true;
  // every rank reaches this point with the same sequence of calls behind it, so the epochs agree from here on
  { const char* x = getenv("MAGIC_DP_XCHG"); xchg_ok_ = !(x && atoi(x) == 0); }
  return "";
}

// Scatter variants of the two 40000-wide weight gradients (97 % of the gradient bytes): same GEMMs, epilogue
// EPI_SCATTER_F32 -> the owner's staging slot of this rank.  Never split along K (plain stores, not red.add).
std::string Engine::build_dp_plans() {
  auto scatter_epi = [&](int which, int layer, float alpha) {
    EpiArgs e;
    memset(&e, 0, sizeof e);
    e.ld_out = lin[layer].ld; e.alpha = alpha; e.aux_split = 1 << 30;
    e.flat_base = lin[layer].off; e.slice_len = slice_len_; e.src_rank = cfg.rank;
    e.world = cfg.world; e.gran_shift = kDpGranShift;
    for (int r = 0; r < cfg.world; ++r) e.peer[r] = peer_stage_[r] + (which == 2 ? alloc_len_ : 0);
    if (const char* d = getenv("MAGIC_EPI_DBG")) e.dbg = atoi(d);   // timing experiments only: results are wrong
    return e;
  };
  {
    GemmProblem p;
    p.kind = GEMM_WGRAD; p.epi = EPI_SCATTER_F32; p.M = cfg.dec[1] + 1; p.N = lin[iOut].Nv; p.K = B;
    add_segments(p, 0, amat(D2_, false, 0, B, p.M), amat(D2_, true, 0, B, p.M), amat(dA_, false, 0, B, p.N),
                 amat(dA_, true, 0, B, p.N), false);
    p.e[0] = scatter_epi(1, iOut, inv_bg);
    p.force_splitk = 1;
    std::string err = gemm_plan(p, num_sms, &pl_wg_out_sc_);
    if (!err.empty()) return "dp wgrad dec out: " + err;
  }
  for (int alt = 0; alt < 2; ++alt) {
    const Act& x = in_[alt].X;
    const Act& dy = dH_[0];
    GemmProblem p;
    p.kind = GEMM_WGRAD; p.epi = EPI_SCATTER_F32; p.ngroups = 2; p.M = D + 1; p.N = lin[0].Nv; p.K = 2 * B;
    add_segments(p, 0, amat(x, false, 0, 2LL * B, p.M), amat(x, true, 0, 2LL * B, p.M), amat(dy, false, 0, 2LL * B, p.N),
                 amat(dy, true, 0, 2LL * B, p.N), true);
    add_segments(p, 1, amat(x, false, B, 2LL * B, p.M), amat(x, true, B, 2LL * B, p.M),
                 amat(dy, false, 2LL * B, 2LL * B, p.N), amat(dy, true, 2LL * B, 2LL * B, p.N), true);
    p.e[0] = scatter_epi(1, 0, 1.f);
    p.e[1] = scatter_epi(2, 0, 1.f);
    p.force_splitk = 1;
    if (bits_a_) {
      p.a_bits = true;
      p.abits[0].base = in_[alt].xb3; p.abits[0].pitch = xb_ld_; p.abits[0].rows = 2 * B;
      p.abits[1].base = in_[alt].xb3 + (size_t)B * xb_ld_; p.abits[1].pitch = xb_ld_; p.abits[1].rows = 2 * B;
      p.kmask[0] = in_[alt].kmask + km_wg_off_[0]; p.kmask[1] = in_[alt].kmask + km_wg_off_[1]; p.kmask_words = km_wg_words_;
    }
    // The small layers' weight gradients ride in this launch as appended whole-K problems, each scattering its tiles to
    // the owners like the main problem (one rank does the same with plain stores: pl_wg_enc0_fused_), instead of six
    // split-K launches accumulating locally, a window clear before and two push_range launches after them (0.15 ms of the
    // step): 2 GPUs 2.131 -> 2.049 ms per step (profiles/r03u_dp_probe_n2.txt).  MAGIC_DP_FUSE_SMALL=0 keeps them apart.
    // (While the packed-operand kernels could not share an SM with the generator this LOST 0.06 ms: the shorter backward
    // pass left more of the generator to block this launch -- profiles/r03k_dp_probe_n2.txt.)
    GemmProblem pf = p;
    size_t nx = 0;
    bool want = !(getenv("MAGIC_DP_FUSE_SMALL") && atoi(getenv("MAGIC_DP_FUSE_SMALL")) == 0);
    for (auto& sp : small_wg_) { nx += sp.ngroups; want = want && sp.nseg == p.nseg; }
    if (want && nx <= (size_t)kMaxExtra) {
      for (size_t i = 0; i < small_wg_.size(); ++i) {
        const GemmProblem& sp = small_wg_[i];
        for (int g = 0; g < sp.ngroups; ++g) {
          GemmExtra x;
          for (int sg = 0; sg < sp.nseg; ++sg) { x.a[sg] = sp.a[g][sg]; x.b[sg] = sp.b[g][sg]; }
          x.e = scatter_epi(sp.ngroups == 2 ? g + 1 : 1, small_wg_layer_[i], sp.e[g].alpha);
          x.M = sp.M; x.N = sp.N; x.K = sp.K;
          pf.extra.push_back(x);
        }
      }
      GemmPlan tmp;
      if (gemm_plan(pf, num_sms, &tmp).empty()) { pl_wg_enc0_sc_[alt] = tmp; dp_small_fused_ = true; continue; }
    }
    if (alt == 1 && dp_small_fused_) return "dp wgrad enc0: the second input buffer's plan does not take the appended problems";
    dp_small_fused_ = false;
    std::string err = gemm_plan(p, num_sms, &pl_wg_enc0_sc_[alt]);
    if (!err.empty()) return "dp wgrad enc0: " + err;
  }
  if (getenv("MAGIC_VERBOSE")) fprintf(stderr, "[magic] data parallel: small layers' wgrads %s\n", dp_small_fused_ ? "in the scatter launch" : "on their own");
  return "";
}

// Sharded state (fp32 master weights, optimiser slots, reduced gradients) lives only on the owner of each granule
// between steps; host accessors assemble it first: zero what this rank does not own, all-reduce(sum).  Collective:
// every rank must make the same call -- true for checkpointing and for the parity tests.
std::string Engine::assemble(float* buf, long long n) {
  mask_unowned(buf, n, cfg.world, cfg.rank, kDpGranShift, stream);
  return allreduce(buf, (size_t)n, stream);
}
std::string Engine::gather_master() {
  if (!dp_fused_ || !master_stale_) return "";
  for (float* b : {p_, m1_, v1_}) MAGIC_TRY(assemble(b, n_total));
  for (float* b : {m2_, v2_}) MAGIC_TRY(assemble(b, n_both));
  MAGIC_CUDA(cudaStreamSynchronize(stream));
  master_stale_ = false;
  return "";
}

// ------------------------------------------------------------------ parameter windows
std::string Engine::copy_param(int i, float* h_dst, const float* h_src, const float* base) {
  if (i < 0 || i >= (int)views.size()) return "parameter index out of range";
  const ParamView& v = views[i];
  const Linear& l = lin[v.layer];
  float* win = const_cast<float*>(base) + l.off + (long long)v.row0 * l.ld + v.col0;
  MAGIC_CUDA(cudaStreamSynchronize(stream));
  if (h_dst)
    MAGIC_CUDA(cudaMemcpy2D(h_dst, sizeof(float) * v.ncols, win, sizeof(float) * l.ld, sizeof(float) * v.ncols, v.nrows,
                            cudaMemcpyDeviceToHost));
  else
    MAGIC_CUDA(cudaMemcpy2D(win, sizeof(float) * l.ld, h_src, sizeof(float) * v.ncols, sizeof(float) * v.ncols, v.nrows,
                            cudaMemcpyHostToDevice));
  return "";
}
std::string Engine::get_param(int i, float* h) {
  MAGIC_TRY(gather_master());
  return copy_param(i, h, nullptr, p_);
}
std::string Engine::set_param(int i, const float* h) {
  if (!h) return "set_param: null source";
  MAGIC_TRY(gather_master());
  shadow_dirty_ = true;
  return copy_param(i, nullptr, h, p_);
}
std::string Engine::get_grad(int which, int i, float* h) {
  if (which != 1 && which != 2) return "get_grad: which must be 1 or 2";
  if (i < 0 || i >= (int)views.size()) return "parameter index out of range";
  if (which == 2 && !(views[i].optimisers & 2)) return "get_grad: optimiser 2 does not train this variable";
  if (dp_fused_ && grads_sharded_) {        // reduced gradients exist on their owners only: gather (collective)
    MAGIC_TRY(assemble(g1_, n_total));
    MAGIC_TRY(assemble(g2_, n_both));
    grads_sharded_ = false;
  }
  return copy_param(i, h, nullptr, which == 1 ? g1_ : g2_);
}
std::string Engine::slot(int which, int sl, int i, float* h_dst, const float* h_src) {
  if (which != 1 && which != 2) return "slot: which must be 1 or 2";
  if (i < 0 || i >= (int)views.size()) return "parameter index out of range";
  if (which == 2 && !(views[i].optimisers & 2)) return "slot: optimiser 2 does not train this variable";
  MAGIC_TRY(gather_master());
  const float* base = which == 1 ? (sl == MAGIC_SLOT_M ? m1_ : v1_) : (sl == MAGIC_SLOT_M ? m2_ : v2_);
  return copy_param(i, h_dst, h_src, base);
}

}  // namespace magic
