// extern "C" boundary (include/magic_b200.h): plain pointers and sizes, error codes, no exceptions.
#include "engine.h"
#include <cstring>
#include <new>

namespace magic {
std::string nccl_unique_id(void* out128);
}

struct magic_handle {
  magic::Engine eng;
};

static thread_local std::string g_last_error;

static int fail(const std::string& msg) {
  g_last_error = msg;
  return 1;
}
static int done(const std::string& err) {
  if (err.empty()) return 0;
  return fail(err);
}
#define CHECK_HANDLE(h) \
  if (!(h)) return fail("null handle")

extern "C" {

const char* magic_last_error(void) { return g_last_error.c_str(); }
int magic_version(void) { return 100; }

int magic_create(const magic_config* cfg, void* stream, magic_handle** out) {
  if (!cfg || !out) return fail("magic_create: null argument");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(std::string("magic_create: no CUDA device (") + cudaGetErrorString(e) +
                "); this library has no CPU path");
  magic_handle* h = new (std::nothrow) magic_handle();
  if (!h) return fail("magic_create: out of host memory");
  std::string err = h->eng.init(*cfg, static_cast<cudaStream_t>(stream));
  if (!err.empty()) {
    delete h;
    cudaGetLastError();
    return fail("magic_create: " + err);
  }
  *out = h;
  return 0;
}

int magic_destroy(magic_handle* h) {
  CHECK_HANDLE(h);
  delete h;
  return 0;
}

int magic_sync(magic_handle* h) {
  CHECK_HANDLE(h);
  cudaError_t e = cudaStreamSynchronize(h->eng.stream);
  if (e != cudaSuccess) return fail(std::string("magic_sync: ") + cudaGetErrorString(e));
  return 0;
}

int magic_get_stats(magic_handle* h, magic_stats* out) {
  CHECK_HANDLE(h);
  if (!out) return fail("null output");
  *out = h->eng.stats;
  return 0;
}

int magic_param_count(magic_handle* h, int32_t* n) {
  CHECK_HANDLE(h);
  if (!n) return fail("null output");
  *n = (int32_t)h->eng.views.size();
  return 0;
}

int magic_param_info(magic_handle* h, int32_t i, const char** name, int64_t shape[2], int32_t* optimisers) {
  CHECK_HANDLE(h);
  if (i < 0 || i >= (int32_t)h->eng.views.size()) return fail("parameter index out of range");
  const magic::ParamView& v = h->eng.views[i];
  if (name) *name = v.name.c_str();
  if (shape) {
    shape[0] = v.is_bias ? v.ncols : v.nrows;
    shape[1] = v.is_bias ? 0 : v.ncols;
  }
  if (optimisers) *optimisers = v.optimisers;
  return 0;
}

int magic_set_param(magic_handle* h, int32_t i, const float* h_src) {
  CHECK_HANDLE(h);
  return done(h->eng.set_param(i, h_src));
}
int magic_get_param(magic_handle* h, int32_t i, float* h_dst) {
  CHECK_HANDLE(h);
  if (!h_dst) return fail("null output");
  return done(h->eng.get_param(i, h_dst));
}
int magic_get_grad(magic_handle* h, int32_t which, int32_t i, float* h_dst) {
  CHECK_HANDLE(h);
  if (!h_dst) return fail("null output");
  return done(h->eng.get_grad(which, i, h_dst));
}
int magic_get_slot(magic_handle* h, int32_t which, int32_t slot, int32_t i, float* h_dst) {
  CHECK_HANDLE(h);
  if (!h_dst) return fail("null output");
  return done(h->eng.slot(which, slot, i, h_dst, nullptr));
}
int magic_set_slot(magic_handle* h, int32_t which, int32_t slot, int32_t i, const float* h_src) {
  CHECK_HANDLE(h);
  if (!h_src) return fail("null source");
  return done(h->eng.slot(which, slot, i, nullptr, h_src));
}
int magic_get_step(magic_handle* h, int64_t* t) {
  CHECK_HANDLE(h);
  if (!t) return fail("null output");
  *t = h->eng.t;
  return 0;
}
int magic_set_step(magic_handle* h, int64_t t) {
  CHECK_HANDLE(h);
  if (t < 0) return fail("negative step");
  h->eng.t = t;
  h->eng.stats.steps = t;
  return 0;
}

int magic_load_interleaved_u8(magic_handle* h, const uint8_t* d_x) {
  CHECK_HANDLE(h);
  return done(h->eng.load_interleaved_u8(d_x));
}
int magic_partial_fit_host_u8(magic_handle* h, const uint8_t* h_x, const float* h_labels, const float* h_eps3,
                              float h_losses5[5], float* h_distance) {
  CHECK_HANDLE(h);
  return done(h->eng.partial_fit_host(h_x, true, h_labels, h_eps3, h_losses5, h_distance));
}
int magic_load_interleaved(magic_handle* h, const float* d_x) {
  CHECK_HANDLE(h);
  return done(h->eng.load_interleaved(d_x));
}
int magic_load_planar_u8(magic_handle* h, const uint8_t* d_lock, const uint8_t* d_rot, const uint8_t* d_key) {
  CHECK_HANDLE(h);
  return done(h->eng.load_planar_u8(d_lock, d_rot, d_key));
}
int magic_load_pairs_u8(magic_handle* h, const uint8_t* d_lock, const uint8_t* d_key, uint64_t seed, uint64_t step) {
  CHECK_HANDLE(h);
  return done(h->eng.load_pairs_u8(d_lock, d_key, seed, step));
}
int magic_load_pairs_bits(magic_handle* h, const uint32_t* d_lock_bits, const uint32_t* d_key_bits, int32_t words_per_row,
                          uint64_t seed, uint64_t step) {
  CHECK_HANDLE(h);
  return done(h->eng.load_pairs_bits(d_lock_bits, d_key_bits, words_per_row, seed, step));
}
int magic_feed_pairs_bits_gather(magic_handle* h, const uint32_t* h_lock_all, const uint32_t* h_key_all, int32_t words_per_row,
                                 int64_t n_rows, const int32_t* h_idx, const float* h_labels, uint64_t seed, uint64_t step,
                                 int32_t activate) {
  CHECK_HANDLE(h);
  return done(h->eng.feed_pairs_bits_gather(h_lock_all, h_key_all, words_per_row, n_rows, h_idx, h_labels, seed, step, activate != 0));
}
int magic_feed_pairs_bits_host(magic_handle* h, const uint32_t* h_lock_bits, const uint32_t* h_key_bits, const float* h_labels,
                               int32_t words_per_row, uint64_t seed, uint64_t step, int32_t activate) {
  CHECK_HANDLE(h);
  return done(h->eng.feed_pairs_bits_host(h_lock_bits, h_key_bits, h_labels, words_per_row, seed, step, activate != 0));
}
int magic_generate_batch(magic_handle* h, uint64_t seed, uint64_t step) {
  CHECK_HANDLE(h);
  return done(h->eng.generate(seed, step));
}
int magic_swap_batch(magic_handle* h) {
  CHECK_HANDLE(h);
  return done(h->eng.swap_batch());
}
int magic_set_option(magic_handle* h, const char* name, int32_t value) {
  CHECK_HANDLE(h);
  return done(h->eng.set_option(name, value));
}
int magic_input_density(magic_handle* h, double* out2) {
  CHECK_HANDLE(h);
  return done(h->eng.input_density(out2));
}
int magic_export_rotation(magic_handle* h, int32_t* d_cs_sn) {
  CHECK_HANDLE(h);
  return done(h->eng.export_rotation(d_cs_sn));
}
int magic_export_latents(magic_handle* h, float* d_z3) {
  CHECK_HANDLE(h);
  return done(h->eng.export_latents(d_z3));
}
int magic_export_batch(magic_handle* h, float* d_x, float* d_labels) {
  CHECK_HANDLE(h);
  return done(h->eng.export_batch(d_x, d_labels));
}

int magic_step(magic_handle* h, const float* d_labels, const float* d_eps3, float* d_losses5, float* d_distance) {
  CHECK_HANDLE(h);
  return done(h->eng.step(d_labels, d_eps3, d_losses5, d_distance, true));
}
int magic_read_losses(magic_handle* h, float h_losses5[5]) {
  CHECK_HANDLE(h);
  return done(h->eng.read_losses(h_losses5));
}
int magic_step_grads_only(magic_handle* h, const float* d_labels, const float* d_eps3, float* d_losses5,
                          float* d_distance) {
  CHECK_HANDLE(h);
  return done(h->eng.step(d_labels, d_eps3, d_losses5, d_distance, false));
}
int magic_partial_fit_host(magic_handle* h, const float* h_x, const float* h_labels, const float* h_eps3,
                           float h_losses5[5], float* h_distance) {
  CHECK_HANDLE(h);
  return done(h->eng.partial_fit_host(h_x, false, h_labels, h_eps3, h_losses5, h_distance));
}
int magic_forward(magic_handle* h, int32_t what, const float* d_eps3, float* d_out) {
  CHECK_HANDLE(h);
  return done(h->eng.forward(what, d_eps3, d_out));
}
int magic_decode(magic_handle* h, const float* d_z, float* d_out) {
  CHECK_HANDLE(h);
  return done(h->eng.decode(d_z, d_out));
}

int magic_cosine_distance(const float* h_a, const float* h_b, int32_t B, int32_t L, float* h_out) {
  if (!h_a || !h_b || !h_out) return fail("magic_cosine_distance: null argument");
  if (B < 1 || L < 1) return fail("magic_cosine_distance: B and L must be >= 1");
  const size_t n = sizeof(float) * (size_t)B * L;
  float *d_a = nullptr, *d_b = nullptr, *d_ws = nullptr, *d_out = nullptr;
  cudaError_t e = cudaMalloc((void**)&d_a, n);
  if (e == cudaSuccess) e = cudaMalloc((void**)&d_b, n);
  if (e == cudaSuccess) e = cudaMalloc((void**)&d_ws, sizeof(float) * 2 * (size_t)L);
  if (e == cudaSuccess) e = cudaMalloc((void**)&d_out, sizeof(float) * (size_t)B);
  if (e == cudaSuccess) e = cudaMemcpy(d_a, h_a, n, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(d_b, h_b, n, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    magic::cosine_distance_rows(d_a, d_b, B, L, d_ws, d_out, nullptr);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(h_out, d_out, sizeof(float) * (size_t)B, cudaMemcpyDeviceToHost);
  cudaFree(d_a); cudaFree(d_b); cudaFree(d_ws); cudaFree(d_out);
  if (e != cudaSuccess) return fail(std::string("magic_cosine_distance: ") + cudaGetErrorString(e));
  return 0;
}

int magic_cosine_distance_noalloc(const float* d_a, const float* d_b, int32_t B, int32_t L, float* d_out, void* stream) {
  if (!d_a || !d_b || !d_out) return fail("magic_cosine_distance_noalloc: null argument");
  if (B < 1 || L < 1) return fail("magic_cosine_distance_noalloc: B and L must be >= 1");
  if (L > magic::kCosineStaticMaxL)
    return fail("magic_cosine_distance_noalloc: L = " + std::to_string(L) + " exceeds the static scratch (maximum " +
                std::to_string(magic::kCosineStaticMaxL) + " latent columns); use magic_cosine_distance");
  const char* err = magic::cosine_distance_rows_static(d_a, d_b, B, L, d_out, static_cast<cudaStream_t>(stream));
  if (err) return fail(std::string("magic_cosine_distance_noalloc: ") + err);
  return 0;
}

int magic_comm_unique_id(void* h_id128) {
  if (!h_id128) return fail("null output");
  return done(magic::nccl_unique_id(h_id128));
}
int magic_comm_init(magic_handle* h, const void* h_nccl_unique_id128) {
  CHECK_HANDLE(h);
  if (!h_nccl_unique_id128) return fail("null id");
  return done(h->eng.comm_init(h_nccl_unique_id128));
}

int magic_comm_export_peer(magic_handle* h, void* h_blob256) {
  CHECK_HANDLE(h);
  if (!h_blob256) return fail("null output");
  return done(h->eng.comm_export_peer(h_blob256));
}
int magic_comm_import_peers(magic_handle* h, const void* h_blobs) {
  CHECK_HANDLE(h);
  if (!h_blobs) return fail("null input");
  return done(h->eng.comm_import_peers(h_blobs));
}
int magic_profile_enable(magic_handle* h, int32_t on) {
  CHECK_HANDLE(h);
  h->eng.profile_enable(on != 0);
  return 0;
}
int magic_profile_report(magic_handle* h, char* h_buf, int64_t cap) {
  CHECK_HANDLE(h);
  if (!h_buf || cap < 1) return fail("null buffer");
  const std::string r = h->eng.profile_report();
  const size_t n = r.size() < (size_t)cap - 1 ? r.size() : (size_t)cap - 1;
  memcpy(h_buf, r.data(), n);
  h_buf[n] = 0;
  return 0;
}

int magic_test_gemm(int32_t kind, int32_t M, int32_t N, int32_t K, const void* d_a, int64_t lda, const void* d_b,
                    int64_t ldb, float* d_out, int64_t ldo, int32_t splitk, void* stream) {
  using namespace magic;
  if (kind < 0 || kind > 2 || !d_a || !d_b || !d_out) return fail("magic_test_gemm: bad argument");
  std::string err = gemm_init();
  if (!err.empty()) return fail(err);
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  GemmProblem p;
  p.kind = kind; p.epi = EPI_STORE_F32; p.M = M; p.N = N; p.K = K; p.force_splitk = splitk;
  const bf16* a = static_cast<const bf16*>(d_a);
  const bf16* b = static_cast<const bf16*>(d_b);
  p.a[0][0] = kind == GEMM_WGRAD ? Mat(a, K, M, lda) : Mat(a, M, K, lda);
  p.b[0][0] = kind == GEMM_DGRAD ? Mat(b, N, K, ldb) : Mat(b, K, N, ldb);
  EpiArgs e;
  memset(&e, 0, sizeof e);
  e.out = d_out; e.ld_out = ldo; e.alpha = 1.f; e.aux_split = 1 << 30;
  p.e[0] = e;
  GemmPlan pl;
  err = gemm_plan(p, sms, &pl);
  if (!err.empty()) return fail(err);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (pl.args.splitk > 1) cudaMemsetAsync(d_out, 0, sizeof(float) * (size_t)M * ldo, s);
  if (gemm_launch(pl, s, &err) < 0) return fail(err);
  return 0;
}

}  // extern "C"
