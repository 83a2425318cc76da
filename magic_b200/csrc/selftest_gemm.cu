// Stand-alone self-test + micro-benchmark of the tcgen05 GEMM against a host double reference.
// Build: see magic_b200/build.py (target selftest_gemm).  Run on a B200: ./selftest_gemm [bench]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>
#include <algorithm>
#include <string>
#include "gemm.h"
#include "kernels.h"

using namespace magic;

#define CK(x)                                                                          \
  do {                                                                                 \
    cudaError_t e_ = (x);                                                              \
    if (e_ != cudaSuccess) {                                                           \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);  \
      exit(2);                                                                         \
    }                                                                                  \
  } while (0)

static uint32_t rng_state = 12345u;
static float frand() {
  rng_state = rng_state * 1664525u + 1013904223u;
  return ((rng_state >> 8) & 0xFFFF) / 65536.0f - 0.5f;
}
static float bf16r(float x) { return __bfloat162float(__float2bfloat16(x)); }

struct HostMat {
  std::vector<float> v;   // bf16-rounded values
  long long rows, cols, ld;
  __nv_bfloat16* d = nullptr;
  void init(long long r, long long c, long long l, float scale, bool binary = false) {
    rows = r; cols = c; ld = l;
    v.assign(r * l, 0.f);
    std::vector<__nv_bfloat16> h(r * l);
    for (long long i = 0; i < r; ++i)
      for (long long j = 0; j < l; ++j) {
        float x = j < c ? (binary ? (frand() > 0.3f ? 1.f : 0.f) : bf16r(frand() * scale)) : 7.0f;  // junk in the pad
        v[i * l + j] = x;
        h[i * l + j] = __float2bfloat16(x);
      }
    CK(cudaMalloc(&d, sizeof(__nv_bfloat16) * r * l));
    CK(cudaMemcpy(d, h.data(), sizeof(__nv_bfloat16) * r * l, cudaMemcpyHostToDevice));
  }
  // zero whole 128 x 64 blocks (r = the 128-aligned axis, c = the 64-aligned axis) so that zero-stage maps have work:
  // two blocks of three, plus every block of the 128-row band `dead_band`
  void punch(bool r_is_row, int dead_band) {
    std::vector<__nv_bfloat16> h(rows * ld);
    for (long long i = 0; i < rows; ++i)
      for (long long j = 0; j < ld; ++j) {
        const long long a = r_is_row ? i / 128 : j / 128, b = r_is_row ? j / 64 : i / 64;
        if (j < cols && ((a + b) % 3 != 0 || a == dead_band)) v[i * ld + j] = 0.f;
        h[i * ld + j] = __float2bfloat16(v[i * ld + j]);
      }
    CK(cudaMemcpy(d, h.data(), sizeof(__nv_bfloat16) * rows * ld, cudaMemcpyHostToDevice));
  }
  Mat mat() const { return Mat(d, rows, cols, ld); }
  float at(long long r, long long c) const { return v[r * ld + c]; }
};

static int g_fail = 0;
static void report(const char* name, double maxerr, double tol) {
  printf("%-46s max_err %.3e  (tol %.1e)  %s\n", name, maxerr, tol, maxerr <= tol ? "OK" : "FAIL");
  if (!(maxerr <= tol)) ++g_fail;
}

// reference D[M,N] for the three kinds, one group, nseg segments
static void host_ref(int kind, int M, int N, int K, int nseg, const HostMat* a, const HostMat* b, std::vector<double>& D) {
  D.assign((size_t)M * N, 0.0);
  for (int s = 0; s < nseg; ++s)
    for (int m = 0; m < M; ++m)
      for (int n = 0; n < N; ++n) {
        double acc = 0;
        for (int k = 0; k < K; ++k) {
          double av = kind == GEMM_WGRAD ? a[s].at(k, m) : a[s].at(m, k);
          double bv = kind == GEMM_DGRAD ? b[s].at(n, k) : b[s].at(k, n);
          acc += av * bv;
        }
        D[(size_t)m * N + n] += acc;
      }
}

static int g_cluster = 1;   // cluster_m applied to the problems built below
static bool g_pair = false; // cta_pair applied to the problems built below (CTA-pair kernels where the variant exists)
static int g_bn = 0;        // block_n applied to the problems built below (0 = the planner's choice)
static int g_bits = 0;      // bench(): A operand as packed bits (all zero)
static float g_density = 1.f;   // bench() with g_bits: fraction of operand stages marked non-zero in a random zero-stage map

static void test_store(int kind, int M, int N, int K, int nseg, int split, int sms) {
  HostMat a[3], b[3];
  for (int s = 0; s < nseg; ++s) {
    if (kind == GEMM_WGRAD) a[s].init(K, M, (M + 63) / 64 * 64, 1.f); else a[s].init(M, K, (K + 63) / 64 * 64, 1.f);
    if (kind == GEMM_DGRAD) b[s].init(N, K, (K + 63) / 64 * 64, 1.f); else b[s].init(K, N, (N + 63) / 64 * 64, 1.f);
  }
  long long ldo = (N + 3) / 4 * 4;
  float* dout;
  CK(cudaMalloc(&dout, sizeof(float) * M * ldo));
  CK(cudaMemset(dout, 0, sizeof(float) * M * ldo));
  GemmProblem p;
  p.kind = kind; p.epi = EPI_STORE_F32; p.nseg = nseg; p.M = M; p.N = N; p.K = K; p.force_splitk = split;
  p.cluster_m = g_cluster; p.cta_pair = g_pair;
  p.block_n = g_bn;
  for (int s = 0; s < nseg; ++s) { p.a[0][s] = a[s].mat(); p.b[0][s] = b[s].mat(); }
  EpiArgs e = {};
  e.out = dout; e.ld_out = ldo; e.alpha = 0.5f;
  p.e[0] = e;
  GemmPlan pl;
  std::string err = gemm_plan(p, sms, &pl);
  if (!err.empty()) { printf("plan error: %s\n", err.c_str()); ++g_fail; return; }
  if (gemm_launch(pl, 0, &err) < 0) { printf("launch error: %s\n", err.c_str()); ++g_fail; return; }
  CK(cudaDeviceSynchronize());
  std::vector<float> h((size_t)M * ldo);
  CK(cudaMemcpy(h.data(), dout, sizeof(float) * M * ldo, cudaMemcpyDeviceToHost));
  std::vector<double> D;
  host_ref(kind, M, N, K, nseg, a, b, D);
  double me = 0;
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) me = fmax(me, fabs(h[(size_t)m * ldo + n] - 0.5 * D[(size_t)m * N + n]));
  char name[128];
  snprintf(name, sizeof name, "store kind=%d M=%d N=%d K=%d nseg=%d split=%d(%d)", kind, M, N, K, nseg, split, pl.args.splitk);
  report(name, me, 2e-3 * sqrt((double)K * nseg) * 0.1 + 1e-4);
  for (int s = 0; s < nseg; ++s) { cudaFree(a[s].d); cudaFree(b[s].d); }
  cudaFree(dout);
}

static void test_wgrad_groups(int M, int N, int K, int sms) {
  HostMat a[2], b[2];
  float* dout[2];
  long long ldo = (N + 3) / 4 * 4;
  GemmProblem p;
  p.kind = GEMM_WGRAD; p.epi = EPI_STORE_F32; p.ngroups = 2; p.nseg = 1; p.M = M; p.N = N; p.K = K; p.cta_pair = g_pair;
  for (int g = 0; g < 2; ++g) {
    a[g].init(K, M, (M + 63) / 64 * 64, 1.f, true);
    b[g].init(K, N, (N + 63) / 64 * 64, 1.f);
    CK(cudaMalloc(&dout[g], sizeof(float) * M * ldo));
    CK(cudaMemset(dout[g], 0, sizeof(float) * M * ldo));
    p.a[g][0] = a[g].mat(); p.b[g][0] = b[g].mat();
    EpiArgs e = {};
    e.out = dout[g]; e.ld_out = ldo; e.alpha = 1.f;
    p.e[g] = e;
  }
  GemmPlan pl;
  std::string err = gemm_plan(p, sms, &pl);
  if (!err.empty()) { printf("plan error: %s\n", err.c_str()); ++g_fail; return; }
  if (gemm_launch(pl, 0, &err) < 0) { printf("launch error: %s\n", err.c_str()); ++g_fail; return; }
  CK(cudaDeviceSynchronize());
  double me = 0;
  for (int g = 0; g < 2; ++g) {
    std::vector<float> h((size_t)M * ldo);
    CK(cudaMemcpy(h.data(), dout[g], sizeof(float) * M * ldo, cudaMemcpyDeviceToHost));
    std::vector<double> D;
    host_ref(GEMM_WGRAD, M, N, K, 1, &a[g], &b[g], D);
    for (int m = 0; m < M; ++m)
      for (int n = 0; n < N; ++n) me = fmax(me, fabs(h[(size_t)m * ldo + n] - D[(size_t)m * N + n]));
  }
  char name[128];
  snprintf(name, sizeof name, "wgrad 2 groups M=%d N=%d K=%d split=%d", M, N, K, pl.args.splitk);
  report(name, me, 2e-3 * sqrt((double)K) * 0.1 + 1e-4);
}

// A two-group wgrad with problems of other shapes appended to the same launch (GemmProblem::extra)
static void test_wgrad_extras(int sms) {
  const int M = 700, N = 500, K = 512;
  const int xs[4][3] = {{501, 500, 256}, {201, 400, 128}, {130, 70, 512}, {64, 256, 64}};   // {M, N, K}
  HostMat a[2], b[2], xa[4], xb[4];
  float* dout[2];
  float* xout[4];
  const long long ldo = 512;
  GemmProblem p;
  p.kind = GEMM_WGRAD; p.epi = EPI_STORE_F32; p.ngroups = 2; p.nseg = 1; p.M = M; p.N = N; p.K = K; p.block_n = g_bn; p.cta_pair = g_pair;
  for (int g = 0; g < 2; ++g) {
    a[g].init(K, M, (M + 63) / 64 * 64, 1.f, true);
    b[g].init(K, N, (N + 63) / 64 * 64, 1.f);
    CK(cudaMalloc(&dout[g], sizeof(float) * M * ldo));
    CK(cudaMemset(dout[g], 0xFF, sizeof(float) * M * ldo));
    p.a[g][0] = a[g].mat(); p.b[g][0] = b[g].mat();
    EpiArgs e = {};
    e.out = dout[g]; e.ld_out = ldo; e.alpha = 1.f;
    p.e[g] = e;
  }
  for (int i = 0; i < 4; ++i) {
    const int m = xs[i][0], n = xs[i][1], k = xs[i][2];
    xa[i].init(k, m, (m + 63) / 64 * 64, 1.f, i % 2 == 0);
    xb[i].init(k, n, (n + 63) / 64 * 64, 1.f);
    CK(cudaMalloc(&xout[i], sizeof(float) * m * ldo));
    CK(cudaMemset(xout[i], 0xFF, sizeof(float) * m * ldo));
    GemmExtra x;
    x.a[0] = xa[i].mat(); x.b[0] = xb[i].mat(); x.M = m; x.N = n; x.K = k;
    EpiArgs e = {};
    e.out = xout[i]; e.ld_out = ldo; e.alpha = 0.5f + i;
    x.e = e;
    p.extra.push_back(x);
  }
  GemmPlan pl;
  std::string err = gemm_plan(p, sms, &pl);
  if (!err.empty()) { printf("plan error: %s\n", err.c_str()); ++g_fail; return; }
  if (gemm_launch(pl, 0, &err) < 0) { printf("launch error: %s\n", err.c_str()); ++g_fail; return; }
  CK(cudaDeviceSynchronize());
  double me = 0;
  for (int g = 0; g < 2; ++g) {
    std::vector<float> h((size_t)M * ldo);
    CK(cudaMemcpy(h.data(), dout[g], sizeof(float) * M * ldo, cudaMemcpyDeviceToHost));
    std::vector<double> D;
    host_ref(GEMM_WGRAD, M, N, K, 1, &a[g], &b[g], D);
    for (int m = 0; m < M; ++m)
      for (int n = 0; n < N; ++n) me = fmax(me, fabs(h[(size_t)m * ldo + n] - D[(size_t)m * N + n]));
  }
  for (int i = 0; i < 4; ++i) {
    const int m = xs[i][0], n = xs[i][1], k = xs[i][2];
    std::vector<float> h((size_t)m * ldo);
    CK(cudaMemcpy(h.data(), xout[i], sizeof(float) * m * ldo, cudaMemcpyDeviceToHost));
    std::vector<double> D;
    host_ref(GEMM_WGRAD, m, n, k, 1, &xa[i], &xb[i], D);
    for (int r = 0; r < m; ++r)
      for (int c = 0; c < n; ++c) me = fmax(me, fabs(h[(size_t)r * ldo + c] - (0.5 + i) * D[(size_t)r * n + c]));
  }
  char name[128];
  snprintf(name, sizeof name, "wgrad 2 groups + 4 appended problems, bn=%d items=%lld", pl.args.bn, pl.items);
  report(name, me, 2e-3 * sqrt((double)K) * 0.1 * 4 + 1e-4);
}

// ---------------------------------------------------------------- A operand as packed bits (A_BITS kernel variants)
// Packs a binary HostMat (rows x cols) into bits on the device: element (r, c) -> bit c & 31 of word r * pitch + (c >> 5).
static uint32_t* pack_bits(const HostMat& m, long long pitch) {
  std::vector<uint32_t> h((size_t)m.rows * pitch, 0u);
  for (long long r = 0; r < m.rows; ++r)
    for (long long c = 0; c < m.cols; ++c)
      if (m.at(r, c) != 0.f) h[(size_t)r * pitch + (c >> 5)] |= 1u << (c & 31);
  uint32_t* d;
  CK(cudaMalloc(&d, sizeof(uint32_t) * h.size()));
  CK(cudaMemcpy(d, h.data(), sizeof(uint32_t) * h.size(), cudaMemcpyHostToDevice));
  return d;
}
static long long bits_pitch(int kind, int M, int K) {
  // (+ 1 tile: a CTA pair may run an empty partner tile behind an odd last M tile)
  const long long need = kind == GEMM_WGRAD ? 4LL * ((M + 127) / 128 + 1) : 2LL * ((K + 63) / 64);
  return (need + 7) / 8 * 8;
}

// The bits variant must reproduce the bf16-operand kernel bit for bit (same MMAs on the same shared-memory image) and
// both must match the host reference.  FWD with nseg segments and split-K, fp32 store.
static void test_bits_fwd_store(int M, int N, int K, int nseg, int split, int sms, bool sparse = false) {
  HostMat a, b[3];
  a.init(M, K, (K + 63) / 64 * 64, 1.f, true);
  if (sparse) a.punch(true, 1);
  for (int s = 0; s < nseg; ++s) b[s].init(K, N, (N + 63) / 64 * 64, 1.f);
  const long long pitch = bits_pitch(GEMM_FWD, M, K);
  uint32_t* dbits = pack_bits(a, pitch);
  const long long ldo = (N + 3) / 4 * 4;
  std::vector<float> h[2];
  int used_split = 0;
  // zero-stage map (sparse runs): built by the library's own kernel from the packed rows
  const int kblocks = (K + 63) / 64, mwords = (kblocks + 31) / 32, tiles = (M + 127) / 128;
  uint32_t* dmask = nullptr;
  if (sparse) {
    CK(cudaMalloc(&dmask, sizeof(uint32_t) * tiles * mwords));
    build_kmask_fwd(dbits, (int)pitch, M, kblocks, dmask, mwords, 0);
    std::vector<uint32_t> hm((size_t)tiles * mwords);
    CK(cudaMemcpy(hm.data(), dmask, sizeof(uint32_t) * hm.size(), cudaMemcpyDeviceToHost));
    long long set = 0, bad = 0;
    for (int t = 0; t < tiles; ++t)
      for (int kk = 0; kk < kblocks; ++kk) {
        bool any = false;
        for (int m = t * 128; m < std::min(M, t * 128 + 128) && !any; ++m)
          for (int k = kk * 64; k < std::min(K, kk * 64 + 64) && !any; ++k) any = a.at(m, k) != 0.f;
        const bool bit = (hm[(size_t)t * mwords + (kk >> 5)] >> (kk & 31)) & 1u;
        set += bit; bad += bit != any;
      }
    char nm[96];
    snprintf(nm, sizeof nm, "   fwd zero-stage map (%lld of %d set)", set, tiles * kblocks);
    report(nm, (double)bad, 0.0);
  }
  for (int v = 0; v < 2; ++v) {
    float* dout;
    CK(cudaMalloc(&dout, sizeof(float) * M * ldo));
    CK(cudaMemset(dout, 0, sizeof(float) * M * ldo));
    GemmProblem p;
    p.kind = GEMM_FWD; p.epi = EPI_STORE_F32; p.nseg = nseg; p.M = M; p.N = N; p.K = K; p.force_splitk = split; p.block_n = 256; p.cta_pair = g_pair;
    for (int s = 0; s < nseg; ++s) { p.a[0][s] = a.mat(); p.b[0][s] = b[s].mat(); }
    if (v == 1) { p.a_bits = true; p.abits[0].base = dbits; p.abits[0].pitch = pitch; p.abits[0].rows = M; }
    if (v == 1 && sparse) { p.kmask[0] = dmask; p.kmask_words = mwords; }
    EpiArgs e = {};
    e.out = dout; e.ld_out = ldo; e.alpha = 0.5f;
    p.e[0] = e;
    GemmPlan pl;
    std::string err = gemm_plan(p, sms, &pl);
    if (!err.empty()) { printf("plan error: %s\n", err.c_str()); ++g_fail; return; }
    gemm_schedule(pl, 0);            // plans with a map: cost-balanced work order (no-op otherwise)
    if (gemm_launch(pl, 0, &err) < 0) { printf("launch error: %s\n", err.c_str()); ++g_fail; return; }
    CK(cudaDeviceSynchronize());
    used_split = pl.args.splitk;
    h[v].resize((size_t)M * ldo);
    CK(cudaMemcpy(h[v].data(), dout, sizeof(float) * M * ldo, cudaMemcpyDeviceToHost));
    cudaFree(dout);
  }
  HostMat as[3] = {a, a, a};
  std::vector<double> D;
  host_ref(GEMM_FWD, M, N, K, nseg, as, b, D);
  double me = 0, diff = 0;
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) {
      me = fmax(me, fabs(h[1][(size_t)m * ldo + n] - 0.5 * D[(size_t)m * N + n]));
      diff = fmax(diff, fabs((double)h[1][(size_t)m * ldo + n] - (double)h[0][(size_t)m * ldo + n]));
    }
  char name[128];
  snprintf(name, sizeof name, "bits%s fwd store M=%d N=%d K=%d nseg=%d split=%d", sparse ? "+skip" : "", M, N, K, nseg, used_split);
  report(name, me, 2e-3 * sqrt((double)K * nseg) * 0.1 + 1e-4);
  // split-K sums partial tiles with red.add in arrival order: exact equality only without a split
  report("   bits vs bf16 operand", diff, used_split > 1 ? 1e-3 : 0.0);
  cudaFree(a.d); cudaFree(dbits);
  if (dmask) cudaFree(dmask);
  for (int s = 0; s < nseg; ++s) cudaFree(b[s].d);
}

// FWD with the fused tanh epilogue (the unsplit first layer at small batch)
static void test_bits_fwd_act(int M, int Nv, int K, int sms) {
  const long long ldo = (Nv + 1 + 63) / 64 * 64;
  HostMat a, b;
  a.init(M, K, (K + 63) / 64 * 64, 1.f, true);
  b.init(K, Nv, (Nv + 63) / 64 * 64, 0.05f);
  const long long pitch = bits_pitch(GEMM_FWD, M, K);
  uint32_t* dbits = pack_bits(a, pitch);
  std::vector<__nv_bfloat16> h[2];
  for (int v = 0; v < 2; ++v) {
    __nv_bfloat16* dout;
    CK(cudaMalloc(&dout, 2 * M * ldo));
    CK(cudaMemset(dout, 0xFF, 2 * M * ldo));
    GemmProblem p;
    p.kind = GEMM_FWD; p.epi = EPI_ACT; p.M = M; p.K = K; p.N = (int)ldo; p.force_splitk = 1; p.block_n = 256; p.cta_pair = g_pair;
    p.a[0][0] = a.mat(); p.b[0][0] = b.mat();
    if (v == 1) { p.a_bits = true; p.abits[0].base = dbits; p.abits[0].pitch = pitch; p.abits[0].rows = M; }
    EpiArgs e = {};
    e.out = dout; e.ld_out = ldo; e.alpha = 1.f; e.n_valid = Nv; e.act = ACT_TANH; e.precise = 1; e.aux_split = 1 << 30;
    p.e[0] = e;
    GemmPlan pl;
    std::string err = gemm_plan(p, sms, &pl);
    if (!err.empty()) { printf("plan error: %s\n", err.c_str()); ++g_fail; return; }
    if (gemm_launch(pl, 0, &err) < 0) { printf("launch error: %s\n", err.c_str()); ++g_fail; return; }
    CK(cudaDeviceSynchronize());
    h[v].resize((size_t)M * ldo);
    CK(cudaMemcpy(h[v].data(), dout, 2 * M * ldo, cudaMemcpyDeviceToHost));
    cudaFree(dout);
  }
  double me = 0, diff = 0;
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < (int)ldo; ++n) {
      double acc = 0;
      if (n < Nv) for (int k = 0; k < K; ++k) acc += (double)a.at(m, k) * b.at(k, n);
      const double want = n < Nv ? tanh(acc) : (n == Nv ? 1.0 : 0.0);
      me = fmax(me, fabs((double)__bfloat162float(h[1][(size_t)m * ldo + n]) - want));
      diff = fmax(diff, fabs((double)__bfloat162float(h[1][(size_t)m * ldo + n]) - (double)__bfloat162float(h[0][(size_t)m * ldo + n])));
    }
  char name[128];
  snprintf(name, sizeof name, "bits fwd tanh M=%d Nv=%d K=%d", M, Nv, K);
  report(name, me, 4e-3);
  report("   bits vs bf16 operand", diff, 0.0);
  cudaFree(a.d); cudaFree(b.d); cudaFree(dbits);
}

// WGRAD, two groups reading overlapping row ranges of one packed matrix (the engine's [rot; lock; key] stack), with and
// without appended problems that keep their bf16 A operand
static void test_bits_wgrad(int M, int N, int Bp, bool extras, int sms, bool sparse = false) {
  // X [3 Bp, M] binary; group 0 = rows [0, 2 Bp), group 1 = rows [Bp, 3 Bp); dY [4 Bp, N]: group 0 rows [0, 2 Bp), group 1 rows [2 Bp, 4 Bp)
  const int K = 2 * Bp;
  HostMat x, dy, xa[2], xb[2];
  x.init(3LL * Bp, M, (M + 63) / 64 * 64, 1.f, true);
  if (sparse) x.punch(false, 1);
  dy.init(4LL * Bp, N, (N + 63) / 64 * 64, 1.f);
  const long long pitch = bits_pitch(GEMM_WGRAD, M, K);
  uint32_t* dbits = pack_bits(x, pitch);
  const long long ldo = 512;
  const int xs[2][3] = {{501, 500, 256}, {130, 70, 192}};
  if (extras)
    for (int i = 0; i < 2; ++i) {
      xa[i].init(xs[i][2], xs[i][0], (xs[i][0] + 63) / 64 * 64, 1.f);
      xb[i].init(xs[i][2], xs[i][1], (xs[i][1] + 63) / 64 * 64, 1.f);
    }
  std::vector<float> h[2][2], hx[2][2];
  const int tiles_m = (M + 127) / 128, wg_kb = (K + 63) / 64, mwords = (wg_kb + 31) / 32;
  uint32_t* dmask[2] = {nullptr, nullptr};
  if (sparse)
    for (int g = 0; g < 2; ++g) {
      CK(cudaMalloc(&dmask[g], sizeof(uint32_t) * ((size_t)tiles_m * mwords + 4)));
      build_kmask_wgrad(dbits + (long long)g * Bp * pitch, (int)pitch, K, tiles_m, dmask[g], mwords, 0);
      std::vector<uint32_t> hm((size_t)tiles_m * mwords);
      CK(cudaMemcpy(hm.data(), dmask[g], sizeof(uint32_t) * hm.size(), cudaMemcpyDeviceToHost));
      long long set = 0, bad = 0;
      for (int t = 0; t < tiles_m; ++t)
        for (int kb = 0; kb < wg_kb; ++kb) {
          bool any = false;
          for (int k = kb * 64; k < std::min(K, kb * 64 + 64) && !any; ++k)
            for (int m = t * 128; m < std::min(M, t * 128 + 128) && !any; ++m) any = x.at((long long)g * Bp + k, m) != 0.f;
          const bool bit = (hm[(size_t)t * mwords + (kb >> 5)] >> (kb & 31)) & 1u;
          set += bit; bad += bit != any;
        }
      char nm[96];
      snprintf(nm, sizeof nm, "   wgrad zero-stage map, group %d (%lld of %d set)", g, set, tiles_m * wg_kb);
      report(nm, (double)bad, 0.0);
    }
  for (int v = 0; v < 2; ++v) {
    float* dout[2];
    float* xout[2] = {nullptr, nullptr};
    GemmProblem p;
    p.kind = GEMM_WGRAD; p.epi = EPI_STORE_F32; p.ngroups = 2; p.nseg = 1; p.M = M; p.N = N; p.K = K; p.block_n = 256; p.cta_pair = g_pair;
    p.force_splitk = 1;
    for (int g = 0; g < 2; ++g) {
      CK(cudaMalloc(&dout[g], sizeof(float) * M * ldo));
      CK(cudaMemset(dout[g], 0xFF, sizeof(float) * M * ldo));
      p.a[g][0] = Mat(x.d + (long long)g * Bp * x.ld, K, M, x.ld);
      p.b[g][0] = Mat(dy.d + (long long)g * 2 * Bp * dy.ld, K, N, dy.ld);
      if (v == 1) { p.a_bits = true; p.abits[g].base = dbits + (long long)g * Bp * pitch; p.abits[g].pitch = pitch; p.abits[g].rows = K; }
      if (v == 1 && sparse) { p.kmask[g] = dmask[g]; p.kmask_words = mwords; }
      EpiArgs e = {};
      e.out = dout[g]; e.ld_out = ldo; e.alpha = 1.f;
      p.e[g] = e;
    }
    if (extras)
      for (int i = 0; i < 2; ++i) {
        CK(cudaMalloc(&xout[i], sizeof(float) * xs[i][0] * ldo));
        CK(cudaMemset(xout[i], 0xFF, sizeof(float) * xs[i][0] * ldo));
        GemmExtra e2;
        e2.a[0] = xa[i].mat(); e2.b[0] = xb[i].mat(); e2.M = xs[i][0]; e2.N = xs[i][1]; e2.K = xs[i][2];
        EpiArgs e = {};
        e.out = xout[i]; e.ld_out = ldo; e.alpha = 2.f;
        e2.e = e;
        p.extra.push_back(e2);
      }
    GemmPlan pl;
    std::string err = gemm_plan(p, sms, &pl);
    if (!err.empty()) { printf("plan error: %s\n", err.c_str()); ++g_fail; return; }
    gemm_schedule(pl, 0);
    if (pl.order) {                  // the order must be a permutation of the work items
      std::vector<int> ho((size_t)pl.items);
      CK(cudaMemcpy(ho.data(), pl.order.get(), sizeof(int) * ho.size(), cudaMemcpyDeviceToHost));
      std::vector<int> seen((size_t)pl.items, 0);
      int bad = 0;
      for (int o : ho) { if (o < 0 || o >= (int)pl.items || seen[o]++) ++bad; }
      report("   work order is a permutation", (double)bad, 0.0);
    }
    if (gemm_launch(pl, 0, &err) < 0) { printf("launch error: %s\n", err.c_str()); ++g_fail; return; }
    CK(cudaDeviceSynchronize());
    for (int g = 0; g < 2; ++g) {
      h[v][g].resize((size_t)M * ldo);
      CK(cudaMemcpy(h[v][g].data(), dout[g], sizeof(float) * M * ldo, cudaMemcpyDeviceToHost));
      cudaFree(dout[g]);
    }
    if (extras)
      for (int i = 0; i < 2; ++i) {
        hx[v][i].resize((size_t)xs[i][0] * ldo);
        CK(cudaMemcpy(hx[v][i].data(), xout[i], sizeof(float) * xs[i][0] * ldo, cudaMemcpyDeviceToHost));
        cudaFree(xout[i]);
      }
  }
  double me = 0, diff = 0;
  for (int g = 0; g < 2; ++g)
    for (int m = 0; m < M; ++m)
      for (int n = 0; n < N; ++n) {
        double acc = 0;
        for (int k = 0; k < K; ++k) acc += (double)x.at((long long)g * Bp + k, m) * dy.at((long long)g * 2 * Bp + k, n);
        me = fmax(me, fabs(h[1][g][(size_t)m * ldo + n] - acc));
        diff = fmax(diff, fabs((double)h[1][g][(size_t)m * ldo + n] - (double)h[0][g][(size_t)m * ldo + n]));
      }
  if (extras)
    for (int i = 0; i < 2; ++i) {
      std::vector<double> D;
      host_ref(GEMM_WGRAD, xs[i][0], xs[i][1], xs[i][2], 1, &xa[i], &xb[i], D);
      for (int r = 0; r < xs[i][0]; ++r)
        for (int c = 0; c < xs[i][1]; ++c) {
          me = fmax(me, fabs(hx[1][i][(size_t)r * ldo + c] - 2.0 * D[(size_t)r * xs[i][1] + c]));
          diff = fmax(diff, fabs((double)hx[1][i][(size_t)r * ldo + c] - (double)hx[0][i][(size_t)r * ldo + c]));
        }
    }
  char name[128];
  snprintf(name, sizeof name, "bits%s wgrad 2 groups M=%d N=%d K=%d%s", sparse ? "+skip" : "", M, N, K, extras ? " + 2 appended" : "");
  report(name, me, 2e-3 * sqrt((double)K) * 0.1 * 2 + 1e-4);
  report("   bits vs bf16 operand", diff, 0.0);
  cudaFree(x.d); cudaFree(dy.d); cudaFree(dbits);
  for (int g = 0; g < 2; ++g) if (dmask[g]) cudaFree(dmask[g]);
  if (extras) for (int i = 0; i < 2; ++i) { cudaFree(xa[i].d); cudaFree(xb[i].d); }
}

static double sigm(double x) { return 1.0 / (1.0 + exp(-x)); }

// ACT (fwd, optional two-pass), DACT (dgrad), BCE (fwd)
static void test_fused(int epi, int M, int Nv, int K, int split, int precise, int sms) {
  const int kind = epi == EPI_DACT ? GEMM_DGRAD : GEMM_FWD;
  const long long ldo = (Nv + 1 + 63) / 64 * 64;
  HostMat a, b, aux;
  a.init(M, K, (K + 63) / 64 * 64, 1.f);
  if (kind == GEMM_DGRAD) b.init(Nv + 1, K, (K + 63) / 64 * 64, 0.5f); else b.init(K, Nv, (Nv + 63) / 64 * 64, 0.5f);
  aux.init(M, ldo, ldo, 1.f, epi == EPI_BCE);
  __nv_bfloat16 *dout, *dlo;
  CK(cudaMalloc(&dout, 2 * M * ldo));
  CK(cudaMalloc(&dlo, 2 * M * ldo));
  CK(cudaMemset(dout, 0xFF, 2 * M * ldo));
  CK(cudaMemset(dlo, 0, 2 * M * ldo));
  // BCE target as packed bits (bit n & 31 of word n >> 5 of row m)
  const int xb_ld = (int)((ldo + 31) / 32 + 7) / 8 * 8;
  uint32_t* dbits = nullptr;
  if (epi == EPI_BCE) {
    std::vector<uint32_t> hb((size_t)M * xb_ld, 0u);
    for (int m = 0; m < M; ++m)
      for (int n = 0; n < Nv; ++n)
        if (aux.at(m, n) != 0.f) hb[(size_t)m * xb_ld + (n >> 5)] |= 1u << (n & 31);
    CK(cudaMalloc(&dbits, sizeof(uint32_t) * hb.size()));
    CK(cudaMemcpy(dbits, hb.data(), sizeof(uint32_t) * hb.size(), cudaMemcpyHostToDevice));
  }
  float* scalar;
  CK(cudaMalloc(&scalar, 4));
  CK(cudaMemset(scalar, 0, 4));
  float* ws = nullptr;
  CK(cudaMalloc(&ws, sizeof(float) * M * ldo));
  GemmProblem p;
  p.kind = kind; p.epi = epi; p.M = M; p.K = K; p.force_splitk = split;
  p.cluster_m = g_cluster; p.cta_pair = g_pair;
  p.block_n = g_bn;
  p.N = epi == EPI_BCE ? Nv : (int)ldo;
  p.a[0][0] = a.mat(); p.b[0][0] = b.mat();
  p.ws = ws; p.ws_ld = ldo;
  EpiArgs e = {};
  e.out = dout; e.out_lo = (epi == EPI_BCE && !precise) ? nullptr : dlo; e.ld_out = ldo;
  e.alpha = epi == EPI_BCE ? 0.25f : 1.f; e.n_valid = Nv;
  e.xbits = dbits; e.xb_ld = xb_ld;
  e.act = ACT_TANH; e.precise = precise; e.aux = aux.d; e.ld_aux = ldo; e.aux_split = M / 2; e.aux_shift = epi == EPI_DACT ? M / 4 : 0;
  e.scalar = scalar;
  if (epi != EPI_DACT) { e.aux_split = 1 << 30; }
  p.e[0] = e;
  GemmPlan pl;
  std::string err = gemm_plan(p, sms, &pl);
  if (!err.empty()) { printf("plan error: %s\n", err.c_str()); ++g_fail; return; }
  if (gemm_launch(pl, 0, &err) < 0) { printf("launch error: %s\n", err.c_str()); ++g_fail; return; }
  CK(cudaDeviceSynchronize());
  std::vector<__nv_bfloat16> h(M * ldo), hl(M * ldo);
  CK(cudaMemcpy(h.data(), dout, 2 * M * ldo, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(hl.data(), dlo, 2 * M * ldo, cudaMemcpyDeviceToHost));
  float hs = 0;
  CK(cudaMemcpy(&hs, scalar, 4, cudaMemcpyDeviceToHost));
  double me = 0, me_pad = 0, loss = 0;
  const int Ncols = epi == EPI_BCE ? Nv : (int)ldo;
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < Ncols; ++n) {
      double acc = 0;
      if (n < (kind == GEMM_DGRAD ? Nv + 1 : Nv))
        for (int k = 0; k < K; ++k) acc += (double)a.at(m, k) * (kind == GEMM_DGRAD ? b.at(n, k) : b.at(k, n));
      double want;
      if (epi == EPI_ACT) want = n < Nv ? tanh(acc) : (n == Nv ? 1.0 : 0.0);
      else if (epi == EPI_DACT) {
        int ar = m < M / 2 ? m : m - M / 4;
        double y = aux.at(ar, n);
        want = n < Nv ? acc * (1 - y * y) : 0.0;
      } else {
        double x = aux.at(m, n);
        want = sigm(acc) - x;
        loss += log1p(exp(-fabs(acc))) + fmax(acc, 0.0) - x * acc;
      }
      double got = (double)__bfloat162float(h[m * ldo + n]) + (double)__bfloat162float(hl[m * ldo + n]);
      double err1 = fabs(got - want);
      if (n < Nv) me = fmax(me, err1); else me_pad = fmax(me_pad, err1);
    }
  char name[128];
  snprintf(name, sizeof name, "fused epi=%d M=%d Nv=%d K=%d split=%d(%d,2pass=%d) prec=%d", epi, M, Nv, K, split,
           pl.args.splitk, (int)pl.two_pass, precise);
  report(name, me, precise ? 3e-4 : 4e-3);
  if (epi != EPI_BCE) report("   pad/ones columns", me_pad, 0.0);
  if (epi == EPI_BCE) {
    double rel = fabs(hs - 0.25 * loss) / fabs(0.25 * loss);
    report("   bce loss sum (relative)", rel, precise ? 1e-5 : 1e-4);
  }
  cudaFree(a.d); cudaFree(b.d); cudaFree(aux.d); cudaFree(dout); cudaFree(dlo); cudaFree(scalar); cudaFree(ws);
  if (dbits) cudaFree(dbits);
}

static void bench(int kind, int M, int N, int K, int split, int sms, const char* label, int epi = EPI_STORE_F32) {
  __nv_bfloat16 *a, *b;
  size_t na = (size_t)(kind == GEMM_WGRAD ? K : M) * (kind == GEMM_WGRAD ? (M + 63) / 64 * 64 : (K + 63) / 64 * 64);
  size_t nb = (size_t)(kind == GEMM_DGRAD ? N : K) * (kind == GEMM_DGRAD ? (K + 63) / 64 * 64 : (N + 63) / 64 * 64);
  CK(cudaMalloc(&a, na * 2)); CK(cudaMalloc(&b, nb * 2));
  CK(cudaMemset(a, 0, na * 2)); CK(cudaMemset(b, 0, nb * 2));
  float* out;
  long long ldo = (N + 3) / 4 * 4;
  CK(cudaMalloc(&out, sizeof(float) * M * ldo));
  CK(cudaMemset(out, 0, sizeof(float) * M * ldo));
  GemmProblem p;
  p.kind = kind; p.epi = epi; p.M = M; p.N = N; p.K = K; p.force_splitk = split;
  p.cluster_m = g_cluster; p.cta_pair = g_pair;
  p.block_n = g_bn;
  p.a[0][0] = kind == GEMM_WGRAD ? Mat(a, K, M, (M + 63) / 64 * 64) : Mat(a, M, K, (K + 63) / 64 * 64);
  p.b[0][0] = kind == GEMM_DGRAD ? Mat(b, N, K, (K + 63) / 64 * 64) : Mat(b, K, N, (N + 63) / 64 * 64);
  EpiArgs e = {};
  e.out = out; e.ld_out = ldo; e.alpha = 1.f;
  if (epi == EPI_BCE) {   // bf16 gradient into the same buffer, all-zero packed targets, loss scalar
    uint32_t* bits;
    float* scalar;
    const int xb_ld = (N + 31) / 32;
    CK(cudaMalloc(&bits, sizeof(uint32_t) * (size_t)M * xb_ld));
    CK(cudaMemset(bits, 0, sizeof(uint32_t) * (size_t)M * xb_ld));
    CK(cudaMalloc(&scalar, 4));
    CK(cudaMemset(scalar, 0, 4));
    e.xbits = bits; e.xb_ld = xb_ld; e.scalar = scalar; e.n_valid = N; e.aux_split = 1 << 30;
    e.ld_out = (N + 7) / 8 * 8;
  }
  if (epi == EPI_ACT || epi == EPI_DACT) {   // bf16 activation out (hi plane only), tanh; DACT reads `b` as the stored y
    e.act = ACT_TANH; e.n_valid = N - 1; e.ld_out = (N + 63) / 64 * 64; e.aux_split = 1 << 30;
    e.aux = reinterpret_cast<const __nv_bfloat16*>(out); e.ld_aux = e.ld_out;
    p.N = (int)e.ld_out;
  }
  if (const char* d = getenv("MAGIC_EPI_DBG")) e.dbg = atoi(d);
  p.e[0] = e;
  uint32_t* abits = nullptr;
  if (g_bits) {
    const long long pitch = bits_pitch(kind, M, K), rows = kind == GEMM_WGRAD ? K : M;
    CK(cudaMalloc(&abits, sizeof(uint32_t) * rows * pitch));
    CK(cudaMemset(abits, 0x5A, sizeof(uint32_t) * rows * pitch));
    p.a_bits = true; p.abits[0].base = abits; p.abits[0].pitch = pitch; p.abits[0].rows = (int)rows; p.block_n = 256;
    if (g_density < 1.f) {     // a random zero-stage map with this fraction of set stages (the bits stay as they are: timing only)
      const int tiles = (M + 127) / 128, kbl = (K + 63) / 64, mw = (kbl + 31) / 32;
      std::vector<uint32_t> hm((size_t)tiles * mw, 0u);
      for (int t = 0; t < tiles; ++t)
        for (int kk = 0; kk < kbl; ++kk)
          if (frand() + 0.5f < g_density) hm[(size_t)t * mw + (kk >> 5)] |= 1u << (kk & 31);
      uint32_t* dm;
      CK(cudaMalloc(&dm, sizeof(uint32_t) * hm.size()));
      CK(cudaMemcpy(dm, hm.data(), sizeof(uint32_t) * hm.size(), cudaMemcpyHostToDevice));
      p.kmask[0] = dm; p.kmask_words = mw;
    }
  }
  GemmPlan pl;
  std::string err = gemm_plan(p, sms, &pl);
  if (!err.empty()) { printf("plan error: %s\n", err.c_str()); return; }
  gemm_schedule(pl, 0);
  if (getenv("MAGIC_TRACE")) {   // per-CTA phase stamps of one warm launch
    long long* tr;
    CK(cudaMalloc(&tr, sizeof(long long) * 64 * pl.grid));
    for (int i = 0; i < 3; ++i) gemm_launch(pl, 0, &err);
    CK(cudaDeviceSynchronize());
    CK(cudaMemset(tr, 0, sizeof(long long) * 64 * pl.grid));
    pl.args.trace = tr;
    gemm_launch(pl, 0, &err);
    CK(cudaDeviceSynchronize());
    pl.args.trace = nullptr;
    std::vector<long long> h((size_t)64 * pl.grid);
    CK(cudaMemcpy(h.data(), tr, sizeof(long long) * h.size(), cudaMemcpyDeviceToHost));
    const char* names[11] = {"entry", "prologue done", "dep wait done", "first TMA issued", "first stage landed",
                             "tile 0 MMAs issued", "all MMAs issued", "epi: tile 0 acc ready", "epi: loop done",
                             "epi: stores drained", "exit"};
    long long gmin = h[11], gmax = h[11];
    for (int c = 0; c < pl.grid; ++c) { gmin = std::min(gmin, h[(size_t)c * 64 + 11]); gmax = std::max(gmax, h[(size_t)c * 64 + 11]); }
    printf("trace %-28s grid=%d, CTA start spread %.2f us; mean / max cycles since CTA entry:\n", label, pl.grid, (gmax - gmin) * 1e-3);
    for (int sidx = 1; sidx <= 10; ++sidx) {
      double sum = 0; long long mx = 0; int n = 0;
      for (int c = 0; c < pl.grid; ++c) {
        long long v = h[(size_t)c * 64 + sidx], b = h[(size_t)c * 64];
        if (v == 0) continue;
        sum += (double)(v - b); mx = std::max(mx, v - b); ++n;
      }
      if (n) printf("   %-24s %9.0f %9lld   (%.2f / %.2f us at 1.9 GHz)\n", names[sidx], sum / n, mx, sum / n / 1900.0, mx / 1900.0);
    }
    // pipeline of the first tiles of CTA 0 (cycles since its entry)
    printf("   CTA 0 tiles: [MMA start, MMA all issued | epi: acc ready, acc released, tile done]\n");
    for (int it = 0; it < 8; ++it) {
      const long long* t = &h[16 + it * 5];
      if (t[0] == 0) break;
      printf("     tile %d: %8lld %8lld | %8lld %8lld %8lld\n", it, t[0] - h[0], t[1] - h[0], t[2] - h[0], t[3] - h[0], t[4] - h[0]);
    }
    CK(cudaFree(tr));
  }
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i) gemm_launch(pl, 0, &err);
  CK(cudaDeviceSynchronize());
  const int iters = 10;
  cudaEventRecord(e0);
  for (int i = 0; i < iters; ++i) gemm_launch(pl, 0, &err);
  cudaEventRecord(e1);
  CK(cudaDeviceSynchronize());
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  ms /= iters;
  printf("bench %-28s M=%d N=%d K=%d split=%d grid=%d: %.3f ms  %.1f TFLOP/s\n", label, M, N, K, pl.args.splitk, pl.grid,
         ms, 2.0 * M * N * K / ms * 1e-9);
  cudaFree(a); cudaFree(b); cudaFree(out);
  if (abits) cudaFree(abits);
}

int main(int argc, char** argv) {
  int dev = 0;
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, dev));
  printf("device: %s, %d SMs, cc %d.%d\n", prop.name, prop.multiProcessorCount, prop.major, prop.minor);
  const int sms = prop.multiProcessorCount;
  setenv("MAGIC_CG2", "1", 0);       // the CTA-pair variants are opt-in in the library; the self-test always checks them
  std::string err = gemm_init();
  if (!err.empty()) { printf("gemm_init: %s\n", err.c_str()); return 2; }
  const bool do_bench = argc > 1 && std::string(argv[1]) == "bench";
  const bool pair_only = argc > 1 && std::string(argv[1]) == "pair";    // the CTA-pair checks and their timings only
  auto pair_bench = [&]() {
    for (int pr = 0; pr < 2; ++pr) {
      g_pair = pr != 0;
      const char* tag = pr ? " [CTA pairs]" : "";
      char lbl[96];
      snprintf(lbl, sizeof lbl, "dec out fwd, BCE%s", tag); bench(GEMM_FWD, 4096, 40000, 512, 1, sms, lbl, EPI_BCE);
      snprintf(lbl, sizeof lbl, "dec out dgrad (auto)%s", tag); bench(GEMM_DGRAD, 4096, 512, 40000, 0, sms, lbl);
      snprintf(lbl, sizeof lbl, "dec out wgrad%s", tag); bench(GEMM_WGRAD, 512, 40000, 4096, 1, sms, lbl);
      snprintf(lbl, sizeof lbl, "enc L0 wgrad, bf16 A%s", tag); bench(GEMM_WGRAD, 40064, 512, 8192, 1, sms, lbl);
      snprintf(lbl, sizeof lbl, "square 8192 dgrad%s", tag); bench(GEMM_DGRAD, 8192, 8192, 8192, 1, sms, lbl);
      g_bits = 1;
      snprintf(lbl, sizeof lbl, "enc L0 fwd, bits (auto)%s", tag); bench(GEMM_FWD, 12288, 512, 40064, 0, sms, lbl);
      snprintf(lbl, sizeof lbl, "enc L0 wgrad, bits%s", tag); bench(GEMM_WGRAD, 40064, 512, 8192, 1, sms, lbl);
      g_bits = 0;
    }
    g_pair = false;
  };
  if (!pair_only) {
  // plain GEMMs, all three operand layouts, K tails, M/N tails, segments, split-K
  test_store(GEMM_FWD, 128, 256, 64, 1, 1, sms);
  test_store(GEMM_FWD, 128, 256, 256, 1, 1, sms);
  test_store(GEMM_DGRAD, 128, 256, 64, 1, 1, sms);
  test_store(GEMM_WGRAD, 128, 256, 64, 1, 1, sms);
  test_store(GEMM_FWD, 300, 500, 200, 1, 1, sms);
  test_store(GEMM_FWD, 300, 500, 1000, 2, 3, sms);
  test_store(GEMM_DGRAD, 300, 501, 840, 1, 1, sms);
  test_store(GEMM_DGRAD, 300, 501, 840, 3, 4, sms);
  test_store(GEMM_WGRAD, 501, 500, 700, 1, 1, sms);
  test_store(GEMM_WGRAD, 501, 40, 1024, 2, 0, sms);
  test_store(GEMM_FWD, 1000, 4, 501, 1, 0, sms);
  test_wgrad_groups(1100, 500, 512, sms);
  g_bn = 256; test_wgrad_extras(sms); g_bn = 0;      // the engine's case: 256-wide tiles (first-layer wgrad fills the grid)
  test_wgrad_extras(sms);
  // fused epilogues
  test_fused(EPI_ACT, 300, 500, 501, 1, 1, sms);
  test_fused(EPI_ACT, 300, 500, 501, 1, 0, sms);
  test_fused(EPI_ACT, 300, 500, 2000, 4, 1, sms);
  test_fused(EPI_DACT, 512, 500, 400, 1, 1, sms);
  test_fused(EPI_DACT, 512, 500, 4000, 5, 1, sms);
  test_fused(EPI_BCE, 300, 1000, 501, 1, 1, sms);
  test_fused(EPI_BCE, 300, 1000, 501, 1, 0, sms);
  // thread-block clusters along M with the B tile multicast (FWD / WGRAD kinds)
  for (int cm = 2; cm <= 4; cm *= 2) {
    g_cluster = cm;
    test_store(GEMM_FWD, 1024, 1000, 520, 1, 1, sms);
    test_store(GEMM_FWD, 1000, 300, 2000, 2, 3, sms);
    test_store(GEMM_WGRAD, 1500, 500, 700, 1, 1, sms);
    test_fused(EPI_BCE, 1024, 3000, 501, 1, 1, sms);
    test_fused(EPI_BCE, 2000, 1000, 501, 1, 0, sms);
    test_fused(EPI_ACT, 2048, 500, 501, 1, 1, sms);
  }
  g_cluster = 1;
  }
  // CTA pairs (tcgen05 cta_group::2): every instantiated variant, odd and even numbers of M tiles, split-K, segments, groups,
  // appended problems, packed A
  g_pair = true; g_bn = 256;
  printf("-- CTA pairs\n");
  test_store(GEMM_DGRAD, 128, 256, 64, 1, 1, sms);
  test_store(GEMM_DGRAD, 300, 501, 840, 1, 1, sms);
  test_store(GEMM_DGRAD, 1300, 501, 8400, 3, 4, sms);
  test_store(GEMM_DGRAD, 4096, 500, 4000, 1, 0, sms);
  test_store(GEMM_WGRAD, 128, 256, 64, 1, 1, sms);
  test_store(GEMM_WGRAD, 501, 500, 700, 1, 1, sms);
  test_store(GEMM_WGRAD, 1500, 1000, 2048, 2, 1, sms);
  test_store(GEMM_WGRAD, 501, 140, 2048, 2, 0, sms);
  test_wgrad_groups(1100, 500, 512, sms);
  test_wgrad_extras(sms);
  test_fused(EPI_BCE, 300, 1000, 501, 1, 1, sms);
  test_fused(EPI_BCE, 300, 1000, 501, 1, 0, sms);
  test_fused(EPI_BCE, 2048, 3000, 501, 1, 0, sms);
  test_bits_fwd_store(128, 256, 64, 1, 1, sms);
  test_bits_fwd_store(300, 500, 1601, 2, 1, sms);
  test_bits_fwd_store(1000, 500, 3000, 2, 3, sms);
  test_bits_fwd_act(300, 500, 1601, sms);
  test_bits_wgrad(700, 500, 96, false, sms);
  test_bits_wgrad(1601, 120, 100, false, sms);
  test_bits_wgrad(1601, 500, 256, true, sms);
  g_pair = false; g_bn = 0;
  if (pair_only) {
    printf("%s (%d failures)\n", g_fail ? "SELFTEST FAILED" : "SELFTEST PASSED", g_fail);
    pair_bench();
    return g_fail ? 1 : 0;
  }
  // 128-wide tiles (6 pipeline stages, 4 accumulator stages), every kind / epilogue instantiated at that width
  g_bn = 128;
  test_store(GEMM_FWD, 128, 128, 64, 1, 1, sms);
  test_store(GEMM_FWD, 300, 500, 200, 1, 1, sms);
  test_store(GEMM_FWD, 2000, 500, 1000, 2, 3, sms);
  test_store(GEMM_DGRAD, 300, 501, 840, 1, 1, sms);
  test_store(GEMM_DGRAD, 1300, 501, 840, 3, 4, sms);
  test_store(GEMM_WGRAD, 501, 500, 700, 1, 1, sms);
  test_store(GEMM_WGRAD, 501, 140, 2048, 2, 0, sms);
  test_wgrad_groups(1100, 500, 512, sms);
  test_wgrad_extras(sms);
  test_fused(EPI_ACT, 300, 500, 501, 1, 1, sms);
  test_fused(EPI_ACT, 3000, 500, 501, 1, 0, sms);
  test_fused(EPI_ACT, 300, 500, 2000, 4, 1, sms);
  test_fused(EPI_DACT, 512, 500, 400, 1, 1, sms);
  test_fused(EPI_DACT, 5120, 500, 512, 1, 1, sms);
  test_fused(EPI_DACT, 512, 500, 4000, 5, 1, sms);
  g_bn = 0;
  // A operand as packed bits: every A_BITS instantiation except the peer scatter (engine-level DP test)
  test_bits_fwd_store(128, 256, 64, 1, 1, sms);
  test_bits_fwd_store(300, 500, 1000, 1, 1, sms);
  test_bits_fwd_store(300, 500, 1601, 2, 1, sms);
  test_bits_fwd_store(1000, 500, 3000, 2, 3, sms);
  test_bits_fwd_act(300, 500, 1601, sms);
  test_bits_wgrad(700, 500, 96, false, sms);
  test_bits_wgrad(1601, 120, 100, false, sms);
  test_bits_wgrad(1601, 500, 256, true, sms);
  // ... with zero-stage maps: skipped stages, interleaved split-K slices, items without any set stage
  test_bits_fwd_store(600, 500, 1601, 1, 1, sms, true);
  test_bits_fwd_store(600, 500, 1601, 2, 1, sms, true);
  test_bits_fwd_store(1000, 500, 3000, 2, 3, sms, true);
  test_bits_fwd_store(300, 256, 5000, 1, 7, sms, true);
  test_bits_wgrad(700, 500, 96, false, sms, true);
  test_bits_wgrad(1601, 500, 256, true, sms, true);
  printf("%s (%d failures)\n", g_fail ? "SELFTEST FAILED" : "SELFTEST PASSED", g_fail);
  if (do_bench) {
    pair_bench();
    bench(GEMM_FWD, 12288, 512, 40064, 1, sms, "enc L0 fwd");
    bench(GEMM_FWD, 12288, 512, 40064, 0, sms, "enc L0 fwd (auto split)");
    bench(GEMM_FWD, 4096, 40000, 512, 1, sms, "dec out fwd");
    bench(GEMM_FWD, 4096, 40000, 512, 1, sms, "dec out fwd, BCE epilogue", EPI_BCE);
    g_cluster = 2; bench(GEMM_FWD, 4096, 40000, 512, 1, sms, "dec out fwd, cluster 2");
    g_cluster = 4; bench(GEMM_FWD, 4096, 40000, 512, 1, sms, "dec out fwd, cluster 4");
    g_cluster = 4; bench(GEMM_WGRAD, 40064, 512, 8192, 1, sms, "enc L0 wgrad, cluster 4");
    g_cluster = 4; bench(GEMM_FWD, 12288, 512, 40064, 1, sms, "enc L0 fwd, cluster 4");
    g_cluster = 1;
    bench(GEMM_DGRAD, 4096, 512, 40000, 0, sms, "dec out dgrad (auto split)");
    bench(GEMM_WGRAD, 512, 40000, 4096, 1, sms, "dec out wgrad");
    bench(GEMM_WGRAD, 40064, 512, 8192, 1, sms, "enc L0 wgrad");
    g_bits = 1;
    bench(GEMM_FWD, 12288, 512, 40064, 1, sms, "enc L0 fwd, A as bits");
    bench(GEMM_FWD, 12288, 512, 40064, 0, sms, "enc L0 fwd, A as bits (auto split)");
    bench(GEMM_WGRAD, 40064, 512, 8192, 1, sms, "enc L0 wgrad, A as bits");
    for (float dn : {0.9999f, 0.75f, 0.5f, 0.25f}) {
      g_density = dn;
      char lbl[64];
      snprintf(lbl, sizeof lbl, "enc L0 fwd, bits, %.0f%% stages", dn * 100);
      bench(GEMM_FWD, 12288, 512, 40064, 0, sms, lbl);
      snprintf(lbl, sizeof lbl, "enc L0 wgrad, bits, %.0f%% stages", dn * 100);
      bench(GEMM_WGRAD, 40064, 512, 8192, 1, sms, lbl);
    }
    g_density = 1.f;
    g_bits = 0;
    bench(GEMM_FWD, 12288, 512, 512, 1, sms, "hidden fwd");
    bench(GEMM_FWD, 12288, 512, 512, 1, sms, "hidden fwd, tanh epilogue", EPI_ACT);
    bench(GEMM_DGRAD, 16384, 512, 512, 1, sms, "hidden dgrad, tanh' epilogue", EPI_DACT);
    g_bn = 256;
    bench(GEMM_FWD, 12288, 512, 512, 1, sms, "hidden fwd, tanh, BN=256", EPI_ACT);
    bench(GEMM_DGRAD, 16384, 512, 512, 1, sms, "hidden dgrad, tanh', BN=256", EPI_DACT);
    bench(GEMM_FWD, 4096, 512, 512, 1, sms, "dec hidden fwd, tanh, BN=256", EPI_ACT);
    g_bn = 128;
    bench(GEMM_FWD, 4096, 512, 512, 1, sms, "dec hidden fwd, tanh, BN=128", EPI_ACT);
    g_bn = 0;
    bench(GEMM_WGRAD, 512, 512, 8192, 0, sms, "hidden wgrad (auto split)");
    bench(GEMM_FWD, 8192, 8192, 8192, 1, sms, "square 8192");
    bench(GEMM_FWD, 4096, 40000, 256, 1, sms, "dec-out shape, K=256");
    bench(GEMM_FWD, 4096, 40000, 1024, 1, sms, "dec-out shape, K=1024");
    bench(GEMM_FWD, 4096, 40000, 2048, 1, sms, "dec-out shape, K=2048");
    bench(GEMM_FWD, 4096, 40000, 4096, 1, sms, "dec-out shape, K=4096");
    bench(GEMM_DGRAD, 4096, 40000, 512, 1, sms, "dec-out shape, K=512, B K-major");
    bench(GEMM_WGRAD, 4096, 40000, 512, 1, sms, "dec-out shape, K=512, A,B MN-major");
  }
  return g_fail ? 1 : 0;
}
