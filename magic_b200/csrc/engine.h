// The training engine behind the C ABI: parameter layout, activations, GEMM plans, step orchestration.
#pragma once
#include <string>
#include <vector>
#include "magic_b200.h"
#include "gemm.h"
#include "kernels.h"

namespace magic {

typedef __nv_bfloat16 bf16;

// One dense layer y = [x | 1] . Wp, with the bias stored as the extra row Kv of the padded weight matrix
// (bias-as-row: biases and bias gradients ride through the GEMMs).
struct Linear {
  int Kv = 0, Nv = 0;          // valid fan-in / fan-out
  long long ld = 0;            // row pitch of Wp (elements), multiple of 64
  long long off = 0;           // offset of Wp in the flat parameter buffers
  long long rows() const { return Kv + 1; }
  long long size() const { return rows() * ld; }
};

// A reference variable (8c/vae.py:89-147) as a rectangular window of a Linear's padded matrix.
struct ParamView {
  std::string name;
  int layer = 0;               // index into Engine::lin
  int row0 = 0, nrows = 0;     // window rows
  int col0 = 0, ncols = 0;     // window cols
  bool is_bias = false;
  int optimisers = 1;          // bitmask: 1 = cost optimiser, 2 = metric optimiser
};

struct Act {                   // bf16 activation buffer (+ low part in the fp32 mode)
  bf16* hi = nullptr;
  bf16* lo = nullptr;
  long long rows = 0, ld = 0;
};

struct NcclApi;

class Engine {
 public:
  Engine() {}
  ~Engine();
  std::string init(const magic_config& cfg, cudaStream_t stream);

  std::string load_interleaved(const float* d_x);
  std::string load_planar_u8(const uint8_t* l, const uint8_t* r, const uint8_t* k);
  std::string generate(uint64_t seed, uint64_t step);
  std::string export_batch(float* d_x, float* d_labels);
  std::string export_rotation(int32_t* d_cs_sn);
  std::string export_latents(float* d_z3);
  std::string swap_batch();
  std::string load_pairs_u8(const uint8_t* d_lock, const uint8_t* d_key, uint64_t seed, uint64_t step_idx);
  std::string load_pairs_bits(const uint32_t* d_lock, const uint32_t* d_key, int wpr, uint64_t seed, uint64_t step_idx);
  std::string feed_pairs_bits_host(const uint32_t* h_lock, const uint32_t* h_key, const float* h_labels, int wpr, uint64_t seed,
                                   uint64_t step_idx, bool activate);
  std::string feed_pairs_bits_gather(const uint32_t* h_lock_all, const uint32_t* h_key_all, int wpr, long long n_rows,
                                     const int32_t* h_idx, const float* h_labels, uint64_t seed, uint64_t step_idx, bool activate);
  std::string step(const float* d_labels, const float* d_eps3, float* d_losses5, float* d_distance, bool apply);
  std::string load_interleaved_u8(const uint8_t* d_x);
  std::string partial_fit_host(const void* h_x, bool x_is_u8, const float* h_labels, const float* h_eps3,
                               float* h_losses5, float* h_distance);
  std::string forward(int what, const float* d_eps3, float* d_out);
  std::string decode(const float* d_z, float* d_out);
  std::string set_option(const char* name, int value);
  std::string read_losses(float* h_losses5);
  std::string input_density(double* out2);    // fraction of operand stages with a set pixel: [forward, weight gradient]
  std::string comm_init(const void* uid);
  // Fused data parallelism over peer memory: every rank exports CUDA IPC handles of its gradient
  // staging buffer and shadow weights (kPeerBlobBytes each), the host framework all-gathers them, every rank imports.
  static constexpr int kPeerBlobBytes = 256;
  std::string comm_export_peer(void* h_blob);
  std::string comm_import_peers(const void* h_blobs /* world x kPeerBlobBytes, rank order */);
  // per-launch CUDA-event timing (bench / optimisation aid): tag -> accumulated ms, launches
  void profile_enable(bool on);
  std::string profile_report();     // synchronises; "tag ms count flops\n" lines; resets the accumulators

  std::string copy_param(int i, float* h_dst, const float* h_src, const float* base);   // get (h_dst) or set (h_src)
  std::string get_param(int i, float* h);
  std::string set_param(int i, const float* h);
  std::string get_grad(int which, int i, float* h);
  std::string slot(int which, int slot, int i, float* h_dst, const float* h_src);

  magic_config cfg{};
  cudaStream_t stream = nullptr;
  std::vector<Linear> lin;          // encoder hidden 0..E-1, heads, dec h1, dec h2, dec out
  std::vector<ParamView> views;
  long long n_both = 0, n_total = 0;
  int64_t t = 0;                    // optimiser step counter
  magic_stats stats{};

 private:
  std::string alloc(void** p, size_t bytes);
  std::string alloc_act(Act* a, long long rows, int width_plus_bias);
  std::string build_plans();
  std::string run_forward(const float* d_eps3, bool train, int* launches);
  std::string run_backward(int* launches);
  std::string launch(GemmPlan& pl, int* launches, const char* tag);
  void prof_begin(const char* tag, double flops = 0, cudaStream_t on = nullptr);
  void prof_end(cudaStream_t on = nullptr);
  struct ProfRec { std::string tag; cudaEvent_t a, b; double flops; };
  std::vector<ProfRec> prof_pending_;
  std::vector<int> prof_open_;      // indices of the records whose end has not been recorded yet (innermost last)
  std::vector<cudaEvent_t> prof_pool_;
  struct ProfAcc { std::string tag; double ms = 0; long long n = 0; double flops = 0; };
  std::vector<ProfAcc> prof_acc_;
  bool profile_ = false;
  void refresh_shadow();
  LatentArgs latent_args(const float* d_eps3, const float* d_labels);
  Mat wmat(int layer, bool lo) const;                 // weights as stored: Mat[Kv+1, Nv]
  static Mat amat(const Act& a, bool lo, long long row0, long long rows, long long cols);
  int add_segments(GemmProblem& p, int g, Mat a_hi, Mat a_lo, Mat b_hi, Mat b_lo, bool a_exact);

  int E = 0, iHead = 0, iD1 = 0, iD2 = 0, iOut = 0;
  int B = 0, D = 0, L = 0;
  bool split = false;               // fp32 mode (bf16x3)
  int num_sms = 148;
  float inv_bg = 1.f;

  // flat parameter-shaped buffers
  float *p_ = nullptr, *g1_ = nullptr, *g2_ = nullptr, *m1_ = nullptr, *v1_ = nullptr, *m2_ = nullptr, *v2_ = nullptr;
  bf16 *w_hi_ = nullptr, *w_lo_ = nullptr;
  bool shadow_dirty_ = true;

  // activations
  struct InBuf {                    // one input batch
    Act X;                          // [3B, D+1] bf16 planes, rows [rot; lock; key] (array feeds; MAGIC_BITS_A=0: every path)
    uint32_t* xb3 = nullptr;        // [3B, xb_ld] the same rows packed 1 bit per pixel, bias column set: the first layer's A operand
    uint32_t* xbits = nullptr;      // = xb3 + B * xb_ld: the packed lock plane, the BCE epilogue's target
    bool planes_valid = false;      // X holds this batch too (false when a loader only wrote the packed rows)
    bool kmask_valid = false;       // kmask / k-block lists / work order describe this batch (built while skip_zero_stages is on)
    uint32_t* kmask = nullptr;      // zero-stage maps of xb3: [fwd | wgrad group 0 | wgrad group 1] (GemmArgs.kmask)
    float* labels = nullptr;        // [B] generated overlap labels
    int* rot = nullptr;             // [B, 2] Q14 (cos, sin) of the generator's lock rotation (tests)
    cudaEvent_t ready = nullptr, consumed = nullptr;
    bool ready_pending = false, consumed_pending = false, has_labels = false, filled = false, has_rot = false;
    // feed_pairs_bits_host(activate = 0): the packed planes have been copied to bits_stage_ (on gen_stream_), the kernel
    // that rotates and lays them out is still to be launched -- by the next step, behind its forward pass (flush_expand)
    bool expand_pending = false;
    uint64_t exp_seed = 0, exp_step = 0;
    int exp_wpr = 0;
  };
  void flush_expand(InBuf& b, bool behind_gate);
  InBuf in_[2];                     // the generator fills in_[cur_ ^ 1] on gen_stream_ while the step reads in_[cur_]
  int cur_ = 0;
  int xb_ld_ = 0;
  bool bits_a_ = true;              // first-layer GEMMs read the packed rows (BitsA); MAGIC_BITS_A=0: the bf16 planes
  bool skip_zero_ = false;          // ... and skip operand stages without a set pixel (MAGIC_SKIP_ZERO=1 / set_option); off: all stages
  int km_fwd_tiles_ = 0, km_fwd_words_ = 0, km_fwd_kb_ = 0;     // map geometry: forward (rows of 128 packed rows x k-blocks of 64 pixels)
  int km_wg_tiles_ = 0, km_wg_words_ = 0;                       // weight gradient (tiles of 128 pixels x k-blocks of 64 batch rows)
  size_t km_wg_off_[2] = {0, 0}, km_total_ = 0;                 // word offsets of the two groups' maps; total words
  void build_kmasks(InBuf& b, cudaStream_t s, bool force = false);
  std::string launch_maybe_dense(GemmPlan& pl, int* launches, const char* tag);
  cudaStream_t gen_stream_ = nullptr;
  cudaEvent_t ev_gate_ = nullptr;   // recorded after the forward pass: earliest start of the next batch's generator
  bool gate_pending_ = false;
  int gen_gate_ = 1;                // 0 = no gate, 1 = after the forward pass, 2 = after the last dgrad (tuning knob MAGIC_GEN_GATE)
  int* not_binary_ = nullptr;       // device flag raised by the array feeds when a pixel is neither 0 nor 1
  std::string check_binary(const char* who);
  void wait_ready(InBuf& b);
  void mark_consumed(InBuf& b);
  std::vector<Act> H_;              // encoder hidden outputs [3B, enc_i + 1]
  float* heads_ = nullptr; long long ld_heads_ = 0;      // [3B, 2L] fp32
  float* z_ = nullptr;              // [3B, L]
  Act zdec_, D1_, D2_;              // decoder input / hidden [B, *]
  Act dA_;                          // [B, D]  sigma(a) - x
  Act dD2_, dD1_;                   // decoder pre-activation grads [B, *]
  float* dz_dec_ = nullptr; long long ld_dz_ = 0;        // [B, L] fp32
  Act dheads_;                      // [4B, 2L]
  std::vector<Act> dH_;             // encoder pre-activation grads [4B, enc_i]
  float* ws_ = nullptr; long long ws_ld_ = 0;            // split-K workspace
  float* statbuf_ = nullptr;        // [8 sums | 2L colsq | L colS]
  float *g_row_ = nullptr, *dist_ = nullptr, *losses_dev_ = nullptr;
  void* stage_x_ = nullptr; size_t stage_x_bytes_ = 0;                            // partial_fit_host staging
  float *stage_labels_ = nullptr, *stage_eps_ = nullptr;
  int32_t* gather_idx_ = nullptr;           // [B] dataset row of every pair of the batch being fed (feed_pairs_bits_gather)
  uint32_t* bits_stage_ = nullptr; size_t bits_stage_bytes_ = 0;                     // feed_pairs_bits_host staging
  const float* labels_ext_ = nullptr;
  bool have_labels_ext_ = false;

  // plans
  std::vector<GemmPlan> fwd_enc_;   // E hidden + heads
  GemmPlan pl_d1_, pl_d2_, pl_out_bce_, pl_out_sig_;
  GemmPlan pl_wg_enc0_fused_[2];            // first-layer wgrad + every small layer's wgrad as extra problems (world == 1)
  bool fuse_wg_ = false;
  std::vector<GemmProblem> small_wg_;       // the weight-gradient problems of every layer but the two 40000-wide ones ...
  std::vector<int> small_wg_layer_;         // ... and the layer each belongs to (build_dp_plans appends them to the scatter launch)
  bool dp_small_fused_ = false;             // data parallel: they ride in the first layer's scatter launch (no split-K, no push)
  // ... and its EPI_ADAM variant: the first layer's dual Adam runs in the wgrad launch's epilogue (gemm_fuse_adam); used
  // by applied single-GPU steps without zero-stage lists (MAGIC_FUSE_ADAM=0 turns it off)
  GemmPlan pl_wg_enc0_adam_[2];
  bool fuse_adam_plans_ok_ = false, fuse_adam_ok_ = false, fuse_adam_ = false;
  GemmPlan fwd_enc0_alt_, pl_out_bce_alt_, pl_wg_enc0_alt_;   // the plans that read the input batch, on in_[1]
  GemmPlan pl_wg_out_, pl_dg_out_, pl_wg_h2_, pl_dg_h2_, pl_wg_h1_, pl_dg_h1_;
  GemmPlan pl_wg_head_, pl_dg_head_;
  std::vector<GemmPlan> pl_wg_enc_, pl_dg_enc_;   // index i: wgrad of layer i; dgrad through layer i (i >= 1)
  // hidden-layer chains (chain_sm100.cuh; bf16 mode, 512-wide hidden layers): one launch instead of fwd_enc_[1 .. E-1] /
  // instead of the dgrads through the heads (when they are at most 512 deep) and hidden layers E-1 .. 1
  ChainPlan chain_fwd_, chain_dg_;
  bool chain_dg_has_head_ = false;
  std::string launch_chain_plan(ChainPlan& cp, int* launches, const char* tag);
  struct ZeroRange { float* ptr; size_t bytes; };
  std::vector<ZeroRange> zero_;     // gradient windows written with red.add (split-K) -> cleared every step

  // data parallel
  NcclApi* nccl_ = nullptr;
  void* comm_ = nullptr;
  cudaStream_t comm_stream_ = nullptr;
  // single GPU: the decoder's Adam update runs on aux_stream_ under the encoder's backward pass (HBM-bound next to
  // tensor-bound work); adam_ holds the step's arguments, early_adam_ says the decoder range has been launched
  cudaStream_t aux_stream_ = nullptr;
  cudaEvent_t ev_aux_fork_ = nullptr, ev_aux_join_ = nullptr;
  AdamArgs adam_;
  bool early_adam_ = false, early_adam_ok_ = true;
  // the step's five losses are final after the forward pass: read_losses() waits for ev_losses_ on its own stream, so a
  // caller that wants its cost every step does not drain the backward pass and the update with it
  cudaEvent_t ev_losses_ = nullptr;
  cudaStream_t read_stream_ = nullptr;
  float* h_losses_ = nullptr;               // pinned, 5 floats
  const float* losses_src_ = nullptr;       // device buffer the last step wrote its losses to

  long long early_begin_ = 0;               // first parameter of the early (side-stream) Adam range
  cudaEvent_t ev_bwd_[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ev_comm_ = nullptr;
  std::string allreduce(float* buf, size_t n, cudaStream_t s);
  std::string assemble(float* buf, long long n);   // sharded buffer -> complete on every rank (collective)
  static constexpr int kDpGranShift = 13, kDpGranule = 1 << kDpGranShift;   // ownership granule: 8192 elements
  // fused DP state
  bool dp_fused_ = false;           // wgrad epilogues scatter to the owners' staging, sharded Adam, shadow push
  bool grads_sharded_ = false;      // reduced g1/g2 exist on their owners only (gathered by get_grad)
  bool master_stale_ = false;       // fp32 p/m/v outside the owned slice are out of date (refreshed on demand)
  int slice_len_ = 0;               // elements of the flat parameter space owned per rank (multiple of 1024)
  long long alloc_len_ = 0;         // world * slice_len (>= n_total): size of the parameter-shaped buffers
  float* stage_ = nullptr;          // [2][world][slice_len]: partial g1 / g2 of the owned slice from every rank
  float* peer_stage_[kMaxRanks] = {};
  size_t xchg_off_ = 0;             // floats: the exchange area of peer_allreduce_small sits behind the two staging halves
  unsigned xchg_epoch_ = 0;
  bool xchg_ok_ = false;            // fused path set up and MAGIC_DP_XCHG != 0: small all-reduces go over peer memory, not NCCL
  bf16* peer_whi_[kMaxRanks] = {};
  bf16* peer_wlo_[kMaxRanks] = {};
  float* barrier_buf_ = nullptr;
  GemmPlan pl_wg_out_sc_, pl_wg_enc0_sc_[2];   // scatter variants of the two 40000-wide wgrads
  std::string build_dp_plans();
  std::string gather_master();
};

}  // namespace magic
