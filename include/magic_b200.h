/* magic_b200.h -- C ABI of the B200-native training engine for the "magic" asymmetric autoencoder.
 *
 * The reference (ag8/magic) has no FFI/plugin layer: its boundary is the Python surface of an experiment
 * directory (constants.py / vae.py / main.py).  This header is the C-ABI that a binding for that surface
 * calls; every entry point names the reference interface it replaces (paths relative to the reference
 * root).  magic_b200/tango.py is the ctypes binding that re-creates `TangoEncoder` on top of it; see
 * INTEGRATION.md for the stub a maintainer adds to <exp>/vae.py.
 *
 * Conventions: every function returns 0 on success, non-zero on error (message: magic_last_error());
 * no exception crosses the boundary.  Pointers named d_* are DEVICE pointers, h_* are HOST pointers.
 * The caller owns every buffer it passes in; the handle owns parameters, optimiser state, activations
 * and workspaces.  Calls enqueue on the cudaStream_t given at create time and do not synchronise unless
 * the name ends in _host or the comment says so.  A handle is not re-entrant: one handle per GPU/rank,
 * driven by one thread (the reference drives its session from one Python thread, 8c/main.py:38-84).
 * The library never falls back to a CPU path: without a CUDA device magic_create fails.
 */
#ifndef MAGIC_B200_H
#define MAGIC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MAGIC_MAX_ENC_LAYERS 8

enum { MAGIC_ACT_TANH = 1, MAGIC_ACT_ELU = 2 };                 /* 8c/vae.py:56, 11a/vae.py:58 */
enum { MAGIC_METRIC_COSINE = 0, MAGIC_METRIC_SQDIFF = 1 };      /* 8c/constants.py:5-7,60 */
enum { MAGIC_PRECISION_BF16 = 0, MAGIC_PRECISION_FP32 = 1 };    /* fp32 = bf16x3 split operands, fp32 accumulate */
enum { MAGIC_FWD_DISTANCE = 0, MAGIC_FWD_ZMEAN = 1, MAGIC_FWD_RECON = 2 };
enum { MAGIC_SLOT_M = 0, MAGIC_SLOT_V = 1 };

/* Hyper-parameters the reference hard-codes in TangoEncoder.__init__ (8c/vae.py:30-66) and FLAGS
 * (8c/constants.py:32,52,60). */
typedef struct magic_config {
  int32_t D;                          /* pixels per image: IMAGE_SIZE^2 = 40000 (8c/vae.py:81); multiple of 8 */
  int32_t n_enc;                      /* number of encoder hidden layers (8c/vae.py:30) */
  int32_t enc[MAGIC_MAX_ENC_LAYERS];  /* their widths */
  int32_t dec[2];                     /* decoder hidden widths (8c/vae.py:36) */
  int32_t L;                          /* latent dimensions (8c/vae.py:49) */
  int32_t act;                        /* MAGIC_ACT_* */
  int32_t metric;                     /* MAGIC_METRIC_* */
  int32_t reciprocal;                 /* distance = 1/distance (9a/vae.py:301,303) */
  float def_weight;                   /* 10 (8c/vae.py:287) or 100 (11a/vae.py:293) */
  float lr1, lr2;                     /* 8c/vae.py:63 */
  float beta1, beta2, eps;            /* tf.train.AdamOptimizer defaults 0.9, 0.999, 1e-8 */
  int32_t precision;                  /* MAGIC_PRECISION_* */
  int32_t batch_local;                /* pairs per step on this rank (static, like 8c/vae.py:367) */
  int32_t batch_global;               /* pairs per step over all ranks: every batch mean divides by this */
  int32_t rank, world;                /* data-parallel position; world > 1 needs magic_comm_init */
  uint64_t seed;                      /* seed of the internal counter-based RNG (noise, generator) */
} magic_config;

typedef struct magic_handle magic_handle;

typedef struct magic_stats {
  int64_t steps;                 /* optimiser steps taken */
  int64_t launches_last_step;    /* kernels of this library launched by the last magic_step */
  double gemm_flops_per_step;    /* issued tensor-core FLOPs per step (incl. padding and split terms) */
  double model_flops_per_pair;   /* algorithmic FLOPs per pair, SURVEY 8.4: 2*(11 enc + 3 dec - 4 D H) */
  int64_t n_params1, n_params2;  /* parameters trained by optimiser 1 / 2 (unpadded) */
  int64_t hbm_bytes;             /* device memory owned by the handle */
} magic_stats;

const char* magic_last_error(void);
int magic_version(void);

/* TangoEncoder.__init__ (8c/vae.py:22-326): builds parameters (Xavier-uniform from cfg.seed, zero biases),
 * both optimisers' slots, activations, TMA descriptors and GEMM plans.  `stream` is a cudaStream_t (NULL =
 * legacy default stream). */
int magic_create(const magic_config* cfg, void* stream, magic_handle** out);
int magic_destroy(magic_handle* h);
int magic_sync(magic_handle* h);
int magic_get_stats(magic_handle* h, magic_stats* out);

/* Variables in the reference's creation order (8c/vae.py:89-147) minus the two dead decoder_out_log_sigma_*
 * tensors: weights [in,out] row-major used as y = x.W + b, biases [out].  shape[1] = 0 for a bias vector. */
int magic_param_count(magic_handle* h, int32_t* n);
int magic_param_info(magic_handle* h, int32_t i, const char** name, int64_t shape[2], int32_t* optimisers);
int magic_set_param(magic_handle* h, int32_t i, const float* h_src);   /* synchronises */
int magic_get_param(magic_handle* h, int32_t i, float* h_dst);         /* synchronises */
/* gradient of the last step: which = 1 -> d cost (optimiser 1), 2 -> d training_loss (optimiser 2) */
int magic_get_grad(magic_handle* h, int32_t which, int32_t i, float* h_dst);
/* optimiser slots and step counter (checkpointing; the reference has none from 7c on) */
int magic_get_slot(magic_handle* h, int32_t which, int32_t slot, int32_t i, float* h_dst);
int magic_set_slot(magic_handle* h, int32_t which, int32_t slot, int32_t i, const float* h_src);
int magic_get_step(magic_handle* h, int64_t* t);
int magic_set_step(magic_handle* h, int64_t t);

/* Input of one step.  The reference feeds X[B,120000] fp32, channel-interleaved (ch0 lock, ch1 rotated lock,
 * ch2 key; 8c/vae.py:166-179, 8c/overlap_input.py:195).  Each loader converts into the handle's input
 * buffer: packed rows [rot; lock; key], 1 bit per pixel (pixels are {0,1}: exact), which the first layer's GEMMs read
 * directly; the array loaders go through planar bf16 on the way. */
int magic_load_interleaved(magic_handle* h, const float* d_x);                       /* [B, 3*D] */
/* Same layout as uint8: any non-zero byte is a set pixel.  The reference's feed_dict converts whatever array it is
 * given to the placeholder dtype, so a uint8/bool X is a legal argument of the same partial_fit (8c/vae.py:395); it
 * moves a quarter of the bytes over PCIe. */
int magic_load_interleaved_u8(magic_handle* h, const uint8_t* d_x);                  /* [B, 3*D] */
int magic_load_planar_u8(magic_handle* h, const uint8_t* d_lock, const uint8_t* d_rot, const uint8_t* d_key);
/* Row N2, the dataset path: lock and key planes [B, D] uint8 as decoded from the {N}_L.png / {N}_K.png files
 * (8c/overlap_input.py:21-73,155-170); the rotated-lock plane of 8c/overlap_input.py:181-182 (random_rotation,
 * 8c/utils.py:329-340: angle ~ U[0, 2 pi), nearest neighbour, zero fill) is produced on the device with the
 * generator's Q14 integer rotation for (seed, step, rank); magic_export_rotation returns the (cos, sin) used. */
int magic_load_pairs_u8(magic_handle* h, const uint8_t* d_lock, const uint8_t* d_key, uint64_t seed, uint64_t step);
/* The same pairs bit-packed: [B, words_per_row] 32-bit words per plane, pixel p = bit (p & 31) of word p >> 5
 * (numpy.packbits(..., bitorder='little') viewed as little-endian uint32): 10 KB per pair instead of the 480 KB of
 * the float32 array the queue hands to feed_dict (8c/main.py:44, 8c/vae.py:395).  Pixels are {0,1} after
 * normalisation (8c/overlap_input.py:185-195), so nothing is lost. */
int magic_load_pairs_bits(magic_handle* h, const uint32_t* d_lock_bits, const uint32_t* d_key_bits, int32_t words_per_row,
                          uint64_t seed, uint64_t step);
/* Host-buffer form for the training loop (the QueueRunner's role, 8c/overlap_input.py:76-124, 8c/main.py:40-45):
 * copies the packed planes and the labels [B] (or NULL) from HOST memory and unpacks them on the engine's input
 * stream, asynchronously -- the batch of step n+1 crosses PCIe while step n runs.  activate != 0 makes the batch the
 * loaded one (like every other loader); activate == 0 leaves it staged in the second input buffer until
 * magic_swap_batch.  The host buffers must stay untouched until the next call that synchronises (magic_sync, a
 * *_host call) or until a step that consumed the batch has been waited for; pinned memory keeps the copy
 * asynchronous.  With labels given, magic_step(d_labels = NULL) uses them. */
int magic_feed_pairs_bits_host(magic_handle* h, const uint32_t* h_lock_bits, const uint32_t* h_key_bits, const float* h_labels,
                               int32_t words_per_row, uint64_t seed, uint64_t step, int32_t activate);
/* The same with the shuffle-batch gather (tf.train.shuffle_batch, 8c/overlap_input.py:297-340) done by the device: the
 * caller keeps the WHOLE packed data set in pinned host memory -- h_lock_all / h_key_all: [n_rows, words_per_row], row i =
 * pair i, words_per_row a multiple of 4 -- and names the rows of this batch, h_idx[B] (entries outside [0, n_rows) give
 * empty images).  A thin kernel on the input stream reads those rows over PCIe straight into the staging area; per step
 * the host touches 16 KB of indices and labels instead of gathering 10 KB per pair into a pinned slot.  Everything else
 * (asynchrony, activate, labels, lifetime of the index / label buffers) as magic_feed_pairs_bits_host; the data set
 * arrays must stay pinned and unchanged while batches are being fed from them. */
int magic_feed_pairs_bits_gather(magic_handle* h, const uint32_t* h_lock_all, const uint32_t* h_key_all, int32_t words_per_row,
                                 int64_t n_rows, const int32_t* h_idx, const float* h_labels, uint64_t seed, uint64_t step,
                                 int32_t activate);
/* Replaces overlap_input.inputs() (8c/overlap_input.py:76-124): synthesises B shape pairs + rotated lock +
 * max-overlap labels on the device for (seed, step, rank).  Labels go to an internal buffer used when
 * magic_step is given d_labels = NULL. */
int magic_generate_batch(magic_handle* h, uint64_t seed, uint64_t step);
int magic_export_batch(magic_handle* h, float* d_x /* [B,3*D] or NULL */, float* d_labels /* [B] or NULL */);
/* Q14 fixed-point (cos, sin) of the rotation the generator applied to each lock, [B,2] int32: with them the rotated
 * plane and the label are reproducible bit for bit from the exported lock/key planes (tests/test_generator.py). */
int magic_export_rotation(magic_handle* h, int32_t* d_cs_sn);
/* The sampled latents z = mu + sqrt(exp(log_sigma_sq)) * eps of the last step / forward call (8c/vae.py:363-371), device
 * fp32 [3, B, L] in the stream order of d_eps3: lock, rot, key.  With zero head weights and biases z IS the N(0,1) draw
 * that replaces tf.random_normal (8c/vae.py:367): the statistical test of the device RNG reads it through this call. */
int magic_export_latents(magic_handle* h, float* d_z3);
/* The engine keeps two input batches in HBM (the loaded one and the one before it).  magic_swap_batch makes the
 * previous one the loaded batch again without touching its data -- a held batch re-fed to partial_fit /
 * get_predictions, as 8c/main.py:66-80 does with its test batch; bench.py alternates two resident batches with it. */
int magic_swap_batch(magic_handle* h);

/* The images are binary (8.5 % of the pixels set: README.md:19-27, overlap_micro.zip), and the first encoder layer --
 * 8c/vae.py:347, 3 of the 4 largest matrix products of the step with its weight gradient -- multiplies them.  The engine
 * keeps them packed (1 bit per pixel) and expands them inside the GEMM.  It can also skip every 128 x 64 operand stage
 * without a set pixel (such a stage adds exact zeros: results are bit-identical to processing it).  On the reference's
 * images the shapes sit anywhere on the canvas and 94-97 % of the stages hold a set pixel, so the per-batch maps cost more
 * than they save and the option is OFF by default; a data set whose shapes leave most of the canvas empty switches it on.
 *   magic_set_option(h, "skip_zero_stages", 1)  skip empty stages from the next loaded batch on (default 0)
 *   magic_set_option(h, "fuse_first_layer_adam", 1)  single GPU: run the first layer's two Adam updates (8c/vae.py:314-321,
 *                                                40 % of all parameters) in the epilogue of its weight-gradient GEMM instead
 *                                                of the Adam launch: same arithmetic, same result, that gradient is never
 *                                                written to HBM (magic_get_grad of encoder_h0_* after an APPLIED step then
 *                                                returns older contents; steps with apply = 0 always store it).  Default 0:
 *                                                measured no faster on B200; fails on handles that
 *                                                cannot do it (data parallel, planar first-layer operand)
 *   magic_input_density                          fraction of stages with a set pixel in the loaded batch:
 *                                                out2[0] forward GEMM, out2[1] weight-gradient GEMM (synchronises) */
int magic_set_option(magic_handle* h, const char* name, int32_t value);
int magic_input_density(magic_handle* h, double* out2);

/* TangoEncoder.partial_fit (8c/vae.py:379-405; 6-tuple variant 11a/vae.py:385-411) on the loaded batch:
 * forward, both backward passes at the pre-step weights, both Adam updates (W <- W - d1 - d2).
 *   d_labels : [B] overlap areas, or NULL to use the generated labels
 *   d_eps3   : [3,B,L] N(0,1) noise in stream order (lock, rot, key) (8c/vae.py:367), or NULL -> internal RNG
 *   d_losses5: [5] = cost, training_loss, rec_loss, lat_loss, def_loss (global-batch means)
 *   d_distance: [B] local distances (11a returns them) or NULL */
int magic_step(magic_handle* h, const float* d_labels, const float* d_eps3, float* d_losses5, float* d_distance);
/* Same, without the Adam update (gradients stay readable through magic_get_grad): parity tests. */
int magic_step_grads_only(magic_handle* h, const float* d_labels, const float* d_eps3, float* d_losses5,
                          float* d_distance);
/* The five losses of the last magic_step / magic_step_grads_only that was given a d_losses5, copied to host memory.
 * partial_fit returns its cost to Python on every call (8c/vae.py:395-405, printed by 8c/main.py:60-63); the losses are
 * final once the forward pass is, so on one rank this call waits for THAT point of the step only -- the backward pass and
 * the update keep running while the caller prepares its next step (data parallel: the losses are published after the
 * step's closing all-reduce, and the call waits for that).  Reading d_losses5 on the step's own stream still works and
 * waits for the whole step. */
int magic_read_losses(magic_handle* h, float h_losses5[5]);
/* Host-buffer form of partial_fit: H2D copies of X/labels/eps, the step, D2H of the results, then a stream
 * synchronise.  h_eps3 and h_distance may be NULL. */
int magic_partial_fit_host(magic_handle* h, const float* h_x, const float* h_labels, const float* h_eps3,
                           float h_losses5[5], float* h_distance);

int magic_partial_fit_host_u8(magic_handle* h, const uint8_t* h_x, const float* h_labels, const float* h_eps3,
                              float h_losses5[5], float* h_distance);

/* get_predictions / transform / reconstruct (8c/vae.py:407-434) on the loaded batch:
 *   MAGIC_FWD_DISTANCE -> d_out[B] (needs d_labels = NULL; sampled z: pass d_eps3 or NULL for RNG)
 *   MAGIC_FWD_ZMEAN    -> d_out[B,L]      MAGIC_FWD_RECON -> d_out[B,D] */
int magic_forward(magic_handle* h, int32_t what, const float* d_eps3, float* d_out);
/* generate (8c/vae.py:416-428): decoder-only forward from a fed z [B,L] -> [B,D] */
int magic_decode(magic_handle* h, const float* d_z, float* d_out);

/* cosine_distance(a, b) (8c/vae.py:438-452), the public free function of the reference's vae.py: both [B, L] arrays are
 * l2-normalised over axis 0 (the BATCH axis, eps 1e-12 under the max) and the products summed over the latent axis.
 * Host fp32 in, host fp32 [B] out; runs on the current device, synchronous.  No handle needed. */
int magic_cosine_distance(const float* h_a, const float* h_b, int32_t B, int32_t L, float* h_out);

/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  cosine_distance(a, b) as above, but on
 * DEVICE arrays a, b [B, L] -> out [B] (fp32), enqueued on `stream` (a cudaStream_t, NULL = default stream) without
 * synchronising, and without allocating anything: no cudaMalloc / cudaFree, no allocating library call.  The kernels'
 * scratch (2 L floats) is a static device array of 32 KiB that every GPU which loads the library holds until exit.
 * Limit: L <= MAGIC_COSINE_NOALLOC_MAX_L; a larger L returns an error (use magic_cosine_distance).  All calls on a device
 * share that scratch: each call clears what it uses first, calls that overlap on the GPU (different streams) race. */
#define MAGIC_COSINE_NOALLOC_MAX_L 4096
int magic_cosine_distance_noalloc(const float* d_a, const float* d_b, int32_t B, int32_t L, float* d_out, void* stream);

/* Data parallelism (new; the reference is single-device).  `nccl_unique_id` is the 128-byte ncclUniqueId
 * created on rank 0 and distributed by the host framework.  Afterwards magic_step all-reduces the cosine
 * batch statistics, the two gradients and the loss scalars over NCCL inside the step. */
int magic_comm_unique_id(void* h_id128);
int magic_comm_init(magic_handle* h, const void* h_nccl_unique_id128);
/* Fused gradient exchange over NVLink peer memory (after magic_comm_init, all ranks on one node): every rank exports a
 * 256-byte blob of CUDA IPC handles (its gradient staging buffer and shadow weights), the host framework all-gathers
 * the blobs in rank order, every rank imports.  From then on the wgrad GEMM epilogues store their tiles straight into
 * the owning rank's staging buffer (reduce-scatter fused into the GEMM), each rank runs both Adams on its 1/world
 * slice only and stores the new bf16 weights into every rank's shadow buffer (all-gather fused into Adam); NCCL is
 * left with the tiny cosine-statistic / loss-scalar all-reduces that double as barriers.  Without this call (or with
 * MAGIC_DP_MODE=allreduce) gradients are all-reduced by NCCL in three buckets behind the backward pass.
 * After it, magic_get/set_param, magic_get/set_slot and magic_get_grad are collective (every rank calls them). */
#define MAGIC_PEER_BLOB_BYTES 256
int magic_comm_export_peer(magic_handle* h, void* h_blob256);
int magic_comm_import_peers(magic_handle* h, const void* h_blobs /* world x 256 bytes, rank order */);

/* Per-launch CUDA-event timing of the step's kernels (bench.py's roofline and profiles/).  The report is text,
 * one line per kernel family: "<tag> <ms summed> <launches> <tensor flops summed>".  Reading it synchronises
 * and resets the accumulators. */
int magic_profile_enable(magic_handle* h, int32_t on);
int magic_profile_report(magic_handle* h, char* h_buf, int64_t cap);

/* Native self-test hook used by tests/: runs one GEMM of the given kind through the tcgen05 kernel.
 * kind 0: D = A[M,K].B[K,N]; 1: D = A[M,K].B[N,K]^T; 2: D = A[K,M]^T.B[K,N].  bf16 in, fp32 out [M,N]. */
int magic_test_gemm(int32_t kind, int32_t M, int32_t N, int32_t K, const void* d_a, int64_t lda, const void* d_b,
                    int64_t ldb, float* d_out, int64_t ldo, int32_t splitk, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MAGIC_B200_H */
