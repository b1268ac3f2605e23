/* apvae.h — C ABI of the B200-native evaluation / sampling hot path of
 * chenwenxiao/adversarial-prior-vae.
 *
 * The reference has no FFI: its hot path is a TensorFlow-1.x graph built by the
 * experiment scripts and executed with session.run.  Each entry point below
 * replaces one such graph fragment (file:line under /root/reference/code).
 * Conventions: plain pointers and sizes only; `*_dev` pointers are CUDA device
 * pointers owned by the caller; `stream` is a cudaStream_t (NULL = default);
 * calls are asynchronous on that stream unless stated; every function returns
 * an apvae_status (0 = OK) and apvae_last_error() gives the message.
 * Tensors are float32, NHWC, sample axis first ([K,B,...] flattened as
 * s = k*B + b), exactly as in the reference.
 */
#ifndef APVAE_H_
#define APVAE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    APVAE_OK = 0,
    APVAE_ERR_INVALID = 1,      /* bad argument / shape (reference: ValueError at graph build) */
    APVAE_ERR_CUDA = 2,         /* CUDA runtime / driver error */
    APVAE_ERR_NO_DEVICE = 3,    /* no usable sm_100 GPU: there is NO CPU fallback */
    APVAE_ERR_WEIGHTS = 4,      /* missing / mis-shaped variable */
    APVAE_ERR_UNSUPPORTED = 5,
    APVAE_ERR_STATE = 6         /* e.g. weights not loaded yet */
} apvae_status;

typedef enum { APVAE_ARCH_A = 0, APVAE_ARCH_B = 1, APVAE_ARCH_M = 2 } apvae_arch;

/* Mirrors the ExpConfig fields the path reads
 * (experiment/dmm/dmm_cifar10.py:27-91, dmm_for_sampling.py:28-92,
 * dmm_bernouli_fashion.py ExpConfig). */
/* Kernel-selection / tuning overrides (A-B comparisons, profiling probes, the layer tests' kernel variants).  All-zero =
 * the library's defaults; "switch" fields: 0 = default, -1 = off, +1 = on.  The library reads NO environment variables:
 * these travel in apvae_config.tuning (model handles) and with apvae_plan_create (layer plans). */
typedef struct apvae_tuning {
    int patch;              /* switch (default on): the A-stationary ("patch") tcgen05 kernel; off: every op on the tap-GEMM kernel */
    int patch_macc;         /* accumulators sharing a weight tile: 1, 2 or 4 (0: planner's choice) */
    int patch_cluster;      /* weight-stream multicast across a cluster of 2 / 4 CTAs (0: none) */
    int patch_max_n;        /* widest output (padded cout) the patch kernel takes (0: 256) */
    int patch_phases;       /* switch (default on): stride-2 transposed convs as sub-pixel phase ops */
    int patch_merge;        /* switch (default on): both column phases of a row phase merged into N */
    int patch_s2;           /* switch (default on): stride-2 convs through space-to-depth maps */
    int patch_fold;         /* switch (default on): 8-wide grids as image pairs */
    int patch_tpi;          /* vertical taps per weight item (0: up to 32 KB per item) */
    int patch_2cta;         /* switch (default: by K per output): CTA pairs, tcgen05 cta_group::2 */
    int patch_2cta_min_k;   /* K elements per output from which CTA pairs are used (0: 500) */
    int patch_a_cap, patch_a_stages, patch_b_stages;   /* ring depths (0: planner's choice) */
    int tap_2cta;           /* switch (default off): CTA pairs in the tap-GEMM kernel */
    int fast_epi;           /* switch (default on): TMA-in / TMA-out epilogue of the patch kernel */
    int fast_mask;          /* switch (default on): leaky-relu' mask through that epilogue (adjoint ops) */
    int tiny_conv;          /* switch (default on): <= 4 -> <= 4 channel convs on the CUDA cores (tinyconv.cu) */
    int small_block;        /* switch (default on): D_psi/level_0 as one fused CUDA-core block (block0.cu) */
    int heads_gemm;         /* switch (default off): G_theta's 1x1 heads as a tensor-core op instead of heads.cu */
    int patch_epi2;         /* switch (default: when a tile is epilogue-bound): two epilogue warpgroups in the patch kernel */
    int patch_split;        /* switch (default on): two-term split-precision ops (precise = 2) on the patch kernel */
    int debug;              /* patch kernel probe modes (skip loads / MMAs / epilogue / stores); 0 in any real run */
    int trace;              /* switch (default off): clock stamps of CTA 0, read back by apvae_plan_bench */
    int patch_virtual;      /* switch (default on): grids whose width is within 1/8 below a power of two (7, 14, 28: arch M) on the patch kernel
                             * as tiles of the next power of two (TMA zero-fills / clips the columns and rows past the image) */
} apvae_tuning;

typedef struct apvae_config {
    int arch;                       /* apvae_arch */
    int x_h, x_w, x_c;              /* config.x_shape */
    int z_dim;                      /* config.z_dim */
    int celeba;                     /* arch B: config.dataset == 'celebA' (stride-2 first/last blocks) */
    float pull_back_energy_weight;  /* config.pull_back_energy_weight */
    float min_logstd_of_q;          /* config.min_logstd_of_q */
    float epsilon;                  /* config.epsilon: min of the likelihood log-scale (-20) */
    int uniform_scale;              /* config.uniform_scale: log_scale = beta */
    int use_truncated;              /* config.use_truncated: TruncatedNormal q(z|x) */
    int bernoulli;                  /* arch M likelihood */
    int max_batch;                  /* capacity: images per q_net launch */
    int max_chunk;                  /* capacity: (image,sample) pairs per decoder/energy launch */
    int enable_grad;                /* build the input-gradient programs (MALA, grad norms) */
    float smallest_step;            /* config.smallest_step (default MALA alpha) */
    int len_train;                  /* config.len_train (log-Z offset) */
    int precise;                    /* 1: split-precision G_theta / D_psi (activations and kernels as hi + lo 16-bit pairs,
                                     * three tensor-core products per term, fp32 accumulate): fp32-class energies for the
                                     * sampler at 3x the MMA work.  2: activations as hi + lo pairs, kernels single (two
                                     * products per term): removes the per-element activation rounding, which differs between
                                     * a MALA state and its proposal; the kernels' rounding stays (a fixed perturbation of D
                                     * that cancels in E(z) - E(z')).  0: plain 16-bit operands (evaluation default). */
    int nf_layers;                  /* config.nf_layers when config.use_flow (dmm_bernouli_fashion.py:48,69,468-474): q(z|x) is
                                     * a FlowDistribution, the base Normal / TruncatedNormal draw pushed through that many planar
                                     * normalizing-flow layers (layers/flows/planar_nf.py); weights `posterior_flow/_i/{w,b,u}`
                                     * (`posterior_flow/{w,b,u}` for one layer).  0: no flow (every script's default). */
    int latent_gan;                 /* 1: also build G_omega, the latent generator whose output `gan_z` is the initial_z of the
                                     * MALA chains in plot_samples (dmm_cifar10.py:503-532,623-639,1061-1066; archs A and B):
                                     * weights `G_omega/...`, entry point apvae_latent_gan. */
    int operand_bf16;               /* 16-bit operand format of the tensor-core path (activations and packed kernels; fp32 accumulate):
                                     * 0 (default): fp16 — 11 significant bits, D(G(z)) within 2e-3 of the float64 oracle;
                                     * 1: bf16 — 8 significant bits, fp32's exponent range (for activations beyond +-65504);
                                     * same tcgen05 rate either way (kind::f16). */
    apvae_tuning tuning;            /* kernel-selection overrides, all-zero = defaults (see apvae_tuning) */
} apvae_config;

typedef struct apvae_handle apvae_handle;
typedef struct apvae_plan apvae_plan;

const char* apvae_version(void);
/* Message of the last failing call on `h` (or of the last failing handle-less call when h == NULL). */
const char* apvae_last_error(const apvae_handle* h);

/* ExpConfig defaults of the three script families. */
int apvae_default_config(int arch, int celeba, apvae_config* out);

/* TF variable schema (names as created under global_reuse scopes:
 * conv2d_.py:147-173, dense_.py:112-137, resnet.py:215-246), e.g.
 * "G_theta/level_3/residual/conv_0/kernel".  No GPU needed. */
int apvae_weight_count(const apvae_config* cfg, int* count);
int apvae_weight_info(const apvae_config* cfg, int index, char* name, int name_cap, int64_t shape[4], int* ndim);

/* One handle per GPU / rank; not thread-safe.  Fails with APVAE_ERR_NO_DEVICE without a GPU. */
int apvae_create(const apvae_config* cfg, int device, apvae_handle** out);
int apvae_destroy(apvae_handle* h);

/* Replaces tf.train.Saver restore / variable init: host float32 arrays in the
 * reference layouts (conv [kh,kw,Cin,Cout], deconv [kh,kw,Cout,Cin], dense [in,out]).
 * The caller keeps ownership; the library converts to bf16 and repacks. Synchronous. */
int apvae_load_weights(apvae_handle* h, int n, const char* const* names, const float* const* data,
                       const int64_t* const* shapes, const int* ndims);
/* 'p_net/xi' (pre-sigmoid, dmm_cifar10.py:604-607), 'beta' (:860), 'log_Z' (:458-463). */
int apvae_set_scalars(apvae_handle* h, float xi_raw, float beta, float log_Z);

/* q_net(x) -> (z_mean, clipped z_logstd) [B, z_dim]   (dmm_cifar10.py:418-452) */
int apvae_encode(apvae_handle* h, const float* x_dev, int B, float* mu_dev, float* logstd_dev, void* stream);
/* G_theta(z, return_std) -> x_mean [N,H,W,C] (tanh applied) and x_logstd (may be NULL)   (:466-500) */
int apvae_decode(apvae_handle* h, const float* z_dev, int N, float* x_mean_dev, float* x_logstd_dev, void* stream);
/* D_psi(x) for images (z_dev == NULL) or D_psi(G_theta(z)) (x_dev == NULL) -> [N]   (:535-562, :175) */
int apvae_energy(apvae_handle* h, const float* x_dev, const float* z_dev, int N, float* d_dev, void* stream);

/* Rows of `per_image_dev` ([APVAE_PER_IMAGE_ROWS][B] float32). */
enum {
    APVAE_ROW_LL = 0,       /* is_loglikelihood, plain N(0,I) prior (dmm_cifar10.py:896-898) */
    APVAE_ROW_LB = 1,       /* elbo (:899) */
    APVAE_ROW_ADV_LL = 2,   /* is_loglikelihood with the energy prior (:901-910) */
    APVAE_ROW_ADV_LB = 3,   /* :911 */
    APVAE_ROW_RECON = 4,    /* mean_k log p(x|z_k) (:886-888) */
    APVAE_ROW_DG = 5,       /* mean_k D(G(z_k)) (:914) */
    APVAE_ROW_LZ_MAX = 6,   /* max_k(-energy - log q)          } another_log_Z_compute_op (:922-924) */
    APVAE_ROW_LZ_SUM = 7,   /* sum_k exp(-energy - log q - max) } before the + log(len_train) */
    APVAE_ROW_REAL_E = 8,   /* D_psi(x_origin) (:913) */
    APVAE_ROW_MSE = 9,      /* min_k sum_pixels (round(x_mean 128 + 127.5) - round(x 128 + 127.5))^2: test_mse before the batch mean
                             * (:889-895; arch M: * 255 on the clipped Bernoulli mean, dmm_bernouli_fashion.py:893-899) */
    APVAE_PER_IMAGE_ROWS = 10
};
/* Rows of the optional `per_sample_dev` ([5][K*B]): log p(x|z), log q(z|x), log N(z), D(G(z)), squared pixel error. */
enum { APVAE_SROW_LOG_PX = 0, APVAE_SROW_LOG_Q = 1, APVAE_SROW_LOG_PN = 2, APVAE_SROW_DG = 3, APVAE_SROW_SQ_ERR = 4,
       APVAE_PER_SAMPLE_ROWS = 5 };

/* The test graph of one Evaluator mini-batch (dmm_cifar10.py:880-927, one session.run at
 * trainer/evaluator.py:189-191): q_net once per image, K reparameterised draws, G_theta and
 * D_psi once per (image,sample), likelihood, per-image log-mean-exp.
 * x: likelihood target; x_origin: q_net / real_energy input (same pointer for archs A, B).
 * eps_dev [K,B,z_dim] N(0,1) noise, or NULL to draw in-kernel (Philox, `seed`). */
int apvae_is_nll(apvae_handle* h, const float* x_dev, const float* x_origin_dev, int B, int K,
                 const float* eps_dev, uint64_t seed, float* per_image_dev, float* per_sample_dev, void* stream);

/* q_net(x, n_z=K)['z'] and its log-densities (dmm_cifar10.py:418-452): z = mu + exp(logstd) * eps, [K,B,z_dim];
 * log q(z|x) and log N(z;0,I), [K,B] each (either may be NULL).  eps_dev NULL: Philox draw with `seed`. */
int apvae_sample_z(apvae_handle* h, const float* x_origin_dev, int B, int K, const float* eps_dev, uint64_t seed,
                   float* z_dev, float* log_q_dev, float* log_pn_dev, void* stream);

/* Shard-local metric accumulation, replacing Evaluator.run's host np.average (trainer/evaluator.py:249-254)
 * and the host logsumexp of the log-Z loop (dmm_cifar10.py:1215-1227).  accum_dev: double[APVAE_ACCUM_SLOTS],
 * zero-initialised by the caller.  Slots 0-5, 8, 11: sums over images of the per-image rows (11 = APVAE_ROW_MSE);
 * slots 6, 7: running (max, sum exp) over BATCHES of another_log_Z_compute_op's batch value (before + log(len_train));
 * slot 9: images seen; slot 10: batches seen; slots 12-15: sums over batches of B x the prior branch's batch scalars
 * (`prior4_dev`, may be NULL: then slots 12-16 are untouched), slot 16: their weight.  One sum-allreduce of the additive
 * slots over ranks gives the dataset metrics; (6, 7) combine by log-sum-exp. */
enum { APVAE_ACCUM_WEIGHT = 9, APVAE_ACCUM_BATCHES = 10, APVAE_ACCUM_MSE = 11, APVAE_ACCUM_PN_ENERGY = 12,
       APVAE_ACCUM_PD_ENERGY = 13, APVAE_ACCUM_KL_ADV = 14, APVAE_ACCUM_LOG_Z_OP = 15, APVAE_ACCUM_PRIOR_WEIGHT = 16,
       APVAE_ACCUM_SLOTS = 20 };
int apvae_accumulate_metrics(apvae_handle* h, const float* per_image_dev, const float* prior4_dev, int B, int K,
                             double* accum_dev, void* stream);

/* The one collective of a sharded evaluation (BASELINE north_star: "per-shard NLL/energy statistics are combined with a
 * single NCCL allreduce over NVLink"): every rank passes the accumulator apvae_accumulate_metrics filled for its image
 * shard (the reference combines on the host: np.average, trainer/evaluator.py:249-254; logsumexp, dmm_cifar10.py:1221-1222).
 * ONE ncclAllGather of APVAE_ACCUM_SLOTS doubles per rank into gather_dev (caller-owned, double[nranks * APVAE_ACCUM_SLOTS]),
 * then a device-side merge into accum_dev on every rank: additive slots summed, the log-Z (max, sum-exp) pair (slots 6, 7)
 * rescaled to the common max.  nccl_comm: the caller's ncclComm_t; NCCL is bound at run time (dlopen of the libnccl.so.2
 * already in the process), APVAE_ERR_UNSUPPORTED if there is none.  nranks == 1: no-op.  Asynchronous on `stream`. */
int apvae_allreduce_metrics(void* nccl_comm, int nranks, double* accum_dev, double* gather_dev, void* stream);
/* The merge alone, for accumulators gathered by another transport (MPI, torch.distributed): gathered_dev
 * double[nranks][APVAE_ACCUM_SLOTS] -> accum_dev double[APVAE_ACCUM_SLOTS]. */
int apvae_merge_metrics(const double* gathered_dev, int nranks, double* accum_dev, void* stream);

/* test_pn_net branch (:883,915-927): Kp prior draws z ~ N(0, I) through G_theta and D_psi.
 * d_g_dev / log_pn_dev [Kp]: D(G(z)) and log N(z) per draw (either may be NULL);
 * prior4_dev [4] (may be NULL): the batch scalars { pn_energy (:918), pd_energy (:915-917), kl_adv_and_gaussian (:925-927),
 * log_Z_compute_op (:919-920) }, reduced on the device. */
enum { APVAE_PRIOR_PN_ENERGY = 0, APVAE_PRIOR_PD_ENERGY = 1, APVAE_PRIOR_KL_ADV = 2, APVAE_PRIOR_LOG_Z_OP = 3 };
int apvae_prior_stats(apvae_handle* h, int Kp, const float* eps_dev, uint64_t seed, float* d_g_dev,
                      float* log_pn_dev, float* prior4_dev, void* stream);

/* EnergyDistribution.sample / get_sgld_proposal (:179-252): `steps` MALA iterations on n chains.
 * z_io_dev [n, z_dim] in/out; eps_dev [steps,n,z_dim], u_dev [steps,n] or NULL (Philox);
 * hist_* [steps, n] or NULL; hist_z_dev [steps, n, z_dim] or NULL: history_z, the state BEFORE each step (:200,235 — the
 * reference's sampling driver restarts from history_z[-1], the state before the last accept, :1061).  Requires enable_grad. */
int apvae_mala(apvae_handle* h, float* z_io_dev, int n, int steps, float alpha, const float* eps_dev,
               const float* u_dev, uint64_t seed, float* hist_e_dev, float* hist_pure_e_dev,
               float* hist_ratio_dev, float* hist_z_dev, void* stream);
/* energy, grad_z energy, pure energy at z (get_sgld_proposal :241-247). Any output may be NULL. */
int apvae_energy_grad(apvae_handle* h, const float* z_dev, int n, float* energy_dev, float* grad_dev,
                      float* pure_dev, void* stream);
/* Out-of-distribution indicators of the evaluation scripts (evaluation/dmm_cifar.py:958-959):
 * D_psi(x) [N], ele_grad = d D_psi(x) / dx [N,H,W,C], ele_grad_norm = sum(ele_grad^2) [N].  Any output may be NULL.
 * Requires enable_grad. */
int apvae_image_energy_grad(apvae_handle* h, const float* x_dev, int N, float* d_dev, float* grad_dev,
                            float* grad_sqnorm_dev, void* stream);
/* x_plots (:976-977,990) + np.round: uint8 NHWC images from latents. */
/* gan_z = G_omega(z_omega) (p_omega_net's 'f_z' mean, dmm_cifar10.py:623-635): z_omega_dev fp32 [n, 128] ~ N(0, I), or
 * NULL for in-kernel Philox draws with `seed`; z0_dev fp32 [n, z_dim] receives the flattened [8, 8, z_channel] output,
 * the `initial_z` handed to apvae_mala (:1061-1066).  Needs config.latent_gan = 1. */
int apvae_latent_gan(apvae_handle* h, const float* z_omega_dev, int n, uint64_t seed, float* z0_dev, void* stream);

int apvae_sample_images(apvae_handle* h, const float* z_dev, int n, uint8_t* out_dev, void* stream);

/* ---- the estimator kernels on plain arrays (no handle; current device): what the reference's own unit tests exercise.
 * DiscretizedLogistic(mean, log_scale, bin_size, min_val, max_val, biased_edges, epsilon).log_prob(given, group_ndims)
 * (tfsnippet/distributions/discretized.py:26-58,198-259; tests/distributions/test_discretized.py:232-363): x / mean /
 * log_scale are fp32 device arrays of rows * group elements (already broadcast), out_dev[r] = sum of the `group`
 * consecutive element log-probs of row r (group = 1: elementwise).  has_min / has_max = 0: min_val / max_val is None. */
int apvae_discretized_logistic_log_prob(const float* x_dev, const float* mean_dev, const float* log_scale_dev, int64_t rows,
                                        int64_t group, float bin_size, int has_min, float min_val, int has_max, float max_val,
                                        int biased_edges, float epsilon, float* out_dev, void* stream);
/* The arch-M likelihood (dmm_bernouli_fashion.py:619-626): p = clip(mean, 1e-7, 1 - 1e-7), logits = log p - log1p(-p),
 * Bernoulli(logits).log_prob(x) = x logits - softplus(logits), summed over `group` consecutive elements. */
int apvae_bernoulli_log_prob(const float* x_dev, const float* mean_dev, int64_t rows, int64_t group, float* out_dev,
                             void* stream);
/* Per-image reductions over the K samples (variational/evaluation.py:9-39, objectives.py:9-42, ops/misc.py:71-107):
 * log_px / log_q / log_pn / d_g [K, B] (+ optional sq_err [K, B]) -> per_image_dev [APVAE_PER_IMAGE_ROWS][B] (rows
 * APVAE_ROW_LL .. APVAE_ROW_LZ_SUM and APVAE_ROW_MSE; APVAE_ROW_REAL_E is left untouched). */
int apvae_is_reduce(const float* log_px_dev, const float* log_q_dev, const float* log_pn_dev, const float* d_g_dev,
                    const float* sq_err_dev, int K, int B, int z_dim, float w_xi, float log_Z, float* per_image_dev,
                    void* stream);

/* NO COUNTERPART IN THE REFERENCE — added for callers that cannot call cudaMalloc
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid allocation).  The estimator tail on plain
 * arrays in ONE call that allocates nothing: element log-probs of DiscretizedLogistic(mean, log_scale, bin_size, ...)
 * summed over `group` consecutive elements per (sample, image) row s = k * B + b  (= apvae_discretized_logistic_log_prob),
 * then the per-image reductions over K (= apvae_is_reduce with sq_err NULL) into per_image_dev [APVAE_PER_IMAGE_ROWS][B].
 * The intermediate per-sample row log p(x|z) [K * B] lives in a fixed device array inside the library (256 KiB on every
 * GPU that loads it, held until exit), so K * B <= APVAE_NOALLOC_MAX_SAMPLES — larger calls are refused with
 * APVAE_ERR_INVALID.  No cudaMalloc / cudaFree / allocating library routine is called.  All calls on one GPU share that
 * array and it keeps the previous call's contents (each call overwrites what it reads; nothing needs zeroing): calls that
 * overlap on the GPU (different streams, unordered graph branches) race — serialise them.  x / mean / log_scale are
 * caller-owned fp32 device arrays of K * B * group elements; the current device is used. */
enum { APVAE_NOALLOC_MAX_SAMPLES = 65536 };
int apvae_is_ll_noalloc(const float* x_dev, const float* mean_dev, const float* log_scale_dev, int64_t group,
                        float bin_size, int has_min, float min_val, int has_max, float max_val, int biased_edges,
                        float epsilon, const float* log_q_dev, const float* log_pn_dev, const float* d_g_dev,
                        int K, int B, int z_dim, float w_xi, float log_Z, float* per_image_dev, void* stream);

/* Philox normals / uniforms the NULL-noise paths use (so tests can reproduce them). */
int apvae_fill_normal(float* out_dev, int64_t n, uint64_t seed, uint64_t offset, int truncated, void* stream);
int apvae_fill_uniform(float* out_dev, int64_t n, uint64_t seed, uint64_t offset, void* stream);

/* Bookkeeping for the benchmark: kernels launched so far, algorithmic MACs per unit. */
int apvae_launch_count(const apvae_handle* h, int64_t* launches);
int apvae_macs(const apvae_handle* h, double* q_per_image, double* g_per_sample, double* d_per_sample);

/* Per-op device timing for bench.py's roofline: net 0 = q_net, 1 = G_theta (+heads), 2 = D_psi, 3 = adjoint program of
 * D_psi, 4 = adjoint program of G_theta (both need enable_grad), 5 = G_omega (needs latent_gan).
 * apvae_op_time launches op `index` on the first `n` units of its (already populated) input
 * buffers 2 + `reps` times and returns the CUDA-event average of the last `reps` launches on
 * `stream` (synchronises that stream). */
int apvae_op_count(const apvae_handle* h, int net, int* count);
int apvae_op_time(apvae_handle* h, int net, int index, int n, int reps, char* name, int name_cap,
                  double* macs_per_unit, float* ms_per_launch, void* stream);

/* Device timing of the HBM-bound CUDA-core kernels of the estimator tail and the sampler (bench.py's "achieved HBM GB/s
 * against peak" table): kernel `which` (APVAE_AUX_*) runs 2 + reps times on the handle's own scratch buffers for `n`
 * units (samples / chains; B images of K samples for APVAE_AUX_IS_REDUCE with n = K * B) and the CUDA-event average of the
 * last `reps` launches is returned together with the kernel's ALGORITHMIC bytes per launch (compulsory reads + writes,
 * stated per kernel in DESIGN.md).  Buffer contents are unspecified afterwards.  The sampler's kernels need enable_grad. */
enum { APVAE_AUX_SAMPLE_Z = 0, APVAE_AUX_IS_REDUCE = 1, APVAE_AUX_SUM_TILES = 2, APVAE_AUX_PACK_IMAGE = 3,
       APVAE_AUX_ROWDOT = 4, APVAE_AUX_FILL_NORMAL = 5, APVAE_AUX_MALA_PROPOSE = 6, APVAE_AUX_MALA_ACCEPT = 7,
       APVAE_AUX_ENERGY_COMBINE = 8, APVAE_AUX_TANH_BACKWARD = 9, APVAE_AUX_QUANTIZE = 10, APVAE_AUX_COUNT = 11 };
int apvae_aux_time(apvae_handle* h, int which, int n, int B, int reps, char* name, int name_cap, double* bytes_per_launch,
                   float* ms_per_launch, void* stream);

/* ---- layer-level interface (tests; mirrors spt.layers.conv2d / deconv2d / dense) ---- */
typedef struct apvae_conv_term {
    int src, transpose, k, stride, cin, cout, in_h, in_w, in_cp;
    const float* kernel;   /* host */
    const float* bias;     /* host, may be NULL */
} apvae_conv_term;

typedef struct apvae_plan_info {
    int kc, npad, nt, n_tiles, cout, num_phases, phase_begin[5], total_steps;
    int bw, bh, bn, rows, wg, hg, tiles_x, tiles_y, ho, wo, os, nsrc, capacity_n;
    uint64_t src_dims[3][5];
    uint64_t src_strides[3][4];
    uint32_t src_box[3][5];
    double macs_per_image;
    /* A-stationary ("patch") plan, when the op is eligible (width 7..128 within 1/8 below a power of two) and not disabled */
    int patch, patch_macc, patch_cluster, patch_rows_tile, patch_groups, patch_items, patch_a_stages, patch_b_stages;
    int patch_prow[2];
    /* activation maps of the patch plan: (source, row phase); patch_s2d = 1: stride-2 conv through the space-to-depth view */
    int patch_s2d, patch_nmaps, patch_map_src[4], patch_map_ph[4], patch_map_prow[4];
    int patch_phase, patch_out_scale, patch_npad;   /* of the selected patch plan (apvae_plan_select_patch): sub-pixel phase 0..3, row phase 4..5, -1 none */
    int patch_fold;   /* 2: 8-pixel-wide images are processed in pairs, side by side in one 16-wide tile row (even capacity) */
    int patch_epi2;   /* 1: the selected patch plan asks for two epilogue warpgroups (epilogue-bound tiles) */
    int patch_wv;     /* tile-row width per image of the patch plan: the grid width, or the next power of two above it (28 -> 32, 14 -> 16, 7 -> 8) */
} apvae_plan_info;

typedef struct apvae_epilogue {
    void* out_raw_dev;  void* out_act_dev;  int ldo;  int bf16_zero_from;
    float* out_f32_dev; int ld_f32; int f32_cols;
    const void* res_dev; int res_mode; int ld_res;
    const void* mask_dev; int ld_mask;
    int tanh_cols;
    /* split precision (apvae_config.precise): low halves v - round16(v) of the 16-bit outputs and of the residual, same
     * layouts as their high halves; NULL = a plain op */
    void* out_raw_lo_dev;  void* out_act_lo_dev;  const void* res_lo_dev;
} apvae_epilogue;

/* Host-only planning (no GPU needed): step table, packed 16-bit weights, TMA view geometry.
 * operand_bf16: the 16-bit format the kernels are packed in and the sources / outputs hold (see apvae_config).
 * tuning: kernel-selection overrides (NULL = defaults), e.g. patch = -1 disables the A-stationary kernel, patch_macc /
 * patch_cluster force its accumulator count / cluster size; kept with the plan and applied by apvae_plan_run / _bench. */
int apvae_plan_create(const apvae_conv_term* terms, int nterms, int capacity_n, int operand_bf16, const apvae_tuning* tuning,
                      apvae_plan** out);
int apvae_plan_get_info(const apvae_plan* p, apvae_plan_info* info);
int apvae_plan_copy_steps(const apvae_plan* p, int32_t* out /* [total_steps][5]: src,dx,dy,ph,c0 */);
int apvae_plan_copy_weights(const apvae_plan* p, uint16_t* out /* bf16 [total_steps][npad][kc] */);
int apvae_plan_copy_bias(const apvae_plan* p, float* out /* [npad] */);
int apvae_plan_copy_patch_groups(const apvae_plan* p, int32_t* out /* [groups][6]: src,dx,py0,k,c0,prow */);
int apvae_plan_copy_patch_weights(const apvae_plan* p, uint16_t* out /* bf16 [items][npad][kc] */);
int apvae_plan_copy_patch_bias(const apvae_plan* p, float* out /* [patch_npad] */);
/* A stride-2 transposed conv is run as four sub-pixel phase plans (which = 0..3: phase (py, px) = (which >> 1, which & 1))
 * or two row-phase plans with both column phases merged into N (which = 4 + py); -1 = the op's own patch plan.  Selects
 * the plan the patch_* fields of apvae_plan_get_info and the apvae_plan_copy_patch_* calls describe (host-side tests). */
int apvae_plan_select_patch(apvae_plan* p, int which);
int apvae_plan_destroy(apvae_plan* p);
/* Runs the plan on `device` (synchronous): src_dev[i] are bf16 NHWC tensors [capacity_n,H,W,Cp]. */
int apvae_plan_run(const apvae_plan* p, int device, const void* const* src_dev, int n_valid,
                   const apvae_epilogue* epi);

/* Like apvae_plan_run, but launches the op 2 + reps times and returns the CUDA-event average of the
 * last `reps` launches (kernel tuning / profiling of one layer in isolation). */
int apvae_plan_bench(const apvae_plan* p, int device, const void* const* src_dev, int n_valid,
                     const apvae_epilogue* epi, int reps, float* ms_per_launch);

#ifdef __cplusplus
}
#endif
#endif /* APVAE_H_ */
