// Fused HBM-bound kernels of the estimator tail and the sampler (launchers).
// All pointers are device pointers; every launcher returns a cudaError_t value.
#pragma once
#include <stdint.h>
#include "likelihood.cuh"

namespace apvae {

// x fp32 [pixels, c] -> bf16 [pixels, cp] raw and leaky_relu(x) (cp = 16 k, zero padded).
// raw_lo / act_lo: optional low halves (v - bf16(v)) for the split-precision mode.
// f16: 16-bit format of the outputs (op16.h), here and in every launcher below that writes operand tensors.
int launch_pack_image(const float* x, long long pixels, int c, int cp, void* out_raw, void* out_act,
                      float alpha, int f16, void* stream, void* raw_lo = nullptr, void* act_lo = nullptr);

// Reparameterised draw + log q(z|x) + log N(z;0,I)   (reference: zd.Normal sample/log_prob via
// tfsnippet/distributions/wrapper.py:112,127-129; truncated_normal.py:142-149; dmm_cifar10.py:449-450,601-603)
//   z[s,j] = mu[b,j] + exp(max(logstd[b,j], min_logstd)) * eps[s,j],  b = (s0 + s) % B
//   (all per-sample pointers are already offset to the chunk; s0 = global index of its first sample)
//   log_q[s], log_pn[s]; z written as fp32 [S,d] and as bf16 into the decoder input layout
//   idx = s*zstride + (j / zc)*cpad + (j % zc)  (raw and, if z_act != null, leaky_relu).
// mu == null: prior draw z = eps (log_q not written).
int launch_sample_z(const float* mu, const float* logstd, const float* eps, int S, int s0, int B, int d,
                    float min_logstd, int truncated, float* z_f32, void* z_raw, void* z_act,
                    long long zstride, int zc, int cpad, float alpha, float* log_q, float* log_pn,
                    int f16, void* stream, void* z_raw_lo = nullptr, void* z_act_lo = nullptr);
// The same draw pushed through `nf` planar normalizing-flow layers (reference: layers/flows/planar_nf.py:108-135,
// distributions/flow.py:118-150): z <- z + u_hat tanh(w.z + b), log q = log q0(z0) - sum_l log|1 + (1 - tanh^2) w.u_hat|.
// flow_w / flow_uhat: [nf][d]; flow_b / flow_wu: [nf].  d <= 512.
int launch_sample_z_flow(const float* mu, const float* logstd, const float* eps, int S, int s0, int B, int d,
                         float min_logstd, int truncated, const float* flow_w, const float* flow_uhat,
                         const float* flow_b, const float* flow_wu, int nf, float* z_f32, void* z_raw, void* z_act,
                         long long zstride, int zc, int cpad, float alpha, float* log_q, float* log_pn, int f16, void* stream,
                         void* z_raw_lo = nullptr, void* z_act_lo = nullptr);

// log p(x|z) summed over pixels from the fp32 heads (reference: distributions/discretized.py:198-259 with the
// dmm_cifar10.py:612-619 parameters; Bernoulli variant dmm_bernouli_fashion.py:619-626).  The hot path computes this
// inside the heads kernel (HeadsParams::lik_*); this pass remains for head layouts that kernel does not run.
//   head fp32 [S*HW, ld]: columns [0,C) = x_mean, [C,2C) = x_logstd (unless uniform_scale / bernoulli)
//   x fp32 [B, HW*C]; local sample s scores image b = (s0 + s) % B.  sq_err (may be null): test_mse's per-sample sum.
int launch_pixel_logprob(const float* head, int ld, const float* x, int S, int s0, int B, int HW, int C,
                         const LikelihoodSpec& spec, int uniform_scale, float beta, float* log_px, float* sq_err,
                         void* stream);
// The likelihood families on plain arrays: out[r] = sum_{j < group} log p(x[r*group + j] | mean, log_scale) (log_scale may
// be null for the Bernoulli).
int launch_elementwise_logprob(const float* x, const float* mean, const float* log_scale, long long rows, long long group,
                               const LikelihoodSpec& spec, float* out, void* stream);

// out[s] = dot(h[s, 0:n], w) + b     (D_psi's level_-1 dense, dmm_cifar10.py:561)
int launch_rowdot(const float* h, int ld, const float* w, const float* b, int n, long long S, float* out,
                  void* stream);

// Per-image importance-sampling reductions over K (reference: variational/evaluation.py:9-39,
// ops/misc.py:71-107, variational/objectives.py:9-42, dmm_cifar10.py:164-177,896-924).
// per_image is [kPerImageRows][B] fp32.
enum PerImageRow {
    ROW_LL = 0,        // log-mean-exp_k (log px + log pN - log q)
    ROW_LB = 1,        // mean_k of the same
    ROW_ADV_LL = 2,    // log-mean-exp_k (log px - energy - max(logZ,-20) - log q)
    ROW_ADV_LB = 3,
    ROW_RECON = 4,     // mean_k log px
    ROW_DG = 5,        // mean_k D(G(z))
    ROW_LZ_MAX = 6,    // max_k (-energy - log q)
    ROW_LZ_SUM = 7,    // sum_k exp(-energy - log q - max)
    ROW_REAL_E = 8,    // D_psi(x_origin): written by the caller, not by is_reduce
    ROW_MSE = 9,       // min_k sum_pixels (round(x_mean 128 + 127.5) - round(x 128 + 127.5))^2 (test_mse, dmm_cifar10.py:889-892)
    kPerImageRows = 10
};
// sq_err [K*B] (may be null -> ROW_MSE = 0): per-sample squared error of the rounded pixels
int launch_is_reduce(const float* log_px, const float* log_q, const float* log_pn, const float* d_g,
                     const float* sq_err, int K, int B, int d, float w_xi, float log_Z, float* per_image,
                     void* stream);

// Shard-local accumulation of the Evaluator metrics (replaces the host np.average of
// trainer/evaluator.py:249-254 and the host logsumexp of dmm_cifar10.py:1221-1222); see elementwise.cu.
constexpr int kAccumSlots = 20;
enum AccumSlot { ACC_WEIGHT = 9, ACC_BATCHES = 10, ACC_MSE = 11, ACC_PN = 12, ACC_PD = 13, ACC_KL = 14, ACC_LOGZ_OP = 15,
                 ACC_PRIOR_WEIGHT = 16 };
int launch_accumulate_metrics(const float* per_image, int B, int K, int rows, const float* prior4, double* accum, void* stream);
// test_pn_net branch scalars from Kp prior draws: out4 = {pn_energy, pd_energy, kl_adv_and_gaussian, log_Z_compute_op}
int launch_prior_reduce(const float* d_g, const float* log_pn, int Kp, int d, float w_xi, float log_Z, float* out4, void* stream);

// Philox4x32-10 + Box-Muller standard normals (timed runs; parity runs feed eps).
// truncated != 0: resample into [-2, 2] (tf.truncated_normal, truncated_normal.py:103).
int launch_fill_normal(float* out, long long n, unsigned long long seed, unsigned long long offset,
                       int truncated, void* stream);
int launch_fill_uniform(float* out, long long n, unsigned long long seed, unsigned long long offset,
                        void* stream);

// MALA (reference: dmm_cifar10.py:197-226,241-252).
//   propose : z' = z - alpha g + sqrt(2 alpha) eps  (fp32 + bf16 decoder-input copies)
//   accept  : ratio, u < ratio select, history rows
int launch_mala_propose(const float* z, const float* g, const float* eps, int n, int d, float alpha,
                        float* z_prime, void* stream);
int launch_mala_accept(float* z, float* e, float* g, float* pure, const float* z_prime, const float* e_prime,
                       const float* g_prime, const float* pure_prime, const float* u, int n, int d,
                       float alpha, float* hist_e, float* hist_pure, float* hist_ratio, void* stream);

// energy E = w_xi * D + 0.5 |z|^2 and grad = w_xi * gz + z
int launch_energy_combine(const float* d_g, const float* gz, const float* z, int n, int d, float w_xi,
                          float* energy, float* grad, void* stream);

// Backward through x_mean = tanh(pre): out[p, c] = g[p, c] * (1 - head[p*ld + c]^2) as bf16 [pixels, cp], zero padded
// (head holds the fp32 tanh OUTPUT; g is the fp32 gradient w.r.t. x_mean, [pixels, C]).
int launch_tanh_backward_pack(const float* g, const float* head, int ld, long long pixels, int C, int cp,
                              void* out_bf16, int f16, void* stream);

// The same backward step fused with the adjoint of the 1x1 mean head (x_mean = tanh(conv 1x1(h)), dmm_cifar10.py:489-492):
//   g_h[p, co] = sum_c g[p, c] (1 - head[p*ld + c]^2) W[co][c],   co < cg, as a 16-bit row of cp_out channels (padding zero).
// C <= 4 image channels, cg <= kHeadGradMaxCg trunk channels: a K = C "GEMM" that is pure HBM streaming (12 B in per pixel,
// cp_out * 2 B out), so it runs on the CUDA cores with the weights in the parameter block.
constexpr int kHeadGradMaxCg = 128, kHeadGradMaxC = 4;
struct HeadGradParams {
    const float* g;  const float* head;  int ld;      // fp32 [pixels, C] and [pixels, ld]
    long long pixels;
    int C, cg, cp_out, f16;
    void* out;                                         // 16-bit [pixels, cp_out]
    float w[kHeadGradMaxCg * kHeadGradMaxC];           // [cg][4]: the head kernel [1,1,cg,C] transposed, zero padded
};
int launch_tanh_backward_head(const HeadGradParams& p, void* stream);

// out[r] = sum_j v[r, j]^2 over `len` contiguous floats per row (ele_grad_norm, evaluation/dmm_cifar.py:958-959)
int launch_row_sqnorm(const float* v, long long len, int rows, float* out, void* stream);

// uint8 = round(clip(256 * x_mean / 2 + 127.5, 0, 255))   (dmm_cifar10.py:976-977,990,1067)
int launch_quantize_images(const float* head, int ld, long long pixels, int C, uint8_t* out, void* stream);

// v = max(v, lo)
int launch_clip_min(float* v, long long n, float lo, void* stream);
// v = max(v, alpha * v)
int launch_leaky_relu_f32(float* v, long long n, float alpha, void* stream);

// copy the first `cols` columns of fp32 [rows, ld] into dense [rows, cols]; optional tanh not applied.
int launch_gather_cols(const float* src, int ld, int col0, int cols, long long rows, float* dst, void* stream);


// G_theta's 1x1 output heads (dmm_cifar10.py:489-496: x_mean = tanh(conv 1x1), x_logstd = conv 1x1) on the CUDA cores:
// nout <= 8 outputs per pixel are an HBM-bound stream over the trunk output, not a tensor-core shape.  One thread per
// pixel; the pixel rows of a 128-pixel tile are staged through shared memory by coalesced 16-byte loads; the weights
// and biases travel in the kernel parameters (constant bank operands, fp32).  Writes the fp32 heads (tanh on the first
// tanh_cols columns) and the bf16 copy of those columns that D_psi reads (raw + leaky_relu, padding columns zero);
// split precision: input = in + in_lo, outputs as hi + lo pairs.
constexpr int kHeadsMaxW = 2048;          // cin * nout fp32 weights in the parameter block
struct HeadsParams {
    const void* in;  const void* in_lo;   // bf16 [pixels, cp]
    int cp, cin, nout, tanh_cols;
    long long pixels_per_image;
    float* out_f32;  int ld_f32;          // [pixels, ld_f32], columns < nout
    void* raw;  void* act;  void* raw_lo;  void* act_lo;   // bf16 [pixels, ldo], columns < tanh_cols, rest zero
    int ldo;
    float alpha;
    int f16;                              // operand format of `in` and of the 16-bit outputs (op16.h)
    // fused likelihood (set per call; lik_part == null: off): pixel p of local sample s scores x[(lik_s0 + s) % lik_B]
    const float* lik_x;                   // fp32 [lik_B, pixels_per_image * C]
    float* lik_part;                      // [samples * tiles_per_image][2]: per-tile sums of log p and of the squared pixel error
    int lik_s0, lik_B, lik_has_std;       // lik_has_std: columns [C, 2C) hold x_logstd
    int lik_uniform;                      // config.uniform_scale: log_scale = lik_beta whatever the columns hold (dmm_for_sampling.py:628)
    float lik_beta;
    LikelihoodSpec lik;
    float bias[8];
    float w[kHeadsMaxW];                  // [cin][nout]
};
int launch_pixel_heads(const HeadsParams& p, int n_images, void* stream);
int heads_tiles_per_image(long long pixels_per_image);
// log_px[s] / sq_err[s] = sum of the sample's tiles_per_image partial sums written by the fused heads kernel
int launch_sum_tiles(const float* part, int tpi, int n, float* log_px, float* sq_err, void* stream);


// A whole resnet_conv2d_block (resnet.py:272-364) with <= 3 input channels on the CUDA cores: D_psi/level_0 of the
// CIFAR models is 3 -> 3 -> 16 channels at 32 x 32 (0.57 MMAC per image).  As implicit GEMMs its two convs pad K from 3
// to 16 per tap and sit on the tensor pipe's per-instruction floor (3 and 18 TFLOP/s measured); here one CTA takes one
// image through shared memory:  r = lrelu(conv_k(lrelu(x)) + b0);  out = conv_k(r) + b1 + x W_s + b_s.
// r is rounded to bf16 (it is stored for the backward pass and that is what the unfused ops fed to conv_1).
// Two kernels: the CIFAR models' block (cout = 16, stride 1, <= 32 x 32: static shared memory, several CTAs per SM) and
// a wider one (cout = 16 or 32, stride 1 or 2 at the block's exit like resnet.py's resize_at_exit, <= 64 x 64, dynamic shared
// memory: arch B's D_psi/level_0 on CIFAR and CelebA shapes), which also takes split-precision tensors (hi + lo pairs).
constexpr int kSmallBlockMaxCin = 3, kSmallBlockMaxCout = 32, kSmallBlockMaxK = 3, kSmallBlockMaxHW = 32, kSmallBlockWideMaxHW = 64;
struct SmallBlockParams {
    const void* in_raw;                   // bf16 [N, H, W, cp_in]
    void* r_act;                          // bf16 [N, H, W, cp_in]: lrelu(conv_0 + b0), padding channels zero
    void* out_raw;  void* out_act;        // bf16 [N, H, W, ldo]: padding channels zero (either may be null)
    int H, W, cp_in, cin, cout, ldo, k;
    int stride;                           // of conv_1 and the shortcut (1 or 2); outputs are [N, ceil(H/stride), ceil(W/stride), ldo]
    float alpha;
    int f16;                              // operand format (op16.h)
    // split precision (wide kernel only): low halves of the input and of every 16-bit output; null = a plain block
    const void* in_raw_lo;  void* r_act_lo;  void* out_raw_lo;  void* out_act_lo;
    float b0[kSmallBlockMaxCin];
    float b1[kSmallBlockMaxCout];         // conv_1 bias + shortcut bias
    float w0[kSmallBlockMaxK * kSmallBlockMaxK * kSmallBlockMaxCin * kSmallBlockMaxCin];      // [k][k][cin][cin]
    float w1[kSmallBlockMaxK * kSmallBlockMaxK * kSmallBlockMaxCin * kSmallBlockMaxCout];     // [k][k][cin][cout]
    float ws[kSmallBlockMaxCin * kSmallBlockMaxCout];                                         // [cin][cout] 1x1 shortcut
};
int launch_small_block(const SmallBlockParams& p, int n, void* stream);

// The adjoint of that block (what tf.gradients(energy, [z]) walks through D_psi/level_0 in get_sgld_proposal,
// dmm_cifar10.py:241-247; model_grad.cu states the algebra): from g_out = dE/d(out) to g_x = dE/dx in one kernel,
//   g_u = conv_1^T(g_out) * lrelu'(r)          (stride-s transpose; kept in shared memory, fp32)
//   g_x = conv_0^T(g_u) * lrelu'(x) + shortcut^T(g_out)   (the 1x1 stride-s transpose lands on pixels (s y, s x))
// with the SAME packed kernels as the forward block.  Replaces 3 (stride 1) or 6 (stride 2: four sub-pixel phases)
// tensor-core ops whose K or N is 3 padded to 16.
struct SmallBlockGradParams {
    const void* g_out;                    // 16-bit [N, H/stride, W/stride, ldg]
    const void* r_act;                    // 16-bit [N, H, W, cp_in]: sign = lrelu'(conv_0 output)
    const void* x_in;                     // 16-bit [N, H, W, cp_in]: the block input (raw or activated: same signs)
    float* gx_f32;  int ld_f32;           // fp32 [N*H*W, ld_f32], cin columns (may be null)
    void* gx_raw;   int ldo;              // 16-bit [N, H, W, ldo], channels >= cin zero (may be null)
    int H, W, cp_in, cin, cout, ldg, stride;
    float alpha;
    int f16;
    float w0[kSmallBlockMaxK * kSmallBlockMaxK * kSmallBlockMaxCin * kSmallBlockMaxCin];      // as SmallBlockParams
    float w1[kSmallBlockMaxK * kSmallBlockMaxK * kSmallBlockMaxCin * kSmallBlockMaxCout];
    float ws[kSmallBlockMaxCin * kSmallBlockMaxCout];
};
int launch_small_block_grad(const SmallBlockGradParams& p, int n, void* stream);


// Stride-1 conv2d / conv2d_transpose with <= 4 input AND <= 4 output channels on the CUDA cores: the first conv of a
// block on the raw image channels (3 -> 3, 1 -> 1) and its adjoint in the sampler's backward pass.  As implicit GEMMs
// these pad K and N from 3 to 16 and sit on the tensor pipe's per-instruction floor (1..4 TFLOP/s measured); here a
// 32 x 8 pixel tile with its halo goes through shared memory and the weights are constant-bank operands.
// Epilogue = the GEMM kernels': v = acc + bias; v *= lrelu'(mask); v += res (same / even pixels); raw, act, fp32 outputs.
constexpr int kTinyMaxC = 4, kTinyMaxK = 5;
struct TinyConvParams {
    const void* in;                       // bf16 [N, H, W, cp_in]
    int H, W, cp_in, cin, cout, k;
    int ntaps;
    int8_t dy[kTinyMaxK * kTinyMaxK], dx[kTinyMaxK * kTinyMaxK];          // source shift of each tap
    float w[kTinyMaxK * kTinyMaxK * kTinyMaxC * kTinyMaxC];              // [tap][ci][co] (cin x cout used, packed to kTinyMaxC)
    float bias[kTinyMaxC];
    void* out_raw;  void* out_act;  int ldo;                             // bf16 [N, H, W, ldo], channels >= cout zero
    float* out_f32;  int ld_f32;                                          // fp32 [N*H*W, ld_f32], cout columns
    const void* res;  int res_mode, ld_res;                              // bf16, RES_SAME: [N,H,W,ld_res]; RES_EVEN: [N,H/2,W/2,ld_res]
    const void* mask;  int ld_mask;                                      // bf16 [N, H, W, ld_mask]
    float alpha;
    int f16;                                                             // operand format (op16.h)
    // split precision (apvae_config.precise): the input and the residual are hi + lo pairs (summed in fp32 on load), every
    // 16-bit output leaves as a pair (v_hi = round16(v), v_lo = round16(v - v_hi)); null = a plain op.  The weights are fp32
    // here anyway (the sum of the hi and lo kernels when both exist).
    const void* in_lo;  const void* res_lo;  void* out_raw_lo;  void* out_act_lo;
};
int launch_tiny_conv(const TinyConvParams& p, int n, void* stream);

}  // namespace apvae
