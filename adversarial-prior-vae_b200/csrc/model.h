// Network programs of the hot path: q_net, G_theta (+ likelihood heads) and
// D_psi of the three script families, lowered to sequences of tap-GEMM ops.
// Reference builders: experiment/dmm/dmm_cifar10.py:418-452,466-500,535-562;
// dmm_for_sampling.py:421-460,476-514,551-576; dmm_bernouli_fashion.py:432-478,492-521,546-573.
#pragma once
#include <map>
#include <memory>
#include <string>
#include <vector>
#include "../../include/apvae.h"
#include "runtime.h"

namespace apvae {

struct HostArr {
    std::vector<float> v;
    std::vector<int64_t> shape;
};

struct BlockSpec {
    const char* scope;
    int cout, k, stride;
};

struct WeightInfo {
    std::string name;
    std::vector<int64_t> shape;
};

// Architecture tables (shared by the schema query and the builder).
std::vector<BlockSpec> q_blocks(const apvae_config& c);
std::vector<BlockSpec> g_blocks(const apvae_config& c);
std::vector<BlockSpec> d_blocks(const apvae_config& c);
std::vector<BlockSpec> omega_blocks(const apvae_config& c);
int omega_base_h(const apvae_config& c);
int omega_base_w(const apvae_config& c);
std::vector<WeightInfo> weight_schema(const apvae_config& c);
bool validate_config(const apvae_config& c, std::string* err);

struct Tensor {
    void* ptr = nullptr;
    int n = 0, h = 0, w = 0, c = 0, cp = 0;
    size_t elems() const { return (size_t)n * h * w * cp; }
};
struct Pair {
    Tensor raw, act;
    Tensor raw_lo, act_lo;   // split-precision mode: v = hi + lo with both halves bf16 (empty otherwise)
    int h() const { return raw.ptr ? raw.h : act.h; }
    int w() const { return raw.ptr ? raw.w : act.w; }
    int c() const { return raw.ptr ? raw.c : act.c; }
};

// Saved forward state of one resnet block (for the input-gradient programs).
struct BlockTape {
    std::string scope;
    bool deconv = false, use_shortcut = false;
    int k = 3, stride = 1, cin = 0, cout = 0;
    Pair in;        // block input (raw + act)
    Tensor r_act;   // leaky_relu(conv_0 output)
    Pair out;
    bool fused = false;   // the forward block is ONE CUDA-core kernel (block0.cu): its adjoint can be one too
};

typedef std::vector<std::unique_ptr<DevOp>> Program;

class Model {
public:
    apvae_config cfg;
    DeviceInfo dev;
    mutable std::string err;
    std::map<std::string, HostArr> weights;
    float xi_raw = 0.f, beta = 0.f, log_Z = 1.f;
    bool built = false;
    int64_t launches = 0;
    double macs_q = 0, macs_g = 0, macs_d = 0;

    ~Model();
    bool build();   // after weights are loaded

    // forward pieces (all on `stream`)
    bool run_encoder(const float* x_dev, int B, void* stream);                  // -> q_mu, q_logstd
    // g_in -> d_in (+ head_f32 when want_f32).  `lik`: also score the samples against x inside the heads kernel
    // (log p(x|z) and the test_mse squared error, per sample) -> log_px_out / sq_err_out [n]
    struct LikArgs { const float* x; int s0, B; float* log_px_out; float* sq_err_out; };
    bool run_decoder_chunk(int n, void* stream, bool want_f32 = true, const LikArgs* lik = nullptr);
    bool run_critic_chunk(int n, float* d_out, void* stream);                   // d_in -> d_out[n]
    // energy E = w xi D(G(z)) + |z|^2 / 2, grad_z E and D(G(z)) for n <= max_chunk latents (fp32 [n, z_dim]);
    // any output may be null (get_sgld_proposal, dmm_cifar10.py:241-247)
    bool energy_grad_chunk(const float* z_dev, int n, float* energy, float* grad, float* pure, void* stream);
    // EnergyDistribution.sample's loop (dmm_cifar10.py:197-238)
    bool mala(float* z_io, int n, int steps, float alpha, const float* eps, const float* u, uint64_t seed,
              float* hist_e, float* hist_pure, float* hist_ratio, float* hist_z, void* stream);
    bool latent_gan_chunk(const float* z_omega_dev, int n, float* z0_dev, void* stream);   // n <= max_chunk
    bool critic_of_images(const float* x_dev, int N, float* d_out, void* stream);
    // D_psi(x), dD/dx [N,H,W,C] and its squared norm per image (evaluation/dmm_cifar.py:958-959); outputs may be null
    bool critic_grad_of_images(const float* x_dev, int N, float* d_out, float* grad_out, float* sqnorm_out, void* stream);
    bool ensure_sample_buffers(size_t samples);

    // geometry
    int H() const { return cfg.x_h; }
    int W() const { return cfg.x_w; }
    int C() const { return cfg.x_c; }
    int HW() const { return cfg.x_h * cfg.x_w; }
    int omega_in_dim() const { return cfg.arch == APVAE_ARCH_M ? cfg.z_dim : 128; }   // z_omega: dmm_cifar10.py:626-629 / dmm_bernouli_fashion.py:636-639
    int head_ld() const { return (cfg.bernoulli || cfg.arch == APVAE_ARCH_M) ? cfg.x_c : 2 * cfg.x_c; }
    int g_base_h() const { return cfg.arch == APVAE_ARCH_B ? 8 : cfg.x_h / 4; }
    int g_base_w() const { return cfg.arch == APVAE_ARCH_B ? 8 : cfg.x_w / 4; }
    int zc() const { return 16 * cfg.z_dim / cfg.x_h / cfg.x_w; }
    // decoder-input layout of z (see kernels.h launch_sample_z)
    long long z_stride() const;
    int z_group() const;
    int z_cpad() const;
    float w_xi() const;
    int f16() const { return cfg.operand_bf16 ? 0 : 1; }   // 16-bit operand format of every tensor of this handle (op16.h)

    // device state
    Program q_prog, g_prog, d_prog;
    Program om_prog;           // G_omega (latent_gan)
    Tensor om_in;              // bf16 [max_chunk, 1, 1, 128]
    float* om_out = nullptr;   // fp32 [max_chunk, z_dim]
    float* om_noise = nullptr; // fp32 [max_chunk, 128] in-kernel z_omega draws
    double macs_omega = 0;
    Program gd_prog, gg_prog;  // input-gradient programs: d D / d x_mean, then d x_mean / d z (enable_grad)
    double macs_grad = 0;
    Tensor d_seed;             // dD/d(flatten) = W64 w_last, one copy per sample (constant)
    float* gx_f32 = nullptr;   // [max_chunk*HW, C] gradient w.r.t. x_mean
    Tensor g_pre;              // gradient w.r.t. the x_mean head's pre-activation, bf16 [max_chunk,H,W,16]
    bool head_grad_fused = false;   // tanh' and the mean head's adjoint as ONE CUDA-core kernel (kernels.h: HeadGradParams)
    HeadGradParams head_grad;
    float* gz_f32 = nullptr;   // [max_chunk, z_dim] d D(G(z)) / d z
    // MALA state (grow-only): current / proposed z, energy, gradient, pure energy, noise
    size_t mala_cap = 0;
    float *mz = nullptr, *mzp = nullptr, *mg = nullptr, *mgp = nullptr, *me = nullptr, *mep = nullptr, *mp = nullptr,
          *mpp = nullptr, *meps = nullptr, *mu = nullptr;
    Pair q_in;                 // [max_batch, H, W, 16]
    float* q_mu = nullptr;     // [max_batch, z_dim]
    float* q_logstd = nullptr;
    // planar posterior flow (nf_layers > 0): w [L][d], u_hat [L][d], b [L], w . u_hat [L]
    float* flow_w = nullptr; float* flow_uhat = nullptr; float* flow_b = nullptr; float* flow_wu = nullptr;
    Pair g_in;                 // decoder input
    float* head_f32 = nullptr; // [max_chunk*HW, head_ld]
    DevOp* heads_op = nullptr; // the CUDA-core heads op at the end of g_prog (null: the heads run as a GEMM op)
    float* lik_part = nullptr; // [max_chunk * tiles_per_image][2] per-tile likelihood sums of the fused heads kernel
    float* prior4 = nullptr;   // the prior branch's four batch scalars (apvae_prior_stats)
    Pair d_in;                 // [max_chunk, H, W, 16]
    float* d64 = nullptr;      // [max_chunk, 64]
    float* d_last_w = nullptr; // level_-1 kernel [64], bias [1]
    float* d_last_b = nullptr;
    std::vector<BlockTape> g_tape, d_tape;
    Tensor g_head_in;          // raw trunk output feeding the heads
    Tensor d_flat;             // raw output of D's last block

    // per-call sample buffers (grow-only)
    size_t sample_cap = 0;
    float *eps_buf = nullptr, *log_px = nullptr, *log_q = nullptr, *log_pn = nullptr, *d_g = nullptr, *sq_err = nullptr;
    float* z_tmp = nullptr;    // [max_chunk, z_dim] scratch

    std::vector<void*> allocs;
    bool fail(const std::string& m) const { err = m; return false; }

    Tensor alloc_act(int n, int h, int w, int c);
    float* alloc_f32(size_t n);
    const HostArr* W_(const std::string& name, std::initializer_list<int64_t> shape);
    // Terms of one (source, kernel) product.  bf16 mode: one term.  Split-precision mode (`split`): activations
    // and kernels are hi + lo bf16 pairs and the product is A_hi W_hi + A_lo W_hi + A_hi W_lo (the lo*lo term,
    // 2^-16 relative, is dropped) accumulated in the same fp32 TMEM accumulator.
    bool push_terms(std::vector<TermSpec>& terms, std::vector<const void*>& srcs, bool split, const Tensor& hi,
                    const Tensor& lo, bool transpose, int k, int stride, int cin, int cout, const std::string& wname,
                    const HostArr* kernel, const HostArr* bias);
    bool split_mode = false;   // set while building G_theta / D_psi of a `precise` handle
    double split_terms() const { return !split_mode ? 1.0 : cfg.precise == 2 ? 2.0 : 3.0; }   // tensor-core terms per algorithmic product
    bool add_op(Program& prog, const std::string& name, const std::vector<TermSpec>& terms,
                const std::vector<const void*>& srcs, const Epilogue& epi, int cap, double* macs);
    bool conv_block(Program& prog, const std::string& root, const BlockSpec& b, const Pair& in, bool need_raw,
                    bool need_act, int cap, Pair* out, double* macs, std::vector<BlockTape>* tape);
    bool deconv_block(Program& prog, const std::string& root, const BlockSpec& b, const Pair& in, bool need_raw,
                      bool need_act, int cap, Pair* out, double* macs, std::vector<BlockTape>* tape);
    bool dense_op(Program& prog, const std::string& name, const Tensor& in_flat, int in_features, int out_features,
                  const Epilogue& epi, int cap, double* macs, const Tensor* in_flat_lo = nullptr);
    bool build_q();
    bool build_g();
    bool build_d();
    bool build_grad();
    bool build_omega();
    bool grad_conv_block(const BlockTape& t, const Tensor& g_out, float* out_f32, int f32_ld, Tensor* g_in_out);
    bool grad_deconv_block(const BlockTape& t, const Tensor& g_out, float* out_f32, int f32_ld, Tensor* g_in_out);
    bool run_program(Program& prog, int n, void* stream);
};

}  // namespace apvae
