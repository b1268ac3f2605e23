// extern "C" surface declared in include/apvae.h.
#include <cuda_runtime.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <mutex>
#include "../../include/apvae.h"
#include "kernels.h"
#include "model.h"
#include "tuning.h"

using namespace apvae;

struct apvae_handle {
    Model m;
};
struct apvae_plan {
    OpPlan plan;
    bool has_patch = false;
    PatchPlan patch;
    std::vector<PatchPlan> phases;   // stride-2 transposed conv: four sub-pixel phase plans (A-stationary kernel)
    int sel = -1;                     // which patch plan the info / copy calls describe (apvae_plan_select_patch)
    std::vector<PatchPlan> row_phases;   // ... or two plans, one per output row phase, both column phases merged into N
    int f16 = 1;                         // operand format (op16.h)
    apvae_tuning tuning = {};            // the options the plan was made with (also steer apvae_plan_run / _bench)
};

static std::string g_err;
static std::mutex g_err_mu;
static int set_global_err(const std::string& m, int code) {
    std::lock_guard<std::mutex> l(g_err_mu);
    g_err = m;
    return code;
}

namespace apvae { int report_global_error(const std::string& m, int code) { return set_global_err(m, code); } }   // noalloc.cu

static_assert((int)kPerImageRows == (int)APVAE_PER_IMAGE_ROWS && (int)ROW_REAL_E == (int)APVAE_ROW_REAL_E &&
              (int)ROW_MSE == (int)APVAE_ROW_MSE, "per-image row tables out of sync");
static_assert((int)ACC_WEIGHT == APVAE_ACCUM_WEIGHT && (int)ACC_BATCHES == APVAE_ACCUM_BATCHES && (int)ACC_MSE == APVAE_ACCUM_MSE &&
              (int)ACC_PN == APVAE_ACCUM_PN_ENERGY && (int)ACC_PD == APVAE_ACCUM_PD_ENERGY && (int)ACC_KL == APVAE_ACCUM_KL_ADV &&
              (int)ACC_LOGZ_OP == APVAE_ACCUM_LOG_Z_OP && (int)ACC_PRIOR_WEIGHT == APVAE_ACCUM_PRIOR_WEIGHT,
              "accumulator layouts out of sync");

extern "C" {

const char* apvae_version(void) { return "apvae-b200 0.1 (sm_100a, tcgen05/TMEM/TMA)"; }

const char* apvae_last_error(const apvae_handle* h) {
    if (h) return h->m.err.c_str();
    std::lock_guard<std::mutex> l(g_err_mu);
    static thread_local std::string copy;
    copy = g_err;
    return copy.c_str();
}

int apvae_default_config(int arch, int celeba, apvae_config* c) {
    if (!c) return set_global_err("null config", APVAE_ERR_INVALID);
    memset(c, 0, sizeof(*c));
    c->arch = arch;
    c->epsilon = -20.f;
    c->max_batch = 64;
    c->max_chunk = 512;
    c->enable_grad = 0;
    switch (arch) {
        case APVAE_ARCH_A:   // dmm_cifar10.py:27-91
            c->x_h = 32; c->x_w = 32; c->x_c = 3; c->z_dim = 1024;
            c->pull_back_energy_weight = 1000.f; c->min_logstd_of_q = -3.f;
            c->smallest_step = 5e-5f; c->len_train = 50000;
            break;
        case APVAE_ARCH_B:   // dmm_for_sampling.py:28-92
            c->celeba = celeba ? 1 : 0;
            c->x_h = c->x_w = celeba ? 64 : 32; c->x_c = 3; c->z_dim = 128;
            c->pull_back_energy_weight = 1000.f; c->min_logstd_of_q = -3.f; c->uniform_scale = 1;
            c->smallest_step = 5e-5f; c->len_train = celeba ? 162770 : 50000;
            break;
        case APVAE_ARCH_M:   // dmm_bernouli_fashion.py ExpConfig
            c->x_h = 28; c->x_w = 28; c->x_c = 1; c->z_dim = 40;
            c->pull_back_energy_weight = 20.f; c->min_logstd_of_q = -5.f; c->use_truncated = 1; c->bernoulli = 1;
            c->smallest_step = 5e-6f; c->len_train = 50000;
            break;
        default:
            return set_global_err("arch must be 0 (A), 1 (B) or 2 (M)", APVAE_ERR_INVALID);
    }
    return APVAE_OK;
}

int apvae_weight_count(const apvae_config* cfg, int* count) {
    std::string e;
    if (!cfg || !count) return set_global_err("null argument", APVAE_ERR_INVALID);
    if (!validate_config(*cfg, &e)) return set_global_err(e, APVAE_ERR_INVALID);
    *count = (int)weight_schema(*cfg).size();
    return APVAE_OK;
}

int apvae_weight_info(const apvae_config* cfg, int index, char* name, int name_cap, int64_t shape[4], int* ndim) {
    std::string e;
    if (!cfg || !name || !shape || !ndim) return set_global_err("null argument", APVAE_ERR_INVALID);
    if (!validate_config(*cfg, &e)) return set_global_err(e, APVAE_ERR_INVALID);
    const std::vector<WeightInfo> s = weight_schema(*cfg);
    if (index < 0 || index >= (int)s.size()) return set_global_err("weight index out of range", APVAE_ERR_INVALID);
    if ((int)s[index].name.size() + 1 > name_cap) return set_global_err("name buffer too small", APVAE_ERR_INVALID);
    strcpy(name, s[index].name.c_str());
    *ndim = (int)s[index].shape.size();
    for (int i = 0; i < 4; ++i) shape[i] = i < *ndim ? s[index].shape[i] : 1;
    return APVAE_OK;
}

int apvae_create(const apvae_config* cfg, int device, apvae_handle** out) {
    std::string e;
    if (!cfg || !out) return set_global_err("null argument", APVAE_ERR_INVALID);
    if (!validate_config(*cfg, &e)) return set_global_err(e, APVAE_ERR_INVALID);
    DeviceInfo dev;
    if (!device_open(device, &dev, &e)) return set_global_err(e, APVAE_ERR_NO_DEVICE);
    apvae_handle* h = new apvae_handle();
    h->m.cfg = *cfg;
    h->m.dev = dev;
    *out = h;
    return APVAE_OK;
}

int apvae_destroy(apvae_handle* h) {
    if (h) {
        cudaSetDevice(h->m.dev.device);
        cudaDeviceSynchronize();
        delete h;
    }
    return APVAE_OK;
}

int apvae_load_weights(apvae_handle* h, int n, const char* const* names, const float* const* data,
                       const int64_t* const* shapes, const int* ndims) {
    if (!h) return set_global_err("null handle", APVAE_ERR_INVALID);
    Model& m = h->m;
    if (m.built) { m.err = "weights already loaded"; return APVAE_ERR_STATE; }
    for (int i = 0; i < n; ++i) {
        HostArr a;
        size_t cnt = 1;
        for (int j = 0; j < ndims[i]; ++j) { a.shape.push_back(shapes[i][j]); cnt *= (size_t)shapes[i][j]; }
        a.v.assign(data[i], data[i] + cnt);
        m.weights[names[i]] = std::move(a);
    }
    for (const WeightInfo& w : weight_schema(m.cfg)) {
        auto it = m.weights.find(w.name);
        if (it == m.weights.end()) { m.err = "missing variable '" + w.name + "'"; return APVAE_ERR_WEIGHTS; }
        if (it->second.shape != w.shape) { m.err = "variable '" + w.name + "' has the wrong shape"; return APVAE_ERR_WEIGHTS; }
    }
    TuningScope tuning_scope(&m.cfg.tuning);          // kernel-selection overrides of this handle's config
    if (!m.build()) return APVAE_ERR_CUDA;
    return APVAE_OK;
}

int apvae_set_scalars(apvae_handle* h, float xi_raw, float beta, float log_Z) {
    if (!h) return set_global_err("null handle", APVAE_ERR_INVALID);
    h->m.xi_raw = xi_raw; h->m.beta = beta; h->m.log_Z = log_Z;
    return APVAE_OK;
}

#define REQUIRE_BUILT(h)                                                              \
    if (!(h)) return set_global_err("null handle", APVAE_ERR_INVALID);                \
    Model& m = (h)->m;                                                                \
    if (!m.built) { m.err = "weights not loaded (apvae_load_weights)"; return APVAE_ERR_STATE; } \
    if (cudaSetDevice(m.dev.device) != cudaSuccess) { m.err = "cudaSetDevice failed"; return APVAE_ERR_CUDA; }

#define LCHECK(call, what)                                                                         \
    do {                                                                                           \
        const int _e = (call);                                                                     \
        ++m.launches;                                                                              \
        if (_e != 0) { m.err = std::string(what) + ": " + cudaGetErrorString((cudaError_t)_e); return APVAE_ERR_CUDA; } \
    } while (0)

int apvae_encode(apvae_handle* h, const float* x_dev, int B, float* mu_dev, float* logstd_dev, void* stream) {
    REQUIRE_BUILT(h);
    if (B <= 0 || !x_dev) { m.err = "encode: empty batch"; return APVAE_ERR_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t per = (size_t)m.HW() * m.C();
    for (int b0 = 0; b0 < B; b0 += m.cfg.max_batch) {
        const int n = std::min(m.cfg.max_batch, B - b0);
        if (!m.run_encoder(x_dev + b0 * per, n, stream)) return APVAE_ERR_CUDA;
        const size_t bytes = (size_t)n * m.cfg.z_dim * sizeof(float);
        if (mu_dev) cudaMemcpyAsync(mu_dev + (size_t)b0 * m.cfg.z_dim, m.q_mu, bytes, cudaMemcpyDeviceToDevice, st);
        if (logstd_dev) cudaMemcpyAsync(logstd_dev + (size_t)b0 * m.cfg.z_dim, m.q_logstd, bytes, cudaMemcpyDeviceToDevice, st);
    }
    return APVAE_OK;
}

// z fp32 [n, d] -> decoder input tensors (no noise: mu = z, eps unused)
static int stage_latents(Model& m, const float* z_dev, int n, void* stream) {
    return launch_sample_z(nullptr, nullptr, z_dev, n, 0, 1, m.cfg.z_dim, 0.f, 0, nullptr, m.g_in.raw.ptr,
                           m.g_in.act.ptr, m.z_stride(), m.z_group(), m.z_cpad(), 0.2f, nullptr, nullptr, m.f16(), stream,
                           m.g_in.raw_lo.ptr, m.g_in.act_lo.ptr);
}

int apvae_decode(apvae_handle* h, const float* z_dev, int N, float* x_mean_dev, float* x_logstd_dev, void* stream) {
    REQUIRE_BUILT(h);
    if (N <= 0 || !z_dev) { m.err = "decode: empty batch"; return APVAE_ERR_INVALID; }
    if (x_logstd_dev && m.head_ld() != 2 * m.C()) { m.err = "this architecture's G_theta has no x_logstd head"; return APVAE_ERR_INVALID; }
    const int Cx = m.C(), ld = m.head_ld();
    for (int s0 = 0; s0 < N; s0 += m.cfg.max_chunk) {
        const int n = std::min(m.cfg.max_chunk, N - s0);
        LCHECK(stage_latents(m, z_dev + (size_t)s0 * m.cfg.z_dim, n, stream), "stage_latents");
        if (!m.run_decoder_chunk(n, stream)) return APVAE_ERR_CUDA;
        const long long rows = (long long)n * m.HW();
        if (x_mean_dev) LCHECK(launch_gather_cols(m.head_f32, ld, 0, Cx, rows, x_mean_dev + (size_t)s0 * m.HW() * Cx, stream), "gather");
        if (x_logstd_dev) LCHECK(launch_gather_cols(m.head_f32, ld, Cx, Cx, rows, x_logstd_dev + (size_t)s0 * m.HW() * Cx, stream), "gather");
    }
    return APVAE_OK;
}

int apvae_energy(apvae_handle* h, const float* x_dev, const float* z_dev, int N, float* d_dev, void* stream) {
    REQUIRE_BUILT(h);
    if (N <= 0 || !d_dev || (!x_dev == !z_dev)) { m.err = "energy: pass exactly one of x_dev / z_dev"; return APVAE_ERR_INVALID; }
    if (x_dev) return m.critic_of_images(x_dev, N, d_dev, stream) ? APVAE_OK : APVAE_ERR_CUDA;
    for (int s0 = 0; s0 < N; s0 += m.cfg.max_chunk) {
        const int n = std::min(m.cfg.max_chunk, N - s0);
        LCHECK(stage_latents(m, z_dev + (size_t)s0 * m.cfg.z_dim, n, stream), "stage_latents");
        if (!m.run_decoder_chunk(n, stream, false)) return APVAE_ERR_CUDA;
        if (!m.run_critic_chunk(n, d_dev + s0, stream)) return APVAE_ERR_CUDA;
    }
    return APVAE_OK;
}

}  // extern "C"

extern "C" {

int apvae_is_nll(apvae_handle* h, const float* x_dev, const float* x_origin_dev, int B, int K,
                 const float* eps_dev, uint64_t seed, float* per_image_dev, float* per_sample_dev, void* stream) {
    REQUIRE_BUILT(h);
    if (B <= 0 || K <= 0 || !x_dev || !per_image_dev) { m.err = "is_nll: bad arguments"; return APVAE_ERR_INVALID; }
    if (B > m.cfg.max_batch) { m.err = "is_nll: B exceeds max_batch"; return APVAE_ERR_INVALID; }
    if (!x_origin_dev) x_origin_dev = x_dev;
    const int d = m.cfg.z_dim;
    const long long S = (long long)K * B;
    if (!m.ensure_sample_buffers((size_t)S)) return APVAE_ERR_CUDA;
    cudaStream_t st = (cudaStream_t)stream;
    // q(z|x): q_net reads input_origin_x in arch M (dmm_bernouli_fashion.py:888), input_x otherwise
    if (!m.run_encoder(m.cfg.arch == APVAE_ARCH_M ? x_origin_dev : x_dev, B, stream)) return APVAE_ERR_CUDA;
    const float* eps = eps_dev;
    if (!eps) {
        LCHECK(launch_fill_normal(m.eps_buf, S * d, seed, 0, m.cfg.use_truncated, stream), "fill_normal");
        eps = m.eps_buf;
    }
    for (long long s0 = 0; s0 < S; s0 += m.cfg.max_chunk) {
        const int n = (int)std::min<long long>(m.cfg.max_chunk, S - s0);
        if (m.cfg.nf_layers > 0)
            LCHECK(launch_sample_z_flow(m.q_mu, m.q_logstd, eps + s0 * d, n, (int)(s0 % B), B, d, m.cfg.min_logstd_of_q,
                                        m.cfg.use_truncated, m.flow_w, m.flow_uhat, m.flow_b, m.flow_wu, m.cfg.nf_layers,
                                        nullptr, m.g_in.raw.ptr, m.g_in.act.ptr, m.z_stride(), m.z_group(), m.z_cpad(), 0.2f,
                                        m.log_q + s0, m.log_pn + s0, m.f16(), stream, m.g_in.raw_lo.ptr, m.g_in.act_lo.ptr),
                   "sample_z_flow");
        else
        LCHECK(launch_sample_z(m.q_mu, m.q_logstd, eps + s0 * d, n, (int)(s0 % B), B, d, m.cfg.min_logstd_of_q,
                               m.cfg.use_truncated, nullptr, m.g_in.raw.ptr, m.g_in.act.ptr, m.z_stride(), m.z_group(),
                               m.z_cpad(), 0.2f, m.log_q + s0, m.log_pn + s0, m.f16(), stream, m.g_in.raw_lo.ptr, m.g_in.act_lo.ptr),
               "sample_z");
        // G_theta, with log p(x|z) and the test_mse squared error scored inside the heads kernel
        const Model::LikArgs lik = {x_dev, (int)(s0 % B), B, m.log_px + s0, m.sq_err + s0};
        if (!m.run_decoder_chunk(n, stream, !m.heads_op, &lik)) return APVAE_ERR_CUDA;
        if (!m.run_critic_chunk(n, m.d_g + s0, stream)) return APVAE_ERR_CUDA;
    }
    LCHECK(launch_is_reduce(m.log_px, m.log_q, m.log_pn, m.d_g, m.sq_err, K, B, d, m.w_xi(), m.log_Z, per_image_dev, stream),
           "is_reduce");
    if (!m.critic_of_images(x_origin_dev, B, per_image_dev + (size_t)APVAE_ROW_REAL_E * B, stream)) return APVAE_ERR_CUDA;
    if (per_sample_dev) {
        const size_t bytes = (size_t)S * sizeof(float);
        cudaMemcpyAsync(per_sample_dev + APVAE_SROW_LOG_PX * S, m.log_px, bytes, cudaMemcpyDeviceToDevice, st);
        cudaMemcpyAsync(per_sample_dev + APVAE_SROW_LOG_Q * S, m.log_q, bytes, cudaMemcpyDeviceToDevice, st);
        cudaMemcpyAsync(per_sample_dev + APVAE_SROW_LOG_PN * S, m.log_pn, bytes, cudaMemcpyDeviceToDevice, st);
        cudaMemcpyAsync(per_sample_dev + APVAE_SROW_DG * S, m.d_g, bytes, cudaMemcpyDeviceToDevice, st);
        cudaMemcpyAsync(per_sample_dev + APVAE_SROW_SQ_ERR * S, m.sq_err, bytes, cudaMemcpyDeviceToDevice, st);
    }
    return APVAE_OK;
}

int apvae_sample_z(apvae_handle* h, const float* x_origin_dev, int B, int K, const float* eps_dev, uint64_t seed,
                   float* z_dev, float* log_q_dev, float* log_pn_dev, void* stream) {
    REQUIRE_BUILT(h);
    if (B <= 0 || K <= 0 || !x_origin_dev || !z_dev) { m.err = "sample_z: empty batch"; return APVAE_ERR_INVALID; }
    if (B > m.cfg.max_batch) { m.err = "sample_z: batch exceeds max_batch"; return APVAE_ERR_INVALID; }
    const long long S = (long long)K * B;
    const int d = m.cfg.z_dim;
    if (!m.run_encoder(x_origin_dev, B, stream)) return APVAE_ERR_CUDA;
    const float* eps = eps_dev;
    if (!eps) {
        if (!m.ensure_sample_buffers((size_t)S)) return APVAE_ERR_CUDA;
        LCHECK(launch_fill_normal(m.eps_buf, S * d, seed, 0, m.cfg.use_truncated, stream), "fill_normal");
        eps = m.eps_buf;
    }
    float* lq = log_q_dev;
    if (!lq) { if (!m.ensure_sample_buffers((size_t)S)) return APVAE_ERR_CUDA; lq = m.log_q; }
    if (m.cfg.nf_layers > 0)
        LCHECK(launch_sample_z_flow(m.q_mu, m.q_logstd, eps, (int)S, 0, B, d, m.cfg.min_logstd_of_q, m.cfg.use_truncated,
                                    m.flow_w, m.flow_uhat, m.flow_b, m.flow_wu, m.cfg.nf_layers, z_dev, nullptr, nullptr, 0, 1,
                                    0, 0.2f, lq, log_pn_dev, m.f16(), stream), "sample_z_flow");
    else
    LCHECK(launch_sample_z(m.q_mu, m.q_logstd, eps, (int)S, 0, B, d, m.cfg.min_logstd_of_q, m.cfg.use_truncated, z_dev,
                           nullptr, nullptr, 0, 1, 0, 0.2f, lq, log_pn_dev, m.f16(), stream), "sample_z");
    return APVAE_OK;
}

int apvae_accumulate_metrics(apvae_handle* h, const float* per_image_dev, const float* prior4_dev, int B, int K,
                             double* accum_dev, void* stream) {
    REQUIRE_BUILT(h);
    static_assert(APVAE_ACCUM_SLOTS == kAccumSlots, "accumulator layouts out of sync");
    if (B <= 0 || K <= 0 || !per_image_dev || !accum_dev) { m.err = "accumulate_metrics: bad arguments"; return APVAE_ERR_INVALID; }
    LCHECK(launch_accumulate_metrics(per_image_dev, B, K, APVAE_PER_IMAGE_ROWS, prior4_dev, accum_dev, stream), "accumulate_metrics");
    return APVAE_OK;
}

int apvae_prior_stats(apvae_handle* h, int Kp, const float* eps_dev, uint64_t seed, float* d_g_dev,
                      float* log_pn_dev, float* prior4_dev, void* stream) {
    REQUIRE_BUILT(h);
    if (Kp <= 0 || (!d_g_dev && !log_pn_dev && !prior4_dev)) { m.err = "prior_stats: bad arguments"; return APVAE_ERR_INVALID; }
    const int d = m.cfg.z_dim;
    if (!m.ensure_sample_buffers((size_t)Kp)) return APVAE_ERR_CUDA;
    const float* eps = eps_dev;
    if (!eps) {
        LCHECK(launch_fill_normal(m.eps_buf, (long long)Kp * d, seed, 1ull << 48, 0, stream), "fill_normal");
        eps = m.eps_buf;
    }
    // per-draw rows land in the caller's arrays when given, else in the handle's per-sample scratch
    float* dg = d_g_dev ? d_g_dev : m.d_g;
    float* lpn = log_pn_dev ? log_pn_dev : m.log_pn;
    for (int s0 = 0; s0 < Kp; s0 += m.cfg.max_chunk) {
        const int n = std::min(m.cfg.max_chunk, Kp - s0);
        LCHECK(launch_sample_z(nullptr, nullptr, eps + (size_t)s0 * d, n, 0, 1, d, 0.f, 0, nullptr, m.g_in.raw.ptr,
                               m.g_in.act.ptr, m.z_stride(), m.z_group(), m.z_cpad(), 0.2f, nullptr, lpn + s0,
                               m.f16(), stream, m.g_in.raw_lo.ptr, m.g_in.act_lo.ptr), "sample_z");
        if (!m.run_decoder_chunk(n, stream, false)) return APVAE_ERR_CUDA;
        if (!m.run_critic_chunk(n, dg + s0, stream)) return APVAE_ERR_CUDA;
    }
    if (prior4_dev) LCHECK(launch_prior_reduce(dg, lpn, Kp, d, m.w_xi(), m.log_Z, prior4_dev, stream), "prior_reduce");
    return APVAE_OK;
}

int apvae_latent_gan(apvae_handle* h, const float* z_omega_dev, int n_total, uint64_t seed, float* z0_dev, void* stream) {
    REQUIRE_BUILT(h);
    if (!m.cfg.latent_gan) { m.err = "latent_gan: create the handle with latent_gan = 1"; return APVAE_ERR_STATE; }
    if (n_total <= 0 || !z0_dev) { m.err = "latent_gan: bad arguments"; return APVAE_ERR_INVALID; }
    for (int s0 = 0; s0 < n_total; s0 += m.cfg.max_chunk) {
        const int n = std::min(m.cfg.max_chunk, n_total - s0);
        const int din = m.omega_in_dim();
        const float* zo = z_omega_dev ? z_omega_dev + (size_t)s0 * din : nullptr;
        if (!zo) {                                       // z ~ N(0, I_128) (dmm_cifar10.py:626-629); arch M: N(0, I_z_dim)
            LCHECK(launch_fill_normal(m.om_noise, (long long)n * din, seed, (uint64_t)s0 * din, 0, stream), "fill_normal");
            zo = m.om_noise;
        }
        if (!m.latent_gan_chunk(zo, n, z0_dev + (size_t)s0 * m.cfg.z_dim, stream)) return APVAE_ERR_CUDA;
    }
    return APVAE_OK;
}

int apvae_sample_images(apvae_handle* h, const float* z_dev, int n_total, uint8_t* out_dev, void* stream) {
    REQUIRE_BUILT(h);
    if (n_total <= 0 || !z_dev || !out_dev) { m.err = "sample_images: bad arguments"; return APVAE_ERR_INVALID; }
    const size_t per = (size_t)m.HW() * m.C();
    for (int s0 = 0; s0 < n_total; s0 += m.cfg.max_chunk) {
        const int n = std::min(m.cfg.max_chunk, n_total - s0);
        LCHECK(stage_latents(m, z_dev + (size_t)s0 * m.cfg.z_dim, n, stream), "stage_latents");
        if (!m.run_decoder_chunk(n, stream)) return APVAE_ERR_CUDA;
        LCHECK(launch_quantize_images(m.head_f32, m.head_ld(), (long long)n * m.HW(), m.C(), out_dev + s0 * per, stream),
               "quantize");
    }
    return APVAE_OK;
}

int apvae_discretized_logistic_log_prob(const float* x_dev, const float* mean_dev, const float* log_scale_dev, int64_t rows,
                                        int64_t group, float bin_size, int has_min, float min_val, int has_max, float max_val,
                                        int biased_edges, float epsilon, float* out_dev, void* stream) {
    if (!x_dev || !mean_dev || !log_scale_dev || !out_dev || rows <= 0 || group <= 0 || !(bin_size > 0.f))
        return set_global_err("discretized_logistic_log_prob: bad arguments", APVAE_ERR_INVALID);
    LikelihoodSpec s;
    memset(&s, 0, sizeof(s));
    s.bernoulli = 0; s.bin_size = bin_size; s.log_bin_size = logf(bin_size);
    s.has_min = has_min ? 1 : 0; s.has_max = has_max ? 1 : 0; s.min_val = min_val; s.max_val = max_val;
    s.biased_edges = biased_edges ? 1 : 0; s.epsilon = epsilon; s.min_log_scale = -INFINITY;
    const int e = launch_elementwise_logprob(x_dev, mean_dev, log_scale_dev, rows, group, s, out_dev, stream);
    return e == 0 ? APVAE_OK : set_global_err(cudaGetErrorString((cudaError_t)e), e == (int)cudaErrorNoDevice || e == (int)cudaErrorInsufficientDriver ? APVAE_ERR_NO_DEVICE : APVAE_ERR_CUDA);
}
int apvae_bernoulli_log_prob(const float* x_dev, const float* mean_dev, int64_t rows, int64_t group, float* out_dev, void* stream) {
    if (!x_dev || !mean_dev || !out_dev || rows <= 0 || group <= 0) return set_global_err("bernoulli_log_prob: bad arguments", APVAE_ERR_INVALID);
    const LikelihoodSpec s = script_likelihood(1, -20.f);
    const int e = launch_elementwise_logprob(x_dev, mean_dev, nullptr, rows, group, s, out_dev, stream);
    return e == 0 ? APVAE_OK : set_global_err(cudaGetErrorString((cudaError_t)e), APVAE_ERR_CUDA);
}
int apvae_is_reduce(const float* log_px_dev, const float* log_q_dev, const float* log_pn_dev, const float* d_g_dev,
                    const float* sq_err_dev, int K, int B, int z_dim, float w_xi, float log_Z, float* per_image_dev, void* stream) {
    if (!log_px_dev || !log_q_dev || !log_pn_dev || !d_g_dev || !per_image_dev || K <= 0 || B <= 0)
        return set_global_err("is_reduce: bad arguments", APVAE_ERR_INVALID);
    const int e = launch_is_reduce(log_px_dev, log_q_dev, log_pn_dev, d_g_dev, sq_err_dev, K, B, z_dim, w_xi, log_Z, per_image_dev, stream);
    return e == 0 ? APVAE_OK : set_global_err(cudaGetErrorString((cudaError_t)e), APVAE_ERR_CUDA);
}

int apvae_fill_normal(float* out_dev, int64_t n, uint64_t seed, uint64_t offset, int truncated, void* stream) {
    const int e = launch_fill_normal(out_dev, n, seed, offset, truncated, stream);
    return e == 0 ? APVAE_OK : set_global_err(cudaGetErrorString((cudaError_t)e), APVAE_ERR_CUDA);
}
int apvae_fill_uniform(float* out_dev, int64_t n, uint64_t seed, uint64_t offset, void* stream) {
    const int e = launch_fill_uniform(out_dev, n, seed, offset, stream);
    return e == 0 ? APVAE_OK : set_global_err(cudaGetErrorString((cudaError_t)e), APVAE_ERR_CUDA);
}

int apvae_launch_count(const apvae_handle* h, int64_t* launches) {
    if (!h || !launches) return set_global_err("null argument", APVAE_ERR_INVALID);
    *launches = h->m.launches;
    return APVAE_OK;
}
int apvae_macs(const apvae_handle* h, double* q, double* g, double* d) {
    if (!h) return set_global_err("null handle", APVAE_ERR_INVALID);
    if (q) *q = h->m.macs_q;
    if (g) *g = h->m.macs_g;
    if (d) *d = h->m.macs_d;
    return APVAE_OK;
}

static Program* pick_net(Model& m, int net) {
    return net == 0 ? &m.q_prog : net == 1 ? &m.g_prog : net == 2 ? &m.d_prog : net == 3 ? &m.gd_prog : net == 4 ? &m.gg_prog :
           net == 5 ? &m.om_prog : nullptr;
}
int apvae_op_count(const apvae_handle* h, int net, int* count) {
    if (!h || !count) return set_global_err("null argument", APVAE_ERR_INVALID);
    Program* p = pick_net(const_cast<Model&>(h->m), net);
    if (!p) return set_global_err("net must be 0 .. 5", APVAE_ERR_INVALID);
    *count = (int)p->size();
    return APVAE_OK;
}
int apvae_op_time(apvae_handle* h, int net, int index, int n, int reps, char* name, int name_cap,
                  double* macs_per_unit, float* ms_per_launch, void* stream) {
    REQUIRE_BUILT(h);
    Program* p = pick_net(m, net);
    if (!p || index < 0 || index >= (int)p->size() || reps <= 0) { m.err = "op_time: bad arguments"; return APVAE_ERR_INVALID; }
    DevOp& op = *(*p)[index];
    if (n <= 0 || n > op.capacity_n) { m.err = "op_time: n exceeds the op's capacity"; return APVAE_ERR_INVALID; }
    if (name && name_cap > 0) {      // op name + which kernel runs it
        const std::string full = op.name + (op.is_tiny ? " [tiny conv, CUDA cores]" : op.is_small_block ? " [fused block, CUDA cores]" : op.is_small_block_grad ? " [fused block adjoint, CUDA cores]" : op.is_heads ? " [heads, CUDA cores]" : op.is_patch ? (op.pparams.fast_epi ? " [patchgemm, TMA epilogue]" : " [patchgemm]") : " [tapgemm]");
        strncpy(name, full.c_str(), name_cap - 1); name[name_cap - 1] = 0;
    }
    if (macs_per_unit) *macs_per_unit = op.macs_per_image;
    cudaStream_t st = (cudaStream_t)stream;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 2; ++i) LCHECK(devop_launch(m.dev, op, n, stream), "op_time warm-up");
    cudaEventRecord(e0, st);
    for (int i = 0; i < reps; ++i) LCHECK(devop_launch(m.dev, op, n, stream), "op_time");
    cudaEventRecord(e1, st);
    const cudaError_t e = cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (e != cudaSuccess) { m.err = std::string("op_time: ") + cudaGetErrorString(e); return APVAE_ERR_CUDA; }
    if (ms_per_launch) *ms_per_launch = ms / reps;
    return APVAE_OK;
}

int apvae_aux_time(apvae_handle* h, int which, int n, int B, int reps, char* name, int name_cap, double* bytes_per_launch,
                   float* ms_per_launch, void* stream) {
    REQUIRE_BUILT(h);
    if (which < 0 || which >= APVAE_AUX_COUNT || n <= 0 || reps <= 0) { m.err = "aux_time: bad arguments"; return APVAE_ERR_INVALID; }
    if (B <= 0) B = 1;
    const int d = m.cfg.z_dim;
    const bool sampler = which >= APVAE_AUX_MALA_PROPOSE && which <= APVAE_AUX_TANH_BACKWARD;
    if (sampler && !m.cfg.enable_grad) { m.err = "aux_time: the sampler's kernels need enable_grad"; return APVAE_ERR_STATE; }
    if (n > m.cfg.max_chunk && which != APVAE_AUX_IS_REDUCE && which != APVAE_AUX_FILL_NORMAL) { m.err = "aux_time: n exceeds max_chunk"; return APVAE_ERR_INVALID; }
    if (!m.ensure_sample_buffers((size_t)std::max(n, m.cfg.max_chunk))) return APVAE_ERR_CUDA;
    if (sampler && (size_t)n > m.mala_cap) {           // sizes the sampler's state buffers (a zero-step run)
        if (!m.mala(m.z_tmp, std::min(n, m.cfg.max_chunk), 0, 1e-3f, nullptr, nullptr, 0, nullptr, nullptr, nullptr, nullptr, stream)) return APVAE_ERR_CUDA;
    }
    const long long px = (long long)n * m.HW();
    const int C3 = m.C();
    const char* nm = "";
    double bytes = 0;
    auto launch = [&]() -> int {
        switch (which) {
            case APVAE_AUX_SAMPLE_Z:          // eps in; z as raw + act operand tensors out (+ log q, log N)
                nm = "sample_z";
                bytes = (double)n * (d * 4.0 + 2.0 * m.z_stride() * 2.0 + 8.0);
                return launch_sample_z(m.q_mu, m.q_logstd, m.eps_buf, n, 0, std::min(B, m.cfg.max_batch), d, m.cfg.min_logstd_of_q,
                                       m.cfg.use_truncated, nullptr, m.g_in.raw.ptr, m.g_in.act.ptr, m.z_stride(), m.z_group(),
                                       m.z_cpad(), 0.2f, m.log_q, m.log_pn, m.f16(), stream);
            case APVAE_AUX_IS_REDUCE: {       // five [K, B] rows in (second pass from L2), ten floats per image out
                nm = "is_reduce";
                const int K = std::max(1, n / B);
                bytes = (double)K * B * 5 * 4.0 + (double)B * kPerImageRows * 4.0;
                return launch_is_reduce(m.log_px, m.log_q, m.log_pn, m.d_g, m.sq_err, K, B, d, m.w_xi(), m.log_Z, m.eps_buf, stream);
            }
            case APVAE_AUX_SUM_TILES: {
                nm = "sum_tiles";
                const int tpi = heads_tiles_per_image(m.HW());
                if (!m.lik_part) return (int)cudaErrorInvalidValue;
                bytes = (double)n * (tpi * 8.0 + 8.0);
                return launch_sum_tiles(m.lik_part, tpi, n, m.log_px, m.sq_err, stream);
            }
            case APVAE_AUX_PACK_IMAGE:        // fp32 image in; raw + act operand rows out
                nm = "pack_image";
                bytes = (double)px * (C3 * 4.0 + 2.0 * m.d_in.raw.cp * 2.0);
                return launch_pack_image(m.head_f32, px, C3, m.d_in.raw.cp, m.d_in.raw.ptr, m.d_in.act.ptr, 0.2f, m.f16(), stream);
            case APVAE_AUX_ROWDOT:
                nm = "rowdot";
                bytes = (double)n * (64 * 4.0 + 4.0);
                return launch_rowdot(m.d64, 64, m.d_last_w, m.d_last_b, 64, n, m.d_g, stream);
            case APVAE_AUX_FILL_NORMAL:
                nm = "fill_normal";
                bytes = (double)n * d * 4.0;
                return launch_fill_normal(m.eps_buf, (long long)n * d, 1234, 0, m.cfg.use_truncated, stream);
            case APVAE_AUX_MALA_PROPOSE:      // z, grad, eps in; z' out
                nm = "mala_propose";
                bytes = (double)n * d * 4.0 * 4;
                return launch_mala_propose(m.z_tmp, m.mg, m.eps_buf, n, d, 1e-3f, m.mzp, stream);
            case APVAE_AUX_MALA_ACCEPT:       // z, z', grad, grad' in; z, grad out for the accepted chains (counted in full)
                nm = "mala_accept";
                bytes = (double)n * (d * 4.0 * 6 + 4.0 * 8);
                return launch_mala_accept(m.z_tmp, m.me, m.mg, m.mp, m.mzp, m.mep, m.mgp, m.mpp, m.log_q, n, d, 1e-3f,
                                          m.log_px, m.log_pn, m.d_g, stream);
            case APVAE_AUX_ENERGY_COMBINE:    // d D/dz, z in; grad out
                nm = "energy_combine";
                bytes = (double)n * (d * 4.0 * 3 + 8.0);
                return launch_energy_combine(m.d_g, m.gz_f32, m.z_tmp, n, d, m.w_xi(), m.me, m.mg, stream);
            case APVAE_AUX_TANH_BACKWARD:     // fp32 gradient + fp32 x_mean in; operand row out
                nm = "tanh_backward_pack";
                bytes = (double)px * (C3 * 4.0 * 2 + m.g_pre.cp * 2.0);
                return launch_tanh_backward_pack(m.gx_f32, m.head_f32, m.head_ld(), px, C3, m.g_pre.cp, m.g_pre.ptr, m.f16(), stream);
            case APVAE_AUX_QUANTIZE:
                nm = "quantize_images";
                bytes = (double)px * C3 * 5.0;
                return launch_quantize_images(m.head_f32, m.head_ld(), px, C3, reinterpret_cast<uint8_t*>(m.d_in.act.ptr), stream);
        }
        return (int)cudaErrorInvalidValue;
    };
    cudaStream_t st = (cudaStream_t)stream;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 2; ++i) LCHECK(launch(), "aux_time warm-up");
    cudaEventRecord(e0, st);
    for (int i = 0; i < reps; ++i) LCHECK(launch(), "aux_time");
    cudaEventRecord(e1, st);
    const cudaError_t e = cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (e != cudaSuccess) { m.err = std::string("aux_time: ") + cudaGetErrorString(e); return APVAE_ERR_CUDA; }
    if (name && name_cap > 0) { strncpy(name, nm, name_cap - 1); name[name_cap - 1] = 0; }
    if (bytes_per_launch) *bytes_per_launch = bytes;
    if (ms_per_launch) *ms_per_launch = ms / reps;
    return APVAE_OK;
}

// ---------------------------------------------------------------- layer-level interface
int apvae_plan_create(const apvae_conv_term* terms, int nterms, int capacity_n, int operand_bf16, const apvae_tuning* tune,
                      apvae_plan** out) {
    if (!terms || nterms <= 0 || !out || capacity_n <= 0) return set_global_err("plan_create: bad arguments", APVAE_ERR_INVALID);
    TuningScope tuning_scope(tune);
    std::vector<TermSpec> ts;
    for (int i = 0; i < nterms; ++i) {
        TermSpec t;
        t.src = terms[i].src; t.transpose = terms[i].transpose; t.k = terms[i].k; t.stride = terms[i].stride;
        t.cin = terms[i].cin; t.cout = terms[i].cout; t.in_h = terms[i].in_h; t.in_w = terms[i].in_w;
        t.in_cp = terms[i].in_cp; t.kernel = terms[i].kernel; t.bias = terms[i].bias;
        t.f16 = operand_bf16 ? 0 : 1;
        ts.push_back(t);
    }
    apvae_plan* p = new apvae_plan();
    p->f16 = operand_bf16 ? 0 : 1;
    if (tune) p->tuning = *tune;
    const int force_macc = tuning().patch_macc, force_cluster = tuning().patch_cluster;
    std::string e;
    if (!plan_op(ts, capacity_n, &p->plan, &e)) { delete p; return set_global_err(e, APVAE_ERR_UNSUPPORTED); }
    bool all_s2t = true;          // stride-2 transposed terms of one geometry: sub-pixel phase plans
    for (const TermSpec& t : ts) all_s2t = all_s2t && t.transpose && t.stride == 2 && t.k == ts[0].k && t.cin == ts[0].cin;
    if (switch_on(tuning().patch, true)) {
        std::string pe;
        p->has_patch = plan_patch_op(ts, capacity_n, force_macc, force_cluster, &p->patch, &pe);
        if (!p->has_patch && switch_on(tuning().patch_phases, true) && all_s2t) {
            p->phases.resize(4);
            for (int ph = 0; ph < 4; ++ph)
                if (!plan_patch_op(ts, capacity_n, force_macc, force_cluster, &p->phases[ph], &pe, ph)) {
                    p->phases.clear();
                    break;
                }
            // both column phases per op (two ops); used by plan_run when the output rows are unpadded
            if (!p->phases.empty() && switch_on(tuning().patch_merge, true)) {
                p->row_phases.resize(2);
                for (int py = 0; py < 2; ++py)
                    if (!plan_patch_op(ts, capacity_n, force_macc, force_cluster, &p->row_phases[py], &pe, 4 + py)) {
                        p->row_phases.clear();
                        break;
                    }
            }
        }
    }
    *out = p;
    return APVAE_OK;
}

static const PatchPlan* selected_patch(const apvae_plan* p) {
    if (p->sel < 0) return p->has_patch ? &p->patch : nullptr;
    if (p->sel < 4) return (size_t)p->sel < p->phases.size() ? &p->phases[p->sel] : nullptr;
    return (size_t)(p->sel - 4) < p->row_phases.size() ? &p->row_phases[p->sel - 4] : nullptr;
}

int apvae_plan_select_patch(apvae_plan* p, int which) {
    if (!p || which < -1 || which > 5) return set_global_err("select_patch: which must be -1 (the op), 0..3 (sub-pixel phase) or 4..5 (row phase)", APVAE_ERR_INVALID);
    const int old = p->sel;
    p->sel = which;
    if (!selected_patch(p)) { p->sel = old; return set_global_err("select_patch: this op has no such patch plan", APVAE_ERR_INVALID); }
    return APVAE_OK;
}

int apvae_plan_get_info(const apvae_plan* p, apvae_plan_info* info) {
    if (!p || !info) return set_global_err("null argument", APVAE_ERR_INVALID);
    const OpPlan& P = p->plan;
    memset(info, 0, sizeof(*info));
    info->kc = P.kc; info->npad = P.npad; info->nt = P.nt; info->n_tiles = P.n_tiles; info->cout = P.cout;
    info->num_phases = P.num_phases;
    for (int i = 0; i < 5; ++i) info->phase_begin[i] = P.phase_begin[i];
    info->total_steps = (int)P.steps.size();
    info->bw = P.bw; info->bh = P.bh; info->bn = P.bn; info->rows = P.rows; info->wg = P.wg; info->hg = P.hg;
    info->tiles_x = P.tiles_x; info->tiles_y = P.tiles_y; info->ho = P.ho; info->wo = P.wo; info->os = P.os;
    info->nsrc = P.nsrc; info->capacity_n = P.capacity_n;
    for (int s = 0; s < P.nsrc; ++s) {
        for (int i = 0; i < 5; ++i) { info->src_dims[s][i] = P.src[s].dims[i]; info->src_box[s][i] = P.src[s].box[i]; }
        for (int i = 0; i < 4; ++i) info->src_strides[s][i] = P.src[s].strides[i];
    }
    info->macs_per_image = P.macs_per_image;
    if (const PatchPlan* qp = selected_patch(p)) {
        const PatchPlan& Q = *qp;
        info->patch_phase = Q.phase; info->patch_out_scale = Q.out_scale; info->patch_npad = Q.npad;
        info->patch = 1; info->patch_macc = Q.macc; info->patch_cluster = Q.cluster; info->patch_rows_tile = Q.rows_tile;
        info->patch_groups = (int)Q.groups.size(); info->patch_items = Q.num_items;
        info->patch_a_stages = Q.a_stages; info->patch_b_stages = Q.b_stages;
        info->patch_prow[0] = Q.prow[0]; info->patch_prow[1] = Q.prow[1];
        info->patch_fold = Q.fold;
        info->patch_epi2 = Q.epi2;
        info->patch_wv = Q.w;
        info->patch_s2d = Q.s2d; info->patch_nmaps = Q.nmaps;
        for (int i = 0; i < 4; ++i) { info->patch_map_src[i] = Q.map_src[i]; info->patch_map_ph[i] = Q.map_ph[i]; info->patch_map_prow[i] = Q.prow[i]; }
    }
    return APVAE_OK;
}

int apvae_plan_copy_steps(const apvae_plan* p, int32_t* out) {
    if (!p || !out) return set_global_err("null argument", APVAE_ERR_INVALID);
    for (size_t i = 0; i < p->plan.steps.size(); ++i) {
        const TapStep& s = p->plan.steps[i];
        out[5 * i + 0] = s.src; out[5 * i + 1] = s.dx; out[5 * i + 2] = s.dy; out[5 * i + 3] = s.ph; out[5 * i + 4] = s.c0;
    }
    return APVAE_OK;
}
int apvae_plan_copy_weights(const apvae_plan* p, uint16_t* out) {
    if (!p || !out) return set_global_err("null argument", APVAE_ERR_INVALID);
    memcpy(out, p->plan.wpack.data(), p->plan.wpack.size() * sizeof(uint16_t));
    return APVAE_OK;
}
int apvae_plan_copy_bias(const apvae_plan* p, float* out) {
    if (!p || !out) return set_global_err("null argument", APVAE_ERR_INVALID);
    memcpy(out, p->plan.bias.data(), p->plan.bias.size() * sizeof(float));
    return APVAE_OK;
}
int apvae_plan_copy_patch_groups(const apvae_plan* p, int32_t* out) {
    const PatchPlan* q = p ? selected_patch(p) : nullptr;
    if (!q || !out) return set_global_err("no patch plan", APVAE_ERR_INVALID);
    for (size_t i = 0; i < q->groups.size(); ++i) {
        const PatchGroup& g = q->groups[i];
        out[6 * i + 0] = g.src; out[6 * i + 1] = g.dx; out[6 * i + 2] = g.py0; out[6 * i + 3] = g.k;
        out[6 * i + 4] = g.c0; out[6 * i + 5] = g.prow;
    }
    return APVAE_OK;
}
int apvae_plan_copy_patch_weights(const apvae_plan* p, uint16_t* out) {
    const PatchPlan* q = p ? selected_patch(p) : nullptr;
    if (!q || !out) return set_global_err("no patch plan", APVAE_ERR_INVALID);
    memcpy(out, q->wpack.data(), q->wpack.size() * sizeof(uint16_t));
    return APVAE_OK;
}
int apvae_plan_copy_patch_bias(const apvae_plan* p, float* out) {
    const PatchPlan* q = p ? selected_patch(p) : nullptr;
    if (!q || !out) return set_global_err("no patch plan", APVAE_ERR_INVALID);
    memcpy(out, q->bias.data(), q->bias.size() * sizeof(float));
    return APVAE_OK;
}
int apvae_plan_destroy(apvae_plan* p) { delete p; return APVAE_OK; }

int apvae_plan_run(const apvae_plan* p, int device, const void* const* src_dev, int n_valid, const apvae_epilogue* epi) {
    return apvae_plan_bench(p, device, src_dev, n_valid, epi, 0, nullptr);
}

int apvae_plan_bench(const apvae_plan* p, int device, const void* const* src_dev, int n_valid,
                     const apvae_epilogue* epi, int reps, float* ms_per_launch) {
    if (!p || !src_dev || !epi) return set_global_err("null argument", APVAE_ERR_INVALID);
    TuningScope tuning_scope(&p->tuning);
    DeviceInfo dev;
    std::string e;
    if (!device_open(device, &dev, &e)) return set_global_err(e, APVAE_ERR_NO_DEVICE);
    Epilogue ep;
    ep.out_raw = epi->out_raw_dev; ep.out_act = epi->out_act_dev; ep.ldo = epi->ldo;
    ep.bf16_zero_from = epi->bf16_zero_from > 0 ? epi->bf16_zero_from : (1 << 30);
    ep.out_f32 = epi->out_f32_dev; ep.ld_f32 = epi->ld_f32; ep.f32_cols = epi->f32_cols;
    ep.res = epi->res_dev; ep.res_mode = epi->res_mode; ep.ld_res = epi->ld_res;
    ep.mask = epi->mask_dev; ep.ld_mask = epi->ld_mask; ep.tanh_cols = epi->tanh_cols;
    ep.out_raw_lo = epi->out_raw_lo_dev; ep.out_act_lo = epi->out_act_lo_dev; ep.res_lo = epi->res_lo_dev;
    ep.f16 = p->f16;
    // the op, or the four phase ops of a stride-2 transposed conv (only with a bf16-only epilogue; else the tap-GEMM op)
    std::vector<std::unique_ptr<DevOp>> ops;
    bool phased = !p->phases.empty();
    if (!p->row_phases.empty()) {            // two merged ops when the epilogue allows it
        bool ok = true;
        for (int py = 0; py < 2 && ok; ++py) {
            std::unique_ptr<DevOp> o(new DevOp());
            o->name = "plan_run/rows" + std::to_string(py);
            std::string e2;
            ok = devop_create_patch(dev, p->row_phases[py], src_dev, ep, o.get(), &e2);
            if (ok) ops.push_back(std::move(o));
        }
        if (ok) phased = false;
        else { ops.clear(); cudaGetLastError(); }
    }
    const bool merged_ok = !ops.empty();
    if (phased) {
        for (int ph = 0; ph < 4 && phased; ++ph) {
            std::unique_ptr<DevOp> o(new DevOp());
            o->name = "plan_run/phase" + std::to_string(ph);
            std::string e2;
            phased = devop_create_patch(dev, p->phases[ph], src_dev, ep, o.get(), &e2);
            if (phased) ops.push_back(std::move(o));
        }
        if (!phased) { ops.clear(); cudaGetLastError(); }
    }
    if (!phased && !merged_ok) {
        std::unique_ptr<DevOp> o(new DevOp());
        o->name = "plan_run";
        const bool use_patch = p->has_patch;
        bool made = use_patch ? devop_create_patch(dev, p->patch, src_dev, ep, o.get(), &e) : devop_create(dev, p->plan, src_dev, ep, o.get(), &e);
        if (!made && use_patch && p->patch.fold > 1) {       // folded op the A-stationary kernel cannot take: tap-GEMM
            cudaGetLastError();
            o.reset(new DevOp());
            o->name = "plan_run";
            made = devop_create(dev, p->plan, src_dev, ep, o.get(), &e);
        }
        if (!made) return set_global_err(e, APVAE_ERR_CUDA);
        ops.push_back(std::move(o));
    }
    auto devop_launch_all = [&](int n) {
        int r = 0;
        for (auto& o : ops) { r = devop_launch(dev, *o, n, nullptr); if (r != 0) break; }
        return r;
    };
    int err = devop_launch_all(n_valid);
    if (err == 0 && reps > 0) {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        err = devop_launch_all(n_valid);
        cudaEventRecord(e0, nullptr);
        for (int i = 0; i < reps && err == 0; ++i) err = devop_launch_all(n_valid);
        cudaEventRecord(e1, nullptr);
        if (err == 0) err = (int)cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms_per_launch) *ms_per_launch = ms / reps;
        cudaEventDestroy(e0); cudaEventDestroy(e1);
    }
    if (err == 0) err = (int)cudaDeviceSynchronize();
    if (err == 0) {
        for (auto& o : ops) {
            if (!o->d_trace) continue;
            std::vector<unsigned long long> t(64 * 16);
            cudaMemcpy(t.data(), o->d_trace, t.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
            const unsigned long long t0 = t[0];
            fprintf(stderr, "trace %s (cycles since the first stamp; 16 slots per tile)\n", o->name.c_str());
            for (int i = 0; i < 64 && t[i * 16] != 0; ++i) {
                fprintf(stderr, "tile %2d:", i);
                for (int j = 0; j < 16; ++j) fprintf(stderr, " %7lld", t[i * 16 + j] ? (long long)(t[i * 16 + j] - t0) : -1LL);
                fprintf(stderr, "\n");
            }
        }
    }
    if (err != 0) return set_global_err(std::string("plan_run: ") + cudaGetErrorString((cudaError_t)err), APVAE_ERR_CUDA);
    return APVAE_OK;
}

// ---------------------------------------------------------------- sampler (input-gradient programs)
int apvae_energy_grad(apvae_handle* h, const float* z_dev, int n, float* energy_dev, float* grad_dev, float* pure_dev,
                      void* stream) {
    REQUIRE_BUILT(h);
    if (!m.cfg.enable_grad) { m.err = "energy_grad: create the handle with enable_grad = 1"; return APVAE_ERR_STATE; }
    if (n <= 0 || !z_dev) { m.err = "energy_grad: empty batch"; return APVAE_ERR_INVALID; }
    const int d = m.cfg.z_dim;
    for (int s0 = 0; s0 < n; s0 += m.cfg.max_chunk) {
        const int c = std::min(m.cfg.max_chunk, n - s0);
        if (!m.energy_grad_chunk(z_dev + (size_t)s0 * d, c, energy_dev ? energy_dev + s0 : nullptr,
                                 grad_dev ? grad_dev + (size_t)s0 * d : nullptr, pure_dev ? pure_dev + s0 : nullptr, stream))
            return APVAE_ERR_CUDA;
    }
    return APVAE_OK;
}
int apvae_image_energy_grad(apvae_handle* h, const float* x_dev, int N, float* d_dev, float* grad_dev, float* grad_sqnorm_dev,
                            void* stream) {
    REQUIRE_BUILT(h);
    if (!m.cfg.enable_grad) { m.err = "image_energy_grad: create the handle with enable_grad = 1"; return APVAE_ERR_STATE; }
    if (N <= 0 || !x_dev) { m.err = "image_energy_grad: empty batch"; return APVAE_ERR_INVALID; }
    return m.critic_grad_of_images(x_dev, N, d_dev, grad_dev, grad_sqnorm_dev, stream) ? APVAE_OK : APVAE_ERR_CUDA;
}
int apvae_mala(apvae_handle* h, float* z_io_dev, int n, int steps, float alpha, const float* eps_dev, const float* u_dev,
               uint64_t seed, float* hist_e_dev, float* hist_pure_e_dev, float* hist_ratio_dev, float* hist_z_dev, void* stream) {
    REQUIRE_BUILT(h);
    if (!m.cfg.enable_grad) { m.err = "mala: create the handle with enable_grad = 1"; return APVAE_ERR_STATE; }
    if (n <= 0 || steps < 0 || !z_io_dev) { m.err = "mala: empty batch"; return APVAE_ERR_INVALID; }
    if (!(alpha > 0.f)) { m.err = "mala: alpha must be positive"; return APVAE_ERR_INVALID; }
    return m.mala(z_io_dev, n, steps, alpha, eps_dev, u_dev, seed, hist_e_dev, hist_pure_e_dev, hist_ratio_dev, hist_z_dev, stream)
               ? APVAE_OK : APVAE_ERR_CUDA;
}

}  // extern "C"
