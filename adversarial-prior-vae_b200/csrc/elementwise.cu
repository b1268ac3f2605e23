// HBM-bound fused kernels: estimator tail, RNG, MALA update.  See kernels.h.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <math.h>
#include "kernels.h"
#include "likelihood.cuh"
#include "op16.h"

namespace apvae {
namespace {

constexpr float kLog2Pi = 1.8378770664093453f;
constexpr float kLogTruncArea = -0.04656529f;   // log(0.95449974), truncated_normal.py:146

__device__ __forceinline__ float lrelu(float v, float a) { return v > 0.f ? v : a * v; }
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// block-wide sum / max for blockDim.x <= 1024; result valid in every thread
__device__ float block_sum(float v) {
    __shared__ float sh[33];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    if (w == 0) {
        float t = (lane < (int)((blockDim.x + 31) >> 5)) ? sh[lane] : 0.f;
        t = warp_sum(t);
        if (lane == 0) sh[32] = t;
    }
    __syncthreads();
    return sh[32];
}
__device__ float block_max(float v) {
    __shared__ float sh[33];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    if (w == 0) {
        float t = (lane < (int)((blockDim.x + 31) >> 5)) ? sh[lane] : -INFINITY;
        t = warp_max(t);
        if (lane == 0) sh[32] = t;
    }
    __syncthreads();
    return sh[32];
}

__global__ void pack_image_kernel(const float* __restrict__ x, long long pixels, int c, int cp,
                                  uint16_t* __restrict__ raw, uint16_t* __restrict__ act, float alpha,
                                  uint16_t* __restrict__ raw_lo, uint16_t* __restrict__ act_lo, int f16) {
    const long long total = pixels * cp;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long p = i / cp;
        const int ch = (int)(i - p * cp);
        const float v = ch < c ? x[p * c + ch] : 0.f;
        if (raw) raw[i] = op16::from_f32(v, f16);
        if (act) act[i] = op16::from_f32(lrelu(v, alpha), f16);
        if (raw_lo) raw_lo[i] = op16::from_f32(v - op16::round(v, f16), f16);
        if (act_lo) { const float a = lrelu(v, alpha); act_lo[i] = op16::from_f32(a - op16::round(a, f16), f16); }
    }
}

// one warp per sample
__global__ void sample_z_kernel(const float* __restrict__ mu, const float* __restrict__ logstd,
                                const float* __restrict__ eps, int S, int s0, int B, int d, float min_logstd,
                                int truncated, float* __restrict__ z_f32, uint16_t* __restrict__ z_raw,
                                uint16_t* __restrict__ z_act, long long zstride, int zc, int cpad,
                                float alpha, float* __restrict__ log_q, float* __restrict__ log_pn,
                                uint16_t* __restrict__ z_raw_lo, uint16_t* __restrict__ z_act_lo, int f16) {
    const int lane = threadIdx.x & 31;
    const int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (s >= S) return;
    const int b = (s0 + s) % B;
    float lq = 0.f, lp = 0.f;
    for (int j = lane; j < d; j += 32) {
        const float e = eps[(size_t)s * d + j];
        float z;
        if (mu != nullptr) {
            const float m = mu[(size_t)b * d + j];
            const float ls = fmaxf(logstd[(size_t)b * d + j], min_logstd);
            z = m + expf(ls) * e;
            // -0.5 log 2pi - logstd - 0.5 exp(-2 logstd) (z - mu)^2, with (z-mu)/std == eps up to rounding;
            // evaluate from z like the reference does.
            const float t = (z - m) * expf(-ls);
            lq += -0.5f * kLog2Pi - ls - 0.5f * t * t - (truncated ? kLogTruncArea : 0.f);
        } else {
            z = e;
        }
        lp += -0.5f * kLog2Pi - 0.5f * z * z;
        if (z_f32) z_f32[(size_t)s * d + j] = z;
        const size_t idx = (size_t)s * zstride + (size_t)(j / zc) * cpad + (j % zc);
        if (z_raw) z_raw[idx] = op16::from_f32(z, f16);
        if (z_act) z_act[idx] = op16::from_f32(lrelu(z, alpha), f16);
        if (z_raw_lo) z_raw_lo[idx] = op16::from_f32(z - op16::round(z, f16), f16);
        if (z_act_lo) { const float a = lrelu(z, alpha); z_act_lo[idx] = op16::from_f32(a - op16::round(a, f16), f16); }
    }
    lq = warp_sum(lq);
    lp = warp_sum(lp);
    if (lane == 0) {
        if (log_q && mu != nullptr) log_q[s] = lq;
        if (log_pn) log_pn[s] = lp;
    }
}

// one warp per sample, z held in registers (PL elements per lane) while the flow layers run
template <int PL>
__global__ void sample_z_flow_kernel(const float* __restrict__ mu, const float* __restrict__ logstd,
                                     const float* __restrict__ eps, int S, int s0, int B, int d, float min_logstd,
                                     int truncated, const float* __restrict__ flow_w, const float* __restrict__ flow_uhat,
                                     const float* __restrict__ flow_b, const float* __restrict__ flow_wu, int nf,
                                     float* __restrict__ z_f32, uint16_t* __restrict__ z_raw,
                                     uint16_t* __restrict__ z_act, long long zstride, int zc, int cpad,
                                     float alpha, float* __restrict__ log_q, float* __restrict__ log_pn,
                                     uint16_t* __restrict__ z_raw_lo, uint16_t* __restrict__ z_act_lo, int f16) {
    const int lane = threadIdx.x & 31;
    const int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (s >= S) return;
    const int b = (s0 + s) % B;
    float z[PL];
    float lq = 0.f;
#pragma unroll
    for (int i = 0; i < PL; ++i) {
        const int j = lane + 32 * i;
        z[i] = 0.f;
        if (j < d) {
            const float e = eps[(size_t)s * d + j];
            const float m = mu[(size_t)b * d + j];
            const float ls = fmaxf(logstd[(size_t)b * d + j], min_logstd);
            z[i] = m + expf(ls) * e;
            const float t = (z[i] - m) * expf(-ls);
            lq += -0.5f * kLog2Pi - ls - 0.5f * t * t - (truncated ? kLogTruncArea : 0.f);
        }
    }
    float logdet = 0.f;
    for (int l = 0; l < nf; ++l) {
        const float* w = flow_w + (size_t)l * d;
        const float* uh = flow_uhat + (size_t)l * d;
        float dot = 0.f;
#pragma unroll
        for (int i = 0; i < PL; ++i) {
            const int j = lane + 32 * i;
            if (j < d) dot += __ldg(w + j) * z[i];
        }
        dot = warp_sum(dot);                                   // every lane holds the total
        const float t = tanhf(dot + __ldg(flow_b + l));
#pragma unroll
        for (int i = 0; i < PL; ++i) {
            const int j = lane + 32 * i;
            if (j < d) z[i] += __ldg(uh + j) * t;
        }
        logdet += logf(fabsf(1.f + (1.f - t * t) * __ldg(flow_wu + l)));
    }
    float lp = 0.f;
#pragma unroll
    for (int i = 0; i < PL; ++i) {
        const int j = lane + 32 * i;
        if (j < d) {
            const float v = z[i];
            lp += -0.5f * kLog2Pi - 0.5f * v * v;
            if (z_f32) z_f32[(size_t)s * d + j] = v;
            const size_t idx = (size_t)s * zstride + (size_t)(j / zc) * cpad + (j % zc);
            if (z_raw) z_raw[idx] = op16::from_f32(v, f16);
            if (z_act) z_act[idx] = op16::from_f32(lrelu(v, alpha), f16);
            if (z_raw_lo) z_raw_lo[idx] = op16::from_f32(v - op16::round(v, f16), f16);
            if (z_act_lo) { const float a = lrelu(v, alpha); z_act_lo[idx] = op16::from_f32(a - op16::round(a, f16), f16); }
        }
    }
    lq = warp_sum(lq);
    lp = warp_sum(lp);
    if (lane == 0) {
        if (log_q) log_q[s] = lq - logdet;
        if (log_pn) log_pn[s] = lp;
    }
}

// one block per sample: the unfused likelihood over the fp32 heads (kept for head layouts the fused heads kernel does not run)
__global__ void pixel_logprob_kernel(const float* __restrict__ head, int ld, const float* __restrict__ x,
                                     int S, int s0, int B, int HW, int C, const LikelihoodSpec spec, int uniform_scale,
                                     float beta, float* __restrict__ log_px, float* __restrict__ sq_err) {
    const int s = blockIdx.x;
    const int b = (s0 + s) % B;
    const int n_x = HW * C;
    float acc = 0.f, se = 0.f;
    for (int i = threadIdx.x; i < n_x; i += blockDim.x) {
        const int p = i / C, c = i - p * C;
        const float* row = head + ((size_t)s * HW + p) * ld;
        const float xv = x[(size_t)b * n_x + i];
        const float mean = row[c];
        const float ls = (spec.bernoulli || uniform_scale) ? beta : row[C + c];
        acc += element_log_prob(spec, xv, mean, ls);
        se += element_sq_err(spec.bernoulli, xv, mean);
    }
    acc = block_sum(acc);
    se = block_sum(se);
    if (threadIdx.x == 0) { log_px[s] = acc; if (sq_err) sq_err[s] = se; }
}

// Stand-alone log_prob of the two likelihood families on plain arrays (apvae_discretized_logistic_log_prob /
// apvae_bernoulli_log_prob): out[r] = sum over `group` consecutive elements; one warp per row.
__global__ void elementwise_logprob_kernel(const float* __restrict__ x, const float* __restrict__ mean,
                                           const float* __restrict__ log_scale, long long rows, long long group,
                                           const LikelihoodSpec spec, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const long long r = blockIdx.x * (long long)(blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= rows) return;
    float acc = 0.f;
    for (long long j = lane; j < group; j += 32) {
        const long long i = r * group + j;
        acc += element_log_prob(spec, x[i], mean[i], log_scale ? log_scale[i] : 0.f);
    }
    acc = warp_sum(acc);
    if (lane == 0) out[r] = acc;
}

__global__ void rowdot_kernel(const float* __restrict__ h, int ld, const float* __restrict__ w,
                              const float* __restrict__ b, int n, long long S, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const long long s = blockIdx.x * (long long)(blockDim.x >> 5) + (threadIdx.x >> 5);
    if (s >= S) return;
    float acc = 0.f;
    for (int j = lane; j < n; j += 32) acc += h[s * ld + j] * w[j];
    acc = warp_sum(acc);
    if (lane == 0) out[s] = acc + b[0];
}

// one block per image
__global__ void is_reduce_kernel(const float* __restrict__ log_px, const float* __restrict__ log_q,
                                 const float* __restrict__ log_pn, const float* __restrict__ d_g,
                                 const float* __restrict__ sq_err, int K, int B,
                                 int d, float w_xi, float log_Z, float* __restrict__ per_image) {
    const int b = blockIdx.x;
    const float lz = fmaxf(log_Z, -20.f);
    float m_p = -INFINITY, m_a = -INFINITY, m_z = -INFINITY;
    float s_p = 0.f, s_a = 0.f, s_px = 0.f, s_dg = 0.f;
    float min_se = INFINITY;
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
        const size_t s = (size_t)k * B + b;
        if (sq_err) min_se = fminf(min_se, sq_err[s]);
        const float lpx = log_px[s], lq = log_q[s], lpn = log_pn[s], dg = d_g[s];
        const float half_z2 = -lpn - 0.5f * d * kLog2Pi;
        const float energy = w_xi * dg + half_z2;
        const float lp = lpx + lpn - lq;
        const float la = lpx - energy - lz - lq;
        m_p = fmaxf(m_p, lp);  m_a = fmaxf(m_a, la);  m_z = fmaxf(m_z, -energy - lq);
        s_p += lp;  s_a += la;  s_px += lpx;  s_dg += dg;
    }
    m_p = block_max(m_p);  m_a = block_max(m_a);  m_z = block_max(m_z);
    min_se = -block_max(-min_se);
    float e_p = 0.f, e_a = 0.f, e_z = 0.f;
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
        const size_t s = (size_t)k * B + b;
        const float lpx = log_px[s], lq = log_q[s], lpn = log_pn[s], dg = d_g[s];
        const float half_z2 = -lpn - 0.5f * d * kLog2Pi;
        const float energy = w_xi * dg + half_z2;
        e_p += expf(lpx + lpn - lq - m_p);
        e_a += expf(lpx - energy - lz - lq - m_a);
        e_z += expf(-energy - lq - m_z);
    }
    e_p = block_sum(e_p);  e_a = block_sum(e_a);  e_z = block_sum(e_z);
    s_p = block_sum(s_p);  s_a = block_sum(s_a);  s_px = block_sum(s_px);  s_dg = block_sum(s_dg);
    if (threadIdx.x == 0) {
        const float invK = 1.f / K;
        per_image[ROW_LL * B + b] = m_p + logf(e_p * invK);
        per_image[ROW_LB * B + b] = s_p * invK;
        per_image[ROW_ADV_LL * B + b] = m_a + logf(e_a * invK);
        per_image[ROW_ADV_LB * B + b] = s_a * invK;
        per_image[ROW_RECON * B + b] = s_px * invK;
        per_image[ROW_DG * B + b] = s_dg * invK;
        per_image[ROW_LZ_MAX * B + b] = m_z;
        per_image[ROW_LZ_SUM * B + b] = e_z;
        per_image[ROW_MSE * B + b] = sq_err ? min_se : 0.f;      // test_mse: min over the K samples (dmm_cifar10.py:892)
    }
}

}  // namespace

namespace {

struct Philox {
    uint32_t c[4];
    uint32_t k[2];
};
__device__ __forceinline__ void philox_round(Philox& s) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    const uint32_t hi0 = __umulhi(M0, s.c[0]), lo0 = M0 * s.c[0];
    const uint32_t hi1 = __umulhi(M1, s.c[2]), lo1 = M1 * s.c[2];
    const uint32_t n0 = hi1 ^ s.c[1] ^ s.k[0], n1 = lo1, n2 = hi0 ^ s.c[3] ^ s.k[1], n3 = lo0;
    s.c[0] = n0; s.c[1] = n1; s.c[2] = n2; s.c[3] = n3;
    s.k[0] += 0x9E3779B9u; s.k[1] += 0xBB67AE85u;
}
__device__ __forceinline__ void philox4x32_10(unsigned long long seed, unsigned long long ctr_lo,
                                              unsigned long long ctr_hi, uint32_t (&out)[4]) {
    Philox s;
    s.c[0] = (uint32_t)ctr_lo; s.c[1] = (uint32_t)(ctr_lo >> 32);
    s.c[2] = (uint32_t)ctr_hi; s.c[3] = (uint32_t)(ctr_hi >> 32);
    s.k[0] = (uint32_t)seed; s.k[1] = (uint32_t)(seed >> 32);
#pragma unroll
    for (int i = 0; i < 10; ++i) philox_round(s);
    out[0] = s.c[0]; out[1] = s.c[1]; out[2] = s.c[2]; out[3] = s.c[3];
}
__device__ __forceinline__ float u01(uint32_t r) { return ((float)(r >> 8) + 0.5f) * (1.0f / 16777216.0f); }
__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float& n0, float& n1) {
    const float r = sqrtf(-2.f * logf(u01(a)));
    float sn, cs;
    sincospif(2.f * u01(b), &sn, &cs);
    n0 = r * cs; n1 = r * sn;
}

__global__ void fill_normal_kernel(float* __restrict__ out, long long n, unsigned long long seed,
                                   unsigned long long offset, int truncated) {
    const long long quads = (n + 3) / 4;
    for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < quads;
         q += (long long)gridDim.x * blockDim.x) {
        uint32_t r[4];
        float v[4];
        philox4x32_10(seed, (unsigned long long)q + offset, 0ull, r);
        box_muller(r[0], r[1], v[0], v[1]);
        box_muller(r[2], r[3], v[2], v[3]);
        if (truncated) {
            for (int i = 0; i < 4; ++i) {
                unsigned long long attempt = 1;
                while (fabsf(v[i]) > 2.f) {
                    uint32_t t[4];
                    float a, b;
                    philox4x32_10(seed, (unsigned long long)q + offset, attempt * 4ull + i, t);
                    box_muller(t[0], t[1], a, b);
                    v[i] = fabsf(a) <= 2.f ? a : b;
                    ++attempt;
                }
            }
        }
        for (int i = 0; i < 4; ++i)
            if (q * 4 + i < n) out[q * 4 + i] = v[i];
    }
}
__global__ void fill_uniform_kernel(float* __restrict__ out, long long n, unsigned long long seed,
                                    unsigned long long offset) {
    const long long quads = (n + 3) / 4;
    for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < quads;
         q += (long long)gridDim.x * blockDim.x) {
        uint32_t r[4];
        philox4x32_10(seed, (unsigned long long)q + offset, 1ull << 40, r);
        for (int i = 0; i < 4; ++i)
            if (q * 4 + i < n) out[q * 4 + i] = u01(r[i]);
    }
}

// one warp per chain
__global__ void mala_propose_kernel(const float* __restrict__ z, const float* __restrict__ g,
                                    const float* __restrict__ eps, int n, int d, float alpha,
                                    float* __restrict__ zp) {
    const int lane = threadIdx.x & 31;
    const int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (s >= n) return;
    const float sc = sqrtf(2.f * alpha);
    for (int j = lane; j < d; j += 32) {
        const size_t i = (size_t)s * d + j;
        const float v = z[i] - alpha * g[i] + sc * eps[i];
        zp[i] = v;
    }
}

__global__ void mala_accept_kernel(float* __restrict__ z, float* __restrict__ e, float* __restrict__ g,
                                   float* __restrict__ pure, const float* __restrict__ zp,
                                   const float* __restrict__ ep, const float* __restrict__ gp,
                                   const float* __restrict__ purep, const float* __restrict__ u, int n, int d,
                                   float alpha, float* __restrict__ hist_e, float* __restrict__ hist_pure,
                                   float* __restrict__ hist_ratio) {
    const int lane = threadIdx.x & 31;
    const int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (s >= n) return;
    float fwd = 0.f, bwd = 0.f;
    for (int j = lane; j < d; j += 32) {
        const size_t i = (size_t)s * d + j;
        const float a = zp[i] - z[i] + alpha * g[i];
        const float b = z[i] - zp[i] + alpha * gp[i];
        fwd += a * a;  bwd += b * b;
    }
    fwd = warp_sum(fwd);  bwd = warp_sum(bwd);
    const float log_q_zprime_z = fwd * (-1.f / (4.f * alpha));
    const float log_q_z_zprime = bwd * (-1.f / (4.f * alpha));
    const float e0 = e[s], e1 = ep[s];
    const float ratio = fminf(fmaxf(expf((-e1 + e0) + (log_q_z_zprime - log_q_zprime_z)), 0.f), 1.f);
    const bool accept = u[s] < ratio;
    if (lane == 0) {
        if (hist_e) hist_e[s] = e0;
        if (hist_pure) hist_pure[s] = pure[s];
        if (hist_ratio) hist_ratio[s] = ratio;
    }
    __syncwarp();
    if (accept) {
        for (int j = lane; j < d; j += 32) {
            const size_t i = (size_t)s * d + j;
            z[i] = zp[i];  g[i] = gp[i];
        }
        if (lane == 0) { e[s] = e1; pure[s] = purep[s]; }
    }
}

__global__ void energy_combine_kernel(const float* __restrict__ d_g, const float* __restrict__ gz,
                                      const float* __restrict__ z, int n, int d, float w_xi,
                                      float* __restrict__ energy, float* __restrict__ grad) {
    const int lane = threadIdx.x & 31;
    const int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (s >= n) return;
    float ss = 0.f;
    for (int j = lane; j < d; j += 32) {
        const size_t i = (size_t)s * d + j;
        const float v = z[i];
        ss += v * v;
        if (grad) grad[i] = w_xi * gz[i] + v;
    }
    ss = warp_sum(ss);
    if (lane == 0 && energy) energy[s] = w_xi * d_g[s] + 0.5f * ss;
}

__global__ void tanh_backward_pack_kernel(const float* __restrict__ g, const float* __restrict__ head, int ld,
                                          long long pixels, int C, int cp, uint16_t* __restrict__ out, int f16) {
    const long long total = pixels * cp;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long p = i / cp;
        const int c = (int)(i - p * cp);
        float v = 0.f;
        if (c < C) {
            const float t = head[p * ld + c];
            v = g[p * C + c] * (1.f - t * t);
        }
        out[i] = op16::from_f32(v, f16);
    }
}

// one thread per (pixel, 8 output channels): 16-byte stores, the pixel's three gradients re-read through L1
template <bool F16>
__global__ void tanh_backward_head_kernel(const __grid_constant__ HeadGradParams p) {
    const int groups = p.cp_out >> 3;
    const long long total = p.pixels * groups;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long px = i / groups;
        const int c0 = (int)(i - px * groups) * 8;
        float gp[kHeadGradMaxC] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int c = 0; c < kHeadGradMaxC; ++c)
            if (c < p.C) { const float t = p.head[px * p.ld + c]; gp[c] = p.g[px * p.C + c] * (1.f - t * t); }
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int co = c0 + j;
            float a = 0.f;
            if (co < p.cg) {
                const float* w = p.w + co * kHeadGradMaxC;
                a = gp[0] * w[0] + gp[1] * w[1] + gp[2] * w[2] + gp[3] * w[3];
            }
            v[j] = a;
        }
        uint4* d = reinterpret_cast<uint4*>(static_cast<uint16_t*>(p.out) + px * p.cp_out + c0);
        *d = make_uint4(op16::pack<F16>(v[0], v[1]), op16::pack<F16>(v[2], v[3]), op16::pack<F16>(v[4], v[5]), op16::pack<F16>(v[6], v[7]));
    }
}

// one block per row
__global__ void row_sqnorm_kernel(const float* __restrict__ v, long long len, float* __restrict__ out) {
    const float* p = v + (size_t)blockIdx.x * len;
    float s = 0.f;
    for (long long i = threadIdx.x; i < len; i += blockDim.x) { const float a = p[i]; s += a * a; }
    s = warp_sum(s);
    __shared__ float part[8];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        float t = threadIdx.x < (blockDim.x >> 5) ? part[threadIdx.x] : 0.f;
        t = warp_sum(t);
        if (threadIdx.x == 0) out[blockIdx.x] = t;
    }
}

__global__ void quantize_kernel(const float* __restrict__ head, int ld, long long pixels, int C,
                                uint8_t* __restrict__ out) {
    const long long total = pixels * C;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long p = i / C;
        const int c = (int)(i - p * C);
        const float v = fminf(fmaxf(256.0f * head[p * ld + c] / 2 + 127.5f, 0.f), 255.f);
        out[i] = (uint8_t)rintf(v);    // np.round = round half to even
    }
}

__global__ void clip_min_kernel(float* v, long long n, float lo) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x)
        v[i] = fmaxf(v[i], lo);
}

__global__ void leaky_relu_f32_kernel(float* v, long long n, float alpha) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x)
        v[i] = fmaxf(v[i], alpha * v[i]);
}

__global__ void gather_cols_kernel(const float* __restrict__ src, int ld, int col0, int cols, long long rows,
                                   float* __restrict__ dst) {
    const long long total = rows * cols;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / cols;
        dst[i] = src[r * ld + col0 + (int)(i - r * cols)];
    }
}

// One block.  accum (double[kAccumSlots]): [r] += sum_b per_image[r][b] for the additive rows (ROW_MSE goes to
// slot ACC_MSE); slots ROW_LZ_MAX / ROW_LZ_SUM hold a running (max, sum exp) over BATCHES of
// lz_b = log-mean-exp over the batch's [K,B] elements (dmm_cifar10.py:922-924,1221-1222);
// slot 9 += B (weight), slot 10 += 1 (batches).  `prior` (4 floats: pn_energy, pd_energy, kl_adv_and_gaussian,
// log_Z_compute_op of this batch's prior branch, or null): slots 12..15 += B * value — the Evaluator weights every
// batch scalar by the batch size (trainer/evaluator.py:249-254).
__global__ void accumulate_metrics_kernel(const float* __restrict__ per_image, int B, int K, int rows,
                                          const float* __restrict__ prior, double* __restrict__ accum) {
    for (int r = 0; r < rows; ++r) {
        if (r == ROW_LZ_MAX || r == ROW_LZ_SUM) continue;
        float acc = 0.f;
        for (int b = threadIdx.x; b < B; b += blockDim.x) acc += per_image[(size_t)r * B + b];
        acc = block_sum(acc);
        if (threadIdx.x == 0) accum[r == ROW_MSE ? ACC_MSE : r] += (double)acc;
    }
    float mx = -INFINITY;
    for (int b = threadIdx.x; b < B; b += blockDim.x) mx = fmaxf(mx, per_image[(size_t)ROW_LZ_MAX * B + b]);
    mx = block_max(mx);
    float sm = 0.f;
    for (int b = threadIdx.x; b < B; b += blockDim.x)
        sm += per_image[(size_t)ROW_LZ_SUM * B + b] * expf(per_image[(size_t)ROW_LZ_MAX * B + b] - mx);
    sm = block_sum(sm);
    if (threadIdx.x == 0) {
        const double lz_b = (double)mx + log((double)sm / ((double)K * (double)B));
        const double nb = accum[ACC_BATCHES];
        if (nb == 0.0) {
            accum[ROW_LZ_MAX] = lz_b;
            accum[ROW_LZ_SUM] = 1.0;
        } else {
            const double m0 = accum[ROW_LZ_MAX], m1 = fmax(m0, lz_b);
            accum[ROW_LZ_SUM] = accum[ROW_LZ_SUM] * exp(m0 - m1) + exp(lz_b - m1);
            accum[ROW_LZ_MAX] = m1;
        }
        accum[ACC_WEIGHT] += (double)B;
        accum[ACC_BATCHES] = nb + 1.0;
        if (prior) {
            for (int i = 0; i < 4; ++i) accum[ACC_PN + i] += (double)B * (double)prior[i];
            accum[ACC_PRIOR_WEIGHT] += (double)B;
        }
    }
}

// The test_pn_net branch's batch scalars from Kp prior draws (dmm_cifar10.py:883,915-927), one block:
//   out[0] pn_energy = mean D(G(z));  out[1] pd_energy = mean[D(G(z)) exp(log_energy_prob - log p_N(z))];
//   out[2] kl_adv_and_gaussian = mean[log p_N(z) - log_energy_prob];  out[3] log_Z_compute_op = log-mean-exp(-energy - log p_N(z))
// with energy = w_xi D + |z|^2 / 2 and log_energy_prob = -energy - max(log_Z, -20) (:164-177), fp32 like the reference.
__global__ void prior_reduce_kernel(const float* __restrict__ d_g, const float* __restrict__ log_pn, int Kp, int d,
                                    float w_xi, float log_Z, float* __restrict__ out) {
    const float lz = fmaxf(log_Z, -20.f);
    float s_pn = 0.f, s_pd = 0.f, s_kl = 0.f, mx = -INFINITY;
    for (int k = threadIdx.x; k < Kp; k += blockDim.x) {
        const float dg = d_g[k], lpn = log_pn[k];
        const float energy = w_xi * dg + (-lpn - 0.5f * d * kLog2Pi);
        const float lep = -energy - lz;
        s_pn += dg;
        s_pd += dg * expf(lep - lpn);
        s_kl += lpn - lep;
        mx = fmaxf(mx, -energy - lpn);
    }
    mx = block_max(mx);
    float se = 0.f;
    for (int k = threadIdx.x; k < Kp; k += blockDim.x) {
        const float dg = d_g[k], lpn = log_pn[k];
        const float energy = w_xi * dg + (-lpn - 0.5f * d * kLog2Pi);
        se += expf(-energy - lpn - mx);
    }
    s_pn = block_sum(s_pn);  s_pd = block_sum(s_pd);  s_kl = block_sum(s_kl);  se = block_sum(se);
    if (threadIdx.x == 0) {
        const float inv = 1.f / Kp;
        out[0] = s_pn * inv;
        out[1] = s_pd * inv;
        out[2] = s_kl * inv;
        out[3] = mx + logf(se * inv);
    }
}

inline int grid_for(long long n, int block, int cap = 148 * 16) {
    long long g = (n + block - 1) / block;
    if (g < 1) g = 1;
    return (int)(g < cap ? g : cap);
}
inline cudaStream_t st(void* s) { return static_cast<cudaStream_t>(s); }

}  // namespace

int launch_pack_image(const float* x, long long pixels, int c, int cp, void* out_raw, void* out_act,
                      float alpha, int f16, void* stream, void* raw_lo, void* act_lo) {
    if (pixels <= 0) return 0;
    pack_image_kernel<<<grid_for(pixels * cp, 256), 256, 0, st(stream)>>>(
        x, pixels, c, cp, (uint16_t*)out_raw, (uint16_t*)out_act, alpha, (uint16_t*)raw_lo, (uint16_t*)act_lo, f16);
    return (int)cudaGetLastError();
}

int launch_sample_z(const float* mu, const float* logstd, const float* eps, int S, int s0, int B, int d,
                    float min_logstd, int truncated, float* z_f32, void* z_raw, void* z_act,
                    long long zstride, int zc, int cpad, float alpha, float* log_q, float* log_pn,
                    int f16, void* stream, void* z_raw_lo, void* z_act_lo) {
    if (S <= 0) return 0;
    sample_z_kernel<<<(S + 7) / 8, 256, 0, st(stream)>>>(mu, logstd, eps, S, s0, B, d, min_logstd, truncated,
                                                        z_f32, (uint16_t*)z_raw, (uint16_t*)z_act,
                                                        zstride, zc, cpad, alpha, log_q, log_pn,
                                                        (uint16_t*)z_raw_lo, (uint16_t*)z_act_lo, f16);
    return (int)cudaGetLastError();
}

int launch_sample_z_flow(const float* mu, const float* logstd, const float* eps, int S, int s0, int B, int d,
                         float min_logstd, int truncated, const float* flow_w, const float* flow_uhat,
                         const float* flow_b, const float* flow_wu, int nf, float* z_f32, void* z_raw, void* z_act,
                         long long zstride, int zc, int cpad, float alpha, float* log_q, float* log_pn, int f16, void* stream,
                         void* z_raw_lo, void* z_act_lo) {
    if (S <= 0) return 0;
    if (d > 512 || !mu || !logstd) return (int)cudaErrorInvalidValue;
#define APVAE_FLOW_LAUNCH(PL) \
    sample_z_flow_kernel<PL><<<(S + 7) / 8, 256, 0, st(stream)>>>(mu, logstd, eps, S, s0, B, d, min_logstd, truncated, \
        flow_w, flow_uhat, flow_b, flow_wu, nf, z_f32, (uint16_t*)z_raw, (uint16_t*)z_act, zstride, zc, cpad, \
        alpha, log_q, log_pn, (uint16_t*)z_raw_lo, (uint16_t*)z_act_lo, f16)
    if (d <= 32) APVAE_FLOW_LAUNCH(1);
    else if (d <= 64) APVAE_FLOW_LAUNCH(2);
    else if (d <= 128) APVAE_FLOW_LAUNCH(4);
    else if (d <= 256) APVAE_FLOW_LAUNCH(8);
    else APVAE_FLOW_LAUNCH(16);
#undef APVAE_FLOW_LAUNCH
    return (int)cudaGetLastError();
}

int launch_pixel_logprob(const float* head, int ld, const float* x, int S, int s0, int B, int HW, int C,
                         const LikelihoodSpec& spec, int uniform_scale, float beta, float* log_px, float* sq_err,
                         void* stream) {
    if (S <= 0) return 0;
    pixel_logprob_kernel<<<S, 256, 0, st(stream)>>>(head, ld, x, S, s0, B, HW, C, spec, uniform_scale, beta, log_px, sq_err);
    return (int)cudaGetLastError();
}

int launch_elementwise_logprob(const float* x, const float* mean, const float* log_scale, long long rows, long long group,
                               const LikelihoodSpec& spec, float* out, void* stream) {
    if (rows <= 0 || group <= 0) return 0;
    elementwise_logprob_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st(stream)>>>(x, mean, log_scale, rows, group, spec, out);
    return (int)cudaGetLastError();
}

int launch_rowdot(const float* h, int ld, const float* w, const float* b, int n, long long S, float* out,
                  void* stream) {
    if (S <= 0) return 0;
    rowdot_kernel<<<(unsigned)((S + 7) / 8), 256, 0, st(stream)>>>(h, ld, w, b, n, S, out);
    return (int)cudaGetLastError();
}

int launch_is_reduce(const float* log_px, const float* log_q, const float* log_pn, const float* d_g,
                     const float* sq_err, int K, int B, int d, float w_xi, float log_Z, float* per_image,
                     void* stream) {
    if (B <= 0) return 0;
    is_reduce_kernel<<<B, 256, 0, st(stream)>>>(log_px, log_q, log_pn, d_g, sq_err, K, B, d, w_xi, log_Z, per_image);
    return (int)cudaGetLastError();
}

int launch_fill_normal(float* out, long long n, unsigned long long seed, unsigned long long offset,
                       int truncated, void* stream) {
    if (n <= 0) return 0;
    fill_normal_kernel<<<grid_for((n + 3) / 4, 256), 256, 0, st(stream)>>>(out, n, seed, offset, truncated);
    return (int)cudaGetLastError();
}
int launch_fill_uniform(float* out, long long n, unsigned long long seed, unsigned long long offset,
                        void* stream) {
    if (n <= 0) return 0;
    fill_uniform_kernel<<<grid_for((n + 3) / 4, 256), 256, 0, st(stream)>>>(out, n, seed, offset);
    return (int)cudaGetLastError();
}

int launch_mala_propose(const float* z, const float* g, const float* eps, int n, int d, float alpha,
                        float* z_prime, void* stream) {
    if (n <= 0) return 0;
    mala_propose_kernel<<<(n + 7) / 8, 256, 0, st(stream)>>>(z, g, eps, n, d, alpha, z_prime);
    return (int)cudaGetLastError();
}
int launch_mala_accept(float* z, float* e, float* g, float* pure, const float* z_prime, const float* e_prime,
                       const float* g_prime, const float* pure_prime, const float* u, int n, int d,
                       float alpha, float* hist_e, float* hist_pure, float* hist_ratio, void* stream) {
    if (n <= 0) return 0;
    mala_accept_kernel<<<(n + 7) / 8, 256, 0, st(stream)>>>(z, e, g, pure, z_prime, e_prime, g_prime, pure_prime,
                                                           u, n, d, alpha, hist_e, hist_pure, hist_ratio);
    return (int)cudaGetLastError();
}
int launch_energy_combine(const float* d_g, const float* gz, const float* z, int n, int d, float w_xi,
                          float* energy, float* grad, void* stream) {
    if (n <= 0) return 0;
    energy_combine_kernel<<<(n + 7) / 8, 256, 0, st(stream)>>>(d_g, gz, z, n, d, w_xi, energy, grad);
    return (int)cudaGetLastError();
}
int launch_tanh_backward_pack(const float* g, const float* head, int ld, long long pixels, int C, int cp,
                              void* out_bf16, int f16, void* stream) {
    if (pixels <= 0) return 0;
    tanh_backward_pack_kernel<<<grid_for(pixels * cp, 256), 256, 0, st(stream)>>>(g, head, ld, pixels, C, cp,
                                                                               (uint16_t*)out_bf16, f16);
    return (int)cudaGetLastError();
}
int launch_tanh_backward_head(const HeadGradParams& p, void* stream) {
    if (p.pixels <= 0) return 0;
    if (p.C < 1 || p.C > kHeadGradMaxC || p.cg < 1 || p.cg > kHeadGradMaxCg || p.cp_out % 8 || p.cp_out < p.cg || !p.g || !p.head || !p.out)
        return (int)cudaErrorInvalidValue;
    const long long total = p.pixels * (p.cp_out >> 3);
    if (p.f16) tanh_backward_head_kernel<true><<<grid_for(total, 256), 256, 0, st(stream)>>>(p);
    else tanh_backward_head_kernel<false><<<grid_for(total, 256), 256, 0, st(stream)>>>(p);
    return (int)cudaGetLastError();
}
int launch_row_sqnorm(const float* v, long long len, int rows, float* out, void* stream) {
    if (rows <= 0) return 0;
    row_sqnorm_kernel<<<rows, 256, 0, st(stream)>>>(v, len, out);
    return (int)cudaGetLastError();
}
int launch_quantize_images(const float* head, int ld, long long pixels, int C, uint8_t* out, void* stream) {
    if (pixels <= 0) return 0;
    quantize_kernel<<<grid_for(pixels * C, 256), 256, 0, st(stream)>>>(head, ld, pixels, C, out);
    return (int)cudaGetLastError();
}
int launch_accumulate_metrics(const float* per_image, int B, int K, int rows, const float* prior, double* accum, void* stream) {
    if (B <= 0) return 0;
    accumulate_metrics_kernel<<<1, 256, 0, st(stream)>>>(per_image, B, K, rows, prior, accum);
    return (int)cudaGetLastError();
}
int launch_prior_reduce(const float* d_g, const float* log_pn, int Kp, int d, float w_xi, float log_Z, float* out4, void* stream) {
    if (Kp <= 0) return 0;
    prior_reduce_kernel<<<1, 256, 0, st(stream)>>>(d_g, log_pn, Kp, d, w_xi, log_Z, out4);
    return (int)cudaGetLastError();
}
int launch_clip_min(float* v, long long n, float lo, void* stream) {
    if (n <= 0) return 0;
    clip_min_kernel<<<grid_for(n, 256), 256, 0, st(stream)>>>(v, n, lo);
    return (int)cudaGetLastError();
}
int launch_leaky_relu_f32(float* v, long long n, float alpha, void* stream) {
    if (n <= 0) return 0;
    leaky_relu_f32_kernel<<<grid_for(n, 256), 256, 0, st(stream)>>>(v, n, alpha);
    return (int)cudaGetLastError();
}
int launch_gather_cols(const float* src, int ld, int col0, int cols, long long rows, float* dst, void* stream) {
    if (rows <= 0) return 0;
    gather_cols_kernel<<<grid_for(rows * cols, 256), 256, 0, st(stream)>>>(src, ld, col0, cols, rows, dst);
    return (int)cudaGetLastError();
}

}  // namespace apvae
