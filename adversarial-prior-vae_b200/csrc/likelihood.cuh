// Per-element log-likelihood of the p(x|z) node, shared by the fused heads kernel (heads.cu) and the stand-alone
// kernels behind apvae_discretized_logistic_log_prob / apvae_bernoulli_log_prob (elementwise.cu).
//
// Reference: tfsnippet/distributions/discretized.py:198-259 (DiscretizedLogistic.log_prob) with the constructor
// arguments of :26-58; Bernoulli: zd.Bernoulli via distributions/univariate.py:65-92 on
// logits = log(p) - log1p(-p), p = clip(x_mean, 1e-7, 1 - 1e-7) (dmm_bernouli_fashion.py:619-626).
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace apvae {

struct LikelihoodSpec {
    int bernoulli;          // 1: Bernoulli on the clipped mean; 0: DiscretizedLogistic
    float bin_size;         // DiscretizedLogistic(bin_size=...)
    float log_bin_size;     // log(bin_size), filled by the host
    int has_min, has_max;   // min_val / max_val given (the edge-bin cases exist only then)
    float min_val, max_val;
    int biased_edges;
    float epsilon;          // threshold of the "very small cdf_delta" branch
    float min_log_scale;    // maybe_clip_value(log_scale, min_val=config.epsilon) (dmm_cifar10.py:614-615); -INFINITY: no clip
};

// the p_net x node of the CIFAR / CelebA scripts (dmm_cifar10.py:612-619)
inline LikelihoodSpec script_likelihood(int bernoulli, float min_log_scale) {
    LikelihoodSpec s;
    s.bernoulli = bernoulli;
    s.bin_size = 2.0f / 256.0f;
    s.log_bin_size = logf(s.bin_size);
    s.has_min = s.has_max = 1;
    s.min_val = -1.0f + 1.0f / 256.0f;
    s.max_val = 1.0f - 1.0f / 256.0f;
    s.biased_edges = 1;
    s.epsilon = 1e-10f;
    s.min_log_scale = min_log_scale;
    return s;
}

#ifdef __CUDACC__
namespace lik {
__device__ __forceinline__ float softplus(float x) { return fmaxf(x, 0.f) + log1pf(expf(-fabsf(x))); }
__device__ __forceinline__ float sigmoid(float x) {
    const float e = expf(-fabsf(x));
    return x >= 0.f ? 1.f / (1.f + e) : e / (1.f + e);
}
}  // namespace lik

// sigmoid with the hardware exponential / reciprocal (ex2.approx, rcp.approx): absolute error ~3e-7, the same class as
// the float32 rounding of a value near 1/2 — log(sigmoid(hi) - sigmoid(lo)) is conditioned by that absolute error
// whichever exponential is used (tests/test_kats_gpu.py states the bound)
__device__ __forceinline__ float fast_sigmoid(float x) {
    const float e = __expf(-fabsf(x));
    const float r = __fdividef(1.f, 1.f + e);
    return x >= 0.f ? r : e * r;
}

// log p(x | mean, log_scale) of one element.  The common case (a middle bin with a representable cdf difference) is
// straight-line code on the hardware transcendentals; the rare cases (edge bins: 2 of 256 pixel values; vanishing cdf
// difference) are real branches on the accurate library functions.
__device__ __forceinline__ float element_log_prob(const LikelihoodSpec& s, float x, float mean, float log_scale) {
    if (s.bernoulli) {
        const float pm = fminf(fmaxf(mean, 1e-7f), 1.0f - 1e-7f);
        const float logits = logf(pm) - log1pf(-pm);
        return x * logits - lik::softplus(logits);        // -sigmoid_cross_entropy_with_logits(labels=x, logits)
    }
    const float ls = fmaxf(log_scale, s.min_log_scale);
    const float inv = expf(-ls);
    const float half_bin = 0.5f * s.bin_size;
    const float hd = half_bin * inv;
    const float xm = (x - mean) * inv;
    const float lo = xm - hd, hi = xm + hd;
    const float delta = fast_sigmoid(hi) - fast_sigmoid(lo);
    float lp = __logf(fmaxf(delta, 1e-12f));
    if (!(delta > s.epsilon)) lp = xm + (s.log_bin_size - ls) - 2.f * lik::softplus(xm);
    if (s.biased_edges) {
        if (s.has_min && x < s.min_val + half_bin) lp = -lik::softplus(-hi);
        if (s.has_max && x >= s.max_val - half_bin) lp = -lik::softplus(lo);
    }
    return lp;
}

// squared error of the rounded pixels, the summand of test_mse (dmm_cifar10.py:889-891:
// (round(mean * 128 + 127.5) - round(x * 128 + 127.5))^2; arch M, dmm_bernouli_fashion.py:893-896: * 255, on the clipped
// Bernoulli mean).  tf.round rounds half to even = rintf.
__device__ __forceinline__ float element_sq_err(int bernoulli, float x, float mean) {
    float a, b;
    if (bernoulli) {
        const float pm = fminf(fmaxf(mean, 1e-7f), 1.0f - 1e-7f);
        // distribution.mean of zd.Bernoulli(logits) = sigmoid(log p - log1p(-p)) = p up to rounding
        a = rintf(pm * 255.0f); b = rintf(x * 255.0f);
    } else {
        a = rintf(mean * 128.0f + 127.5f); b = rintf(x * 128.0f + 127.5f);
    }
    return (a - b) * (a - b);
}
#endif

}  // namespace apvae
