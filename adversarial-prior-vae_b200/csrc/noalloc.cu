// apvae_is_ll_noalloc: the estimator tail (element log-probs -> per-sample log p(x|z) -> per-image IS / ELBO rows) for
// callers that cannot allocate (CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// The tail's only scratch — the per-sample row log p(x|z) [K * B] — lives in a file-scope device array of fixed size
// instead of the handle's cudaMalloc'ed arena.  Every GPU that loads this library holds the array (256 KiB) until
// the process exits.  In host code the array's NAME is only a host-side stand-in, not a device address: the address a
// kernel can use comes from cudaGetSymbolAddress, asked once per device with that device current (below).
//
// Every call shares the one array per device and it keeps the last call's contents: a call overwrites elements
// [0, K * B) before it reads them and never reads beyond, so nothing has to start at zero — but two calls that overlap
// on the same GPU (different streams, or graph branches without an edge between them) race on it.
#include <cuda_runtime.h>
#include <math.h>
#include <string.h>
#include <mutex>
#include <string>
#include "../../include/apvae.h"
#include "kernels.h"

namespace apvae { int report_global_error(const std::string& m, int code); }   // capi.cu

using namespace apvae;

namespace {

constexpr int kMaxDevices = 64;
static_assert(APVAE_NOALLOC_MAX_SAMPLES == 65536, "scratch size is part of the ABI documentation");

// per-sample log p(x|z) of the call in flight (float32 [APVAE_NOALLOC_MAX_SAMPLES])
static __device__ __align__(256) float g_noalloc_log_px[APVAE_NOALLOC_MAX_SAMPLES];

std::mutex g_mu;
float* g_scratch_addr[kMaxDevices] = {};        // device address of g_noalloc_log_px on each device, 0 = not asked yet

// Device address of the scratch array on the CURRENT device (asked from the runtime once per device).
int scratch_for_current_device(float** out, std::string& why) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { why = std::string("cudaGetDevice: ") + cudaGetErrorString(e); return (int)e; }
    if (dev < 0 || dev >= kMaxDevices) { why = "device ordinal beyond the scratch table"; return (int)cudaErrorInvalidDevice; }
    std::lock_guard<std::mutex> l(g_mu);
    if (!g_scratch_addr[dev]) {
        float* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_noalloc_log_px);
        if (e != cudaSuccess || !p) {
            why = std::string("cudaGetSymbolAddress(scratch): ") + cudaGetErrorString(e);
            return e != cudaSuccess ? (int)e : (int)cudaErrorInvalidSymbol;
        }
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return 0;
}

int cuda_status(int e) {
    return e == (int)cudaErrorNoDevice || e == (int)cudaErrorInsufficientDriver ? APVAE_ERR_NO_DEVICE : APVAE_ERR_CUDA;
}

}  // namespace

extern "C" int apvae_is_ll_noalloc(const float* x_dev, const float* mean_dev, const float* log_scale_dev, int64_t group,
                                   float bin_size, int has_min, float min_val, int has_max, float max_val, int biased_edges,
                                   float epsilon, const float* log_q_dev, const float* log_pn_dev, const float* d_g_dev,
                                   int K, int B, int z_dim, float w_xi, float log_Z, float* per_image_dev, void* stream) {
    if (!x_dev || !mean_dev || !log_scale_dev || !log_q_dev || !log_pn_dev || !d_g_dev || !per_image_dev || K <= 0 || B <= 0 ||
        group <= 0 || !(bin_size > 0.f))
        return report_global_error("is_ll_noalloc: bad arguments", APVAE_ERR_INVALID);
    const int64_t rows = (int64_t)K * B;
    if (rows > APVAE_NOALLOC_MAX_SAMPLES)
        return report_global_error("is_ll_noalloc: K * B = " + std::to_string(rows) + " exceeds APVAE_NOALLOC_MAX_SAMPLES = " +
                                   std::to_string((int)APVAE_NOALLOC_MAX_SAMPLES) + " (the fixed device scratch); split the batch or use "
                                   "apvae_discretized_logistic_log_prob + apvae_is_reduce with a buffer of your own", APVAE_ERR_INVALID);
    float* log_px = nullptr;
    std::string why;
    int e = scratch_for_current_device(&log_px, why);
    if (e != 0) return report_global_error("is_ll_noalloc: " + why, cuda_status(e));
    LikelihoodSpec s;
    memset(&s, 0, sizeof(s));
    s.bernoulli = 0; s.bin_size = bin_size; s.log_bin_size = logf(bin_size);
    s.has_min = has_min ? 1 : 0; s.has_max = has_max ? 1 : 0; s.min_val = min_val; s.max_val = max_val;
    s.biased_edges = biased_edges ? 1 : 0; s.epsilon = epsilon; s.min_log_scale = -INFINITY;
    e = launch_elementwise_logprob(x_dev, mean_dev, log_scale_dev, rows, group, s, log_px, stream);
    if (e == 0) e = launch_is_reduce(log_px, log_q_dev, log_pn_dev, d_g_dev, nullptr, K, B, z_dim, w_xi, log_Z, per_image_dev, stream);
    return e == 0 ? APVAE_OK : report_global_error(std::string("is_ll_noalloc: ") + cudaGetErrorString((cudaError_t)e), cuda_status(e));
}
