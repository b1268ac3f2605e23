// apvae_allreduce_metrics: the one collective of a sharded evaluation, on the C ABI.
//
// Reference: the reference evaluates on one device; its metric combine is the host-side np.average of
// trainer/evaluator.py:249-254 and the host logsumexp of the log-Z loop (dmm_cifar10.py:1221-1222).  Here every rank
// holds the accumulator vector apvae_accumulate_metrics filled for its image shard; ONE ncclAllGather of
// APVAE_ACCUM_SLOTS doubles per rank over NVLink, then every rank merges the R vectors on the device: additive slots are
// summed, the running (max, sum-exp) pair of the log-Z batches is merged by rescaling to the common max.  Same
// arithmetic as apvae/evaluation.py::merge_accumulators (the torch.distributed path of the Python host side).
//
// NCCL is bound at run time (dlopen of the libnccl.so.2 already in the process — the one that created the caller's
// ncclComm_t — or the system one), so the library has no link-time NCCL dependency and loads on hosts without it.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <math.h>
#include <mutex>
#include <string>
#include "../../include/apvae.h"
#include "kernels.h"

namespace apvae { int report_global_error(const std::string& m, int code); }   // capi.cu

using namespace apvae;

namespace {

// the slice of nccl.h this file needs (ABI-stable since NCCL 2.0: ncclResult_t 0 = success, ncclFloat64 = 8)
typedef int (*nccl_all_gather_fn)(const void* sendbuff, void* recvbuff, size_t sendcount, int datatype, void* comm, cudaStream_t stream);
typedef const char* (*nccl_error_string_fn)(int result);
constexpr int kNcclFloat64 = 8;

std::mutex g_mu;
void* g_nccl = nullptr;
nccl_all_gather_fn g_all_gather = nullptr;
nccl_error_string_fn g_error_string = nullptr;

bool bind_nccl(std::string& why) {
    std::lock_guard<std::mutex> l(g_mu);
    if (g_all_gather) return true;
    // RTLD_NOLOAD first: reuse the copy the host program (e.g. torch) already mapped, whatever its path
    for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
        g_nccl = dlopen(name, RTLD_NOW | RTLD_NOLOAD);
        if (!g_nccl) g_nccl = dlopen(name, RTLD_NOW | RTLD_LOCAL);
        if (g_nccl) break;
    }
    if (!g_nccl) { why = std::string("cannot load libnccl.so.2: ") + (dlerror() ? dlerror() : "not found"); return false; }
    g_all_gather = (nccl_all_gather_fn)dlsym(g_nccl, "ncclAllGather");
    g_error_string = (nccl_error_string_fn)dlsym(g_nccl, "ncclGetErrorString");
    if (!g_all_gather) { why = "libnccl has no ncclAllGather"; return false; }
    return true;
}

// gathered [R][kAccumSlots] -> accum [kAccumSlots]; one warp
__global__ void merge_accumulators_kernel(const double* __restrict__ gathered, int R, double* __restrict__ accum) {
    const int slot = threadIdx.x;
    double gmax = -INFINITY;
    bool any = false;
    for (int r = 0; r < R; ++r) {
        if (gathered[r * kAccumSlots + ACC_BATCHES] > 0) { any = true; gmax = fmax(gmax, gathered[r * kAccumSlots + ROW_LZ_MAX]); }
    }
    if (slot >= kAccumSlots) return;
    double v = 0.0;
    if (slot == ROW_LZ_MAX) {
        v = any ? gmax : 0.0;
    } else if (slot == ROW_LZ_SUM) {
        for (int r = 0; r < R; ++r)
            if (gathered[r * kAccumSlots + ACC_BATCHES] > 0)
                v += gathered[r * kAccumSlots + ROW_LZ_SUM] * exp(gathered[r * kAccumSlots + ROW_LZ_MAX] - gmax);
    } else {
        for (int r = 0; r < R; ++r) v += gathered[r * kAccumSlots + slot];
    }
    accum[slot] = v;
}

}  // namespace

extern "C" int apvae_allreduce_metrics(void* nccl_comm, int nranks, double* accum_dev, double* gather_dev, void* stream) {
    if (!accum_dev || nranks < 1 || (nranks > 1 && (!nccl_comm || !gather_dev)))
        return report_global_error("allreduce_metrics: bad arguments", APVAE_ERR_INVALID);
    if (nranks == 1) return APVAE_OK;                       // one shard: the accumulator is already the dataset's
    std::string why;
    if (!bind_nccl(why)) return report_global_error("allreduce_metrics: " + why, APVAE_ERR_UNSUPPORTED);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int r = g_all_gather(accum_dev, gather_dev, (size_t)kAccumSlots, kNcclFloat64, nccl_comm, st);
    if (r != 0)
        return report_global_error(std::string("allreduce_metrics: ncclAllGather: ") + (g_error_string ? g_error_string(r) : "error"),
                                   APVAE_ERR_CUDA);
    merge_accumulators_kernel<<<1, 32, 0, st>>>(gather_dev, nranks, accum_dev);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? APVAE_OK : report_global_error(std::string("allreduce_metrics: ") + cudaGetErrorString(e), APVAE_ERR_CUDA);
}

// The merge alone (no NCCL): [nranks][APVAE_ACCUM_SLOTS] gathered by any transport -> accum_dev.
extern "C" int apvae_merge_metrics(const double* gathered_dev, int nranks, double* accum_dev, void* stream) {
    if (!gathered_dev || !accum_dev || nranks < 1) return report_global_error("merge_metrics: bad arguments", APVAE_ERR_INVALID);
    merge_accumulators_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(gathered_dev, nranks, accum_dev);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? APVAE_OK : report_global_error(std::string("merge_metrics: ") + cudaGetErrorString(e),
                                                              e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? APVAE_ERR_NO_DEVICE : APVAE_ERR_CUDA);
}
