// Device-side instantiation of planned ops: uploads step tables / weights,
// encodes TMA tensor maps and launches the tap-GEMM kernel.
#pragma once
#include <stdint.h>
#include <string>
#include <vector>
#include "planner.h"
#include "kernels.h"
#include "tapgemm.h"
#include "patchgemm.h"

namespace apvae {

struct Epilogue {
    void* out_raw = nullptr;
    void* out_act = nullptr;
    int ldo = 0;
    int bf16_zero_from = 1 << 30;
    float* out_f32 = nullptr;
    int ld_f32 = 0, f32_cols = 0;
    const void* res = nullptr;
    int res_mode = RES_NONE, ld_res = 0;
    const void* mask = nullptr;
    int ld_mask = 0;
    int tanh_cols = 0;
    int f16 = 1;                     // 16-bit operand / output format of the op: 1 = fp16, 0 = bf16 (op16.h)
    void* out_raw_lo = nullptr;      // split-precision low halves (see tapgemm.h)
    void* out_act_lo = nullptr;
    const void* res_lo = nullptr;
};

struct DeviceInfo {
    int device = -1;
    int num_sms = 0;
    int cc_major = 0, cc_minor = 0;
    void* encode_fn = nullptr;   // cuTensorMapEncodeTiled, resolved at run time (no libcuda link dependency)
};

// Fails (returns false) when there is no CUDA device of compute capability 10.x.
bool device_open(int device, DeviceInfo* info, std::string* err);

struct DevOp {
    std::string name;
    TapStep* d_steps = nullptr;
    uint16_t* d_w = nullptr;
    float* d_bias = nullptr;
    unsigned long long* d_trace = nullptr;
    TapGemmParams params;
    bool is_patch = false;          // true: the A-stationary kernel (patchgemm.h) runs this op
    PatchGroup* d_groups = nullptr;
    PatchGemmParams pparams;
    bool is_heads = false;          // true: G_theta's 1x1 output heads on the CUDA cores (kernels.h: HeadsParams)
    HeadsParams heads;
    bool is_tiny = false;           // true: a <= 4 -> <= 4 channel stride-1 conv on the CUDA cores (kernels.h: TinyConvParams)
    TinyConvParams tiny;
    bool is_small_block = false;    // true: a whole <= 3-channel-input resnet block on the CUDA cores (kernels.h: SmallBlockParams)
    SmallBlockParams sblock;
    bool is_small_block_grad = false;   // true: the adjoint of such a block (kernels.h: SmallBlockGradParams)
    SmallBlockGradParams sbgrad;
    double macs_per_image = 0;
    int capacity_n = 0;
    ~DevOp();
    DevOp() = default;
    DevOp(const DevOp&) = delete;
    DevOp& operator=(const DevOp&) = delete;
};

// Uploads the plan and encodes the tensor maps against `src_ptrs` (bf16 NHWC device tensors).
bool devop_create(const DeviceInfo& dev, const OpPlan& plan, const void* const* src_ptrs, const Epilogue& epi,
                  DevOp* op, std::string* err);
bool devop_create_patch(const DeviceInfo& dev, const PatchPlan& plan, const void* const* src_ptrs, const Epilogue& epi,
                        DevOp* op, std::string* err);
// Launch for the first n_valid images.
int devop_launch(const DeviceInfo& dev, DevOp& op, int n_valid, void* stream);

#define APVAE_CUDA_OK(expr, errp)                                                            \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess) {                                                              \
            if (errp) *(errp) = std::string(#expr) + ": " + cudaGetErrorString(_e);           \
            return false;                                                                     \
        }                                                                                     \
    } while (0)

}  // namespace apvae
