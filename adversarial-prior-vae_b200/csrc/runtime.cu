#include "runtime.h"
#include "tuning.h"

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdlib.h>
#include <string.h>

namespace apvae {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

bool device_open(int device, DeviceInfo* info, std::string* err) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count <= 0) {
        if (err) *err = std::string("no CUDA device available (") + cudaGetErrorString(e) +
                        "); this library has no CPU fallback";
        cudaGetLastError();
        return false;
    }
    if (device < 0 || device >= count) {
        if (err) *err = "device index out of range";
        return false;
    }
    APVAE_CUDA_OK(cudaSetDevice(device), err);
    cudaDeviceProp prop;
    APVAE_CUDA_OK(cudaGetDeviceProperties(&prop, device), err);
    if (prop.major != 10) {
        if (err) *err = std::string("device '") + prop.name + "' is not sm_100 (Blackwell B200); kernels are sm_100a only";
        return false;
    }
    info->device = device;
    info->num_sms = prop.multiProcessorCount;
    info->cc_major = prop.major;
    info->cc_minor = prop.minor;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    APVAE_CUDA_OK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres), err);
    if (fn == nullptr || qres != cudaDriverEntryPointSuccess) {
        if (err) *err = "cuTensorMapEncodeTiled not available from the driver";
        return false;
    }
    info->encode_fn = fn;
    return true;
}

static CUtensorMapSwizzle swizzle_for(int kc) {
    return kc == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : kc == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
}

DevOp::~DevOp() {
    if (d_steps) cudaFree(d_steps);
    if (d_w) cudaFree(d_w);
    if (d_bias) cudaFree(d_bias);
    if (d_trace) cudaFree(d_trace);
    if (d_groups) cudaFree(d_groups);
}

static void fill_epilogue(EpilogueArgs& e, const Epilogue& epi, const float* bias) {
    e.bias = bias;
    e.out_raw = epi.out_raw; e.out_act = epi.out_act; e.ldo = epi.ldo;
    e.bf16_zero_from = epi.bf16_zero_from;
    e.out_f32 = epi.out_f32; e.ld_f32 = epi.ld_f32; e.f32_cols = epi.f32_cols;
    e.res = epi.res; e.res_mode = epi.res_mode; e.ld_res = epi.ld_res;
    e.mask = epi.mask; e.ld_mask = epi.ld_mask;
    e.tanh_cols = epi.tanh_cols;
    e.lrelu_alpha = 0.2f;
    e.out_raw_lo = epi.out_raw_lo; e.out_act_lo = epi.out_act_lo; e.res_lo = epi.res_lo;
    e.f16 = epi.f16;
}

static bool check_epilogue(const Epilogue& epi, int npad, const std::string& name, std::string* err) {
    if ((epi.out_raw || epi.out_act) && (epi.ldo % 16 != 0 || epi.ldo < npad)) {
        if (err) *err = "epilogue: ldo must be a multiple of 16 and >= padded cout (op " + name + ")";
        return false;
    }
    if (epi.res && (epi.ld_res % 8 != 0 || epi.ld_res < npad)) {
        if (err) *err = "epilogue: residual row too short (op " + name + ")";
        return false;
    }
    if (epi.mask && (epi.ld_mask % 8 != 0 || epi.ld_mask < npad)) {
        if (err) *err = "epilogue: mask row too short (op " + name + ")";
        return false;
    }
    return true;
}

bool devop_create_patch(const DeviceInfo& dev, const PatchPlan& plan, const void* const* src_ptrs, const Epilogue& epi,
                        DevOp* op, std::string* err) {
    EncodeTiledFn encode = reinterpret_cast<EncodeTiledFn>(dev.encode_fn);
    if (!check_epilogue(epi, plan.phase >= 4 ? plan.npad / 2 : plan.npad, op->name, err)) return false;
    const size_t ngroups = plan.groups.size();
    if (ngroups > (size_t)kPatchMaxGroups) { if (err) *err = "patch plan has too many groups"; return false; }
    APVAE_CUDA_OK(cudaMalloc(&op->d_w, plan.wpack.size() * sizeof(uint16_t)), err);
    APVAE_CUDA_OK(cudaMalloc(&op->d_bias, plan.bias.size() * sizeof(float)), err);
    APVAE_CUDA_OK(cudaMemcpy(op->d_w, plan.wpack.data(), plan.wpack.size() * sizeof(uint16_t), cudaMemcpyHostToDevice), err);
    APVAE_CUDA_OK(cudaMemcpy(op->d_bias, plan.bias.data(), plan.bias.size() * sizeof(float), cudaMemcpyHostToDevice), err);
    PatchGemmParams& p = op->pparams;
    memset(&p, 0, sizeof(p));
    const cuuint32_t ones[5] = {1, 1, 1, 1, 1};
    // W: tile-row width per image (box widths); Wr: the grid width in memory (tensor-map dims and strides).  They differ for
    // the virtual widths of planner.cpp (28 -> 32 ...): every box then hangs over the image's right / bottom edge, where
    // the TMA unit zero-fills loads and clips stores
    const cuuint64_t C = plan.in_cp, W = plan.w, Wr = plan.w_real > 0 ? plan.w_real : plan.w, H = plan.h;
    const bool virt = (Wr != W);
    const bool fold = plan.fold > 1;        // 8-wide images in pairs: dim 2 is the image inside the pair, dim 4 the pair
    if (fold && (plan.fold != 2 || W != 8 || (plan.capacity_n & 1))) { if (err) *err = "patch: bad fold geometry for op " + op->name; return false; }
    if (fold && epi.res_mode == RES_EVEN) { if (err) *err = "patch: folded ops take no half-resolution residual (op " + op->name + ")"; return false; }
    const cuuint64_t NP = fold ? (cuuint64_t)plan.capacity_n / 2 : (cuuint64_t)plan.capacity_n;
    // one map per (source, row phase): a stride-2 conv reads the space-to-depth view of its 2H x 2W input, column phase
    // folded into the channel dim (2C), row phase selected by the map's base address (planner.h: PatchPlan::s2d)
    const cuuint64_t cs = plan.s2d ? 2 : 1;
    const cuuint64_t row_b = cs * cs * Wr * C * 2, img_b = H * row_b;      // bytes per grid row (cs input rows) and per image
    if (plan.nmaps < 1 || plan.nmaps > kPatchMaxMaps) { if (err) *err = "patch: bad map count for op " + op->name; return false; }
    for (int s = 0; s < plan.nmaps; ++s) {
        cuuint64_t dims[5] = {cs * C, Wr, fold ? 2u : 1u, H, NP};
        cuuint64_t strides[4] = {cs * C * 2, fold ? img_b : row_b, row_b, (fold ? 2 : 1) * img_b};
        cuuint32_t box[5] = {(cuuint32_t)plan.kc, (cuuint32_t)W, fold ? 2u : 1u, (cuuint32_t)plan.prow[s], 1};
        char* base = static_cast<char*>(const_cast<void*>(src_ptrs[plan.map_src[s]])) + (size_t)plan.map_ph[s] * (cs * Wr * C * 2);
        CUresult r = encode(&p.tm_a[s], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, base, dims, strides,
                            box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(plan.kc),
                            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            if (err) *err = "cuTensorMapEncodeTiled(patch A) failed with CUresult " + std::to_string((int)r) + " for op " + op->name;
            return false;
        }
    }
    for (int s = plan.nmaps; s < kPatchMaxMaps; ++s) p.tm_a[s] = p.tm_a[0];
    {
        cuuint64_t dims[2] = {(cuuint64_t)plan.kc, (cuuint64_t)plan.num_items * plan.npad};
        cuuint64_t strides[1] = {(cuuint64_t)plan.kc * 2};
        cuuint32_t box[2] = {(cuuint32_t)plan.kc, (cuuint32_t)(plan.npad / plan.cluster)};
        CUresult r = encode(&p.tm_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, op->d_w, dims, strides, box, ones,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(plan.kc), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            if (err) *err = "cuTensorMapEncodeTiled(patch B) failed with CUresult " + std::to_string((int)r) + " for op " + op->name;
            return false;
        }
    }
    for (size_t i = 0; i < ngroups; ++i) p.groups[i] = plan.groups[i];
    p.num_groups = (int)ngroups; p.num_items = plan.num_items;
    p.W = plan.w * plan.fold; p.Wr = (int)Wr; p.H = plan.h; p.fold = plan.fold; p.rows_tile = plan.rows_tile; p.macc = plan.macc; p.tiles_y = plan.tiles_y;
    p.n_valid = 0;
    p.Npad = plan.npad; p.NT = plan.npad; p.kc = plan.kc;
    p.tpi = plan.tpi; p.tap_bytes = plan.tap_bytes;
    p.a_stages = plan.a_stages; p.b_stages = plan.b_stages;
    p.a_slot_bytes = plan.a_slot_bytes; p.b_slot_bytes = plan.b_slot_bytes;
    p.cluster = plan.cluster;
    p.two_cta = (plan.two_cta && plan.cluster == 2) ? 1 : 0;
    p.debug = tuning().debug;
    if (switch_on(tuning().trace, false)) {        // tuning only: clock stamps of CTA 0 (read back by apvae_plan_bench)
        APVAE_CUDA_OK(cudaMalloc(&op->d_trace, 64 * 16 * sizeof(unsigned long long)), err);
        APVAE_CUDA_OK(cudaMemset(op->d_trace, 0, 64 * 16 * sizeof(unsigned long long)), err);
        p.trace = op->d_trace;
    }
    fill_epilogue(p.epi, epi, op->d_bias);
    // bf16-only epilogues (the conv blocks' interior layers) go through bulk tensor loads / stores
    p.fast_epi = 0;
    p.res_rows = 32;
    {
        // a leaky-relu' mask (adjoint ops) rides along when the op writes no act output: its box takes the act staging box
        // (a merged-column-phase op needs unpadded mask rows, like its output rows)
        const bool mask_ok = !epi.mask || ((plan.phase < 4 || epi.ld_mask * 2 == plan.npad) && !epi.out_act && epi.out_raw && switch_on(tuning().fast_mask, true));
        // split precision (hi + lo outputs, hi + lo residual): its own epilogue variant, when every output and the residual
        // come as complete pairs and there is no mask
        const bool any_lo = epi.out_raw_lo || epi.out_act_lo || epi.res_lo;
        const bool split_ok = !any_lo || ((!epi.out_raw == !epi.out_raw_lo) && (!epi.out_act == !epi.out_act_lo) &&
                                          (epi.res_mode == RES_NONE ? !epi.res_lo : epi.res_lo != nullptr) && !epi.mask &&
                                          switch_on(tuning().patch_split, true));
        const bool plain = mask_ok && !epi.out_f32 && split_ok && epi.tanh_cols == 0 && epi.bf16_zero_from >= plan.npad &&
                           (epi.out_raw || epi.out_act) && (epi.res_mode == RES_NONE || epi.res != nullptr);
        if (plain && switch_on(tuning().fast_epi, true)) {
            const cuuint64_t pixels = (cuuint64_t)plan.capacity_n * H * Wr;
            bool ok = true;
            auto map2d = [&](CUtensorMap* tm, const void* ptr, cuuint64_t rows, int ld, cuuint32_t box_rows) {
                cuuint64_t dims[2] = {(cuuint64_t)plan.npad, rows};
                cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
                cuuint32_t box[2] = {64, box_rows};
                const CUresult r = encode(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, ones,
                                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                          CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                if (r != CUDA_SUCCESS) ok = false;
            };
            // outputs: 4-D (C, W, H, N) view of the output tensor on the INPUT grid; a sub-pixel phase (py, px) of a stride-2
            // transposed conv writes every second pixel of a 2H x 2W image starting at (py, px)
            const int osc = plan.out_scale;
            // phase 4 + py: both column phases in one op; its 2 * cout outputs per input pixel are output pixels (2x, 2x + 1),
            // contiguous only when the output rows carry no channel padding
            const bool merged = plan.phase >= 4;
            if (merged && epi.ldo * 2 != plan.npad) { if (err) *err = "patch: merged column phases need ldo == cout (op " + op->name + ")"; return false; }
            const int py = merged ? plan.phase - 4 : plan.phase >= 0 ? (plan.phase >> 1) : 0, px = (plan.phase >= 0 && !merged) ? (plan.phase & 1) : 0;
            auto map_grid = [&](CUtensorMap* tm, const void* ptr, int ld_elems, int osc, int py, int px) {
                const cuuint64_t ld = (cuuint64_t)ld_elems * 2;                      // bytes per output pixel
                const cuuint64_t Wo = (cuuint64_t)osc * Wr, Ho = (cuuint64_t)osc * H;
                char* base = static_cast<char*>(const_cast<void*>(ptr)) + ((size_t)py * Wo + px) * ld;
                if (fold) {                 // (C, 8 pixels, 2 images, H, pairs): a block's 32 rows are [2 rows][2 images][8 pixels]
                    cuuint64_t dims5[5] = {(cuuint64_t)plan.npad, Wr, 2, H, NP};
                    cuuint64_t strides4[4] = {osc * ld, Ho * Wo * ld, osc * Wo * ld, 2 * Ho * Wo * ld};
                    cuuint32_t box5[5] = {64, 8, 2, 2, 1};
                    const CUresult r = encode(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, base, dims5, strides4, box5, ones,
                                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                    if (r != CUDA_SUCCESS) ok = false;
                    return;
                }
                cuuint64_t dims[4] = {(cuuint64_t)plan.npad, Wr, H, (cuuint64_t)plan.capacity_n};
                cuuint64_t strides[3] = {osc * ld, osc * Wo * ld, Ho * Wo * ld};
                const cuuint32_t bx = (cuuint32_t)std::min<cuuint64_t>(W, 32);
                cuuint32_t box[4] = {64, bx, 32 / bx, 1};
                const cuuint32_t ones4[4] = {1, 1, 1, 1};
                const CUresult r = encode(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, base, dims, strides, box, ones4,
                                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                          CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                if (r != CUDA_SUCCESS) ok = false;
            };
            auto map_out = [&](CUtensorMap* tm, const void* ptr, int ld_elems) { map_grid(tm, ptr, ld_elems, osc, py, px); };
            const void* any_out = epi.out_raw ? epi.out_raw : epi.out_act;
            map_out(&p.tm_raw, epi.out_raw ? epi.out_raw : any_out, epi.ldo);
            map_out(&p.tm_act, epi.out_act ? epi.out_act : any_out, epi.ldo);
            auto map_res = [&](CUtensorMap* tm, const void* ptr) {
                if (epi.res_mode == RES_EVEN && virt) {
                    // half-resolution residual of a virtual-width grid: (C, Wr / 2, H / 2, N), one (partial) row of it per block
                    p.res_rows = (plan.w >= 32) ? 16 : 8;
                    const cuuint64_t ld = (cuuint64_t)epi.ld_res * 2;
                    cuuint64_t dims[4] = {(cuuint64_t)plan.npad, Wr / 2, H / 2, (cuuint64_t)plan.capacity_n};
                    cuuint64_t strides[3] = {ld, (Wr / 2) * ld, (H / 2) * (Wr / 2) * ld};
                    cuuint32_t box[4] = {64, (cuuint32_t)p.res_rows, 1, 1};
                    const cuuint32_t ones4[4] = {1, 1, 1, 1};
                    const CUresult r = encode(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), dims, strides, box, ones4,
                                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                    if (r != CUDA_SUCCESS || (Wr & 1) || (H & 1)) ok = false;
                } else if (epi.res_mode == RES_EVEN) {
                    p.res_rows = (plan.w >= 32) ? 16 : 8;
                    map2d(tm, ptr, pixels / 4, epi.ld_res, (cuuint32_t)p.res_rows);
                } else if (epi.res_mode == RES_SAME && fold) {
                    const cuuint64_t ld = (cuuint64_t)epi.ld_res * 2;
                    cuuint64_t dims5[5] = {(cuuint64_t)plan.npad, Wr, 2, H, NP};
                    cuuint64_t strides4[4] = {ld, H * Wr * ld, Wr * ld, 2 * H * Wr * ld};
                    cuuint32_t box5[5] = {64, 8, 2, 2, 1};
                    const CUresult r = encode(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(ptr), dims5, strides4,
                                              box5, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                    if (r != CUDA_SUCCESS) ok = false;
                } else if (epi.res_mode == RES_SAME && virt) {
                    map_grid(tm, ptr, epi.ld_res, 1, 0, 0);              // the output's 4-D view: loaded at the store's coordinates
                } else if (epi.res_mode == RES_SAME) {
                    map2d(tm, ptr, pixels, epi.ld_res, 32);
                } else {
                    *tm = p.tm_raw;
                }
            };
            map_res(&p.tm_res, epi.res);
            p.res4d = (virt && !fold) ? 1 : 0;
            p.split = any_lo ? 1 : 0;
            p.epi2 = (plan.epi2 && !any_lo) ? 1 : 0;                   // two epilogue warpgroups: 16-bit epilogue (mask allowed)
            if (any_lo) {
                const void* any_lo_out = epi.out_raw_lo ? epi.out_raw_lo : epi.out_act_lo;
                map_out(&p.tm_raw_lo, epi.out_raw_lo ? epi.out_raw_lo : any_lo_out, epi.ldo);
                map_out(&p.tm_act_lo, epi.out_act_lo ? epi.out_act_lo : any_lo_out, epi.ldo);
                map_res(&p.tm_res_lo, epi.res_lo);
            } else {
                p.tm_raw_lo = p.tm_raw; p.tm_act_lo = p.tm_act; p.tm_res_lo = p.tm_res;
            }
            p.fast_mask = 0;
            if (epi.mask && plan.phase >= 0) {            // the mask of a sub-pixel phase: the same strided view as its output
                map_out(&p.tm_mask, epi.mask, epi.ld_mask);
                p.fast_mask = ok ? (fold ? 1 : 2) : 0;     // 2: 4-D load at the store's coordinates
            } else if (epi.mask && virt && !fold) {
                map_grid(&p.tm_mask, epi.mask, epi.ld_mask, 1, 0, 0);
                p.fast_mask = ok ? 2 : 0;
            } else if (epi.mask) {
                if (fold) {
                    const cuuint64_t ld = (cuuint64_t)epi.ld_mask * 2;
                    cuuint64_t dims5[5] = {(cuuint64_t)plan.npad, Wr, 2, H, NP};
                    cuuint64_t strides4[4] = {ld, H * Wr * ld, Wr * ld, 2 * H * Wr * ld};
                    cuuint32_t box5[5] = {64, 8, 2, 2, 1};
                    const CUresult r = encode(&p.tm_mask, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(epi.mask), dims5, strides4,
                                              box5, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                    if (r != CUDA_SUCCESS) ok = false;
                } else {
                    map2d(&p.tm_mask, epi.mask, pixels, epi.ld_mask, 32);
                }
                p.fast_mask = ok ? 1 : 0;
            } else {
                p.tm_mask = p.tm_raw;
            }
            p.fast_epi = ok ? 1 : 0;
        }
    }
    if (plan.phase >= 0 && (!p.fast_epi || epi.res_mode != RES_NONE)) {
        if (err) *err = "a sub-pixel phase op needs the bf16-only epilogue without residual (op " + op->name + ")";
        return false;
    }
    op->is_patch = true;
    op->macs_per_image = plan.macs_per_image;
    op->capacity_n = plan.capacity_n;
    return true;
}


bool devop_create(const DeviceInfo& dev, const OpPlan& plan, const void* const* src_ptrs, const Epilogue& epi,
                  DevOp* op, std::string* err) {
    EncodeTiledFn encode = reinterpret_cast<EncodeTiledFn>(dev.encode_fn);
    const size_t nsteps = plan.steps.size();
    APVAE_CUDA_OK(cudaMalloc(&op->d_steps, nsteps * sizeof(TapStep)), err);
    APVAE_CUDA_OK(cudaMalloc(&op->d_w, plan.wpack.size() * sizeof(uint16_t)), err);
    APVAE_CUDA_OK(cudaMalloc(&op->d_bias, plan.bias.size() * sizeof(float)), err);
    APVAE_CUDA_OK(cudaMemcpy(op->d_steps, plan.steps.data(), nsteps * sizeof(TapStep), cudaMemcpyHostToDevice), err);
    APVAE_CUDA_OK(cudaMemcpy(op->d_w, plan.wpack.data(), plan.wpack.size() * sizeof(uint16_t), cudaMemcpyHostToDevice), err);
    APVAE_CUDA_OK(cudaMemcpy(op->d_bias, plan.bias.data(), plan.bias.size() * sizeof(float), cudaMemcpyHostToDevice), err);

    TapGemmParams& p = op->params;
    memset(&p, 0, sizeof(p));
    const cuuint32_t ones[5] = {1, 1, 1, 1, 1};
    for (int s = 0; s < plan.nsrc; ++s) {
        const SrcGeom& g = plan.src[s];
        cuuint64_t dims[5], strides[4];
        cuuint32_t box[5];
        for (int i = 0; i < 5; ++i) { dims[i] = g.dims[i]; box[i] = g.box[i]; }
        for (int i = 0; i < 4; ++i) strides[i] = g.strides[i];
        CUresult r = encode(&p.tm_a[s], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(src_ptrs[s]), dims,
                            strides, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(plan.kc),
                            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            if (err) *err = "cuTensorMapEncodeTiled(A) failed with CUresult " + std::to_string((int)r) + " for op " + op->name;
            return false;
        }
    }
    for (int s = plan.nsrc; s < kMaxSrc; ++s) p.tm_a[s] = p.tm_a[0];
    {
        cuuint64_t dims[2] = {(cuuint64_t)plan.kc, (cuuint64_t)nsteps * plan.npad};
        cuuint64_t strides[1] = {(cuuint64_t)plan.kc * 2};
        // CTA pairs (cta_group::2): each CTA keeps half of every weight tile.  Pays off once the weight tile is wide
        // enough for the shared-memory traffic to matter; tuning.tap_2cta = -1 / +1 overrides.
        const bool split_epi = epi.out_raw_lo || epi.out_act_lo || epi.res_lo;
        bool two = false;    // measured: no gain for this kernel (1062 vs 1070 TFLOP/s at 192 wide, slower for short K): it is bound by the
                             // per-stage barrier round trip and the epilogue, not by shared-memory traffic; kept for experiments
        if (tuning().tap_2cta != 0) two = tuning().tap_2cta > 0 && !split_epi;
        p.two_cta = two ? 1 : 0;
        cuuint32_t box[2] = {(cuuint32_t)plan.kc, (cuuint32_t)(two ? plan.nt / 2 : plan.nt)};
        CUresult r = encode(&p.tm_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, op->d_w, dims, strides, box, ones,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(plan.kc), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            if (err) *err = "cuTensorMapEncodeTiled(B) failed with CUresult " + std::to_string((int)r) + " for op " + op->name;
            return false;
        }
    }
    p.steps = op->d_steps;
    p.inline_steps = nsteps <= (size_t)kMaxInlineSteps ? 1 : 0;
    if (p.inline_steps) for (size_t i = 0; i < nsteps; ++i) p.steps_inline[i] = plan.steps[i];
    for (int i = 0; i <= kMaxPhases; ++i) p.phase_begin[i] = plan.phase_begin[i];
    p.num_phases = plan.num_phases;
    p.bw = plan.bw; p.bh = plan.bh; p.bn = plan.bn; p.rows = plan.rows;
    p.Wg = plan.wg; p.Hg = plan.hg; p.tiles_x = plan.tiles_x; p.tiles_y = plan.tiles_y;
    p.n_valid = 0;
    p.Ho = plan.ho; p.Wo = plan.wo; p.os = plan.os;
    p.Npad = plan.npad; p.NT = plan.nt; p.n_tiles = plan.n_tiles;
    p.kc = plan.kc;
    int min_steps = 1 << 30;
    for (int ph = 0; ph < plan.num_phases; ++ph)
        min_steps = std::min(min_steps, plan.phase_begin[ph + 1] - plan.phase_begin[ph]);
    p.num_stages = tapgemm_pick_stages(plan.kc, plan.nt, std::max(2, (int)nsteps));
    p.a_bytes = (uint32_t)plan.rows * plan.kc * 2;
    p.b_bytes = (uint32_t)plan.nt * plan.kc * 2;
    p.stage_bytes = (kTileM * plan.kc * 2 + plan.nt * plan.kc * 2 + 1023u) & ~1023u;
    fill_epilogue(p.epi, epi, op->d_bias);
    op->macs_per_image = plan.macs_per_image;
    op->capacity_n = plan.capacity_n;
    if (!check_epilogue(epi, plan.npad, op->name, err)) return false;
    return true;
}

int devop_launch(const DeviceInfo& dev, DevOp& op, int n_valid, void* stream) {
    if (n_valid > op.capacity_n) return (int)cudaErrorInvalidValue;
    if (op.is_tiny) return launch_tiny_conv(op.tiny, n_valid, stream);
    if (op.is_small_block) return launch_small_block(op.sblock, n_valid, stream);
    if (op.is_small_block_grad) return launch_small_block_grad(op.sbgrad, n_valid, stream);
    if (op.is_heads) return launch_pixel_heads(op.heads, n_valid, stream);
    if (op.is_patch) {
        op.pparams.n_real = n_valid;
        op.pparams.n_valid = op.pparams.fold > 1 ? (n_valid + 1) / 2 : n_valid;
        return launch_patchgemm(op.pparams, dev.num_sms, stream);
    }
    op.params.n_valid = n_valid;
    return launch_tapgemm(op.params, dev.num_sms, stream);
}

}  // namespace apvae
