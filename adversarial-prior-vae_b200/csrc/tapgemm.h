// Tap-GEMM: the one tensor-core kernel of the path.
//
// Every conv2d / conv2d_transpose / dense on the hot path (reference:
// tfsnippet/layers/convolutional/conv2d_.py:178-211,464-471 and
// layers/core/dense_.py:142-147) is executed as
//
//     out[128 pixels, NT channels] = sum over steps  A_step[128, KC] * B_step[NT, KC]^T
//
// where one "step" is (filter tap, KC-channel chunk).  A_step is a TMA box of
// the NHWC bf16 activation tensor shifted by the tap offset (TMA zero-fills
// the SAME-padding halo), B_step is a TMA box of the pre-packed bf16 weights,
// and the accumulator lives in TMEM (tcgen05.mma, M=128, N=NT, K=16).
// Stride-2 convs read the input through a space-to-depth 5-D tensor map;
// stride-2 transposed convs run as 4 sub-pixel phases over the input grid.
//
// This header is shared between the host planner (plain C++) and the kernel.
#pragma once
#include <stdint.h>
#include <cuda.h>

namespace apvae {

struct TapStep {
    int16_t src;     // which A tensor map (0..2)
    int16_t dx, dy;  // box shift in the W' / H' dims of the source map
    int16_t ph;      // coordinate in the phase dim (space-to-depth row phase)
    int32_t c0;      // coordinate in the innermost (channel) dim
};

constexpr int kMaxPhases = 4;
constexpr int kMaxSrc = 3;
constexpr int kTileM = 128;
constexpr int kMaxInlineSteps = 160;   // step tables up to this size travel in the kernel parameters (uniform constant-bank reads)

enum ResMode { RES_NONE = 0, RES_SAME = 1, RES_EVEN = 2 };

// Fused epilogue shared by the tensor-core kernels:
//   v = acc + bias ; v *= lrelu'(mask) ; v += res ; tanh on cols < tanh_cols
struct EpilogueArgs {
    const float* bias;          // [Npad]
    void* out_raw;              // bf16 [pix, ldo]   v
    void* out_act;              // bf16 [pix, ldo]   leaky_relu(v)
    int ldo;
    int bf16_zero_from;         // columns >= this are stored as 0 in the bf16 outputs
    float* out_f32;             // fp32 [pix, ld_f32], columns < f32_cols
    int ld_f32, f32_cols;
    const void* res;            // bf16 residual
    int res_mode, ld_res;
    const void* mask;           // bf16 forward tensor whose sign selects lrelu'
    int ld_mask;
    int tanh_cols;
    float lrelu_alpha;
    // split-precision mode (activations carried as hi + lo bf16 pairs, see model.h `precise`): low halves
    void* out_raw_lo;           // bf16(v - bf16(v))
    void* out_act_lo;           // bf16(leaky_relu(v) - bf16(leaky_relu(v)))
    const void* res_lo;         // low half of the residual (same indexing as res)
    int f16;                    // 16-bit operand / output format of this op: 1 = fp16, 0 = bf16 (op16.h)
};

struct alignas(64) TapGemmParams {
    CUtensorMap tm_a[kMaxSrc];  // 5-D: (C', W', P, H', N)
    CUtensorMap tm_b;           // 2-D: (KC, total_steps * Npad)
    const TapStep* steps;       // device pointer, all phases back to back (used when the table does not fit inline)
    int inline_steps;           // 1: read steps_inline
    TapStep steps_inline[kMaxInlineSteps];
    int phase_begin[kMaxPhases + 1];
    int num_phases;             // 1, or 4 (py = ph >> 1, px = ph & 1)
    // pixel tile = TMA box (bw, bh, bn); rows = bw*bh*bn <= 128
    int bw, bh, bn, rows;
    int Wg, Hg;                 // extent of the tile grid (per image)
    int tiles_x, tiles_y;
    int n_valid;                // images (or samples) to process
    // output grid
    int Ho, Wo, os;             // output H, W; output pixel = grid pixel * os + phase
    int Npad, NT, n_tiles;      // padded Cout, N per tile, tiles along N
    int kc;                     // 16, 32 or 64 channels per step
    int num_stages;
    uint32_t a_bytes, b_bytes, stage_bytes;
    int two_cta;                // 1: launched as 2-CTA clusters, one cta_group::2 MMA per pair of pixel tiles
                                // (each CTA loads its own A tile and HALF of every weight tile; tm_b's box is NT / 2 rows)
    EpilogueArgs epi;
};

// Per-warp shared-memory staging of the epilogue (sm100_ptx.cuh: epilogue_block).
constexpr uint32_t kEpiBlockCols = 64;
constexpr uint32_t kEpiPitch = kEpiBlockCols * 2 + 16;                 // bytes per staged row
constexpr uint32_t kEpiStageBytes = 32 * kEpiPitch;                    // one 32-row staging block
constexpr uint32_t kEpiWarpBytes = 2 * kEpiStageBytes + 1024;          // block A (residual, then raw), block B (act), bias[256]
constexpr uint32_t kEpiSmemBytes = 4 * kEpiWarpBytes;                  // four epilogue warps

// Launches the kernel on `stream`.  Returns a cudaError_t value (0 = ok).
int launch_tapgemm(const TapGemmParams& p, int num_sms, void* stream);

// Dynamic shared memory the kernel needs for (kc, NT, stages).
uint32_t tapgemm_smem_bytes(int kc, int NT, int stages);
// Pick the deepest pipeline that fits.
int tapgemm_pick_stages(int kc, int NT, int total_steps_min);

}  // namespace apvae
