// Host-side planner: turns one conv2d / conv2d_transpose / dense (plus an
// optional fused 1x1 shortcut term) into a tap-GEMM step table, a packed bf16
// weight matrix and the TMA view geometry.  Pure C++ (no CUDA calls), so its
// output can be dumped and checked on a machine without a GPU.
//
// Reference semantics restated here:
//   conv2d  : tfsnippet/layers/convolutional/conv2d_.py:178-211, TF SAME padding
//   deconv2d: conv2d_.py:464-471 (adjoint of the SAME conv), output = in*stride
//             (layers/convolutional/utils.py:78-96)
//   dense   : layers/core/dense_.py:142-147 (a 1x1 conv over a 1x1 image)
#pragma once
#include <stdint.h>
#include <string>
#include <vector>
#include "tapgemm.h"
#include "patchgemm.h"

namespace apvae {

struct TermSpec {
    int src;            // which A source (0..2) this term reads
    int transpose;      // 0: conv2d, kernel [kh,kw,Cin,Cout]; 1: deconv2d, kernel [kh,kw,Cout,Cin]
    int k, stride;      // square kernel, stride 1 or 2
    int cin, cout;      // logical channels
    int in_h, in_w;     // source spatial size
    int in_cp;          // source channels as stored (padded to a multiple of 16)
    const float* kernel;
    const float* bias;  // may be null
    int f16 = 1;        // 16-bit operand format the kernel is packed in: 1 = fp16, 0 = bf16 (op16.h); all terms of an op agree
};

struct SrcGeom {
    uint64_t dims[5];     // (C', W', P, H', N)
    uint64_t strides[4];  // bytes, dims 1..4
    uint32_t box[5];
};

struct OpPlan {
    int kc = 0, npad = 0, nt = 0, n_tiles = 0, cout = 0;
    int num_phases = 1;
    int phase_begin[kMaxPhases + 1] = {0, 0, 0, 0, 0};
    std::vector<TapStep> steps;
    std::vector<uint16_t> wpack;   // bf16 [steps][npad][kc]
    std::vector<float> bias;       // [npad], sum over terms
    int bw = 0, bh = 0, bn = 0, rows = 0;
    int wg = 0, hg = 0, tiles_x = 0, tiles_y = 0;
    int ho = 0, wo = 0, os = 1;
    int nsrc = 0;
    int capacity_n = 0;
    SrcGeom src[kMaxSrc];
    double macs_per_image = 0;     // algorithmic multiply-accumulates (unpadded)
};

// Plan of the A-stationary kernel (patchgemm.h) for stride-1 terms on widths 16..128.
struct PatchPlan {
    int kc = 0, npad = 0, cout = 0;
    int w = 0, h = 0, rows_tile = 0, macc = 1, tiles_y = 0;
    int w_real = 0;                // grid width in memory; w is the tile-row width per image: w_real, or the next power of two above it
                                   // (28 -> 32, 14 -> 16, 7 -> 8: TMA zero-fills / clips columns >= w_real and rows >= h)
    int fold = 1;                  // 2: 8-pixel-wide images, two per tile row (the kernel sees a 16-wide "image" per pair)
    int nsrc = 0;
    // activation tensor maps: one per source, or (stride-2 conv through the space-to-depth view, `s2d`) one per
    // (source, row phase) in use; PatchGroup::src indexes this list
    int nmaps = 0;
    int map_src[kPatchMaxMaps] = {0, 0, 0, 0};
    int map_ph[kPatchMaxMaps] = {0, 0, 0, 0};
    int prow[kPatchMaxMaps] = {0, 0, 0, 0};     // patch rows (TMA box height) per map
    int s2d = 0;                   // 1: stride-2 conv; w, h are the OUTPUT grid, the input is 2w x 2h
    int in_cp = 0;
    std::vector<PatchGroup> groups;
    int num_items = 0;
    std::vector<uint16_t> wpack;   // bf16 [items][npad][kc]
    std::vector<float> bias;       // [npad]
    int tpi = 1;                   // taps per weight item
    uint32_t tap_bytes = 0;
    int a_stages = 0, b_stages = 0;
    uint32_t a_slot_bytes = 0, b_slot_bytes = 0;
    int cluster = 1;
    int two_cta = 0;               // cta_group::2 pairs (cluster == 2): each CTA stores half of every weight tile
    int epi2 = 0;                  // two epilogue warpgroups (patchgemm.h): the rings are sized for kPatchFixedSmem2
    // sub-pixel phase of a stride-2 transposed conv (-1: ordinary stride-1 op).  Phase (py, px) = (phase >> 1, phase & 1)
    // is a stride-1 correlation over the INPUT grid whose results land on output pixels (2y + py, 2x + px).
    int phase = -1;                // 4 + py: both column phases of output rows 2y + py (N = [px 0 | px 1] x cout, see planner.cpp)
    int out_scale = 1;             // output grid = out_scale x input grid
    int capacity_n = 0;
    double macs_per_image = 0;
};

inline int round_up(int v, int m) { return (v + m - 1) / m * m; }
inline int pad16(int c) { return round_up(c, 16); }

// TF SAME padding rule.
void same_padding(int in, int k, int s, int* out, int* lo, int* hi);

uint16_t f32_to_bf16(float f);
float bf16_to_f32(uint16_t h);

// Builds the plan.  All terms must agree on the output grid and on in_cp.
// Returns false and fills *err on unsupported shapes.
bool plan_op(const std::vector<TermSpec>& terms, int capacity_n, OpPlan* plan, std::string* err);

// Returns false (with *err = reason) when the op is not eligible for the patch kernel:
// every term must be stride 1 (or every term a stride-2 conv2d with even input size) on the same W x H output grid with W in {8 (pairs of images, even capacity),16,32,64,128}, <= 2 terms, cout <= 256.
// force_macc / force_cluster: 0 = choose automatically.
// `phase` >= 0: plan that sub-pixel phase of a single stride-2 transposed term (see PatchPlan::phase).
bool plan_patch_op(const std::vector<TermSpec>& terms, int capacity_n, int force_macc, int force_cluster,
                   PatchPlan* plan, std::string* err, int phase = -1);

}  // namespace apvae
