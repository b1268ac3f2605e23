// Patch-GEMM: the A-stationary implicit-GEMM kernel for stride-1 conv2d / conv2d_transpose
// (reference: tfsnippet/layers/convolutional/conv2d_.py:204-211,464-471) on image widths
// W in {16, 32, 64, 128}.
//
// The tap-GEMM kernel (tapgemm.h) re-reads the shifted activation window from L2 once per
// filter tap, which makes L2 -> SM bandwidth (about 6300 B/clk chip-wide) the limiter of the
// wide layers.  Here a tile is `macc` x 128 consecutive output pixels (whole rows) of ONE image.
// For every (channel chunk, horizontal tap offset dx) the producer loads ONE activation patch
//     [rows_tile + k - 1 rows][W pixels][kc channels]
// by TMA (box shifted by dx; the SAME-padding halo is the TMA zero fill).  All k vertical taps
// of that dx then read the SAME patch: tap j's A operand is the patch advanced by j*W rows, i.e.
// the UMMA shared-memory descriptor start address + j*W*kc*2 bytes (a multiple of the swizzle
// period, so no re-swizzling).  Activations are therefore fetched ndx*(rows+k-1)/rows times
// instead of k*ndx times.  Weight tiles stream through their own ring; with cluster size c > 1
// each CTA fetches 1/c of every weight tile and TMA-multicasts it to the whole cluster.
//
// A "group" = (source tensor, channel chunk, dx); it owns k weight tiles, fetched `tpi` taps at a time
// ("items": one ring slot and one barrier round trip per item).
#pragma once
#include <stdint.h>
#include <cuda.h>
#include "tapgemm.h"

namespace apvae {

struct PatchGroup {
    int16_t src;     // which A tensor map (0..kPatchMaxMaps-1)
    int16_t dx;      // horizontal shift of the patch box
    int16_t py0;     // vertical shift of patch row 0 relative to the tile's first output row
    int16_t k;       // weight tiles (vertical taps) that read this patch
    int32_t c0;      // channel coordinate of the chunk
    int32_t prow;    // patch rows of this source (= TMA box height)
};

constexpr int kPatchMaxSrc = 2;
constexpr int kPatchMaxMaps = 4;     // activation tensor maps: (source, row phase of the space-to-depth view)
constexpr int kPatchMaxGroups = 48;
// TMA epilogue (bf16 outputs, optional bf16 residual, no mask / tanh / fp32 output): per epilogue warp two
// residual boxes, one raw and one act staging box of 32 rows x 128 bytes (SWIZZLE_128B), plus the bias row
constexpr uint32_t kFastEpiBoxBytes = 32u * 128u;
constexpr uint32_t kFastEpiWarpBytes = 4u * kFastEpiBoxBytes + 1024u;
constexpr uint32_t kFastEpiSmemBytes = 4u * kFastEpiWarpBytes;
// Two epilogue warpgroups (PatchGemmParams::epi2: warps 4..7 take the even 64-column blocks of a tile, warps 8..11 the odd
// ones; 384 threads): per warp ONE residual box, one raw and one act staging box and the bias row.  For ops whose tile
// costs the epilogue more than the MMAs (few K-elements per output: 3x3 96-wide layers, 1x1 shortcuts, sub-pixel phases,
// everything narrow).
constexpr uint32_t kFastEpi2WarpBytes = 3u * kFastEpiBoxBytes + 1024u;
constexpr uint32_t kPatchFixedSmem2 = 1024u + 1024u + 8u * kFastEpi2WarpBytes;
// shared memory the kernel needs besides the two rings: barriers, realignment slack, epilogue staging
constexpr uint32_t kPatchFixedSmem = 1024u + 1024u + (kFastEpiSmemBytes > kEpiSmemBytes ? kFastEpiSmemBytes : kEpiSmemBytes);

struct alignas(64) PatchGemmParams {
    CUtensorMap tm_a[kPatchMaxMaps]; // 5-D (C, W, 1, H, N), box (kc, W, 1, prow_src, 1); a stride-2 conv reads the
                                     // space-to-depth view (2C, W/2, 1, H/2, N) of one row phase per map
    CUtensorMap tm_b;                // 2-D (kc, items*Npad), box (kc, NT / cluster)
    PatchGroup groups[kPatchMaxGroups];   // in the kernel parameters: read through the constant bank (uniform loads)
    int num_groups, num_items;
    int W, H;                        // tile-row width (a power of two; per image pair when folded) and image height
    int Wr;                          // image width in memory; < W / fold for the virtual widths (28 in 32, 14 in 16, 7 in 8): the TMA
                                     // unit zero-fills / clips what hangs over, the generic epilogue skips those pixels
    int res4d;                       // 1: tm_res is the 4-D view (Npad, Wr, H, N) of a virtual-width grid (RES_EVEN: of its half-resolution grid)
    int rows_tile, macc;             // output rows per tile = macc * 128 / W
    int tiles_y;
    int n_valid;                     // images to process (folded ops: image PAIRS, see `fold`)
    int fold;                        // 1: ordinary op.  2: 8-pixel-wide images, two per 128-pixel tile: the patch is
                                     // [rows][2 images][8 pixels] (W = 16 here), tm_a is (C, 8, 2, H, N/2) and the epilogue
                                     // maps tm_raw / tm_act / tm_res are 5-D (Npad, 8, 2, H, N/2) with boxes (64, 8, 2, 2, 1)
    int n_real;                      // images to process (== n_valid unless folded)
    int Npad, NT;                    // NT == Npad <= 256, macc * NT <= 256
    int kc;                          // 16, 32 or 64 channels per chunk
    int tpi;                         // vertical taps per weight item (one barrier / ring slot per item)
    uint32_t tap_bytes;              // NT * kc * 2: one tap's weight tile inside an item
    int a_stages, b_stages;
    uint32_t a_slot_bytes, b_slot_bytes;
    int cluster;                     // 1, 2 or 4 CTAs sharing the weight stream
    int two_cta;                     // 1 (cluster == 2): cta_group::2 MMAs, each CTA holds half of every weight tile
    int fast_epi;                    // 1: TMA epilogue through tm_raw / tm_act / tm_res
    int res_rows;                    // rows of the residual box (32; 16 or 8 for RES_EVEN)
    CUtensorMap tm_raw, tm_act;      // 4-D (Npad, W, H, N) over the output (strided by 2 for a sub-pixel phase),
                                     // box (64, min(W, 32), 32 / min(W, 32), 1), SWIZZLE_128B
    CUtensorMap tm_res;              // 2-D (Npad, residual pixels), box (64, res_rows)
    CUtensorMap tm_mask;             // fast_epi with epi.mask (adjoint ops, raw output only): the forward activation whose sign
                                     // selects leaky-relu', same view and box as a same-resolution residual; it is loaded
                                     // one block ahead into the (unused) act staging box
    int fast_mask;                   // 1: tm_mask is valid and the TMA epilogue multiplies by leaky-relu'(mask); 2: tm_mask is the
                                     // 4-D strided view of a sub-pixel phase, loaded at the store's coordinates
    int epi2;                        // 1 (fast_epi, no mask, no split): two epilogue warpgroups, 384 threads, kPatchFixedSmem2
    int split;                       // 1 (fast_epi, no mask): split-precision epilogue — every output is written as a hi + lo
                                     // pair (tm_raw / tm_act and tm_raw_lo / tm_act_lo, two store passes through the same
                                     // staging boxes) and the residual is read as hi + lo (tm_res / tm_res_lo, one box each)
    CUtensorMap tm_raw_lo, tm_act_lo, tm_res_lo;
    int debug;                       // tuning only (apvae_tuning.debug): 1 = no loads, 2 = no weight loads, 3 = no patch loads, 4 = no epilogue work, 5 = no output stores, 1x = x plus no loads (results are garbage)
    unsigned long long* trace;       // tuning only (APVAE_PATCH_TRACE): per-tile clock64 stamps of CTA 0, 16 slots per tile, 64 tiles
    EpilogueArgs epi;
};

int launch_patchgemm(const PatchGemmParams& p, int num_sms, void* stream);
uint32_t patchgemm_smem_bytes(const PatchGemmParams& p);

}  // namespace apvae
