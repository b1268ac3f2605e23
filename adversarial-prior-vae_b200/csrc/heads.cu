// G_theta's 1x1 output heads on the CUDA cores (see kernels.h: HeadsParams).
//
// Reference: dmm_cifar10.py:489-496 (x_mean = tanh(conv2d 1x1), x_logstd = conv2d 1x1 on the last resnet block's
// output); dmm_for_sampling.py:503-514 and dmm_bernouli_fashion.py:511-521 have the x_mean head only.
//
// Roofline: HBM.  Per pixel the kernel reads cp bf16 (192 B at arch A) and writes nout fp32 + two 32-byte bf16 rows
// (88 B at arch A); the 2 * cin * nout FLOP per pixel (1152 at arch A) are ~4 FLOP per byte.  As a 128 x 16 x 16
// tensor-core tile this op ran at 4.5x its HBM time (epilogue-bound: 16 useful columns per 128-row accumulator).
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <math.h>
#include <stdint.h>
#include "kernels.h"

namespace apvae {
namespace {

constexpr int kTile = 128;                 // pixels per tile = threads per block

__device__ __forceinline__ float bf_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack2(float a, float b) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&v);
}
__device__ __forceinline__ float bf_round(float a) { return __bfloat162float(__float2bfloat16_rn(a)); }

// NOUT outputs per pixel; USED = ceil(cin / 8) 16-byte chunks per pixel row when known at compile time (the weight
// reads are then constant-bank operands of the FMAs with immediate offsets), 0 = run-time count.
template <int NOUT, int USED>
__global__ void __launch_bounds__(kTile) pixel_heads_kernel(const __grid_constant__ HeadsParams p, long long pixels) {
    extern __shared__ __align__(16) uint8_t sm[];
    const int chunks = p.cp >> 3;                       // 16-byte chunks per pixel row in memory
    // padded row pitch: (cp / 8 + 1) is odd, so the eight rows of a quarter-warp's 16-byte reads fall into eight
    // different bank quadruples (conflict-free LDS.128 with one thread per row)
    const int pitch = p.cp * 2 + 16;
    const bool split = p.in_lo != nullptr;
    uint8_t* const sm_lo = sm + kTile * pitch;
    const int used = USED ? USED : (p.cin + 7) >> 3;
    const int t = threadIdx.x;
    const long long tiles = (pixels + kTile - 1) / kTile;
    for (long long tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const long long p0 = tile * kTile;
        const int np = (int)((pixels - p0) < (long long)kTile ? (pixels - p0) : (long long)kTile);
        const int total = np * chunks;
        {
            const uint4* src = reinterpret_cast<const uint4*>(p.in) + p0 * chunks;
            constexpr int U = USED > 0 ? USED : 1;
            if (USED > 0 && np == kTile && chunks == U) {
                // full tile, compile-time row length: all of this thread's 16-byte loads are issued before the first
                // shared-memory store (U independent loads in flight per thread)
                uint4 v[U];
#pragma unroll
                for (int i = 0; i < U; ++i) v[i] = __ldg(src + t + i * kTile);
#pragma unroll
                for (int i = 0; i < U; ++i) {
                    const int c = t + i * kTile, px = c / U, ch = c - px * U;
                    *reinterpret_cast<uint4*>(sm + px * pitch + ch * 16) = v[i];
                }
            } else
            for (int c = t; c < total; c += kTile) {
                const int px = c / chunks, ch = c - px * chunks;
                if (ch < used) *reinterpret_cast<uint4*>(sm + px * pitch + ch * 16) = __ldg(src + c);
            }
            if (split) {
                const uint4* src_lo = reinterpret_cast<const uint4*>(p.in_lo) + p0 * chunks;
                for (int c = t; c < total; c += kTile) {
                    const int px = c / chunks, ch = c - px * chunks;
                    if (ch < used) *reinterpret_cast<uint4*>(sm_lo + px * pitch + ch * 16) = __ldg(src_lo + c);
                }
            }
        }
        __syncthreads();
        if (t < np) {
            float acc[NOUT];
#pragma unroll
            for (int n = 0; n < NOUT; ++n) acc[n] = p.bias[n];
            const uint8_t* row = sm + t * pitch;
            const uint8_t* row_lo = sm_lo + t * pitch;
            auto chunk = [&](int j) {
                const uint4 v = *reinterpret_cast<const uint4*>(row + j * 16);
                float f[8] = {bf_lo(v.x), bf_hi(v.x), bf_lo(v.y), bf_hi(v.y), bf_lo(v.z), bf_hi(v.z), bf_lo(v.w), bf_hi(v.w)};
                if (split) {
                    const uint4 l = *reinterpret_cast<const uint4*>(row_lo + j * 16);
                    f[0] += bf_lo(l.x); f[1] += bf_hi(l.x); f[2] += bf_lo(l.y); f[3] += bf_hi(l.y);
                    f[4] += bf_lo(l.z); f[5] += bf_hi(l.z); f[6] += bf_lo(l.w); f[7] += bf_hi(l.w);
                }
                const float* w = p.w + j * 8 * NOUT;
#pragma unroll
                for (int i = 0; i < 8; ++i)
#pragma unroll
                    for (int n = 0; n < NOUT; ++n) acc[n] = fmaf(f[i], w[i * NOUT + n], acc[n]);
            };
            if constexpr (USED > 0) {
#pragma unroll
                for (int j = 0; j < USED; ++j) chunk(j);
            } else {
                for (int j = 0; j < used; ++j) chunk(j);
            }
            // v = acc (bias included); tanh on the x_mean columns
            float vb[8];                                 // the bf16 columns (first tanh_cols outputs), rest zero
#pragma unroll
            for (int n = 0; n < 8; ++n) vb[n] = 0.f;
#pragma unroll
            for (int n = 0; n < NOUT; ++n) {
                if (n < p.tanh_cols) { acc[n] = tanhf(acc[n]); vb[n] = acc[n]; }
            }
            const long long pix = p0 + t;
            if (p.out_f32) {
                float* o = p.out_f32 + pix * p.ld_f32;
                if ((NOUT % 2 == 0) && (p.ld_f32 % 2 == 0)) {
#pragma unroll
                    for (int n = 0; n < NOUT / 2 * 2; n += 2) *reinterpret_cast<float2*>(o + n) = make_float2(acc[n], acc[n + 1]);
                } else {
#pragma unroll
                    for (int n = 0; n < NOUT; ++n) o[n] = acc[n];
                }
            }
            const int och = p.ldo >> 3;                  // 16-byte chunks per bf16 output row
            auto store_row = [&](void* base, const float (&v)[8]) {
                if (!base) return;
                uint4* o = reinterpret_cast<uint4*>(static_cast<uint8_t*>(base) + pix * p.ldo * 2);
                o[0] = make_uint4(pack2(v[0], v[1]), pack2(v[2], v[3]), pack2(v[4], v[5]), pack2(v[6], v[7]));
                for (int c = 1; c < och; ++c) o[c] = make_uint4(0u, 0u, 0u, 0u);
            };
            float va[8];
#pragma unroll
            for (int n = 0; n < 8; ++n) va[n] = fmaxf(vb[n], p.alpha * vb[n]);
            store_row(p.raw, vb);
            store_row(p.act, va);
            if (p.raw_lo || p.act_lo) {
                float rl[8], al[8];
#pragma unroll
                for (int n = 0; n < 8; ++n) { rl[n] = vb[n] - bf_round(vb[n]); al[n] = va[n] - bf_round(va[n]); }
                store_row(p.raw_lo, rl);
                store_row(p.act_lo, al);
            }
        }
        __syncthreads();
    }
}

}  // namespace

int launch_pixel_heads(const HeadsParams& p, long long pixels, void* stream) {
    if (pixels <= 0) return 0;
    if (p.cp % 16 || p.cin > p.cp || p.nout < 1 || p.nout > 8 || p.tanh_cols > p.nout || p.ldo % 8 ||
        ((p.cin + 7) / 8) * 8 * p.nout > kHeadsMaxW)
        return (int)cudaErrorInvalidValue;
    const int used = (p.cin + 7) / 8;
    void (*kern)(HeadsParams, long long) = nullptr;
#define APVAE_HEADS_CASE(N, U) if (p.nout == N && used == U) kern = pixel_heads_kernel<N, U>;
    APVAE_HEADS_CASE(6, 12)          // arch A: 96 -> x_mean | x_logstd
    APVAE_HEADS_CASE(3, 4)           // arch B: 32 -> x_mean
    APVAE_HEADS_CASE(3, 8)
    APVAE_HEADS_CASE(6, 4)
    APVAE_HEADS_CASE(1, 4)           // arch M
    APVAE_HEADS_CASE(1, 2)
#undef APVAE_HEADS_CASE
    if (!kern) {
        switch (p.nout) {
            case 1: kern = pixel_heads_kernel<1, 0>; break;
            case 2: kern = pixel_heads_kernel<2, 0>; break;
            case 3: kern = pixel_heads_kernel<3, 0>; break;
            case 4: kern = pixel_heads_kernel<4, 0>; break;
            case 5: kern = pixel_heads_kernel<5, 0>; break;
            case 6: kern = pixel_heads_kernel<6, 0>; break;
            case 7: kern = pixel_heads_kernel<7, 0>; break;
            default: kern = pixel_heads_kernel<8, 0>; break;
        }
    }
    const int pitch = p.cp * 2 + 16;
    const int smem = kTile * pitch * (p.in_lo ? 2 : 1);
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (err != cudaSuccess) return (int)err;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int per_sm = (227 * 1024) / (smem + 1024);
    if (per_sm > 16) per_sm = 16;
    if (per_sm < 1) per_sm = 1;
    const long long tiles = (pixels + kTile - 1) / kTile;
    const long long cap = (long long)sms * per_sm;
    const unsigned grid = (unsigned)(tiles < cap ? tiles : cap);
    kern<<<grid, kTile, smem, static_cast<cudaStream_t>(stream)>>>(p, pixels);
    return (int)cudaGetLastError();
}

}  // namespace apvae
