// Stride-1 conv2d / conv2d_transpose with <= 4 input and <= 4 output channels on the CUDA cores (kernels.h: TinyConvParams).
//
// Reference: tfsnippet/layers/convolutional/conv2d_.py:204-211 (conv2d, SAME) and :464-471 (conv2d_transpose) as
// used by the first conv of a resnet block on the image channels (resnet.py:215-223; D_psi/level_0/conv_0 in
// dmm_for_sampling.py:556-560, dmm_bernouli_fashion.py:557-561) and by tf.gradients through it in get_sgld_proposal
// (dmm_cifar10.py:241-247).
//
// Roofline: HBM.  Per pixel: one bf16 input row read (cp_in * 2 B, reused k*k times from shared memory), the mask /
// residual rows when present, one or two bf16 output rows and cout fp32 values written.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "kernels.h"
#include "op16.h"
#include "tapgemm.h"      // ResMode

namespace apvae {
namespace {

constexpr int kTW = 32, kTH = 8, kHalo = kTinyMaxK - 1;
constexpr int kPW = kTW + kHalo, kPH = kTH + kHalo;

template <int CIN, int COUT, bool F16>
__global__ void __launch_bounds__(kTW * kTH) tiny_conv_kernel(const __grid_constant__ TinyConvParams p, int n_images) {
    __shared__ float tile[CIN][kPH * kPW];
    const int H = p.H, W = p.W;
    const int tiles_x = (W + kTW - 1) / kTW, tiles_y = (H + kTH - 1) / kTH;
    const long long tiles = (long long)n_images * tiles_y * tiles_x;
    const int t = threadIdx.x, tx = t % kTW, ty = t / kTW;
    // the taps' shifts lie in [-r, r]; the shared tile starts r pixels up / left of the output tile
    const int r = p.k / 2;
    for (long long tl = blockIdx.x; tl < tiles; tl += gridDim.x) {
        const int bx = (int)(tl % tiles_x), by = (int)((tl / tiles_x) % tiles_y), n = (int)(tl / ((long long)tiles_x * tiles_y));
        const int x0 = bx * kTW, y0 = by * kTH;
        const uint8_t* src = static_cast<const uint8_t*>(p.in) + (size_t)n * H * W * p.cp_in * 2;
        const uint8_t* src_lo = p.in_lo ? static_cast<const uint8_t*>(p.in_lo) + (size_t)n * H * W * p.cp_in * 2 : nullptr;
        for (int i = t; i < kPH * kPW; i += kTW * kTH) {
            const int py = i / kPW, px = i - py * kPW;
            const int y = y0 + py - r, x = x0 + px - r;
            float c[4] = {0.f, 0.f, 0.f, 0.f};
            if (py < kTH + 2 * r && px < kTW + 2 * r && y >= 0 && y < H && x >= 0 && x < W) {
                const uint2 v = *reinterpret_cast<const uint2*>(src + ((size_t)y * W + x) * p.cp_in * 2);
                c[0] = op16::lo<F16>(v.x); c[1] = op16::hi<F16>(v.x); c[2] = op16::lo<F16>(v.y); c[3] = op16::hi<F16>(v.y);
                if (src_lo) {
                    const uint2 l = *reinterpret_cast<const uint2*>(src_lo + ((size_t)y * W + x) * p.cp_in * 2);
                    c[0] += op16::lo<F16>(l.x); c[1] += op16::hi<F16>(l.x); c[2] += op16::lo<F16>(l.y); c[3] += op16::hi<F16>(l.y);
                }
            }
#pragma unroll
            for (int ci = 0; ci < CIN; ++ci) tile[ci][i] = c[ci];
        }
        __syncthreads();
        const int y = y0 + ty, x = x0 + tx;
        if (y < H && x < W) {
            float acc[COUT];
#pragma unroll
            for (int co = 0; co < COUT; ++co) acc[co] = p.bias[co];
            for (int tp = 0; tp < p.ntaps; ++tp) {
                const int o = (ty + r + p.dy[tp]) * kPW + tx + r + p.dx[tp];
                const float* w = p.w + tp * kTinyMaxC * kTinyMaxC;
#pragma unroll
                for (int ci = 0; ci < CIN; ++ci) {
                    const float a = tile[ci][o];
#pragma unroll
                    for (int co = 0; co < COUT; ++co) acc[co] = fmaf(a, w[ci * kTinyMaxC + co], acc[co]);
                }
            }
            const size_t pix = ((size_t)n * H + y) * W + x;
            if (p.mask) {
                const uint2 m = *reinterpret_cast<const uint2*>(static_cast<const uint8_t*>(p.mask) + pix * p.ld_mask * 2);
                const bool mv[4] = {op16::pos_lo(m.x), op16::pos_hi(m.x), op16::pos_lo(m.y), op16::pos_hi(m.y)};
#pragma unroll
                for (int co = 0; co < COUT; ++co) acc[co] *= mv[co] ? 1.f : p.alpha;
            }
            if (p.res_mode == RES_SAME || (p.res_mode == RES_EVEN && !((y | x) & 1))) {
                const size_t rp = p.res_mode == RES_SAME ? pix : ((size_t)n * (H >> 1) + (y >> 1)) * (W >> 1) + (x >> 1);
                const uint2 rv = *reinterpret_cast<const uint2*>(static_cast<const uint8_t*>(p.res) + rp * p.ld_res * 2);
                const float rr[4] = {op16::lo<F16>(rv.x), op16::hi<F16>(rv.x), op16::lo<F16>(rv.y), op16::hi<F16>(rv.y)};
#pragma unroll
                for (int co = 0; co < COUT; ++co) acc[co] += rr[co];
                if (p.res_lo) {
                    const uint2 rl = *reinterpret_cast<const uint2*>(static_cast<const uint8_t*>(p.res_lo) + rp * p.ld_res * 2);
                    const float r2[4] = {op16::lo<F16>(rl.x), op16::hi<F16>(rl.x), op16::lo<F16>(rl.y), op16::hi<F16>(rl.y)};
#pragma unroll
                    for (int co = 0; co < COUT; ++co) acc[co] += r2[co];
                }
            }
            float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int co = 0; co < COUT; ++co) v[co] = acc[co];
            if (p.out_f32) {
#pragma unroll
                for (int co = 0; co < COUT; ++co) p.out_f32[pix * p.ld_f32 + co] = v[co];
            }
            const int och = p.ldo >> 3;
            if (p.out_raw) {
                uint4* d = reinterpret_cast<uint4*>(static_cast<uint8_t*>(p.out_raw) + pix * p.ldo * 2);
                const uint32_t h0 = op16::pack<F16>(v[0], v[1]), h1 = op16::pack<F16>(v[2], v[3]);
                d[0] = make_uint4(h0, h1, 0u, 0u);
                for (int c = 1; c < och; ++c) d[c] = make_uint4(0u, 0u, 0u, 0u);
                if (p.out_raw_lo) {
                    uint4* dl = reinterpret_cast<uint4*>(static_cast<uint8_t*>(p.out_raw_lo) + pix * p.ldo * 2);
                    dl[0] = make_uint4(op16::pack<F16>(v[0] - op16::lo<F16>(h0), v[1] - op16::hi<F16>(h0)),
                                       op16::pack<F16>(v[2] - op16::lo<F16>(h1), v[3] - op16::hi<F16>(h1)), 0u, 0u);
                    for (int c = 1; c < och; ++c) dl[c] = make_uint4(0u, 0u, 0u, 0u);
                }
            }
            if (p.out_act) {
                float a[4];
#pragma unroll
                for (int co = 0; co < 4; ++co) a[co] = fmaxf(v[co], p.alpha * v[co]);
                uint4* d = reinterpret_cast<uint4*>(static_cast<uint8_t*>(p.out_act) + pix * p.ldo * 2);
                const uint32_t h0 = op16::pack<F16>(a[0], a[1]), h1 = op16::pack<F16>(a[2], a[3]);
                d[0] = make_uint4(h0, h1, 0u, 0u);
                for (int c = 1; c < och; ++c) d[c] = make_uint4(0u, 0u, 0u, 0u);
                if (p.out_act_lo) {
                    uint4* dl = reinterpret_cast<uint4*>(static_cast<uint8_t*>(p.out_act_lo) + pix * p.ldo * 2);
                    dl[0] = make_uint4(op16::pack<F16>(a[0] - op16::lo<F16>(h0), a[1] - op16::hi<F16>(h0)),
                                       op16::pack<F16>(a[2] - op16::lo<F16>(h1), a[3] - op16::hi<F16>(h1)), 0u, 0u);
                    for (int c = 1; c < och; ++c) dl[c] = make_uint4(0u, 0u, 0u, 0u);
                }
            }
        }
        __syncthreads();
    }
}

template <int CIN>
void (*pick_cout(int cout, bool f16))(TinyConvParams, int) {
    switch (cout) {
        case 1: return f16 ? tiny_conv_kernel<CIN, 1, true> : tiny_conv_kernel<CIN, 1, false>;
        case 2: return f16 ? tiny_conv_kernel<CIN, 2, true> : tiny_conv_kernel<CIN, 2, false>;
        case 3: return f16 ? tiny_conv_kernel<CIN, 3, true> : tiny_conv_kernel<CIN, 3, false>;
        default: return f16 ? tiny_conv_kernel<CIN, 4, true> : tiny_conv_kernel<CIN, 4, false>;
    }
}

}  // namespace

int launch_tiny_conv(const TinyConvParams& p, int n, void* stream) {
    if (n <= 0) return 0;
    if (p.cin < 1 || p.cin > kTinyMaxC || p.cout < 1 || p.cout > kTinyMaxC || p.k < 1 || p.k > kTinyMaxK || !(p.k & 1) ||
        p.ntaps < 1 || p.ntaps > p.k * p.k || p.cp_in % 8 || (p.ldo % 8 && (p.out_raw || p.out_act)) || !p.in ||
        (p.res_mode != RES_NONE && (!p.res || p.ld_res % 4)) || (p.mask && p.ld_mask % 4))
        return (int)cudaErrorInvalidValue;
    void (*kern)(TinyConvParams, int) = nullptr;
    switch (p.cin) {
        case 1: kern = pick_cout<1>(p.cout, p.f16 != 0); break;
        case 2: kern = pick_cout<2>(p.cout, p.f16 != 0); break;
        case 3: kern = pick_cout<3>(p.cout, p.f16 != 0); break;
        default: kern = pick_cout<4>(p.cout, p.f16 != 0); break;
    }
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long long tiles = (long long)n * ((p.H + kTH - 1) / kTH) * ((p.W + kTW - 1) / kTW);
    const long long cap = (long long)sms * 8;
    kern<<<(unsigned)(tiles < cap ? tiles : cap), kTW * kTH, 0, static_cast<cudaStream_t>(stream)>>>(p, n);
    return (int)cudaGetLastError();
}

}  // namespace apvae
