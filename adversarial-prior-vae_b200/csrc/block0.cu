// resnet_conv2d_block with <= 3 input channels on the CUDA cores (see kernels.h: SmallBlockParams).
//
// Reference: tfsnippet/layers/convolutional/resnet.py:145-267 (pre-activation residual block: activation, conv_0,
// activation, conv_1, plus the 1x1 shortcut conv when channels change) as used by D_psi/level_0 in
// experiment/dmm/dmm_cifar10.py:547-549.
//
// Roofline: HBM.  Per image the kernel reads H*W*cp_in bf16 (32 KB) and writes r (32 KB) + raw and act outputs
// (2 x 32 KB): 128 KB per image; the 2 * 0.57 MFLOP per image are CUDA-core FMAs with the weights as constant-bank
// operands (kernel parameters).
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "kernels.h"
#include "op16.h"

namespace apvae {
namespace {

constexpr int kThreads = 256;

// CIN input channels, COUT outputs, 3x3 taps, one image per CTA (H, W <= 32).  Shared memory: lrelu(x), x and r on a
// zero-bordered (H+2) x (W+2) grid, channel-planar fp32.
template <int CIN, int COUT, bool F16>
__global__ void __launch_bounds__(kThreads) small_block_kernel(const __grid_constant__ SmallBlockParams p, int n_images) {
    constexpr int PW = kSmallBlockMaxHW + 2;
    __shared__ float xa[CIN][PW * PW];        // lrelu(x), zero border
    __shared__ float xr[CIN][PW * PW];        // x (interior only is read)
    __shared__ float rr[CIN][PW * PW];        // r = bf16(lrelu(conv_0 + b0)), zero border
    const int H = p.H, W = p.W, HW = H * W;
    const int t = threadIdx.x;
    for (int img = blockIdx.x; img < n_images; img += gridDim.x) {
        for (int i = t; i < CIN * PW * PW; i += kThreads) { (&xa[0][0])[i] = 0.f; (&rr[0][0])[i] = 0.f; }
        __syncthreads();
        const uint8_t* src = static_cast<const uint8_t*>(p.in_raw) + (size_t)img * HW * p.cp_in * 2;
        for (int px = t; px < HW; px += kThreads) {
            const uint2 v = *reinterpret_cast<const uint2*>(src + (size_t)px * p.cp_in * 2);     // first 4 channels
            const float c[4] = {op16::lo<F16>(v.x), op16::hi<F16>(v.x), op16::lo<F16>(v.y), op16::hi<F16>(v.y)};
            const int y = px / W, x = px - y * W, o = (y + 1) * PW + x + 1;
#pragma unroll
            for (int ci = 0; ci < CIN; ++ci) { xr[ci][o] = c[ci]; xa[ci][o] = fmaxf(c[ci], p.alpha * c[ci]); }
        }
        __syncthreads();
        // conv_0: CIN -> CIN, r rounded to bf16; also written out (bf16 rows of cp_in channels, padding zero)
        uint8_t* rdst = static_cast<uint8_t*>(p.r_act) + (size_t)img * HW * p.cp_in * 2;
        const int rch = p.cp_in >> 3;
        for (int px = t; px < HW; px += kThreads) {
            const int y = px / W, x = px - y * W, o = y * PW + x;       // top-left of the 3x3 window in the padded grid
            float acc[CIN];
#pragma unroll
            for (int co = 0; co < CIN; ++co) acc[co] = p.b0[co];
#pragma unroll
            for (int di = 0; di < 3; ++di)
#pragma unroll
                for (int dj = 0; dj < 3; ++dj)
#pragma unroll
                    for (int ci = 0; ci < CIN; ++ci) {
                        const float a = xa[ci][o + di * PW + dj];
#pragma unroll
                        for (int co = 0; co < CIN; ++co) acc[co] = fmaf(a, p.w0[((di * 3 + dj) * CIN + ci) * CIN + co], acc[co]);
                    }
            float rv[4] = {0.f, 0.f, 0.f, 0.f};          // CIN <= 3 values, the rest of the first 4 channels zero
#pragma unroll
            for (int co = 0; co < CIN; ++co) {
                rv[co] = op16::round<F16>(fmaxf(acc[co], p.alpha * acc[co]));
                rr[co][o + PW + 1] = rv[co];
            }
            if (p.r_act) {
                uint4* d = reinterpret_cast<uint4*>(rdst + (size_t)px * p.cp_in * 2);
                d[0] = make_uint4(op16::pack<F16>(rv[0], rv[1]), op16::pack<F16>(rv[2], rv[3]), 0u, 0u);
                for (int c = 1; c < rch; ++c) d[c] = make_uint4(0u, 0u, 0u, 0u);
            }
        }
        __syncthreads();
        // conv_1: CIN -> COUT, + 1x1 shortcut on x
        uint8_t* oraw = static_cast<uint8_t*>(p.out_raw) + (size_t)img * HW * p.ldo * 2;
        uint8_t* oact = static_cast<uint8_t*>(p.out_act) + (size_t)img * HW * p.ldo * 2;
        const int och = p.ldo >> 3;
        for (int px = t; px < HW; px += kThreads) {
            const int y = px / W, x = px - y * W, o = y * PW + x;
            float acc[COUT];
#pragma unroll
            for (int co = 0; co < COUT; ++co) acc[co] = p.b1[co];
#pragma unroll
            for (int di = 0; di < 3; ++di)
#pragma unroll
                for (int dj = 0; dj < 3; ++dj)
#pragma unroll
                    for (int ci = 0; ci < CIN; ++ci) {
                        const float a = rr[ci][o + di * PW + dj];
#pragma unroll
                        for (int co = 0; co < COUT; ++co) acc[co] = fmaf(a, p.w1[((di * 3 + dj) * CIN + ci) * COUT + co], acc[co]);
                    }
#pragma unroll
            for (int ci = 0; ci < CIN; ++ci) {
                const float a = xr[ci][o + PW + 1];
#pragma unroll
                for (int co = 0; co < COUT; ++co) acc[co] = fmaf(a, p.ws[ci * COUT + co], acc[co]);
            }
            uint32_t wr[COUT / 2], wa[COUT / 2];
#pragma unroll
            for (int co = 0; co < COUT; co += 2) {
                wr[co / 2] = op16::pack<F16>(acc[co], acc[co + 1]);
                wa[co / 2] = op16::pack<F16>(fmaxf(acc[co], p.alpha * acc[co]), fmaxf(acc[co + 1], p.alpha * acc[co + 1]));
            }
            auto store = [&](uint8_t* base, const uint32_t (&w)[COUT / 2]) {
                uint4* d = reinterpret_cast<uint4*>(base + (size_t)px * p.ldo * 2);
#pragma unroll
                for (int c = 0; c < COUT / 8; ++c) d[c] = make_uint4(w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
                for (int c = COUT / 8; c < och; ++c) d[c] = make_uint4(0u, 0u, 0u, 0u);
            };
            if (p.out_raw) store(oraw, wr);
            if (p.out_act) store(oact, wa);
        }
        __syncthreads();
    }
}


// The wider block: any H, W <= 64 (whole image in dynamic shared memory, fp32 planes with a zero border), COUT 16 or 32,
// conv_1 and the shortcut with stride 1 or 2 (TF SAME: a 3x3 stride-2 conv on an even grid pads bottom / right only, so the
// window of output (y, x) starts at input (2y, 2x)).  Split precision: the input is hi + lo (summed on load), r stays fp32
// in shared memory and every 16-bit output leaves as a pair.
constexpr int kWideThreads = 512;

template <int CIN, int COUT, bool F16>
__global__ void __launch_bounds__(kWideThreads) small_block_wide_kernel(const __grid_constant__ SmallBlockParams p, int n_images) {
    extern __shared__ float sm[];
    const int H = p.H, W = p.W, HW = H * W, PW = W + 2, plane = (H + 2) * PW, s = p.stride;
    const int Ho = (H + s - 1) / s, Wo = (W + s - 1) / s, HWo = Ho * Wo;
    float* const xa = sm;                     // lrelu(x), zero border          [CIN][plane]
    float* const xr = sm + CIN * plane;       // x (interior only is read)
    float* const rr = sm + 2 * CIN * plane;   // r = lrelu(conv_0 + b0), zero border
    const bool split = p.in_raw_lo != nullptr;
    const int t = threadIdx.x;
    const int o1 = (s == 1) ? 0 : PW + 1;     // window origin of conv_1 in the padded grid, relative to (s y, s x)
    for (int img = blockIdx.x; img < n_images; img += gridDim.x) {
        for (int i = t; i < CIN * plane; i += kWideThreads) { xa[i] = 0.f; rr[i] = 0.f; }
        __syncthreads();
        const uint8_t* src = static_cast<const uint8_t*>(p.in_raw) + (size_t)img * HW * p.cp_in * 2;
        const uint8_t* src_lo = split ? static_cast<const uint8_t*>(p.in_raw_lo) + (size_t)img * HW * p.cp_in * 2 : nullptr;
        for (int px = t; px < HW; px += kWideThreads) {
            const uint2 v = *reinterpret_cast<const uint2*>(src + (size_t)px * p.cp_in * 2);
            float c[4] = {op16::lo<F16>(v.x), op16::hi<F16>(v.x), op16::lo<F16>(v.y), op16::hi<F16>(v.y)};
            if (split) {
                const uint2 l = *reinterpret_cast<const uint2*>(src_lo + (size_t)px * p.cp_in * 2);
                c[0] += op16::lo<F16>(l.x); c[1] += op16::hi<F16>(l.x); c[2] += op16::lo<F16>(l.y); c[3] += op16::hi<F16>(l.y);
            }
            const int y = px / W, x = px - y * W, o = (y + 1) * PW + x + 1;
#pragma unroll
            for (int ci = 0; ci < CIN; ++ci) { xr[ci * plane + o] = c[ci]; xa[ci * plane + o] = fmaxf(c[ci], p.alpha * c[ci]); }
        }
        __syncthreads();
        // conv_0: CIN -> CIN at the input resolution
        const int rch = p.cp_in >> 3;
        for (int px = t; px < HW; px += kWideThreads) {
            const int y = px / W, x = px - y * W, o = y * PW + x;
            float acc[CIN];
#pragma unroll
            for (int co = 0; co < CIN; ++co) acc[co] = p.b0[co];
#pragma unroll
            for (int di = 0; di < 3; ++di)
#pragma unroll
                for (int dj = 0; dj < 3; ++dj)
#pragma unroll
                    for (int ci = 0; ci < CIN; ++ci) {
                        const float a = xa[ci * plane + o + di * PW + dj];
#pragma unroll
                        for (int co = 0; co < CIN; ++co) acc[co] = fmaf(a, p.w0[((di * 3 + dj) * CIN + ci) * CIN + co], acc[co]);
                    }
            float rv[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int co = 0; co < CIN; ++co) {
                const float a = fmaxf(acc[co], p.alpha * acc[co]);
                rv[co] = split ? a : op16::round<F16>(a);           // plain block: what the unfused ops feed to conv_1
                rr[co * plane + o + PW + 1] = rv[co];
            }
            if (p.r_act) {
                const uint32_t h0 = op16::pack<F16>(rv[0], rv[1]), h1 = op16::pack<F16>(rv[2], rv[3]);
                uint4* d = reinterpret_cast<uint4*>(static_cast<uint8_t*>(p.r_act) + ((size_t)img * HW + px) * p.cp_in * 2);
                d[0] = make_uint4(h0, h1, 0u, 0u);
                for (int c = 1; c < rch; ++c) d[c] = make_uint4(0u, 0u, 0u, 0u);
                if (p.r_act_lo) {
                    uint4* dl = reinterpret_cast<uint4*>(static_cast<uint8_t*>(p.r_act_lo) + ((size_t)img * HW + px) * p.cp_in * 2);
                    dl[0] = make_uint4(op16::pack<F16>(rv[0] - op16::lo<F16>(h0), rv[1] - op16::hi<F16>(h0)),
                                       op16::pack<F16>(rv[2] - op16::lo<F16>(h1), rv[3] - op16::hi<F16>(h1)), 0u, 0u);
                    for (int c = 1; c < rch; ++c) dl[c] = make_uint4(0u, 0u, 0u, 0u);
                }
            }
        }
        __syncthreads();
        // conv_1: CIN -> COUT with stride s, + the 1x1 stride-s shortcut on x
        const int och = p.ldo >> 3;
        for (int px = t; px < HWo; px += kWideThreads) {
            const int y = px / Wo, x = px - y * Wo, o = (s * y) * PW + s * x + o1;
            float acc[COUT];
#pragma unroll
            for (int co = 0; co < COUT; ++co) acc[co] = p.b1[co];
#pragma unroll
            for (int di = 0; di < 3; ++di)
#pragma unroll
                for (int dj = 0; dj < 3; ++dj)
#pragma unroll
                    for (int ci = 0; ci < CIN; ++ci) {
                        const float a = rr[ci * plane + o + di * PW + dj];
#pragma unroll
                        for (int co = 0; co < COUT; ++co) acc[co] = fmaf(a, p.w1[((di * 3 + dj) * CIN + ci) * COUT + co], acc[co]);
                    }
#pragma unroll
            for (int ci = 0; ci < CIN; ++ci) {
                const float a = xr[ci * plane + (s * y + 1) * PW + s * x + 1];
#pragma unroll
                for (int co = 0; co < COUT; ++co) acc[co] = fmaf(a, p.ws[ci * COUT + co], acc[co]);
            }
            const size_t off = ((size_t)img * HWo + px) * p.ldo * 2;
            auto store = [&](void* hi_base, void* lo_base, bool act) {
                uint32_t wh[COUT / 2], wl[COUT / 2];
#pragma unroll
                for (int co = 0; co < COUT; co += 2) {
                    const float a0 = act ? fmaxf(acc[co], p.alpha * acc[co]) : acc[co];
                    const float a1 = act ? fmaxf(acc[co + 1], p.alpha * acc[co + 1]) : acc[co + 1];
                    wh[co / 2] = op16::pack<F16>(a0, a1);
                    wl[co / 2] = op16::pack<F16>(a0 - op16::lo<F16>(wh[co / 2]), a1 - op16::hi<F16>(wh[co / 2]));
                }
                uint4* d = reinterpret_cast<uint4*>(static_cast<uint8_t*>(hi_base) + off);
#pragma unroll
                for (int c = 0; c < COUT / 8; ++c) d[c] = make_uint4(wh[4 * c], wh[4 * c + 1], wh[4 * c + 2], wh[4 * c + 3]);
                for (int c = COUT / 8; c < och; ++c) d[c] = make_uint4(0u, 0u, 0u, 0u);
                if (lo_base) {
                    uint4* dl = reinterpret_cast<uint4*>(static_cast<uint8_t*>(lo_base) + off);
#pragma unroll
                    for (int c = 0; c < COUT / 8; ++c) dl[c] = make_uint4(wl[4 * c], wl[4 * c + 1], wl[4 * c + 2], wl[4 * c + 3]);
                    for (int c = COUT / 8; c < och; ++c) dl[c] = make_uint4(0u, 0u, 0u, 0u);
                }
            };
            if (p.out_raw) store(p.out_raw, p.out_raw_lo, false);
            if (p.out_act) store(p.out_act, p.out_act_lo, true);
        }
        __syncthreads();
    }
}

template <int COUT>
void (*pick_wide(int cin, bool f16))(SmallBlockParams, int) {
    switch (cin) {
        case 1: return f16 ? small_block_wide_kernel<1, COUT, true> : small_block_wide_kernel<1, COUT, false>;
        case 2: return f16 ? small_block_wide_kernel<2, COUT, true> : small_block_wide_kernel<2, COUT, false>;
        default: return f16 ? small_block_wide_kernel<3, COUT, true> : small_block_wide_kernel<3, COUT, false>;
    }
}


// Adjoint of the block (kernels.h: SmallBlockGradParams).  One image per CTA; g_u on a zero-bordered fp32 grid in dynamic
// shared memory.  Pixels are walked as [row][column parity][column / 2], so that with stride 2 the taps a warp needs
// (di = y mod 2 (+2), dj = x mod 2 (+2)) are warp-uniform and the weights stay uniform constant-bank loads.
template <int CIN, int COUT, bool F16>
__global__ void __launch_bounds__(kWideThreads) small_block_grad_kernel(const __grid_constant__ SmallBlockGradParams p, int n_images) {
    extern __shared__ float sm[];
    const int H = p.H, W = p.W, HW = H * W, PW = W + 2, plane = (H + 2) * PW, s = p.stride;
    const int Ho = H / s, Wo = W / s, HWo = Ho * Wo, halfW = W >> 1;
    float* const gu = sm;                                       // [CIN][plane]
    const int t = threadIdx.x;
    const int p1 = (s == 1) ? 1 : 0;                            // pad_lo of conv_1 (TF SAME)
    auto load_row = [&](const uint8_t* row, float (&g)[COUT]) {
#pragma unroll
        for (int c = 0; c < COUT / 8; ++c) {
            const uint4 v = *reinterpret_cast<const uint4*>(row + c * 16);
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) { g[c * 8 + 2 * j] = op16::lo<F16>(w[j]); g[c * 8 + 2 * j + 1] = op16::hi<F16>(w[j]); }
        }
    };
    for (int img = blockIdx.x; img < n_images; img += gridDim.x) {
        for (int i = t; i < CIN * plane; i += kWideThreads) gu[i] = 0.f;
        __syncthreads();
        const uint8_t* go = static_cast<const uint8_t*>(p.g_out) + (size_t)img * HWo * p.ldg * 2;
        const uint8_t* rm = static_cast<const uint8_t*>(p.r_act) + (size_t)img * HW * p.cp_in * 2;
        const uint8_t* xm = static_cast<const uint8_t*>(p.x_in) + (size_t)img * HW * p.cp_in * 2;
        // g_u = conv_1^T(g_out) * lrelu'(r)
        for (int q = t; q < HW; q += kWideThreads) {
            const int y = q / W, rem = q - y * W, par = rem / halfW, x = 2 * (rem - par * halfW) + par;
            float acc[CIN];
#pragma unroll
            for (int ci = 0; ci < CIN; ++ci) acc[ci] = 0.f;
            // output (oy, ox) reads input (s oy + di - p1, s ox + dj - p1): this pixel receives from the taps with
            // (y + p1 - di) and (x + p1 - dj) divisible by s
            for (int di = (s == 1) ? 0 : ((y + p1) & 1); di < 3; di += s) {
                const int ty = y + p1 - di;
                if (ty < 0 || ty >= H) continue;
                for (int dj = (s == 1) ? 0 : ((x + p1) & 1); dj < 3; dj += s) {
                    const int tx = x + p1 - dj;
                    if (tx < 0 || tx >= W) continue;
                    float g[COUT];
                    load_row(go + ((size_t)(ty / s) * Wo + tx / s) * p.ldg * 2, g);
                    const float* w = p.w1 + (di * 3 + dj) * CIN * COUT;
#pragma unroll
                    for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
                        for (int co = 0; co < COUT; ++co) acc[ci] = fmaf(g[co], w[ci * COUT + co], acc[ci]);
                }
            }
            const uint2 m = *reinterpret_cast<const uint2*>(rm + ((size_t)y * W + x) * p.cp_in * 2);
            const bool pos[4] = {op16::pos_lo(m.x), op16::pos_hi(m.x), op16::pos_lo(m.y), op16::pos_hi(m.y)};
#pragma unroll
            for (int ci = 0; ci < CIN; ++ci) gu[ci * plane + (y + 1) * PW + x + 1] = acc[ci] * (pos[ci] ? 1.f : p.alpha);
        }
        __syncthreads();
        // g_x = conv_0^T(g_u) * lrelu'(x) + shortcut^T(g_out)
        const int och = p.ldo >> 3;
        for (int q = t; q < HW; q += kWideThreads) {
            const int y = q / W, x = q - y * W;
            float acc[CIN];
#pragma unroll
            for (int ci = 0; ci < CIN; ++ci) acc[ci] = 0.f;
            // u(y', x', co) reads a(y' + di - 1, x' + dj - 1, ci): a(y, x) receives from u(y - di + 1, x - dj + 1)
#pragma unroll
            for (int di = 0; di < 3; ++di)
#pragma unroll
                for (int dj = 0; dj < 3; ++dj)
#pragma unroll
                    for (int co = 0; co < CIN; ++co) {
                        const float g = gu[co * plane + (y - di + 2) * PW + x - dj + 2];
#pragma unroll
                        for (int ci = 0; ci < CIN; ++ci) acc[ci] = fmaf(g, p.w0[((di * 3 + dj) * CIN + ci) * CIN + co], acc[ci]);
                    }
            const uint2 m = *reinterpret_cast<const uint2*>(xm + ((size_t)y * W + x) * p.cp_in * 2);
            const bool pos[4] = {op16::pos_lo(m.x), op16::pos_hi(m.x), op16::pos_lo(m.y), op16::pos_hi(m.y)};
#pragma unroll
            for (int ci = 0; ci < CIN; ++ci) acc[ci] *= pos[ci] ? 1.f : p.alpha;
            if (!(y % s) && !(x % s)) {
                float g[COUT];
                load_row(go + ((size_t)(y / s) * Wo + x / s) * p.ldg * 2, g);
#pragma unroll
                for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
                    for (int co = 0; co < COUT; ++co) acc[ci] = fmaf(g[co], p.ws[ci * COUT + co], acc[ci]);
            }
            const size_t pix = (size_t)img * HW + q;
            if (p.gx_f32) {
#pragma unroll
                for (int ci = 0; ci < CIN; ++ci) p.gx_f32[pix * p.ld_f32 + ci] = acc[ci];
            }
            if (p.gx_raw) {
                float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int ci = 0; ci < CIN; ++ci) v[ci] = acc[ci];
                uint4* d = reinterpret_cast<uint4*>(static_cast<uint8_t*>(p.gx_raw) + pix * p.ldo * 2);
                d[0] = make_uint4(op16::pack<F16>(v[0], v[1]), op16::pack<F16>(v[2], v[3]), 0u, 0u);
                for (int c = 1; c < och; ++c) d[c] = make_uint4(0u, 0u, 0u, 0u);
            }
        }
        __syncthreads();
    }
}

template <int COUT>
void (*pick_grad(int cin, bool f16))(SmallBlockGradParams, int) {
    switch (cin) {
        case 1: return f16 ? small_block_grad_kernel<1, COUT, true> : small_block_grad_kernel<1, COUT, false>;
        case 2: return f16 ? small_block_grad_kernel<2, COUT, true> : small_block_grad_kernel<2, COUT, false>;
        default: return f16 ? small_block_grad_kernel<3, COUT, true> : small_block_grad_kernel<3, COUT, false>;
    }
}

}  // namespace

int launch_small_block(const SmallBlockParams& p, int n, void* stream) {
    if (n <= 0) return 0;
    const int stride = p.stride ? p.stride : 1;
    const bool split = p.in_raw_lo || p.r_act_lo || p.out_raw_lo || p.out_act_lo;
    if (p.k != 3 || p.cp_in % 8 || p.ldo % 8 || p.cin < 1 || p.cin > kSmallBlockMaxCin || (p.cout != 16 && p.cout != 32) ||
        p.ldo < p.cout || !p.in_raw || (stride != 1 && stride != 2) || p.H > kSmallBlockWideMaxHW || p.W > kSmallBlockWideMaxHW)
        return (int)cudaErrorInvalidValue;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (p.cout == 16 && stride == 1 && !split && p.H <= kSmallBlockMaxHW && p.W <= kSmallBlockMaxHW) {
        void (*kern)(SmallBlockParams, int) = nullptr;
        switch (p.cin) {
            case 1: kern = p.f16 ? small_block_kernel<1, 16, true> : small_block_kernel<1, 16, false>; break;
            case 2: kern = p.f16 ? small_block_kernel<2, 16, true> : small_block_kernel<2, 16, false>; break;
            default: kern = p.f16 ? small_block_kernel<3, 16, true> : small_block_kernel<3, 16, false>; break;
        }
        const int cap = sms * 4;
        kern<<<n < cap ? n : cap, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(p, n);
        return (int)cudaGetLastError();
    }
    if (split && !p.in_raw_lo) return (int)cudaErrorInvalidValue;
    SmallBlockParams q = p;
    q.stride = stride;
    void (*kern)(SmallBlockParams, int) = p.cout == 16 ? pick_wide<16>(p.cin, p.f16 != 0) : pick_wide<32>(p.cin, p.f16 != 0);
    const size_t smem = (size_t)3 * p.cin * (p.H + 2) * (p.W + 2) * sizeof(float);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    const int per_sm = smem > 110 * 1024 ? 1 : smem > 72 * 1024 ? 2 : smem > 54 * 1024 ? 3 : 4;
    const int cap = sms * per_sm;
    kern<<<n < cap ? n : cap, kWideThreads, smem, static_cast<cudaStream_t>(stream)>>>(q, n);
    return (int)cudaGetLastError();
}

int launch_small_block_grad(const SmallBlockGradParams& p, int n, void* stream) {
    if (n <= 0) return 0;
    if (p.cp_in % 8 || p.ldg % 8 || p.cin < 1 || p.cin > kSmallBlockMaxCin || (p.cout != 16 && p.cout != 32) || p.ldg < p.cout ||
        !p.g_out || !p.r_act || !p.x_in || (p.stride != 1 && p.stride != 2) || p.H > kSmallBlockWideMaxHW || p.W > kSmallBlockWideMaxHW ||
        (p.H % p.stride) || (p.W % 2) || (p.gx_raw && p.ldo % 8) || (!p.gx_f32 && !p.gx_raw))
        return (int)cudaErrorInvalidValue;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    void (*kern)(SmallBlockGradParams, int) = p.cout == 16 ? pick_grad<16>(p.cin, p.f16 != 0) : pick_grad<32>(p.cin, p.f16 != 0);
    const size_t smem = (size_t)p.cin * (p.H + 2) * (p.W + 2) * sizeof(float);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    const int per_sm = smem > 110 * 1024 ? 1 : smem > 72 * 1024 ? 2 : smem > 54 * 1024 ? 3 : 4;
    const int cap = sms * per_sm;
    kern<<<n < cap ? n : cap, kWideThreads, smem, static_cast<cudaStream_t>(stream)>>>(p, n);
    return (int)cudaGetLastError();
}

}  // namespace apvae
