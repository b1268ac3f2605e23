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

}  // namespace

int launch_small_block(const SmallBlockParams& p, int n, void* stream) {
    if (n <= 0) return 0;
    if (p.k != 3 || p.H > kSmallBlockMaxHW || p.W > kSmallBlockMaxHW || p.cp_in % 8 || p.ldo % 8 || p.cin < 1 ||
        p.cin > kSmallBlockMaxCin || p.cout != 16 || p.ldo < p.cout || !p.in_raw)
        return (int)cudaErrorInvalidValue;
    void (*kern)(SmallBlockParams, int) = nullptr;
    switch (p.cin) {
        case 1: kern = p.f16 ? small_block_kernel<1, 16, true> : small_block_kernel<1, 16, false>; break;
        case 2: kern = p.f16 ? small_block_kernel<2, 16, true> : small_block_kernel<2, 16, false>; break;
        default: kern = p.f16 ? small_block_kernel<3, 16, true> : small_block_kernel<3, 16, false>; break;
    }
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int cap = sms * 4;
    kern<<<n < cap ? n : cap, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(p, n);
    return (int)cudaGetLastError();
}

}  // namespace apvae
