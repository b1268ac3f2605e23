// G_theta's 1x1 output heads on the CUDA cores (see kernels.h: HeadsParams).
//
// Reference: dmm_cifar10.py:489-496 (x_mean = tanh(conv2d 1x1), x_logstd = conv2d 1x1 on the last resnet block's
// output); dmm_for_sampling.py:503-514 and dmm_bernouli_fashion.py:511-521 have the x_mean head only.
//
// Roofline: HBM.  Per pixel the kernel reads cp 16-bit channels (192 B at arch A) and writes D_psi's 16-bit input row(s)
// (32 B each); the 2 * cin * nout FLOP per pixel (1152 at arch A) are ~4 FLOP per byte.  As a 128 x 16 x 16
// tensor-core tile this op ran at 4.5x its HBM time (epilogue-bound: 16 useful columns per 128-row accumulator).
//
// Fused likelihood (HeadsParams::lik_part != null): x_mean / x_logstd never leave the registers.  The thread that owns
// a pixel scores it against x[b] (distributions/discretized.py:198-259 | Bernoulli), the tile's sum of log p and of the
// rounded-pixel squared error (test_mse, dmm_cifar10.py:889-891) go to lik_part[tile] and a tiny kernel adds the tiles
// of each sample in a fixed order (deterministic).  The fp32 heads are only written when out_f32 is set (decode /
// sample_images / the sampler's tanh').
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <math.h>
#include <stdint.h>
#include <type_traits>
#include "kernels.h"
#include "likelihood.cuh"
#include "op16.h"

namespace apvae {
namespace {

constexpr int kTile = 128;                 // pixels per tile = threads per block


// NOUT outputs per pixel; USED = ceil(cin / 8) 16-byte chunks per pixel row when known at compile time (the weight
// reads are then constant-bank operands of the FMAs with immediate offsets), 0 = run-time count.
template <int NOUT, int USED, bool F16>
__global__ void __launch_bounds__(kTile) pixel_heads_kernel(const __grid_constant__ HeadsParams p, int n_images) {
    extern __shared__ __align__(16) uint8_t sm[];
    const int chunks = p.cp >> 3;                       // 16-byte chunks per pixel row in memory
    // padded row pitch: (cp / 8 + 1) is odd, so the eight rows of a quarter-warp's 16-byte reads fall into eight
    // different bank quadruples (conflict-free LDS.128 with one thread per row)
    const int pitch = p.cp * 2 + 16;
    const bool split = p.in_lo != nullptr;
    uint8_t* const sm_lo = sm + kTile * pitch;
    const int used = USED ? USED : (p.cin + 7) >> 3;
    const int t = threadIdx.x;
    // tiles never straddle two images (a sample's likelihood is a sum over ITS tiles): tpi tiles per image, the last
    // one ragged when H * W is not a multiple of 128 (28 x 28)
    const int HW = (int)p.pixels_per_image;
    const int tpi = (HW + kTile - 1) / kTile;
    const long long tiles = (long long)n_images * tpi;
    __shared__ float red[2][kTile / 32];
    for (long long tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int img = (int)(tile / tpi), ti = (int)(tile - (long long)img * tpi);
        const long long p0 = (long long)img * HW + ti * kTile;
        const int np = (HW - ti * kTile) < kTile ? (HW - ti * kTile) : kTile;
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
        float lp_sum = 0.f, se_sum = 0.f;
        if (t < np) {
            float acc[NOUT];
#pragma unroll
            for (int n = 0; n < NOUT; ++n) acc[n] = p.bias[n];
            const uint8_t* row = sm + t * pitch;
            const uint8_t* row_lo = sm_lo + t * pitch;
            auto chunk = [&](int j) {
                const uint4 v = *reinterpret_cast<const uint4*>(row + j * 16);
                const float2 a0 = op16::unpack<F16>(v.x), a1 = op16::unpack<F16>(v.y), a2 = op16::unpack<F16>(v.z), a3 = op16::unpack<F16>(v.w);
                float f[8] = {a0.x, a0.y, a1.x, a1.y, a2.x, a2.y, a3.x, a3.y};
                if (split) {
                    const uint4 l = *reinterpret_cast<const uint4*>(row_lo + j * 16);
                    const float2 l0 = op16::unpack<F16>(l.x), l1 = op16::unpack<F16>(l.y), l2 = op16::unpack<F16>(l.z), l3 = op16::unpack<F16>(l.w);
                    f[0] += l0.x; f[1] += l0.y; f[2] += l1.x; f[3] += l1.y;
                    f[4] += l2.x; f[5] += l2.y; f[6] += l3.x; f[7] += l3.y;
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
                o[0] = make_uint4(op16::pack<F16>(v[0], v[1]), op16::pack<F16>(v[2], v[3]), op16::pack<F16>(v[4], v[5]), op16::pack<F16>(v[6], v[7]));
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
                for (int n = 0; n < 8; ++n) { rl[n] = vb[n] - op16::round<F16>(vb[n]); al[n] = va[n] - op16::round<F16>(va[n]); }
                store_row(p.raw_lo, rl);
                store_row(p.act_lo, al);
            }
            if (p.lik_part) {                            // log p(x | z) of this pixel, x_mean / x_logstd still in registers
                const int b = (p.lik_s0 + img) % p.lik_B;
                auto score = [&](auto has_std_c) {
                    constexpr bool HS = decltype(has_std_c)::value;
                    constexpr int CC = HS ? NOUT / 2 : NOUT;
                    const float* xr = p.lik_x + ((size_t)b * HW + ti * kTile + t) * CC;
#pragma unroll
                    for (int c = 0; c < CC; ++c) {
                        const float xv = __ldg(xr + c);
                        const float ls = (HS && !p.lik_uniform) ? acc[(HS ? CC : 0) + c] : p.lik_beta;
                        lp_sum += element_log_prob(p.lik, xv, acc[c], ls);
                        se_sum += element_sq_err(p.lik.bernoulli, xv, acc[c]);
                    }
                };
                if (p.lik_has_std) { if constexpr (NOUT % 2 == 0) score(std::true_type{}); }
                else score(std::false_type{});
            }
        }
        if (p.lik_part) {                                // the tile's sums: warp shuffles, then four partials through shared memory
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                lp_sum += __shfl_xor_sync(0xffffffffu, lp_sum, o);
                se_sum += __shfl_xor_sync(0xffffffffu, se_sum, o);
            }
            if ((t & 31) == 0) { red[0][t >> 5] = lp_sum; red[1][t >> 5] = se_sum; }
        }
        __syncthreads();
        if (p.lik_part && t == 0) {
            float a = 0.f, c = 0.f;
#pragma unroll
            for (int w = 0; w < kTile / 32; ++w) { a += red[0][w]; c += red[1][w]; }
            p.lik_part[2 * tile] = a;
            p.lik_part[2 * tile + 1] = c;
        }
    }
}

// log_px[s] = sum over the sample's tiles (fixed order), sq_err[s] likewise
__global__ void sum_tiles_kernel(const float* __restrict__ part, int tpi, int n, float* __restrict__ log_px,
                                 float* __restrict__ sq_err) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    float a = 0.f, c = 0.f;
    for (int i = 0; i < tpi; ++i) { a += part[2 * ((size_t)s * tpi + i)]; c += part[2 * ((size_t)s * tpi + i) + 1]; }
    if (log_px) log_px[s] = a;
    if (sq_err) sq_err[s] = c;
}

}  // namespace

int heads_tiles_per_image(long long pixels_per_image) { return (int)((pixels_per_image + kTile - 1) / kTile); }

int launch_sum_tiles(const float* part, int tpi, int n, float* log_px, float* sq_err, void* stream) {
    if (n <= 0) return 0;
    sum_tiles_kernel<<<(n + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(part, tpi, n, log_px, sq_err);
    return (int)cudaGetLastError();
}

int launch_pixel_heads(const HeadsParams& p, int n_images, void* stream) {
    if (n_images <= 0) return 0;
    if (p.cp % 16 || p.cin > p.cp || p.nout < 1 || p.nout > 8 || p.tanh_cols > p.nout || p.ldo % 8 ||
        ((p.cin + 7) / 8) * 8 * p.nout > kHeadsMaxW)
        return (int)cudaErrorInvalidValue;
    if (p.lik_part) {      // fused likelihood: the x channels are the tanh columns; x_logstd columns follow them when present
        const int cc = p.lik_has_std ? p.nout / 2 : p.nout;
        if (!p.lik_x || p.lik_B <= 0 || cc != p.tanh_cols || (p.lik_has_std && (p.nout & 1))) return (int)cudaErrorInvalidValue;
    }
    const int used = (p.cin + 7) / 8;
    void (*kern)(HeadsParams, int) = nullptr;
    const bool f16 = p.f16 != 0;
#define APVAE_HEADS_CASE(N, U) if (p.nout == N && used == U) kern = f16 ? pixel_heads_kernel<N, U, true> : pixel_heads_kernel<N, U, false>;
    APVAE_HEADS_CASE(6, 12)          // arch A: 96 -> x_mean | x_logstd
    APVAE_HEADS_CASE(3, 4)           // arch B: 32 -> x_mean
    APVAE_HEADS_CASE(1, 2)           // arch M: 16 -> x_mean
#undef APVAE_HEADS_CASE
    if (!kern) {
#define APVAE_HEADS_ANY(N) case N: kern = f16 ? pixel_heads_kernel<N, 0, true> : pixel_heads_kernel<N, 0, false>; break;
        switch (p.nout) {
            APVAE_HEADS_ANY(1) APVAE_HEADS_ANY(2) APVAE_HEADS_ANY(3) APVAE_HEADS_ANY(4)
            APVAE_HEADS_ANY(5) APVAE_HEADS_ANY(6) APVAE_HEADS_ANY(7)
            default: kern = f16 ? pixel_heads_kernel<8, 0, true> : pixel_heads_kernel<8, 0, false>; break;
        }
#undef APVAE_HEADS_ANY
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
    const long long tiles = (long long)n_images * heads_tiles_per_image(p.pixels_per_image);
    const long long cap = (long long)sms * per_sm;
    const unsigned grid = (unsigned)(tiles < cap ? tiles : cap);
    kern<<<grid, kTile, smem, static_cast<cudaStream_t>(stream)>>>(p, n_images);
    return (int)cudaGetLastError();
}

}  // namespace apvae
