// Device-side sm_100a primitives shared by the tensor-core kernels: mbarrier, TMA
// (cp.async.bulk.tensor), tcgen05 (MMA / commit / TMEM load), cluster helpers, UMMA descriptors
// and the common epilogue (bias -> lrelu' mask -> residual -> tanh -> fp32 / bf16 stores).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdio.h>
#include "tapgemm.h"

namespace apvae {
namespace ptx {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Bounded wait: a pipeline bug must become a trap, never a hung GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    long long t0 = 0;
    for (uint32_t it = 0;; ++it) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (done) return;
        if (it == 64) t0 = clock64();
        if (it > 64 && (it & 1023) == 0 && clock64() - t0 > 4000000000LL) {
            printf("apvae: mbarrier timeout block %d thread %d bar %u parity %u\n", (int)blockIdx.x,
                   (int)threadIdx.x, bar, parity);
            __trap();
        }
    }
}
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void tma_load_5d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                            int c2, int c3, int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
// Multicast: the box lands at the same smem offset in every CTA of `mask`, and completes the
// transaction on the mbarrier at the same offset in each of them.
__device__ __forceinline__ void tma_load_2d_mc(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                               uint16_t mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
        " [%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "h"(mask) : "memory");
}
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// Arrives on the mbarrier at this offset in every CTA of `mask` once the prior MMAs retire.
__device__ __forceinline__ void umma_commit_mc(uint32_t bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"(mask) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// K-major shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
// start>>4 [0,14) | LBO>>4 [16,30) (unused: one swizzle atom along K) |
// SBO>>4 [32,46) = 8 rows * row bytes | version=1 [46,48) | layout [61,64).
template <int KC>
__device__ __forceinline__ uint64_t make_kmajor_desc(uint32_t saddr) {
    constexpr uint64_t row_bytes = KC * 2;
    constexpr uint64_t layout = (KC == 64) ? 2 : (KC == 32) ? 4 : 6;   // SW128 / SW64 / SW32
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= ((8 * row_bytes) >> 4) << 32;
    d |= 1ull << 46;
    d |= layout << 61;
    return d;
}
// kind::f16 instruction descriptor: D = f32, A = B = bf16, both K-major, M x N.
__device__ __forceinline__ uint32_t make_idesc_bf16(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xFFFF0000u); }

// Epilogue for 16 accumulator columns [col, col+16) of one output pixel:
//   v = acc + bias ; v *= lrelu'(mask) ; v += res ; tanh on cols < tanh_cols ; fp32 / bf16 stores.
__device__ __forceinline__ void epilogue16(const EpilogueArgs& e, const uint32_t (&v)[16], int col, size_t opix,
                                           size_t rpix, bool has_res) {
    float f[16];
    const float4* b4 = reinterpret_cast<const float4*>(e.bias + col);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float4 b = __ldg(b4 + i);
        f[4 * i + 0] = __uint_as_float(v[4 * i + 0]) + b.x;
        f[4 * i + 1] = __uint_as_float(v[4 * i + 1]) + b.y;
        f[4 * i + 2] = __uint_as_float(v[4 * i + 2]) + b.z;
        f[4 * i + 3] = __uint_as_float(v[4 * i + 3]) + b.w;
    }
    if (e.mask != nullptr) {
        const uint4* m4 = reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(e.mask) + opix * e.ld_mask + col);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const uint4 m = __ldg(m4 + i);
            const uint32_t w[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                f[8 * i + 2 * j + 0] *= (bf16_lo(w[j]) > 0.f) ? 1.f : e.lrelu_alpha;
                f[8 * i + 2 * j + 1] *= (bf16_hi(w[j]) > 0.f) ? 1.f : e.lrelu_alpha;
            }
        }
    }
    if (has_res) {
        const uint4* r4 = reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(e.res) + rpix * e.ld_res + col);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const uint4 m = __ldg(r4 + i);
            const uint32_t w[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                f[8 * i + 2 * j + 0] += bf16_lo(w[j]);
                f[8 * i + 2 * j + 1] += bf16_hi(w[j]);
            }
        }
    }
    if (col < e.tanh_cols) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if (col + i < e.tanh_cols) f[i] = tanhf(f[i]);
    }
    if (e.out_f32 != nullptr && col < e.f32_cols) {
        float* o = e.out_f32 + opix * e.ld_f32 + col;
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if (col + i < e.f32_cols) o[i] = f[i];
    }
    if (col + 16 > e.bf16_zero_from) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if (col + i >= e.bf16_zero_from) f[i] = 0.f;
    }
    if (e.out_raw != nullptr) {
        uint4* o = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(e.out_raw) + opix * e.ldo + col);
        o[0] = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
        o[1] = make_uint4(pack_bf16(f[8], f[9]), pack_bf16(f[10], f[11]), pack_bf16(f[12], f[13]), pack_bf16(f[14], f[15]));
    }
    if (e.out_act != nullptr) {
#pragma unroll
        for (int i = 0; i < 16; ++i) f[i] = f[i] > 0.f ? f[i] : e.lrelu_alpha * f[i];
        uint4* o = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(e.out_act) + opix * e.ldo + col);
        o[0] = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
        o[1] = make_uint4(pack_bf16(f[8], f[9]), pack_bf16(f[10], f[11]), pack_bf16(f[12], f[13]), pack_bf16(f[14], f[15]));
    }
}

}  // namespace ptx
}  // namespace apvae
