// tcgen05 / TMEM / TMA implicit-GEMM kernel for sm_100a.  See tapgemm.h.
//
// Warp roles (256 threads, 1 CTA per SM, persistent over tiles):
//   warp 0      TMA producer  (lane 0 issues cp.async.bulk.tensor for A and B)
//   warp 1      MMA issuer    (lane 0 issues tcgen05.mma, commits to mbarriers)
//   warp 2      TMEM allocator
//   warps 4..7  epilogue      (tcgen05.ld -> bias/mask/residual/act -> global)
// Pipelines: smem full/empty ring (TMA <-> MMA) and a 2-deep TMEM accumulator
// ring (MMA <-> epilogue), all on mbarriers.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdio.h>
#include "tapgemm.h"

namespace apvae {

namespace {

constexpr int kThreads = 256;
constexpr int kAccStages = 2;
constexpr int kAccColStride = 256;     // TMEM columns between accumulator stages
constexpr int kTmemCols = 512;
constexpr uint32_t kSmemBudget = 227u * 1024u;
constexpr uint32_t kBarrierBytes = 1024;

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
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
                 : "memory");
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
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) return;
        if (it == 64) t0 = clock64();
        if (it > 64 && (it & 1023) == 0 && clock64() - t0 > 8000000000LL) {
            printf("apvae tapgemm: mbarrier timeout block %d thread %d bar %u parity %u\n",
                   (int)blockIdx.x, (int)threadIdx.x, bar, parity);
            __trap();
        }
    }
}

__device__ __forceinline__ void tma_load_5d(uint32_t dst, const CUtensorMap* map, uint32_t bar,
                                            int c0, int c1, int c2, int c3, int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar,
                                            int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                          uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                 ::"r"(bar) : "memory");
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
          "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]),
          "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

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

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xFFFF0000u); }

struct TileCoord {
    int n0, y0, x0, ph, nt;
};

__device__ __forceinline__ TileCoord decode_tile(const TapGemmParams& p, int tile) {
    TileCoord t;
    t.nt = tile % p.n_tiles;  tile /= p.n_tiles;
    t.ph = tile % p.num_phases;  tile /= p.num_phases;
    t.x0 = (tile % p.tiles_x) * p.bw;  tile /= p.tiles_x;
    t.y0 = (tile % p.tiles_y) * p.bh;  tile /= p.tiles_y;
    t.n0 = tile * p.bn;
    return t;
}

}  // namespace

template <int KC>
__global__ void __launch_bounds__(kThreads, 1)
tapgemm_kernel(const __grid_constant__ TapGemmParams p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // Dynamic smem is only guaranteed 16 B aligned by the ABI: realign to 1024 B
    // (swizzle atoms of the UMMA descriptors assume it).
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + p.num_stages * p.stage_bytes;
    // barrier layout: full[S], empty[S], tmem_full[2], tmem_empty[2], tmem_ptr
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (p.num_stages + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * p.num_stages + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * p.num_stages + kAccStages + a); };
    const uint32_t tmem_ptr_smem = bar_base + 8u * (2 * p.num_stages + 2 * kAccStages);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&p.tm_a[0]);
        prefetch_tmap(&p.tm_b);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < p.num_stages; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), 1);
        }
        for (int a = 0; a < kAccStages; ++a) {
            mbar_init(tfull_bar(a), 1);
            mbar_init(tempty_bar(a), 4);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(tmem_ptr_smem), "r"((uint32_t)kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_ptr_smem));

    const int tiles_n = (p.n_valid + p.bn - 1) / p.bn;
    const int total_tiles = tiles_n * p.tiles_y * p.tiles_x * p.num_phases * p.n_tiles;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                const TileCoord t = decode_tile(p, tile);
                const int s_begin = p.phase_begin[t.ph], s_end = p.phase_begin[t.ph + 1];
                for (int s = s_begin; s < s_end; ++s) {
                    const TapStep st = p.steps[s];
                    mbar_wait(empty_bar(stage), phase ^ 1);
                    const uint32_t a_dst = smem_base + stage * p.stage_bytes;
                    const uint32_t b_dst = a_dst + kTileM * KC * 2;
                    mbar_expect_tx(full_bar(stage), p.a_bytes + p.b_bytes);
                    tma_load_5d(a_dst, &p.tm_a[st.src], full_bar(stage), st.c0, t.x0 + st.dx,
                                st.ph, t.y0 + st.dy, t.n0);
                    tma_load_2d(b_dst, &p.tm_b, full_bar(stage), 0, s * p.Npad + t.nt * p.NT);
                    if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) |
                                   ((uint32_t)(p.NT >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                const TileCoord t = decode_tile(p, tile);
                const int s_begin = p.phase_begin[t.ph], s_end = p.phase_begin[t.ph + 1];
                mbar_wait(tempty_bar(acc), acc_phase ^ 1);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * kAccColStride;
                for (int s = s_begin; s < s_end; ++s) {
                    mbar_wait(full_bar(stage), phase);
                    tc_fence_after();
                    const uint32_t a_src = smem_base + stage * p.stage_bytes;
                    const uint32_t b_src = a_src + kTileM * KC * 2;
                    const uint64_t da = make_kmajor_desc<KC>(a_src);
                    const uint64_t db = make_kmajor_desc<KC>(b_src);
#pragma unroll
                    for (int k = 0; k < KC / 16; ++k) {
                        // +32 B per K=16 slice inside the swizzle atom (start address is >>4)
                        umma_bf16(tmem_d, da + 2 * k, db + 2 * k, idesc, (s > s_begin || k > 0) ? 1u : 0u);
                    }
                    umma_commit(empty_bar(stage));
                    if (s == s_end - 1) umma_commit(tfull_bar(acc));
                    if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
                }
                if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
            }
        }
        __syncwarp();
    } else if (warp >= 4) {
        // ===================== epilogue =====================
        const int q = warp & 3;                 // TMEM lane quarter this warp may read
        const int r = q * 32 + lane;            // accumulator row = pixel in tile
        const int px_x = r % p.bw;
        const int px_y = (r / p.bw) % p.bh;
        const int px_n = r / (p.bw * p.bh);
        __nv_bfloat16* out_raw = static_cast<__nv_bfloat16*>(p.out_raw);
        __nv_bfloat16* out_act = static_cast<__nv_bfloat16*>(p.out_act);
        const __nv_bfloat16* res = static_cast<const __nv_bfloat16*>(p.res);
        const __nv_bfloat16* msk = static_cast<const __nv_bfloat16*>(p.mask);
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            const TileCoord t = decode_tile(p, tile);
            const int n = t.n0 + px_n, gy = t.y0 + px_y, gx = t.x0 + px_x;
            const bool valid = (r < p.rows) && (n < p.n_valid) && (gy < p.Hg) && (gx < p.Wg);
            const int Y = gy * p.os + (p.num_phases == 4 ? (t.ph >> 1) : 0);
            const int X = gx * p.os + (p.num_phases == 4 ? (t.ph & 1) : 0);
            const size_t opix = ((size_t)n * p.Ho + Y) * p.Wo + X;
            size_t rpix = opix;
            bool has_res = valid && (p.res_mode != RES_NONE);
            if (p.res_mode == RES_EVEN) {
                has_res = has_res && ((Y & 1) == 0) && ((X & 1) == 0);
                rpix = ((size_t)n * (p.Ho >> 1) + (Y >> 1)) * (p.Wo >> 1) + (X >> 1);
            }
            mbar_wait(tfull_bar(acc), acc_phase);
            tc_fence_after();
            const uint32_t taddr0 = tmem_base + acc * kAccColStride + ((uint32_t)(q * 32) << 16);
            for (int c = 0; c < p.NT; c += 16) {
                uint32_t v[16];
                tmem_ld16(taddr0 + c, v);
                tmem_ld_wait();
                if (valid) {
                    const int col = t.nt * p.NT + c;
                    float f[16];
                    const float4* b4 = reinterpret_cast<const float4*>(p.bias + col);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float4 b = __ldg(b4 + i);
                        f[4 * i + 0] = __uint_as_float(v[4 * i + 0]) + b.x;
                        f[4 * i + 1] = __uint_as_float(v[4 * i + 1]) + b.y;
                        f[4 * i + 2] = __uint_as_float(v[4 * i + 2]) + b.z;
                        f[4 * i + 3] = __uint_as_float(v[4 * i + 3]) + b.w;
                    }
                    if (msk != nullptr) {
                        const uint4* m4 = reinterpret_cast<const uint4*>(msk + opix * p.ld_mask + col);
#pragma unroll
                        for (int i = 0; i < 2; ++i) {
                            const uint4 m = __ldg(m4 + i);
                            const uint32_t w[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                f[8 * i + 2 * j + 0] *= (bf16_lo(w[j]) > 0.f) ? 1.f : p.lrelu_alpha;
                                f[8 * i + 2 * j + 1] *= (bf16_hi(w[j]) > 0.f) ? 1.f : p.lrelu_alpha;
                            }
                        }
                    }
                    if (has_res) {
                        const uint4* r4 = reinterpret_cast<const uint4*>(res + rpix * p.ld_res + col);
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
                    if (col < p.tanh_cols) {
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            if (col + i < p.tanh_cols) f[i] = tanhf(f[i]);
                    }
                    if (p.out_f32 != nullptr && col < p.f32_cols) {
                        float* o = p.out_f32 + opix * p.ld_f32 + col;
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            if (col + i < p.f32_cols) o[i] = f[i];
                    }
                    if (col + 16 > p.bf16_zero_from) {
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            if (col + i >= p.bf16_zero_from) f[i] = 0.f;
                    }
                    if (out_raw != nullptr) {
                        uint4* o = reinterpret_cast<uint4*>(out_raw + opix * p.ldo + col);
                        o[0] = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]),
                                          pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
                        o[1] = make_uint4(pack_bf16(f[8], f[9]), pack_bf16(f[10], f[11]),
                                          pack_bf16(f[12], f[13]), pack_bf16(f[14], f[15]));
                    }
                    if (out_act != nullptr) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) f[i] = f[i] > 0.f ? f[i] : p.lrelu_alpha * f[i];
                        uint4* o = reinterpret_cast<uint4*>(out_act + opix * p.ldo + col);
                        o[0] = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]),
                                          pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
                        o[1] = make_uint4(pack_bf16(f[8], f[9]), pack_bf16(f[10], f[11]),
                                          pack_bf16(f[12], f[13]), pack_bf16(f[14], f[15]));
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(acc));
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;"
                     ::"r"(tmem_base), "r"((uint32_t)kTmemCols) : "memory");
    }
}

uint32_t tapgemm_smem_bytes(int kc, int NT, int stages) {
    uint32_t stage = kTileM * kc * 2 + NT * kc * 2;
    stage = (stage + 1023u) & ~1023u;
    return stages * stage + kBarrierBytes + 1024u /* realignment slack */;
}

int tapgemm_pick_stages(int kc, int NT, int max_useful) {
    int s = 8;
    while (s > 2 && tapgemm_smem_bytes(kc, NT, s) > kSmemBudget) --s;
    if (max_useful >= 2 && s > max_useful) s = max_useful;
    return s;
}

int launch_tapgemm(const TapGemmParams& p, int num_sms, void* stream) {
    const int tiles_n = (p.n_valid + p.bn - 1) / p.bn;
    const long long total = (long long)tiles_n * p.tiles_y * p.tiles_x * p.num_phases * p.n_tiles;
    if (total <= 0) return 0;
    const int grid = (int)(total < num_sms ? total : num_sms);
    const uint32_t smem = tapgemm_smem_bytes(p.kc, p.NT, p.num_stages);
    cudaError_t err;
    void (*kern)(TapGemmParams) = nullptr;
    switch (p.kc) {
        case 16: kern = tapgemm_kernel<16>; break;
        case 32: kern = tapgemm_kernel<32>; break;
        case 64: kern = tapgemm_kernel<64>; break;
        default: return (int)cudaErrorInvalidValue;
    }
    err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return (int)err;
    kern<<<grid, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(p);
    return (int)cudaGetLastError();
}

}  // namespace apvae
