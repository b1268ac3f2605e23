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
#include "sm100_ptx.cuh"

namespace apvae {

using namespace ptx;

namespace {

constexpr int kThreads = 256;
constexpr int kAccStages = 2;
constexpr int kAccColStride = 256;     // TMEM columns between accumulator stages
constexpr int kTmemCols = 512;
constexpr uint32_t kSmemBudget = 227u * 1024u;
constexpr uint32_t kBarrierBytes = 1024;

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

// Pair mode: round g of a CTA pair = (N tile, phase, pair of pixel tiles); rank r takes pixel tile 2*pp + r.
// A pixel tile beyond the grid (odd count) decodes to n0 >= n_valid: its loads are harmless, its rows are masked.
__device__ __forceinline__ TileCoord decode_pair_tile(const TapGemmParams& p, int g, int rank, int tiles_n) {
    TileCoord t;
    t.nt = g % p.n_tiles;  g /= p.n_tiles;
    t.ph = g % p.num_phases;  g /= p.num_phases;
    int tile = 2 * g + rank;
    t.x0 = (tile % p.tiles_x) * p.bw;  tile /= p.tiles_x;
    t.y0 = (tile % p.tiles_y) * p.bh;  tile /= p.tiles_y;
    t.n0 = tile * p.bn;                                   // tile >= tiles_n: out of range
    (void)tiles_n;
    return t;
}

}  // namespace

template <int KC, bool kSplit, bool kTwo, bool kF16>
__global__ void __launch_bounds__(kThreads, 1)
tapgemm_kernel(const __grid_constant__ TapGemmParams p) {
    constexpr bool two = kTwo;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // Dynamic smem is only guaranteed 16 B aligned by the ABI: realign to 1024 B
    // (swizzle atoms of the UMMA descriptors assume it).
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t epi_base = smem_base + p.num_stages * p.stage_bytes;
    const uint32_t bar_base = epi_base + kEpiSmemBytes;
    // barrier layout: full[S], empty[S], tmem_full[2], tmem_empty[2], tmem_ptr
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (p.num_stages + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * p.num_stages + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * p.num_stages + kAccStages + a); };
    const uint32_t tmem_ptr_smem = bar_base + 8u * (2 * p.num_stages + 2 * kAccStages);

    const int warp = uniform_warp_idx();
    const int lane = threadIdx.x & 31;

    if (warp == 0 && elect_one_sync()) {
        prefetch_tmap(&p.tm_a[0]);
        prefetch_tmap(&p.tm_b);
    }
    if (warp == 1 && elect_one_sync()) {
        for (int s = 0; s < p.num_stages; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), 1);
        }
        for (int a = 0; a < kAccStages; ++a) {
            mbar_init(tfull_bar(a), 1);
            mbar_init(tempty_bar(a), two ? 8 : 4);
        }
        fence_barrier_init();
    }
    if (warp == 2) {
        if constexpr (two) tmem_alloc_2cta(tmem_ptr_smem, kTmemCols); else tmem_alloc(tmem_ptr_smem, kTmemCols);
    }
    tc_fence_before();
    __syncthreads();
    if constexpr (two) cluster_sync_all();      // the peer's barriers exist before anything is signalled across the pair
    tc_fence_after();
    const uint32_t rank = two ? cluster_ctarank() : 0u;
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_ptr_smem));
    tmem_base = __shfl_sync(0xffffffffu, tmem_base, 0);

    const int tiles_n = (p.n_valid + p.bn - 1) / p.bn;
    const int pix_tiles = tiles_n * p.tiles_y * p.tiles_x;
    // work items of this CTA: tiles (one per CTA) or, in pair mode, rounds of a CTA pair
    const int total_tiles = (two ? (pix_tiles + 1) / 2 : pix_tiles) * p.num_phases * p.n_tiles;
    const int w_first = two ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
    const int w_step = two ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    auto decode = [&](int w) { return two ? decode_pair_tile(p, w, (int)rank, tiles_n) : decode_tile(p, w); };

    // The role loops run with the whole warp converged; one elected lane issues the TMA / tcgen05
    // instructions (see elect_one_sync in sm100_ptx.cuh).
    if (warp == 0) {
        // ===================== TMA producer =====================
        int stage = 0;
        uint32_t phase = 0;
        for (int tile = w_first; tile < total_tiles; tile += w_step) {
            const TileCoord t = decode(tile);
            const int s_begin = p.phase_begin[t.ph], s_end = p.phase_begin[t.ph + 1];
            for (int s = s_begin; s < s_end; ++s) {
                TapStep st;
                if (p.inline_steps) st = p.steps_inline[s];
                else st = p.steps[s];
                mbar_wait(empty_bar(stage), phase ^ 1);
                if (elect_one_sync()) {
                    const uint32_t a_dst = smem_base + stage * p.stage_bytes;
                    const uint32_t b_dst = a_dst + kTileM * KC * 2;
                    if constexpr (!two) {
                        mbar_expect_tx(full_bar(stage), p.a_bytes + p.b_bytes);
                        tma_load_5d(a_dst, &p.tm_a[st.src], full_bar(stage), st.c0, t.x0 + st.dx, st.ph, t.y0 + st.dy, t.n0);
                        tma_load_2d(b_dst, &p.tm_b, full_bar(stage), 0, s * p.Npad + t.nt * p.NT);
                    } else {       // both CTAs' boxes complete on the leader's barrier; each CTA keeps half of the weight tile
                        const uint32_t bar = mapa_shared(full_bar(stage), 0);
                        if (rank == 0) mbar_expect_tx(full_bar(stage), 2 * p.a_bytes + p.b_bytes);
                        tma_load_5d_2cta(a_dst, &p.tm_a[st.src], bar, st.c0, t.x0 + st.dx, st.ph, t.y0 + st.dy, t.n0);
                        tma_load_2d_2cta(b_dst, &p.tm_b, bar, 0, s * p.Npad + t.nt * p.NT + (int)rank * (p.NT / 2));
                    }
                }
                __syncwarp();
                if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1 && (!two || rank == 0)) {
        // ===================== MMA issuer (pair mode: the leader CTA only) =====================
        const uint32_t idesc = make_idesc_16(two ? 2 * kTileM : kTileM, p.NT, kF16);
        int stage = 0, acc = 0;
        uint32_t phase = 0, acc_phase = 0;
        for (int tile = w_first; tile < total_tiles; tile += w_step) {
            const TileCoord t = decode(tile);
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
                if (elect_one_sync()) {
#pragma unroll
                    for (int k = 0; k < KC / 16; ++k) {
                        // +32 B per K=16 slice inside the swizzle atom (start address is >>4)
                        if constexpr (two) umma_bf16_2cta(tmem_d, da + 2 * k, db + 2 * k, idesc, (s > s_begin || k > 0) ? 1u : 0u);
                        else umma_bf16(tmem_d, da + 2 * k, db + 2 * k, idesc, (s > s_begin || k > 0) ? 1u : 0u);
                    }
                    if constexpr (two) {
                        umma_commit_2cta(empty_bar(stage), 3);
                        if (s == s_end - 1) umma_commit_2cta(tfull_bar(acc), 3);
                    } else {
                        umma_commit(empty_bar(stage));
                        if (s == s_end - 1) umma_commit(tfull_bar(acc));
                    }
                }
                __syncwarp();
                if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
            }
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
    } else if (warp >= 4) {
        // ===================== epilogue =====================
        const int q = warp & 3;                 // TMEM lane quarter this warp may read
        const int r = q * 32 + lane;            // accumulator row = pixel in tile
        const int px_x = r % p.bw;
        const int px_y = (r / p.bw) % p.bh;
        const int px_n = r / (p.bw * p.bh);
        const uint32_t wsm = epi_base + q * kEpiWarpBytes;
        const uint32_t bias_s = wsm + 2 * kEpiStageBytes;
        int acc = 0, bias_col = -1;
        uint32_t acc_phase = 0;
        auto make_row = [&](const TileCoord& t) {
            const int n = t.n0 + px_n, gy = t.y0 + px_y, gx = t.x0 + px_x;
            const bool valid = (r < p.rows) && (n < p.n_valid) && (gy < p.Hg) && (gx < p.Wg);
            const int Y = gy * p.os + (p.num_phases == 4 ? (t.ph >> 1) : 0);
            const int X = gx * p.os + (p.num_phases == 4 ? (t.ph & 1) : 0);
            EpiRow row;
            row.opix = ((size_t)n * p.Ho + Y) * p.Wo + X;
            row.rpix = row.opix;
            row.valid = valid;
            row.has_res = valid && (p.epi.res_mode != RES_NONE);
            if (p.epi.res_mode == RES_EVEN) {
                row.has_res = row.has_res && ((Y & 1) == 0) && ((X & 1) == 0);
                row.rpix = ((size_t)n * (p.Ho >> 1) + (Y >> 1)) * (p.Wo >> 1) + (X >> 1);
            }
            return row;
        };
        uint4 rv[8];
        bool pre = false;
        if (w_first < total_tiles) {
            const TileCoord t0 = decode(w_first);
            const EpiRow row0 = make_row(t0);
            if (__any_sync(0xffffffffu, row0.has_res)) {       // first tile's residual: before the accumulator is ready
                epilogue_res_issue(p.epi, row0, p.NT, t0.nt * p.NT, 0, lane, rv);
                pre = true;
            }
        }
        for (int tile = w_first; tile < total_tiles; tile += w_step) {
            const TileCoord t = decode(tile);
            if (bias_col != t.nt * p.NT) { bias_col = t.nt * p.NT; epilogue_load_bias(p.epi, bias_s, p.NT, bias_col, lane); }
            const EpiRow row = make_row(t);
            const bool more = tile + w_step < total_tiles;
            const TileCoord tn = decode(more ? tile + w_step : tile);
            const EpiRow nrow = make_row(tn);
            mbar_wait(tfull_bar(acc), acc_phase);
            tc_fence_after();
            const uint32_t taddr0 = tmem_base + acc * kAccColStride + ((uint32_t)(q * 32) << 16);
            pre = epilogue_block<kSplit, kF16>(p.epi, taddr0, p.NT, t.nt * p.NT, row, wsm, bias_s, lane, rv, pre, more ? &nrow : nullptr, tn.nt * p.NT);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) { if constexpr (two) mbar_arrive_cluster(mapa_shared(tempty_bar(acc), 0)); else mbar_arrive(tempty_bar(acc)); }
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if constexpr (two) cluster_sync_all();
    if (warp == 2) {
        if constexpr (two) tmem_dealloc_2cta(tmem_base, kTmemCols); else tmem_dealloc(tmem_base, kTmemCols);
    }
}

uint32_t tapgemm_smem_bytes(int kc, int NT, int stages) {
    uint32_t stage = kTileM * kc * 2 + NT * kc * 2;
    stage = (stage + 1023u) & ~1023u;
    return stages * stage + kEpiSmemBytes + kBarrierBytes + 1024u /* realignment slack */;
}

int tapgemm_pick_stages(int kc, int NT, int max_useful) {
    int s = 8;
    while (s > 2 && tapgemm_smem_bytes(kc, NT, s) > kSmemBudget) --s;
    if (max_useful >= 2 && s > max_useful) s = max_useful;
    return s;
}

int launch_tapgemm(const TapGemmParams& p, int num_sms, void* stream) {
    const int tiles_n = (p.n_valid + p.bn - 1) / p.bn;
    const long long pix_tiles = (long long)tiles_n * p.tiles_y * p.tiles_x;
    const bool two = p.two_cta != 0;
    const long long total = (two ? (pix_tiles + 1) / 2 : pix_tiles) * p.num_phases * p.n_tiles;
    if (total <= 0) return 0;
    const uint32_t smem = tapgemm_smem_bytes(p.kc, p.NT, p.num_stages);
    cudaError_t err;
    void (*kern)(TapGemmParams) = nullptr;
    const bool split = p.epi.out_raw_lo != nullptr || p.epi.out_act_lo != nullptr || p.epi.res_lo != nullptr;
    if (two && split) return (int)cudaErrorInvalidValue;
    const bool f16 = p.epi.f16 != 0;
#define APVAE_TAP_PICK(KC)                                                                                            \
    kern = two ? (f16 ? tapgemm_kernel<KC, false, true, true> : tapgemm_kernel<KC, false, true, false>)               \
         : split ? (f16 ? tapgemm_kernel<KC, true, false, true> : tapgemm_kernel<KC, true, false, false>)             \
                 : (f16 ? tapgemm_kernel<KC, false, false, true> : tapgemm_kernel<KC, false, false, false>)
    switch (p.kc) {
        case 16: APVAE_TAP_PICK(16); break;
        case 32: APVAE_TAP_PICK(32); break;
        case 64: APVAE_TAP_PICK(64); break;
        default: return (int)cudaErrorInvalidValue;
    }
#undef APVAE_TAP_PICK
    err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return (int)err;
    if (!two) {
        const int grid = (int)(total < num_sms ? total : num_sms);
        kern<<<grid, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(p);
        return (int)cudaGetLastError();
    }
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cfg.gridDim = dim3((num_sms / 2) * 2);
    static int cached[6] = {0, 0, 0, 0, 0, 0};               // [kc index, format] -> max co-resident pairs
    const int ki = (p.kc == 16 ? 0 : p.kc == 32 ? 1 : 2) + (f16 ? 3 : 0);
    if (cached[ki] == 0) {
        int nc = 0;
        if (cudaOccupancyMaxActiveClusters(&nc, kern, &cfg) == cudaSuccess && nc > 0) cached[ki] = nc;
        else { cached[ki] = num_sms / 2; cudaGetLastError(); }
    }
    const long long pairs = total < cached[ki] ? total : cached[ki];
    cfg.gridDim = dim3((unsigned)(pairs * 2));
    err = cudaLaunchKernelEx(&cfg, kern, p);
    return (int)err;
}

}  // namespace apvae
