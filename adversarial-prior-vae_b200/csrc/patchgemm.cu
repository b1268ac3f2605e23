// A-stationary tcgen05 / TMEM / TMA implicit GEMM for stride-1 convolutions.  See patchgemm.h.
//
// Warp roles (256 threads, 1 CTA per SM, persistent over tiles, optional cluster of 2 / 4):
//   warp 0      activation-patch producer (TMA 5-D box, zero-filled halo)
//   warp 1      MMA issuer (tcgen05.mma; macc accumulators of 128 rows each)
//   warp 2      TMEM allocator
//   warp 3      weight-tile producer (TMA 2-D; 1/cluster of each tile, multicast to the cluster)
//   warps 4..7  epilogue (tcgen05.ld -> bias/mask/residual/act -> global)
// Rings: patches (full/empty), weight tiles (full/empty; empty collects one arrival per cluster CTA
// through tcgen05.commit multicast), and a 2-deep TMEM accumulator ring.
#include <cuda_runtime.h>
#include <string.h>
#include "patchgemm.h"
#include "sm100_ptx.cuh"

namespace apvae {

using namespace ptx;

namespace {

constexpr int kThreads = 256;
constexpr int kAccStages = 2;
constexpr int kAccColStride = 256;
constexpr int kTmemCols = 512;
constexpr uint32_t kSmemBudget = 227u * 1024u;
constexpr int kMaxStages = 12;

}  // namespace

template <int KC>
__global__ void __launch_bounds__(kThreads, 1)
patchgemm_kernel(const __grid_constant__ PatchGemmParams p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // identical offset in every CTA of the cluster (same kernel, same static layout)
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_base = smem_base;
    const uint32_t b_base = a_base + p.a_stages * p.a_slot_bytes;
    const uint32_t epi_base = b_base + p.b_stages * p.b_slot_bytes;
    const uint32_t bar_base = epi_base + kEpiSmemBytes;
    auto fullA = [&](int s) { return bar_base + 8u * s; };
    auto emptyA = [&](int s) { return bar_base + 8u * (kMaxStages + s); };
    auto fullB = [&](int s) { return bar_base + 8u * (2 * kMaxStages + s); };
    auto emptyB = [&](int s) { return bar_base + 8u * (3 * kMaxStages + s); };
    auto tfull = [&](int a) { return bar_base + 8u * (4 * kMaxStages + a); };
    auto tempty = [&](int a) { return bar_base + 8u * (4 * kMaxStages + kAccStages + a); };
    const uint32_t tmem_ptr_smem = bar_base + 8u * (4 * kMaxStages + 2 * kAccStages);

    const int warp = uniform_warp_idx();
    const int lane = threadIdx.x & 31;
    const int csize = p.cluster;
    const uint32_t rank = csize > 1 ? cluster_ctarank() : 0u;
    const uint16_t cmask = (uint16_t)((1u << csize) - 1u);

    if (warp == 0 && elect_one_sync()) {
        prefetch_tmap(&p.tm_a[0]);
        prefetch_tmap(&p.tm_a[1]);
        prefetch_tmap(&p.tm_b);
    }
    if (warp == 1 && elect_one_sync()) {
        for (int s = 0; s < p.a_stages; ++s) { mbar_init(fullA(s), 1); mbar_init(emptyA(s), 1); }
        for (int s = 0; s < p.b_stages; ++s) { mbar_init(fullB(s), 1); mbar_init(emptyB(s), csize); }
        for (int a = 0; a < kAccStages; ++a) { mbar_init(tfull(a), 1); mbar_init(tempty(a), 4); }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_ptr_smem, kTmemCols);
    tc_fence_before();
    __syncthreads();
    if (csize > 1) cluster_sync_all();      // peers' barriers are initialised before any multicast
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_ptr_smem));
    tmem_base = __shfl_sync(0xffffffffu, tmem_base, 0);

    const int total_tiles = p.n_valid * p.tiles_y;
    const int num_clusters = gridDim.x / csize;
    const int cluster_id = blockIdx.x / csize;
    const int rounds_total = (total_tiles + csize - 1) / csize;   // tile groups; every CTA of a cluster runs the same rounds
    constexpr uint32_t row_bytes = KC * 2;
    const bool skipA = (p.debug == 1 || p.debug == 3), skipB = (p.debug == 1 || p.debug == 2);

    if (warp == 0) {
        // ===================== activation-patch producer (whole warp, one lane issues) =====================
        if (!skipA) {
            int sa = 0;
            uint32_t pa = 0;
            for (int g = cluster_id; g < rounds_total; g += num_clusters) {
                const int tile = g * csize + (int)rank;
                const int n = tile / p.tiles_y;                  // tile >= total_tiles -> n >= n_valid: zero-filled / masked
                const int y0 = (tile % p.tiles_y) * p.rows_tile;
                for (int gi = 0; gi < p.num_groups; ++gi) {
                    const PatchGroup gr = p.groups[gi];
                    mbar_wait(emptyA(sa), pa ^ 1);
                    if (elect_one_sync()) {
                        mbar_expect_tx(fullA(sa), (uint32_t)gr.prow * p.W * row_bytes);
                        tma_load_5d(a_base + sa * p.a_slot_bytes, &p.tm_a[gr.src], fullA(sa), gr.c0, gr.dx, 0,
                                    y0 + gr.py0, n);
                    }
                    __syncwarp();
                    if (++sa == p.a_stages) { sa = 0; pa ^= 1; }
                }
            }
        }
    } else if (warp == 3) {
        // ===================== weight-tile producer =====================
        if (!skipB) {
            int sb = 0;
            uint32_t pb = 0;
            const int slice_rows = p.NT / csize;
            const uint32_t slice_bytes = (uint32_t)slice_rows * row_bytes;
            for (int g = cluster_id; g < rounds_total; g += num_clusters) {
                for (int it = 0; it < p.num_items; ++it) {
                    mbar_wait(emptyB(sb), pb ^ 1);               // every CTA of the cluster released this slot
                    if (elect_one_sync()) {
                        mbar_expect_tx(fullB(sb), (uint32_t)p.NT * row_bytes);
                        const uint32_t dst = b_base + sb * p.b_slot_bytes + rank * slice_bytes;
                        if (csize > 1)
                            tma_load_2d_mc(dst, &p.tm_b, fullB(sb), 0, it * p.Npad + (int)rank * slice_rows, cmask);
                        else
                            tma_load_2d(dst, &p.tm_b, fullB(sb), 0, it * p.Npad);
                    }
                    __syncwarp();
                    if (++sb == p.b_stages) { sb = 0; pb ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (whole warp runs the loops, one lane issues) =====================
        const uint32_t idesc = make_idesc_bf16(kTileM, p.NT);
        int sa = 0, sb = 0, acc = 0;
        uint32_t pa = 0, pb = 0, acc_phase = 0;
        for (int g = cluster_id; g < rounds_total; g += num_clusters) {
            mbar_wait(tempty(acc), acc_phase ^ 1);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + acc * kAccColStride;
            uint32_t accumulate = 0;
            for (int gi = 0; gi < p.num_groups; ++gi) {
                const int k = p.groups[gi].k;
                if (!skipA) mbar_wait(fullA(sa), pa);
                const uint32_t a_src = a_base + sa * p.a_slot_bytes;
                for (int j = 0; j < k; ++j) {
                    if (!skipB) mbar_wait(fullB(sb), pb);
                    tc_fence_after();
                    const uint64_t db = make_kmajor_desc<KC>(b_base + sb * p.b_slot_bytes);
                    if (elect_one_sync()) {
                        for (int m = 0; m < p.macc; ++m) {
                            // tap j reads the patch advanced by j image rows; accumulator m by 128 pixels
                            const uint64_t da = make_kmajor_desc<KC>(a_src + (uint32_t)(j * p.W + m * kTileM) * row_bytes);
#pragma unroll
                            for (int kk = 0; kk < KC / 16; ++kk)
                                umma_bf16(tmem_d + m * p.NT, da + 2 * kk, db + 2 * kk, idesc, (kk == 0) ? accumulate : 1u);
                        }
                        if (skipB) {}
                        else if (csize > 1) umma_commit_mc(emptyB(sb), cmask);
                        else umma_commit(emptyB(sb));
                    }
                    __syncwarp();
                    accumulate = 1;
                    if (++sb == p.b_stages) { sb = 0; pb ^= 1; }
                }
                if (!skipA && elect_one_sync()) umma_commit(emptyA(sa));
                __syncwarp();
                if (++sa == p.a_stages) { sa = 0; pa ^= 1; }
            }
            if (elect_one_sync()) umma_commit(tfull(acc));
            __syncwarp();
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
    } else if (warp >= 4) {
        // ===================== epilogue =====================
        const int q = warp & 3;
        const int r = q * 32 + lane;
        const uint32_t wsm = epi_base + q * kEpiWarpBytes;
        const uint32_t bias_s = wsm + 2 * kEpiStageBytes;
        epilogue_load_bias(p.epi, bias_s, p.NT, 0, lane);
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int g = cluster_id; g < rounds_total; g += num_clusters) {
            const int tile = g * csize + (int)rank;
            const int n = tile / p.tiles_y;
            const int y0 = (tile % p.tiles_y) * p.rows_tile;
            mbar_wait(tfull(acc), acc_phase);
            tc_fence_after();
            for (int m = 0; m < p.macc; ++m) {
                const int pix = m * kTileM + r;
                const int y = y0 + pix / p.W, x = pix % p.W;
                const bool valid = (n < p.n_valid) && (y < p.H);
                const size_t opix = ((size_t)n * p.H + y) * p.W + x;
                EpiRow row;
                row.opix = opix; row.rpix = opix; row.valid = valid; row.has_res = valid && (p.epi.res_mode != RES_NONE);
                const uint32_t taddr0 = tmem_base + acc * kAccColStride + m * p.NT + ((uint32_t)(q * 32) << 16);
                epilogue_block(p.epi, taddr0, p.NT, 0, row, wsm, bias_s, lane);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty(acc));
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (csize > 1) cluster_sync_all();      // no CTA exits while a peer may still multicast into it
    if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

uint32_t patchgemm_smem_bytes(const PatchGemmParams& p) {
    return p.a_stages * p.a_slot_bytes + p.b_stages * p.b_slot_bytes + kPatchFixedSmem;
}

int launch_patchgemm(const PatchGemmParams& p, int num_sms, void* stream) {
    const long long total = (long long)p.n_valid * p.tiles_y;
    if (total <= 0) return 0;
    const int c = p.cluster;
    if (p.a_stages > kMaxStages || p.b_stages > kMaxStages || p.num_groups > kPatchMaxGroups || (c != 1 && c != 2 && c != 4)) return (int)cudaErrorInvalidValue;
    const long long rounds = (total + c - 1) / c;
    const uint32_t smem = patchgemm_smem_bytes(p);
    if (smem > kSmemBudget) return (int)cudaErrorInvalidValue;
    void (*kern)(PatchGemmParams) = nullptr;
    switch (p.kc) {
        case 16: kern = patchgemm_kernel<16>; break;
        case 32: kern = patchgemm_kernel<32>; break;
        case 64: kern = patchgemm_kernel<64>; break;
        default: return (int)cudaErrorInvalidValue;
    }
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return (int)err;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((num_sms / c) * c);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = c;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    // one persistent cluster per co-resident slot (GPC boundaries can leave a few SMs without a full cluster)
    int max_clusters = num_sms / c;
    if (c > 1) {
        static int cached[3][5] = {{0}};                     // [kc index][cluster] -> max active clusters
        const int ki = p.kc == 16 ? 0 : p.kc == 32 ? 1 : 2;
        if (cached[ki][c] == 0) {
            int nc = 0;
            if (cudaOccupancyMaxActiveClusters(&nc, kern, &cfg) == cudaSuccess && nc > 0) cached[ki][c] = nc;
            else { cached[ki][c] = max_clusters; cudaGetLastError(); }
        }
        if (cached[ki][c] < max_clusters) max_clusters = cached[ki][c];
    }
    const int clusters = (int)(rounds < max_clusters ? rounds : max_clusters);
    cfg.gridDim = dim3(clusters * c);
    err = cudaLaunchKernelEx(&cfg, kern, p);
    return (int)err;
}

}  // namespace apvae
