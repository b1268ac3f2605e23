// A-stationary tcgen05 / TMEM / TMA implicit GEMM for stride-1 convolutions.  See patchgemm.h.
//
// Warp roles (256 threads, 1 CTA per SM, persistent over tiles, optional cluster of 2 / 4):
//   warp 0      activation-patch producer (TMA 5-D box, zero-filled halo)
//   warp 1      MMA issuer (tcgen05.mma; macc accumulators of 128 rows each)
//   warp 2      TMEM allocator
//   warp 3      weight-tile producer (TMA 2-D; 1/cluster of each tile, multicast to the cluster)
//   warps 4..7  epilogue (tcgen05.ld -> bias/mask/residual/act -> global); bf16-only epilogues take the
//               TMA path: residual boxes prefetched two blocks ahead by bulk tensor loads, outputs staged
//               swizzled in shared memory and written by bulk tensor stores
// Rings: patches (full/empty), weight tiles (full/empty; empty collects one arrival per cluster CTA
// through tcgen05.commit multicast), and a 2-deep TMEM accumulator ring.
#include <cuda_runtime.h>
#include <type_traits>
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

template <int KC, bool kTwo, bool kMask, bool kF16, bool kSplit, bool kEpi2>
__global__ void __launch_bounds__(kEpi2 ? kThreads + 128 : kThreads, 1)
patchgemm_kernel(const __grid_constant__ PatchGemmParams p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // identical offset in every CTA of the cluster (same kernel, same static layout)
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_base = smem_base;
    const uint32_t b_base = a_base + p.a_stages * p.a_slot_bytes;
    const uint32_t epi_base = b_base + p.b_stages * p.b_slot_bytes;
    const uint32_t bar_base = epi_base + ((kEpi2 ? kPatchFixedSmem2 : kPatchFixedSmem) - 2048u);
    auto fullA = [&](int s) { return bar_base + 8u * s; };
    auto emptyA = [&](int s) { return bar_base + 8u * (kMaxStages + s); };
    auto fullB = [&](int s) { return bar_base + 8u * (2 * kMaxStages + s); };
    auto emptyB = [&](int s) { return bar_base + 8u * (3 * kMaxStages + s); };
    auto tfull = [&](int a) { return bar_base + 8u * (4 * kMaxStages + a); };
    auto tempty = [&](int a) { return bar_base + 8u * (4 * kMaxStages + kAccStages + a); };
    const uint32_t tmem_ptr_smem = bar_base + 8u * (4 * kMaxStages + 2 * kAccStages);
    auto resbar = [&](int q, int b) { return bar_base + 8u * (4 * kMaxStages + 2 * kAccStages + 1 + 2 * q + b); };
    auto maskbar = [&](int q) { return bar_base + 8u * (4 * kMaxStages + 2 * kAccStages + 1 + 8 + q); };
    auto resbar2 = [&](int e) { return bar_base + 8u * (4 * kMaxStages + 2 * kAccStages + 1 + 8 + 4 + e); };   // kEpi2: one per epilogue warp
    auto maskbar2 = [&](int e) { return bar_base + 8u * (4 * kMaxStages + 2 * kAccStages + 1 + 8 + 4 + 8 + e); };

    const int warp = uniform_warp_idx();
    const int lane = threadIdx.x & 31;
    const int csize = p.cluster;
    const uint32_t rank = csize > 1 ? cluster_ctarank() : 0u;
    const uint16_t cmask = (uint16_t)((1u << csize) - 1u);
    constexpr bool two = kTwo;                  // CTA pair: one cta_group::2 MMA per pair, issued by rank 0 (its own
                                                // instantiation: a kernel containing cta_group::2 code cannot be launched
                                                // without a cluster of 2)

    if (warp == 0 && elect_one_sync()) {
        prefetch_tmap(&p.tm_a[0]);
        prefetch_tmap(&p.tm_a[1]);
        prefetch_tmap(&p.tm_b);
    }
    if (warp == 1 && elect_one_sync()) {
        for (int s = 0; s < p.a_stages; ++s) { mbar_init(fullA(s), 1); mbar_init(emptyA(s), 1); }
        for (int s = 0; s < p.b_stages; ++s) { mbar_init(fullB(s), 1); mbar_init(emptyB(s), two ? 1 : csize); }
        for (int a = 0; a < kAccStages; ++a) { mbar_init(tfull(a), 1); mbar_init(tempty(a), (two ? 8 : 4) * (kEpi2 ? 2 : 1)); }
        if constexpr (kEpi2) { for (int e = 0; e < 8; ++e) { mbar_init(resbar2(e), 1); mbar_init(maskbar2(e), 1); } }
        for (int q = 0; q < 4; ++q) { mbar_init(resbar(q, 0), 1); mbar_init(resbar(q, 1), 1); mbar_init(maskbar(q), 1); }
        fence_barrier_init();
    }
    if (warp == 2) { if constexpr (two) tmem_alloc_2cta(tmem_ptr_smem, kTmemCols); else tmem_alloc(tmem_ptr_smem, kTmemCols); }
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
    // tuning only: clock stamps of CTA 0's first 64 tiles, slot layout in tools/trace_layer.py
    unsigned long long* const trace = (blockIdx.x == 0) ? p.trace : nullptr;
    auto stamp = [&](int it, int slot) { if (trace && it < 64) trace[it * 16 + slot] = (unsigned long long)clock64(); };
    const bool skipA = (p.debug == 1 || p.debug == 3 || p.debug >= 10), skipB = (p.debug == 1 || p.debug == 2 || p.debug >= 10);

    if (warp == 0) {
        // ===================== activation-patch producer (whole warp, one lane issues) =====================
        if (!skipA) {
            int sa = 0;
            uint32_t pa = 0;
            int it = 0;
            for (int g = cluster_id; g < rounds_total; g += num_clusters, ++it) {
                const int tile = g * csize + (int)rank;
                const int n = tile / p.tiles_y;                  // tile >= total_tiles -> n >= n_valid: zero-filled / masked
                const int y0 = (tile % p.tiles_y) * p.rows_tile;
                for (int gi = 0; gi < p.num_groups; ++gi) {
                    const PatchGroup gr = p.groups[gi];
                    mbar_wait(emptyA(sa), pa ^ 1);
                    if (elect_one_sync()) {
                        if (gi == 0) stamp(it, 14);
                        if (gi == p.num_groups - 1) stamp(it, 15);
                        const uint32_t bytes = (uint32_t)gr.prow * p.W * row_bytes;
                        if constexpr (!two) {
                            mbar_expect_tx(fullA(sa), bytes);
                            tma_load_5d(a_base + sa * p.a_slot_bytes, &p.tm_a[gr.src], fullA(sa), gr.c0, gr.dx, 0,
                                        y0 + gr.py0, n);
                        } else {                                 // both patches of the pair complete on the leader's barrier
                            if (rank == 0) mbar_expect_tx(fullA(sa), 2 * bytes);
                            tma_load_5d_2cta(a_base + sa * p.a_slot_bytes, &p.tm_a[gr.src], mapa_shared(fullA(sa), 0),
                                             gr.c0, gr.dx, 0, y0 + gr.py0, n);
                        }
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
            // pair mode: each CTA keeps only ITS half of every weight tile (at the same offset in both CTAs)
            const uint32_t tap_b = two ? p.tap_bytes / 2 : p.tap_bytes;
            for (int g = cluster_id; g < rounds_total; g += num_clusters) {
                int tap0 = 0;                                    // global tap index = row block of the packed weights
                for (int gi = 0; gi < p.num_groups; ++gi) {
                    const int k = p.groups[gi].k;
                    for (int j0 = 0; j0 < k; j0 += p.tpi) {
                        const int nt = min(p.tpi, k - j0);       // taps in this item
                        mbar_wait(emptyB(sb), pb ^ 1);           // every CTA of the cluster released this slot
                        if (elect_one_sync()) {
                            if (!two || rank == 0) mbar_expect_tx(fullB(sb), (uint32_t)nt * p.tap_bytes);
                            const uint32_t bar = two ? mapa_shared(fullB(sb), 0) : fullB(sb);
                            for (int t = 0; t < nt; ++t) {
                                const int row0 = (tap0 + j0 + t) * p.Npad + (int)rank * slice_rows;
                                if constexpr (two) {
                                    tma_load_2d_2cta(b_base + sb * p.b_slot_bytes + t * tap_b, &p.tm_b, bar, 0, row0);
                                } else {
                                    const uint32_t dst = b_base + sb * p.b_slot_bytes + t * p.tap_bytes + rank * slice_bytes;
                                    if (csize > 1) tma_load_2d_mc(dst, &p.tm_b, bar, 0, row0, cmask);
                                    else tma_load_2d(dst, &p.tm_b, bar, 0, row0);
                                }
                            }
                        }
                        __syncwarp();
                        if (++sb == p.b_stages) { sb = 0; pb ^= 1; }
                    }
                    tap0 += k;
                }
            }
        }
    } else if (warp == 1 && (!two || rank == 0)) {
        // ===================== MMA issuer (whole warp runs the loops, one lane issues) =====================
        const uint32_t idesc = make_idesc_16(two ? 2 * kTileM : kTileM, p.NT, kF16);
        // loop invariants in registers (kernel parameters would be re-read from the constant bank after every
        // barrier wait); descriptors advance by plain adds: the low 14 bits hold (smem address >> 4)
        const int macc = p.macc, tpi = p.tpi, num_groups = p.num_groups, a_stages = p.a_stages, b_stages = p.b_stages;
        const uint32_t NT = (uint32_t)p.NT, a_slot = p.a_slot_bytes, b_slot = p.b_slot_bytes;
        const uint64_t a_tap_step = (uint64_t)(((uint32_t)p.W * row_bytes) >> 4);     // one image row of the patch
        const uint64_t a_m_step = (uint64_t)((kTileM * row_bytes) >> 4);               // next 128-pixel accumulator
        const uint64_t b_tap_step = (uint64_t)((two ? p.tap_bytes / 2 : p.tap_bytes) >> 4);
        int sa = 0, sb = 0, acc = 0;
        uint32_t pa = 0, pb = 0, acc_phase = 0;
        int it = 0;
        for (int g = cluster_id; g < rounds_total; g += num_clusters, ++it) {
            if (lane == 0) stamp(it, 0);
            mbar_wait(tempty(acc), acc_phase ^ 1);
            tc_fence_after();
            if (lane == 0) stamp(it, 1);
            const uint32_t tmem_d = tmem_base + acc * kAccColStride;
            uint32_t accumulate = 0;
            for (int gi = 0; gi < num_groups; ++gi) {
                const int k = p.groups[gi].k;
                if (!skipA) mbar_wait(fullA(sa), pa);
                if (gi == 0 && lane == 0) stamp(it, 2);
                uint64_t da_row = make_kmajor_desc<KC>(a_base + sa * a_slot);          // vertical tap 0
                for (int j0 = 0; j0 < k; j0 += tpi) {
                    const int nt = min(tpi, k - j0);
                    if (!skipB) mbar_wait(fullB(sb), pb);
                    tc_fence_after();
                    const uint64_t db0 = make_kmajor_desc<KC>(b_base + sb * b_slot);
                    if (elect_one_sync()) {
                        // tap j reads the patch advanced by j image rows; accumulator m by 128 pixels
                        auto issue = [&](auto macc_c) {
                            constexpr int M = decltype(macc_c)::value;
                            uint64_t da_t = da_row, db = db0;
                            for (int t = 0; t < nt; ++t) {
#pragma unroll
                                for (int m = 0; m < M; ++m) {
#pragma unroll
                                    for (int kk = 0; kk < KC / 16; ++kk) {
                                        if constexpr (two) umma_bf16_2cta(tmem_d + m * NT, da_t + m * a_m_step + 2 * kk, db + 2 * kk, idesc,
                                                                (kk == 0) ? accumulate : 1u);
                                        else umma_bf16(tmem_d + m * NT, da_t + m * a_m_step + 2 * kk, db + 2 * kk, idesc,
                                                       (kk == 0) ? accumulate : 1u);
                                    }
                                }
                                accumulate = 1;
                                da_t += a_tap_step;
                                db += b_tap_step;
                            }
                        };
                        if (p.debug == 6 || p.debug == 16) {}
                        else if (macc == 1) issue(std::integral_constant<int, 1>{});
                        else if (macc == 2) issue(std::integral_constant<int, 2>{});
                        else issue(std::integral_constant<int, 4>{});
                        if (skipB) {}
                        else if constexpr (two) umma_commit_2cta(emptyB(sb), cmask);
                        else if (csize > 1) umma_commit_mc(emptyB(sb), cmask);
                        else umma_commit(emptyB(sb));
                    }
                    __syncwarp();
                    accumulate = 1;
                    da_row += (uint64_t)nt * a_tap_step;
                    if (++sb == b_stages) { sb = 0; pb ^= 1; }
                }
                if (!skipA && elect_one_sync()) { if constexpr (two) umma_commit_2cta(emptyA(sa), cmask); else umma_commit(emptyA(sa)); }
                __syncwarp();
                if (++sa == a_stages) { sa = 0; pa ^= 1; }
            }
            if (lane == 0) stamp(it, 3);
            if (elect_one_sync()) { if constexpr (two) umma_commit_2cta(tfull(acc), cmask); else umma_commit(tfull(acc)); }
            __syncwarp();
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
    } else if (kEpi2 && warp >= 4) {
        // ===================== epilogue, TMA in / TMA out, TWO warpgroups =====================
        // Warps 4..7 (grp 0) take the even 64-column blocks of a tile, warps 8..11 (grp 1) the odd ones; both warps of a
        // lane quarter q read the same TMEM lanes.  Per warp: one residual box (fetched one of ITS blocks ahead), one
        // (raw, act) staging pair; the value goes chunk by chunk from the TMEM registers to the staging boxes.
        const int e = warp - 4, q = e & 3, grp = e >> 2;
        const uint32_t wbase = epi_base + e * kFastEpi2WarpBytes;
        const uint32_t rawbuf = wbase + kFastEpiBoxBytes, actbuf = wbase + 2 * kFastEpiBoxBytes, bias_s = wbase + 3 * kFastEpiBoxBytes;
        epilogue_load_bias(p.epi, bias_s, p.NT, 0, lane);
        const bool leader = elect_one_sync();
        const bool want_raw = p.epi.out_raw != nullptr, want_act = p.epi.out_act != nullptr;
        const int res_mode = p.epi.res_mode;
        const bool do_store = (p.debug != 5 && p.debug != 15), do_work = (p.debug != 4 && p.debug != 14);
        const uint32_t res_bytes = (uint32_t)p.res_rows * 128u;
        const float alpha = p.epi.lrelu_alpha;
        const uint32_t sw = (uint32_t)(lane & 7);
        const int W = p.W, H = p.H, NT = p.NT, macc = p.macc, tiles_y = p.tiles_y, rows_tile = p.rows_tile, n_valid = p.n_valid;
        const int lgW = 31 - __clz(W);
        const bool fold = p.fold > 1;
        const int rsh = (res_mode == RES_EVEN) ? 1 : 0;      // 4-D residual view (virtual widths): half-resolution coordinates
        const int tile_step = num_clusters * csize;
        const int dn = tile_step / tiles_y, dty = tile_step - dn * tiles_y;
        const int nb_acc = (NT + 63) >> 6, nb_tile = macc * nb_acc;      // 64-column blocks per accumulator / per tile
        const bool any_mine = nb_tile > grp;                             // a one-block tile leaves grp 1 without work
        const int bw_last = NT - ((NT - 1) & ~63);
        struct Pos { int g, n, ty, bi; };                                // a block of mine: round, image, row tile, block in tile
        auto advance = [&](Pos& s) {
            s.bi += 2;
            if (s.bi >= nb_tile) {
                s.bi = grp; s.g += num_clusters; s.n += dn; s.ty += dty;
                if (s.ty >= tiles_y) { s.ty -= tiles_y; ++s.n; }
            }
        };
        auto pos_ok = [&](const Pos& s) { return s.g < rounds_total && s.n < n_valid; };
        auto issue_res = [&](const Pos& s) {
            if (res_mode == RES_NONE || !pos_ok(s)) return false;
            const int m = s.bi / nb_acc, cb = (s.bi - m * nb_acc) << 6;
            const int po = m * kTileM + q * 32;
            const int gy = s.n * H + s.ty * rows_tile + (po >> lgW), x0 = po & (W - 1);
            int rp0 = gy * W + x0;
            if (res_mode == RES_EVEN) {
                if (gy & 1) return false;
                rp0 = (gy >> 1) * (W >> 1) + (x0 >> 1);
            }
            if (leader) {
                mbar_expect_tx(resbar2(e), res_bytes);
                if (fold) tma_load_5d(wbase, &p.tm_res, resbar2(e), cb, 0, 0, s.ty * rows_tile + (po >> lgW), s.n);
                else if (p.res4d) tma_load_4d(wbase, &p.tm_res, resbar2(e), cb, x0 >> rsh, (s.ty * rows_tile + (po >> lgW)) >> rsh, s.n);
                else tma_load_2d(wbase, &p.tm_res, resbar2(e), cb, rp0);
            }
            return true;
        };
        // leaky-relu' mask (adjoint ops, raw output only): the forward activation box of one of MY blocks, into the act box
        auto issue_mask = [&](const Pos& s) {
            if (!kMask || !pos_ok(s)) return false;
            const int m = s.bi / nb_acc, cb = (s.bi - m * nb_acc) << 6;
            const int po = m * kTileM + q * 32;
            if (leader) {
                mbar_expect_tx(maskbar2(e), 32u * 128u);
                if (fold) tma_load_5d(actbuf, &p.tm_mask, maskbar2(e), cb, 0, 0, s.ty * rows_tile + (po >> lgW), s.n);
                else if (p.fast_mask == 2) tma_load_4d(actbuf, &p.tm_mask, maskbar2(e), cb, po & (W - 1), s.ty * rows_tile + (po >> lgW), s.n);
                else tma_load_2d(actbuf, &p.tm_mask, maskbar2(e), cb, (s.n * H + s.ty * rows_tile + (po >> lgW)) * W + (po & (W - 1)));
            }
            return true;
        };
        int rrow = lane;
        bool lane_res = true;
        if (res_mode == RES_EVEN) {
            const int pix = q * 32 + lane;
            const int x = pix & (W - 1);
            lane_res = !(x & 1) && !((pix >> lgW) & 1);
            rrow = (W >= 32) ? (lane >> 1) : ((lane & 15) >> 1);
        }
        const uint32_t res_off = (uint32_t)rrow * 128u, rsw = (uint32_t)(rrow & 7);
        const uint32_t st_off = (uint32_t)lane * 128u;
        Pos cur;
        {
            const int tile0 = cluster_id * csize + (int)rank;
            cur.g = cluster_id; cur.n = tile0 / tiles_y; cur.ty = tile0 - cur.n * tiles_y; cur.bi = grp;
        }
        Pos pre = cur;
        bool iss = any_mine && issue_res(pre);
        bool miss = any_mine && issue_mask(pre);
        if (any_mine) advance(pre);
        uint32_t rph = 0, mph = 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        auto release_acc = [&]() {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) { if constexpr (two) mbar_arrive_cluster(mapa_shared(tempty(acc), 0)); else mbar_arrive(tempty(acc)); }
        };
        auto block2 = [&](auto bw_c, uint32_t taddr, int cb, int m, bool last_of_tile) {
            constexpr int BW = decltype(bw_c)::value;
            constexpr int NCH = BW / 8;
            uint32_t v[(BW + 31) / 32 * 32];
            {
                uint32_t (&v0)[32] = *reinterpret_cast<uint32_t (*)[32]>(&v[0]);
                if constexpr (BW >= 32) tmem_ld32(taddr, v0);
                else { uint32_t (&l0)[16] = *reinterpret_cast<uint32_t (*)[16]>(&v[0]); tmem_ld16(taddr, l0); }
                if constexpr (BW == 64) { uint32_t (&v1)[32] = *reinterpret_cast<uint32_t (*)[32]>(&v[32]); tmem_ld32(taddr + 32, v1); }
                if constexpr (BW == 48) { uint32_t (&l1)[16] = *reinterpret_cast<uint32_t (*)[16]>(&v[32]); tmem_ld16(taddr + 32, l1); }
            }
            const bool has_res = iss;
            if (has_res) { mbar_wait(resbar2(e), rph); rph ^= 1; }
            const bool use_mask = kMask && miss;
            if (use_mask) { mbar_wait(maskbar2(e), mph); mph ^= 1; }
            if (leader) tma_store_wait_read();             // my previous block's stores have left the staging pair
            tmem_ld_wait();
            if (last_of_tile) release_acc();               // my part of the accumulator is in registers
            __syncwarp();
            const uint32_t rbuf = wbase + res_off;
            const uint32_t bias_b = bias_s + cb * 4;
#pragma unroll
            for (int ch = 0; ch < NCH; ++ch) {
                float f[8];
                const float4 b0 = ld_shared_f4(bias_b + ch * 32);
                const float4 b1 = ld_shared_f4(bias_b + ch * 32 + 16);
                f[0] = __uint_as_float(v[ch * 8 + 0]) + b0.x; f[1] = __uint_as_float(v[ch * 8 + 1]) + b0.y;
                f[2] = __uint_as_float(v[ch * 8 + 2]) + b0.z; f[3] = __uint_as_float(v[ch * 8 + 3]) + b0.w;
                f[4] = __uint_as_float(v[ch * 8 + 4]) + b1.x; f[5] = __uint_as_float(v[ch * 8 + 5]) + b1.y;
                f[6] = __uint_as_float(v[ch * 8 + 6]) + b1.z; f[7] = __uint_as_float(v[ch * 8 + 7]) + b1.w;
                if (use_mask) {                            // v *= leaky-relu'(forward activation)
                    const uint4 mv = ld_shared_v4(actbuf + st_off + (((uint32_t)ch ^ sw) << 4));
                    f[0] *= op16::pos_lo(mv.x) ? 1.f : alpha; f[1] *= op16::pos_hi(mv.x) ? 1.f : alpha;
                    f[2] *= op16::pos_lo(mv.y) ? 1.f : alpha; f[3] *= op16::pos_hi(mv.y) ? 1.f : alpha;
                    f[4] *= op16::pos_lo(mv.z) ? 1.f : alpha; f[5] *= op16::pos_hi(mv.z) ? 1.f : alpha;
                    f[6] *= op16::pos_lo(mv.w) ? 1.f : alpha; f[7] *= op16::pos_hi(mv.w) ? 1.f : alpha;
                }
                if (has_res) {
                    const uint4 rv = ld_shared_v4(rbuf + (((uint32_t)ch ^ rsw) << 4));
                    if (lane_res) {
                        const float2 r0 = op16::unpack<kF16>(rv.x), r1 = op16::unpack<kF16>(rv.y), r2 = op16::unpack<kF16>(rv.z), r3 = op16::unpack<kF16>(rv.w);
                        f[0] += r0.x; f[1] += r0.y; f[2] += r1.x; f[3] += r1.y;
                        f[4] += r2.x; f[5] += r2.y; f[6] += r3.x; f[7] += r3.y;
                    }
                }
                if (want_raw)
                    st_shared_v4(rawbuf + st_off + (((uint32_t)ch ^ sw) << 4), op16::pack<kF16>(f[0], f[1]), op16::pack<kF16>(f[2], f[3]),
                                 op16::pack<kF16>(f[4], f[5]), op16::pack<kF16>(f[6], f[7]));
                if (want_act)
                    st_shared_v4(actbuf + st_off + (((uint32_t)ch ^ sw) << 4),
                                 op16::pack<kF16>(fmaxf(f[0], alpha * f[0]), fmaxf(f[1], alpha * f[1])),
                                 op16::pack<kF16>(fmaxf(f[2], alpha * f[2]), fmaxf(f[3], alpha * f[3])),
                                 op16::pack<kF16>(fmaxf(f[4], alpha * f[4]), fmaxf(f[5], alpha * f[5])),
                                 op16::pack<kF16>(fmaxf(f[6], alpha * f[6]), fmaxf(f[7], alpha * f[7])));
            }
            __syncwarp();                                  // every lane has read the residual (and mask) box: fetch my next block's
            iss = issue_res(pre);
            miss = issue_mask(pre);
            advance(pre);
            fence_proxy_async_smem();
            __syncwarp();
            if (leader && do_store && cur.n < n_valid) {
                const int po = m * kTileM + q * 32;
                const int oyy = cur.ty * rows_tile + (po >> lgW), ox0 = po & (W - 1);
                if (fold) {
                    if (want_raw) tma_store_5d(&p.tm_raw, rawbuf, cb, 0, 0, oyy, cur.n);
                    if (want_act) tma_store_5d(&p.tm_act, actbuf, cb, 0, 0, oyy, cur.n);
                } else {
                    if (want_raw) tma_store_4d(&p.tm_raw, rawbuf, cb, ox0, oyy, cur.n);
                    if (want_act) tma_store_4d(&p.tm_act, actbuf, cb, ox0, oyy, cur.n);
                }
                tma_store_commit();
            }
        };
        for (int g = cluster_id; g < rounds_total; g += num_clusters) {
            mbar_wait(tfull(acc), acc_phase);
            tc_fence_after();
            if (!any_mine || !do_work) {
                release_acc();
            } else {
                for (int bi = grp; bi < nb_tile; bi += 2) {
                    const int m = bi / nb_acc, cbi = bi - m * nb_acc, cb = cbi << 6;
                    const bool last = bi + 2 >= nb_tile;
                    const int bw = (cbi + 1 == nb_acc) ? bw_last : 64;
                    const uint32_t taddr = tmem_base + acc * kAccColStride + m * NT + cb + ((uint32_t)(q * 32) << 16);
                    if (bw == 64) block2(std::integral_constant<int, 64>{}, taddr, cb, m, last);
                    else if (bw == 32) block2(std::integral_constant<int, 32>{}, taddr, cb, m, last);
                    else if (bw == 16) block2(std::integral_constant<int, 16>{}, taddr, cb, m, last);
                    else block2(std::integral_constant<int, 48>{}, taddr, cb, m, last);
                }
            }
            // my next tile
            cur.g += num_clusters; cur.n += dn; cur.ty += dty;
            if (cur.ty >= tiles_y) { cur.ty -= tiles_y; ++cur.n; }
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
        if (leader) tma_store_wait_all();
        __syncwarp();
    } else if (warp >= 4 && warp < 8 && p.fast_epi) {
        // ===================== epilogue, TMA in / TMA out =====================
        // Work unit = "block": 32 rows (this warp's TMEM lane quarter) x up to 64 columns of one accumulator.
        // Everything that positions a block is warp-uniform and advances by adds / shifts (W is a power of two);
        // the block body is compiled once per column count so that it is straight-line code.
        const int q = warp & 3;
        const uint32_t wbase = epi_base + q * kFastEpiWarpBytes;
        const uint32_t rawbuf = wbase + 2 * kFastEpiBoxBytes, actbuf = wbase + 3 * kFastEpiBoxBytes;
        const uint32_t bias_s = wbase + 4 * kFastEpiBoxBytes;
        epilogue_load_bias(p.epi, bias_s, p.NT, 0, lane);
        const bool leader = elect_one_sync();              // the one thread that owns this warp's bulk-store groups
        const bool want_raw = p.epi.out_raw != nullptr, want_act = p.epi.out_act != nullptr;
        const int res_mode = p.epi.res_mode;
        const bool dbl = (res_mode == RES_NONE);
        const bool do_store = (p.debug != 5 && p.debug != 15), do_work = (p.debug != 4 && p.debug != 14);
        const uint32_t res_bytes = (uint32_t)p.res_rows * 128u;
        const float alpha = p.epi.lrelu_alpha;
        const uint32_t sw = (uint32_t)(lane & 7);
        const int W = p.W, H = p.H, NT = p.NT, macc = p.macc, tiles_y = p.tiles_y, rows_tile = p.rows_tile, n_valid = p.n_valid;
        const int lgW = 31 - __clz(W);
        const bool fold = p.fold > 1;                      // two 8-wide images per tile row (tiles_y == 1): 5-D epilogue maps
        const int rsh = (res_mode == RES_EVEN) ? 1 : 0;    // 4-D residual view (virtual widths): half-resolution coordinates
        const int tile_step = num_clusters * csize;
        const int dn = tile_step / tiles_y, dty = tile_step - dn * tiles_y;
        // a block's place in the work sequence: round g, accumulator m, first column cb; (n, ty) = image and row-tile of
        // this CTA's tile in round g, advanced incrementally
        struct Seq { int g, m, cb, n, ty; };
        auto seq_next = [&](Seq& s) {
            s.cb += 64;
            if (s.cb >= NT) {
                s.cb = 0;
                if (++s.m == macc) {
                    s.m = 0; s.g += num_clusters; s.n += dn; s.ty += dty;
                    if (s.ty >= tiles_y) { s.ty -= tiles_y; ++s.n; }
                }
            }
        };
        auto seq_ok = [&](const Seq& s) { return s.g < rounds_total && s.n < n_valid; };
        // bulk load of the residual box of block s into buffer b; returns whether one was issued (warp-uniform)
        auto issue_res = [&](const Seq& s, int b) {
            if (!seq_ok(s)) return false;
            const int po = s.m * kTileM + q * 32;          // first pixel of this warp's rows inside the tile
            const int gy = s.n * H + s.ty * rows_tile + (po >> lgW), x0 = po & (W - 1);
            int rp0 = gy * W + x0;
            if (res_mode == RES_EVEN) {                    // half-resolution residual on even pixels only (H is even)
                if (gy & 1) return false;
                rp0 = (gy >> 1) * (W >> 1) + (x0 >> 1);
            }
            if (leader) {
                mbar_expect_tx(resbar(q, b), res_bytes);
                if (fold) tma_load_5d(wbase + b * kFastEpiBoxBytes, &p.tm_res, resbar(q, b), s.cb, 0, 0, s.ty * rows_tile + (po >> lgW), s.n);
                else if (p.res4d) tma_load_4d(wbase + b * kFastEpiBoxBytes, &p.tm_res, resbar(q, b), s.cb, x0 >> rsh, (s.ty * rows_tile + (po >> lgW)) >> rsh, s.n);
                else tma_load_2d(wbase + b * kFastEpiBoxBytes, &p.tm_res, resbar(q, b), s.cb, rp0);
            }
            return true;
        };
        // split precision: the residual of block s as a hi + lo pair into boxes 0 and 1 behind ONE barrier (single buffered)
        auto issue_res_split = [&](const Seq& s) {
            if (res_mode == RES_NONE || !seq_ok(s)) return false;
            const int po = s.m * kTileM + q * 32;
            const int gy = s.n * H + s.ty * rows_tile + (po >> lgW), x0 = po & (W - 1);
            int rp0 = gy * W + x0;
            if (res_mode == RES_EVEN) {
                if (gy & 1) return false;
                rp0 = (gy >> 1) * (W >> 1) + (x0 >> 1);
            }
            if (leader) {
                mbar_expect_tx(resbar(q, 0), 2 * res_bytes);
                if (fold) {
                    tma_load_5d(wbase, &p.tm_res, resbar(q, 0), s.cb, 0, 0, s.ty * rows_tile + (po >> lgW), s.n);
                    tma_load_5d(wbase + kFastEpiBoxBytes, &p.tm_res_lo, resbar(q, 0), s.cb, 0, 0, s.ty * rows_tile + (po >> lgW), s.n);
                } else if (p.res4d) {
                    tma_load_4d(wbase, &p.tm_res, resbar(q, 0), s.cb, x0 >> rsh, (s.ty * rows_tile + (po >> lgW)) >> rsh, s.n);
                    tma_load_4d(wbase + kFastEpiBoxBytes, &p.tm_res_lo, resbar(q, 0), s.cb, x0 >> rsh, (s.ty * rows_tile + (po >> lgW)) >> rsh, s.n);
                } else {
                    tma_load_2d(wbase, &p.tm_res, resbar(q, 0), s.cb, rp0);
                    tma_load_2d(wbase + kFastEpiBoxBytes, &p.tm_res_lo, resbar(q, 0), s.cb, rp0);
                }
            }
            return true;
        };
        // leaky-relu' mask (adjoint ops): the forward activation box of block s, one block ahead, into the act staging box
        // (such ops write no act output); returns whether one was issued (warp-uniform)
        constexpr bool has_mask = kMask;               // its own instantiation: the forward ops' epilogue carries none of this
        auto issue_mask = [&](const Seq& s) {
            if (!has_mask || !seq_ok(s)) return false;
            const int po = s.m * kTileM + q * 32;
            if (leader) {
                mbar_expect_tx(maskbar(q), 32u * 128u);
                if (fold) tma_load_5d(actbuf, &p.tm_mask, maskbar(q), s.cb, 0, 0, s.ty * rows_tile + (po >> lgW), s.n);
                else if (p.fast_mask == 2) tma_load_4d(actbuf, &p.tm_mask, maskbar(q), s.cb, po & (W - 1), s.ty * rows_tile + (po >> lgW), s.n);
                else tma_load_2d(actbuf, &p.tm_mask, maskbar(q), s.cb, (s.n * H + s.ty * rows_tile + (po >> lgW)) * W + (po & (W - 1)));
            }
            return true;
        };
        // this lane's residual row inside the box, and whether this lane's pixel takes a residual
        int rrow = lane;
        bool lane_res = true;
        if (res_mode == RES_EVEN) {
            const int pix = q * 32 + lane;                 // (m * 128 does not change x parity or the row split)
            const int x = pix & (W - 1);
            lane_res = !(x & 1) && !((pix >> lgW) & 1);
            rrow = (W >= 32) ? (lane >> 1) : ((lane & 15) >> 1);
        }
        const uint32_t res_off = (uint32_t)rrow * 128u, rsw = (uint32_t)(rrow & 7);
        const uint32_t st_off = (uint32_t)lane * 128u;
        Seq cur;
        {
            const int tile0 = cluster_id * csize + (int)rank;
            cur.g = cluster_id; cur.m = 0; cur.cb = 0; cur.n = tile0 / tiles_y; cur.ty = tile0 - cur.n * tiles_y;
        }
        Seq pre = cur;
        Seq mseq = cur;
        bool miss = issue_mask(mseq);                      // mask of the first block
        seq_next(mseq);
        uint32_t mph = 0;
        bool iss0 = false, iss1 = false;
        if constexpr (kSplit) {
            iss0 = issue_res_split(pre); seq_next(pre);
        } else if (!dbl) {
            iss0 = issue_res(pre, 0); seq_next(pre);
            iss1 = issue_res(pre, 1); seq_next(pre);
        }
        uint32_t ph0 = 0, ph1 = 0, nblk = 0;
        int kb = 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        int tit = 0, tblk = 0;                             // trace only

        // one block of BW columns: TMEM -> registers -> (+bias, +residual, leaky-relu) -> swizzled staging -> bulk stores
        auto block = [&](auto bw_c, uint32_t taddr, bool last_of_tile) {
            constexpr int BW = decltype(bw_c)::value;
            constexpr int NCH = BW / 8;                    // 16-byte chunks per row
            uint32_t v[(BW + 31) / 32 * 32];
            {
                uint32_t (&v0)[32] = *reinterpret_cast<uint32_t (*)[32]>(&v[0]);
                if constexpr (BW >= 32) tmem_ld32(taddr, v0);
                else { uint32_t (&l0)[16] = *reinterpret_cast<uint32_t (*)[16]>(&v[0]); tmem_ld16(taddr, l0); }
                if constexpr (BW == 64) { uint32_t (&v1)[32] = *reinterpret_cast<uint32_t (*)[32]>(&v[32]); tmem_ld32(taddr + 32, v1); }
                if constexpr (BW == 48) { uint32_t (&l1)[16] = *reinterpret_cast<uint32_t (*)[16]>(&v[32]); tmem_ld16(taddr + 32, l1); }
            }
            const bool has_res = kb ? iss1 : iss0;
            if (has_res) {
                if (kb) { mbar_wait(resbar(q, 1), ph1); ph1 ^= 1; }
                else { mbar_wait(resbar(q, 0), ph0); ph0 ^= 1; }
            }
            const bool use_mask = has_mask && miss;
            if (use_mask) { mbar_wait(maskbar(q), mph); mph ^= 1; }
            tmem_ld_wait();
            if (q == 0 && lane == 0 && tblk == 0) stamp(tit, 6);
            if (last_of_tile) {                            // the accumulator is in registers: hand it back to the MMA warp now
                tc_fence_before();
                __syncwarp();
                if (lane == 0) { if constexpr (two) mbar_arrive_cluster(mapa_shared(tempty(acc), 0)); else mbar_arrive(tempty(acc)); }
            }
            const uint32_t rbuf = wbase + kb * kFastEpiBoxBytes + res_off;
            const uint32_t bias_b = bias_s + cur.cb * 4;
            uint32_t rr[NCH * 4], aa[NCH * 4];             // packed bf16 pairs: raw and activated
#pragma unroll
            for (int ch = 0; ch < NCH; ++ch) {
                float f[8];
                const float4 b0 = ld_shared_f4(bias_b + ch * 32);
                const float4 b1 = ld_shared_f4(bias_b + ch * 32 + 16);
                f[0] = __uint_as_float(v[ch * 8 + 0]) + b0.x; f[1] = __uint_as_float(v[ch * 8 + 1]) + b0.y;
                f[2] = __uint_as_float(v[ch * 8 + 2]) + b0.z; f[3] = __uint_as_float(v[ch * 8 + 3]) + b0.w;
                f[4] = __uint_as_float(v[ch * 8 + 4]) + b1.x; f[5] = __uint_as_float(v[ch * 8 + 5]) + b1.y;
                f[6] = __uint_as_float(v[ch * 8 + 6]) + b1.z; f[7] = __uint_as_float(v[ch * 8 + 7]) + b1.w;
                if (use_mask) {                            // v *= leaky-relu'(forward activation): 1 where it is positive, else alpha
                    const uint4 mv = ld_shared_v4(actbuf + st_off + (((uint32_t)ch ^ sw) << 4));
                    f[0] *= op16::pos_lo(mv.x) ? 1.f : alpha; f[1] *= op16::pos_hi(mv.x) ? 1.f : alpha;
                    f[2] *= op16::pos_lo(mv.y) ? 1.f : alpha; f[3] *= op16::pos_hi(mv.y) ? 1.f : alpha;
                    f[4] *= op16::pos_lo(mv.z) ? 1.f : alpha; f[5] *= op16::pos_hi(mv.z) ? 1.f : alpha;
                    f[6] *= op16::pos_lo(mv.w) ? 1.f : alpha; f[7] *= op16::pos_hi(mv.w) ? 1.f : alpha;
                }
                if (has_res) {
                    const uint4 rv = ld_shared_v4(rbuf + (((uint32_t)ch ^ rsw) << 4));
                    if (lane_res) {
                        const float2 r0 = op16::unpack<kF16>(rv.x), r1 = op16::unpack<kF16>(rv.y), r2 = op16::unpack<kF16>(rv.z), r3 = op16::unpack<kF16>(rv.w);
                        f[0] += r0.x; f[1] += r0.y; f[2] += r1.x; f[3] += r1.y;
                        f[4] += r2.x; f[5] += r2.y; f[6] += r3.x; f[7] += r3.y;
                    }
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    rr[ch * 4 + i] = op16::pack<kF16>(f[2 * i], f[2 * i + 1]);
                    aa[ch * 4 + i] = op16::pack<kF16>(fmaxf(f[2 * i], alpha * f[2 * i]), fmaxf(f[2 * i + 1], alpha * f[2 * i + 1]));
                }
            }
            if (q == 0 && lane == 0 && tblk == 0) stamp(tit, 7);
            if (!dbl) {
                __syncwarp();                              // every lane is done with the residual box:
                const bool is = issue_res(pre, kb);        // refill it with the residual of the block two ahead
                if (kb) iss1 = is; else iss0 = is;
                seq_next(pre);
            }
            if (has_mask) {
                __syncwarp();                              // ... and with the mask box: fetch the next block's mask
                miss = issue_mask(mseq);
                seq_next(mseq);
            }
            // staging boxes: without a residual the two residual boxes serve as a second (raw, act) pair, so only
            // the stores of the block before the previous one must have left shared memory
            const bool alt = dbl && (nblk & 1);
            const uint32_t rawb = (alt ? wbase : rawbuf), actb = (alt ? wbase + kFastEpiBoxBytes : actbuf);
            ++nblk;
            if (leader) { if (dbl) tma_store_wait_read1(); else tma_store_wait_read(); }
            __syncwarp();
            if (q == 0 && lane == 0 && tblk == 0) stamp(tit, 8);
            if (want_raw) {
#pragma unroll
                for (int ch = 0; ch < NCH; ++ch)
                    st_shared_v4(rawb + st_off + (((uint32_t)ch ^ sw) << 4), rr[ch * 4], rr[ch * 4 + 1], rr[ch * 4 + 2], rr[ch * 4 + 3]);
            }
            if (want_act) {
#pragma unroll
                for (int ch = 0; ch < NCH; ++ch)
                    st_shared_v4(actb + st_off + (((uint32_t)ch ^ sw) << 4), aa[ch * 4], aa[ch * 4 + 1], aa[ch * 4 + 2], aa[ch * 4 + 3]);
            }
            if (q == 0 && lane == 0 && tblk == 0) stamp(tit, 9);
            fence_proxy_async_smem();
            __syncwarp();
            if (q == 0 && lane == 0 && tblk == 0) stamp(tit, 10);
            if (leader && do_store && cur.n < n_valid) {
                const int po = cur.m * kTileM + q * 32;
                const int oyy = cur.ty * rows_tile + (po >> lgW), ox0 = po & (W - 1);
                if (fold) {
                    if (want_raw) tma_store_5d(&p.tm_raw, rawb, cur.cb, 0, 0, oyy, cur.n);
                    if (want_act) tma_store_5d(&p.tm_act, actb, cur.cb, 0, 0, oyy, cur.n);
                } else {
                    if (want_raw) tma_store_4d(&p.tm_raw, rawb, cur.cb, ox0, oyy, cur.n);
                    if (want_act) tma_store_4d(&p.tm_act, actb, cur.cb, ox0, oyy, cur.n);
                }
                tma_store_commit();
            }
            if (q == 0 && lane == 0) stamp(tit, tblk == 0 ? 11 : (tblk == 1 ? 12 : 13));
            ++tblk;
            kb ^= 1;
        };

        // Split-precision block: the value v = acc + bias + (res_hi + res_lo) leaves as v_hi = round16(v), v_lo = round16(v - v_hi)
        // (and the same pair of leaky_relu(v)).  Two store passes through the ONE (raw, act) staging pair — hi, then lo; the
        // value is recomputed per 8-column chunk in each pass (TMEM registers, bias row and residual boxes stay put), so
        // the register footprint is that of the plain block.  The residual boxes are refilled when pass 2 has read them.
        auto block_split = [&](auto bw_c, uint32_t taddr, bool last_of_tile) {
            constexpr int BW = decltype(bw_c)::value;
            constexpr int NCH = BW / 8;
            uint32_t v[(BW + 31) / 32 * 32];
            {
                uint32_t (&v0)[32] = *reinterpret_cast<uint32_t (*)[32]>(&v[0]);
                if constexpr (BW >= 32) tmem_ld32(taddr, v0);
                else { uint32_t (&l0)[16] = *reinterpret_cast<uint32_t (*)[16]>(&v[0]); tmem_ld16(taddr, l0); }
                if constexpr (BW == 64) { uint32_t (&v1)[32] = *reinterpret_cast<uint32_t (*)[32]>(&v[32]); tmem_ld32(taddr + 32, v1); }
                if constexpr (BW == 48) { uint32_t (&l1)[16] = *reinterpret_cast<uint32_t (*)[16]>(&v[32]); tmem_ld16(taddr + 32, l1); }
            }
            const bool has_res = iss0;
            if (has_res) { mbar_wait(resbar(q, 0), ph0); ph0 ^= 1; }
            tmem_ld_wait();
            if (last_of_tile) {
                tc_fence_before();
                __syncwarp();
                if (lane == 0) { if constexpr (two) mbar_arrive_cluster(mapa_shared(tempty(acc), 0)); else mbar_arrive(tempty(acc)); }
            }
            const uint32_t rbuf = wbase + res_off;
            const uint32_t bias_b = bias_s + cur.cb * 4;
            auto value = [&](int ch, float (&f)[8]) {
                const float4 b0 = ld_shared_f4(bias_b + ch * 32);
                const float4 b1 = ld_shared_f4(bias_b + ch * 32 + 16);
                f[0] = __uint_as_float(v[ch * 8 + 0]) + b0.x; f[1] = __uint_as_float(v[ch * 8 + 1]) + b0.y;
                f[2] = __uint_as_float(v[ch * 8 + 2]) + b0.z; f[3] = __uint_as_float(v[ch * 8 + 3]) + b0.w;
                f[4] = __uint_as_float(v[ch * 8 + 4]) + b1.x; f[5] = __uint_as_float(v[ch * 8 + 5]) + b1.y;
                f[6] = __uint_as_float(v[ch * 8 + 6]) + b1.z; f[7] = __uint_as_float(v[ch * 8 + 7]) + b1.w;
                if (has_res) {
                    const uint4 rh = ld_shared_v4(rbuf + (((uint32_t)ch ^ rsw) << 4));
                    const uint4 rl = ld_shared_v4(rbuf + kFastEpiBoxBytes + (((uint32_t)ch ^ rsw) << 4));
                    if (lane_res) {
                        const uint32_t wh[4] = {rh.x, rh.y, rh.z, rh.w}, wl[4] = {rl.x, rl.y, rl.z, rl.w};
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float2 a = op16::unpack<kF16>(wh[i]), b = op16::unpack<kF16>(wl[i]);
                            f[2 * i] += a.x + b.x; f[2 * i + 1] += a.y + b.y;
                        }
                    }
                }
            };
#pragma unroll
            for (int pass = 0; pass < 2; ++pass) {
                if (leader) tma_store_wait_read();         // the staging pair is free again (previous pass / previous block)
                __syncwarp();
#pragma unroll
                for (int ch = 0; ch < NCH; ++ch) {
                    float f[8];
                    value(ch, f);
                    uint32_t r4[4], a4[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float a0 = fmaxf(f[2 * i], alpha * f[2 * i]), a1 = fmaxf(f[2 * i + 1], alpha * f[2 * i + 1]);
                        const uint32_t rh = op16::pack<kF16>(f[2 * i], f[2 * i + 1]), ah = op16::pack<kF16>(a0, a1);
                        if (pass == 0) { r4[i] = rh; a4[i] = ah; }
                        else {
                            r4[i] = op16::pack<kF16>(f[2 * i] - op16::lo<kF16>(rh), f[2 * i + 1] - op16::hi<kF16>(rh));
                            a4[i] = op16::pack<kF16>(a0 - op16::lo<kF16>(ah), a1 - op16::hi<kF16>(ah));
                        }
                    }
                    if (want_raw) st_shared_v4(rawbuf + st_off + (((uint32_t)ch ^ sw) << 4), r4[0], r4[1], r4[2], r4[3]);
                    if (want_act) st_shared_v4(actbuf + st_off + (((uint32_t)ch ^ sw) << 4), a4[0], a4[1], a4[2], a4[3]);
                }
                if (pass == 1) {
                    __syncwarp();                          // every lane has read the residual pair: fetch the next block's
                    iss0 = issue_res_split(pre);
                    seq_next(pre);
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (leader && do_store && cur.n < n_valid) {
                    const int po = cur.m * kTileM + q * 32;
                    const int oyy = cur.ty * rows_tile + (po >> lgW), ox0 = po & (W - 1);
                    const CUtensorMap* mr = pass ? &p.tm_raw_lo : &p.tm_raw;
                    const CUtensorMap* ma = pass ? &p.tm_act_lo : &p.tm_act;
                    if (fold) {
                        if (want_raw) tma_store_5d(mr, rawbuf, cur.cb, 0, 0, oyy, cur.n);
                        if (want_act) tma_store_5d(ma, actbuf, cur.cb, 0, 0, oyy, cur.n);
                    } else {
                        if (want_raw) tma_store_4d(mr, rawbuf, cur.cb, ox0, oyy, cur.n);
                        if (want_act) tma_store_4d(ma, actbuf, cur.cb, ox0, oyy, cur.n);
                    }
                    tma_store_commit();
                }
            }
            ++tblk;
        };

        const int bw_last = NT - ((NT - 1) & ~63);          // columns of the last block of an accumulator: 16, 32, 48 or 64
        for (int g = cluster_id; g < rounds_total; g += num_clusters, ++tit) {
            if (q == 0 && lane == 0) stamp(tit, 4);
            mbar_wait(tfull(acc), acc_phase);
            tc_fence_after();
            if (q == 0 && lane == 0) stamp(tit, 5);
            tblk = 0;
            if (do_work) {
                for (int m = 0; m < macc; ++m) {
                    const uint32_t tbase = tmem_base + acc * kAccColStride + m * NT + ((uint32_t)(q * 32) << 16);
                    for (int cb = 0; cb < NT; cb += 64) {
                        const bool last_cb = cb + 64 >= NT;
                        const bool last = last_cb && (m + 1 == macc);
                        const int bw = last_cb ? bw_last : 64;
                        if constexpr (kSplit) {
                            if (bw == 64) block_split(std::integral_constant<int, 64>{}, tbase + cb, last);
                            else if (bw == 32) block_split(std::integral_constant<int, 32>{}, tbase + cb, last);
                            else if (bw == 16) block_split(std::integral_constant<int, 16>{}, tbase + cb, last);
                            else block_split(std::integral_constant<int, 48>{}, tbase + cb, last);
                        } else {
                            if (bw == 64) block(std::integral_constant<int, 64>{}, tbase + cb, last);
                            else if (bw == 32) block(std::integral_constant<int, 32>{}, tbase + cb, last);
                            else if (bw == 16) block(std::integral_constant<int, 16>{}, tbase + cb, last);
                            else block(std::integral_constant<int, 48>{}, tbase + cb, last);
                        }
                        seq_next(cur);
                    }
                }
            } else {
                tc_fence_before();
                __syncwarp();
                if (lane == 0) { if constexpr (two) mbar_arrive_cluster(mapa_shared(tempty(acc), 0)); else mbar_arrive(tempty(acc)); }
            }
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
        if (leader) tma_store_wait_all();
        __syncwarp();
    } else if (warp >= 4 && warp < 8) {
        // ===================== epilogue =====================
        const int q = warp & 3;
        const int r = q * 32 + lane;
        const uint32_t wsm = epi_base + q * kEpiWarpBytes;
        const uint32_t bias_s = wsm + 2 * kEpiStageBytes;
        epilogue_load_bias(p.epi, bias_s, p.NT, 0, lane);
        int acc = 0;
        uint32_t acc_phase = 0;
        const bool split_epi = p.epi.out_raw_lo != nullptr || p.epi.out_act_lo != nullptr || p.epi.res_lo != nullptr;
        auto make_row = [&](int tile, int m) {
            const int n = tile / p.tiles_y;
            const int y0 = (tile % p.tiles_y) * p.rows_tile;
            const int pix = m * kTileM + r;
            int y = y0 + pix / p.W, x = pix % p.W;
            int ni = n;
            const int wr = p.Wr;                        // row width in memory (< the tile-row width for the virtual widths)
            bool valid = (n < p.n_valid) && (y < p.H) && (x < wr);
            if (p.fold > 1) {                           // x = (image of the pair, pixel): image 2 n + x / 8, <= 8 pixels per row
                ni = 2 * n + (x >> 3); x &= 7;
                valid = (ni < p.n_real) && (y < p.H) && (x < wr);
            }
            const size_t opix = ((size_t)ni * p.H + y) * wr + x;
            EpiRow row;
            row.opix = opix; row.rpix = opix; row.valid = valid; row.has_res = valid && (p.epi.res_mode != RES_NONE);
            if (p.epi.res_mode == RES_EVEN) {           // 1x1 stride-2 transposed shortcut: x W lands on even output pixels only
                row.has_res = row.has_res && !((y | x) & 1);
                row.rpix = ((size_t)ni * (p.H >> 1) + (y >> 1)) * (wr >> 1) + (x >> 1);
            }
            return row;
        };
        uint4 rv[8];
        bool pre = false;
        EpiRow row = make_row(cluster_id * csize + (int)rank, 0);
        if (cluster_id < rounds_total && __any_sync(0xffffffffu, row.has_res)) {     // residual of the first tile: before the accumulator is ready
            epilogue_res_issue(p.epi, row, p.NT, 0, 0, lane, rv);
            pre = true;
        }
        for (int g = cluster_id; g < rounds_total; g += num_clusters) {
            const int tile = g * csize + (int)rank;
            mbar_wait(tfull(acc), acc_phase);
            tc_fence_after();
            for (int m = 0; m < p.macc; ++m) {
                const bool last_m = (m + 1 == p.macc);
                const bool more = !last_m || (g + num_clusters < rounds_total);
                const EpiRow nrow = last_m ? make_row((g + num_clusters) * csize + (int)rank, 0) : make_row(tile, m + 1);
                const uint32_t taddr0 = tmem_base + acc * kAccColStride + m * p.NT + ((uint32_t)(q * 32) << 16);
                // split precision (hi + lo outputs / residual): the same block body with the low halves written from registers
                pre = split_epi ? epilogue_block<true, kF16>(p.epi, taddr0, p.NT, 0, row, wsm, bias_s, lane, rv, pre, more ? &nrow : nullptr, 0)
                                : epilogue_block<false, kF16>(p.epi, taddr0, p.NT, 0, row, wsm, bias_s, lane, rv, pre, more ? &nrow : nullptr, 0);
                row = nrow;
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) { if constexpr (two) mbar_arrive_cluster(mapa_shared(tempty(acc), 0)); else mbar_arrive(tempty(acc)); }
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (csize > 1) cluster_sync_all();      // no CTA exits while a peer may still multicast into it
    if (warp == 2) { if constexpr (two) tmem_dealloc_2cta(tmem_base, kTmemCols); else tmem_dealloc(tmem_base, kTmemCols); }
}

uint32_t patchgemm_smem_bytes(const PatchGemmParams& p) {
    return p.a_stages * p.a_slot_bytes + p.b_stages * p.b_slot_bytes + (p.epi2 && p.fast_epi ? kPatchFixedSmem2 : kPatchFixedSmem);
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
    const bool two = p.two_cta != 0;
    if (two && c != 2) return (int)cudaErrorInvalidValue;
    const bool mk = p.fast_epi && p.fast_mask;
    const bool f16 = p.epi.f16 != 0;
    const bool sp = p.fast_epi && p.split;
    if (sp && mk) return (int)cudaErrorInvalidValue;          // the split-precision TMA epilogue carries no mask
    const bool e2 = p.fast_epi && p.epi2;
    if (e2 && sp) return (int)cudaErrorInvalidValue;          // two epilogue warpgroups: 16-bit epilogue, with or without mask
#define APVAE_PATCH_PICK2(KC, TWO, MK, SP) kern = f16 ? patchgemm_kernel<KC, TWO, MK, true, SP, false> : patchgemm_kernel<KC, TWO, MK, false, SP, false>
#define APVAE_PATCH_PICK_E2(KC, TWO, MK) kern = f16 ? patchgemm_kernel<KC, TWO, MK, true, false, true> : patchgemm_kernel<KC, TWO, MK, false, false, true>
#define APVAE_PATCH_PICK(KC)                                                                                           \
    do {                                                                                                               \
        if (e2) { if (two) { if (mk) APVAE_PATCH_PICK_E2(KC, true, true); else APVAE_PATCH_PICK_E2(KC, true, false); }  \
                  else { if (mk) APVAE_PATCH_PICK_E2(KC, false, true); else APVAE_PATCH_PICK_E2(KC, false, false); } }  \
        else if (two) { if (sp) APVAE_PATCH_PICK2(KC, true, false, true); else if (mk) APVAE_PATCH_PICK2(KC, true, true, false);     \
                   else APVAE_PATCH_PICK2(KC, true, false, false); }                                                   \
        else { if (sp) APVAE_PATCH_PICK2(KC, false, false, true); else if (mk) APVAE_PATCH_PICK2(KC, false, true, false);       \
               else APVAE_PATCH_PICK2(KC, false, false, false); }                                                      \
    } while (0)
    switch (p.kc) {
        case 16: APVAE_PATCH_PICK(16); break;
        case 32: APVAE_PATCH_PICK(32); break;
        case 64: APVAE_PATCH_PICK(64); break;
        default: return (int)cudaErrorInvalidValue;
    }
#undef APVAE_PATCH_PICK
#undef APVAE_PATCH_PICK2
#undef APVAE_PATCH_PICK_E2
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return (int)err;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((num_sms / c) * c);
    cfg.blockDim = dim3(e2 ? kThreads + 128 : kThreads);
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
        static int cached[72][5] = {{0}};                     // [kc index, pair mode, epilogue variant, format][cluster] -> max active clusters
        const int ki = (p.kc == 16 ? 0 : p.kc == 32 ? 1 : 2) + (two ? 3 : 0) + (e2 ? (mk ? 24 : 18) : mk ? 6 : sp ? 12 : 0) + (f16 ? 30 : 0);
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
