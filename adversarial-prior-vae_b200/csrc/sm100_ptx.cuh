// Device-side sm_100a primitives shared by the tensor-core kernels: mbarrier, TMA
// (cp.async.bulk.tensor), tcgen05 (MMA / commit / TMEM load), cluster helpers, UMMA descriptors
// and the common epilogue (bias -> lrelu' mask -> residual -> tanh -> fp32 / bf16 stores).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdio.h>
#include "tapgemm.h"
#include "op16.h"

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
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
// Bulk tensor store shared -> global (box clipped at the tensor extents).  Bulk async-groups are per
// thread: the thread that issues and commits is the one that waits.
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(map), "r"(src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2, int c3) {
    asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
                 ::"l"(map), "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_store_5d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2, int c3, int c4) {
    asm volatile("cp.async.bulk.tensor.5d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];"
                 ::"l"(map), "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all committed groups have finished READING shared memory (the staging buffer may be rewritten)
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// ... all but the most recent committed group (double-buffered staging)
__device__ __forceinline__ void tma_store_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// generic-proxy writes to shared memory become visible to the async proxy (TMA)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
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
// ---- cta_group::2: the CTA pair of a 2-CTA cluster runs ONE 256-row MMA; the leader (rank 0) issues, each CTA
// holds its own 128 rows of A, its half of B (N/2 rows) and its 128 accumulator rows.
__device__ __forceinline__ void umma_bf16_2cta(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                               uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit_2cta(uint32_t bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"(mask) : "memory");
}
__device__ __forceinline__ void tmem_alloc_2cta(uint32_t dst_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2cta(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
// shared::cluster address of `local` (a shared::cta address) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_shared(uint32_t local, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster_addr) : "memory");
}
// TMA loads of a CTA pair: data lands in THIS CTA's shared memory, the transaction completes on `bar`, a
// shared::cluster address that may belong to the peer (the leader's full barrier)
__device__ __forceinline__ void tma_load_5d_2cta(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2,
                                                 int c3, int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4) : "memory");
}
__device__ __forceinline__ void tma_load_2d_2cta(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
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

// One lane of the (converged) warp; the role loops run warp-uniformly and only the tcgen05 / TMA
// instruction itself is issued under this predicate, so that ptxas keeps descriptors, coordinates
// and barrier addresses in uniform registers (a lane-0 branch around the whole loop makes every
// operand "divergent" and each UTCHMMA / UTMALDG gets wrapped in an ELECT + R2UR waterfall loop).
__device__ __forceinline__ bool elect_one_sync() {
    uint32_t pred = 0;
    asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "elect.sync _|P1, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, P1;\n\t}"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ int uniform_warp_idx() { return __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0); }

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
// kind::f16 instruction descriptor: D = f32, A = B = fp16 (format 0) or bf16 (format 1), both K-major, M x N.
__device__ __forceinline__ uint32_t make_idesc_16(int M, int N, bool f16) {
    const uint32_t fmt = f16 ? 0u : 1u;
    return (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xFFFF0000u); }

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
    uint4 r;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr) : "memory");
    return r;
}
__device__ __forceinline__ float4 ld_shared_f4(uint32_t addr) {
    float4 r;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr) : "memory");
    return r;
}

// ---------------------------------------------------------------------------------------------
// Warp-level epilogue of one 32-row x NT-column accumulator block (this warp's TMEM lane quarter).
//   v = acc + bias ; v *= lrelu'(mask) ; v += res ; tanh on cols < tanh_cols ; fp32 / bf16 stores.
// bf16 outputs (and the bf16 residual) go through a per-warp shared-memory staging block of
// 32 rows x 64 columns so that global accesses are 128-byte row segments instead of one 32-byte
// piece per lane: lane r owns accumulator row r, but 8 consecutive lanes move one row segment.
// Staging row pitch = 64*2 + 16 bytes: the 16-byte lane accesses of a quarter-warp hit distinct banks.
// (kEpiBlockCols, kEpiPitch, kEpiStageBytes, kEpiWarpBytes, kEpiSmemBytes: tapgemm.h)

struct EpiRow {
    size_t opix;     // output pixel of this lane's row
    size_t rpix;     // residual pixel
    bool valid;      // row is written
    bool has_res;    // residual is added to this row
};

// Issues (does not wait for) the coalesced loads of one 64-column residual block: lane l fetches the
// 16-byte segments l, l + 32, ... of the 32 rows x bw columns block.  Called one block ahead of its use
// (and for the first block before the accumulator is even ready), so the L2 latency hides behind the
// previous block's arithmetic.
__device__ __forceinline__ void epilogue_res_issue(const EpilogueArgs& e, const EpiRow& row, int NT, int col_base, int cb,
                                                   int lane, uint4 (&rv)[8]) {
    const int bw = min((int)kEpiBlockCols, NT - cb);
    const int segs = bw >> 3;
    const int sh = (segs == 8) ? 3 : (segs == 4) ? 2 : (segs == 2) ? 1 : -1;      // segs is even; 6 (48 columns) divides
    const __nv_bfloat16* res = static_cast<const __nv_bfloat16*>(e.res) + col_base + cb;
    const uint32_t rmask = __ballot_sync(0xffffffffu, row.has_res);
    const uint32_t my_rp = (uint32_t)row.rpix;
    // all addresses first, then all loads back to back: a shuffle between two loads would wait on the scoreboard the
    // previous load holds and serialise the eight L2 round trips
    const uint4* ptr[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) {
        ptr[t] = nullptr;
        if (t < segs) {                                                // idx = lane + 32 t < 32 segs  (warp-uniform)
            const int idx = lane + 32 * t;
            const int r = sh >= 0 ? (idx >> sh) : (idx / segs);
            const int sg = idx - r * segs;
            const uint32_t rp = __shfl_sync(0xffffffffu, my_rp, r);
            if ((rmask >> r) & 1u) ptr[t] = reinterpret_cast<const uint4*>(res + (size_t)rp * e.ld_res) + sg;
        }
    }
#pragma unroll
    for (int t = 0; t < 8; ++t) {
        rv[t] = make_uint4(0, 0, 0, 0);
        if (t < segs && ptr[t] != nullptr) rv[t] = __ldg(ptr[t]);
    }
}

// `wsm`: this warp's kEpiWarpBytes of shared memory.  `bias_s`: smem address of bias[col_base ...] (NT floats).
// `rv` / `preloaded`: residual registers, already holding block 0 of `row` when `preloaded`.  When `next` is
// given, its residual block 0 (at next_col_base) is issued during this tile's last block and the function returns true
// (pass that as the next call's `preloaded`).
template <bool kSplit, bool kF16>
__device__ __forceinline__ bool epilogue_block(const EpilogueArgs& e, uint32_t taddr0, int NT, int col_base,
                                               const EpiRow& row, uint32_t wsm, uint32_t bias_s, int lane,
                                               uint4 (&rv)[8], bool preloaded, const EpiRow* next, int next_col_base) {
    const uint32_t stA = wsm, stB = wsm + kEpiStageBytes;
    const bool any_res = __any_sync(0xffffffffu, row.has_res);
    bool have = preloaded, next_issued = false;
    for (int cb = 0; cb < NT; cb += (int)kEpiBlockCols) {
        const int bw = min((int)kEpiBlockCols, NT - cb);               // columns in this block (multiple of 16)
        const int segs = bw >> 3;                                      // 16-byte segments per row
        const int total = 32 * segs;
        if (any_res) {                                                 // coalesced residual block -> staging A
            if (!have) epilogue_res_issue(e, row, NT, col_base, cb, lane, rv);
#pragma unroll
            for (int t = 0; t < 8; ++t) {
                if (t < segs) {
                    const int idx = lane + 32 * t;
                    const int r = (segs == 8) ? (idx >> 3) : (segs == 4) ? (idx >> 2) : (segs == 2) ? (idx >> 1) : (idx / segs);
                    const int sg = idx - r * segs;
                    st_shared_v4(stA + r * kEpiPitch + sg * 16, rv[t].x, rv[t].y, rv[t].z, rv[t].w);
                }
            }
            __syncwarp();
            have = false;
            if (cb + (int)kEpiBlockCols < NT) { epilogue_res_issue(e, row, NT, col_base, cb + (int)kEpiBlockCols, lane, rv); have = true; }
        }
        if (cb + (int)kEpiBlockCols >= NT && next != nullptr && e.res_mode != RES_NONE && !have) {
            if (__any_sync(0xffffffffu, next->has_res)) { epilogue_res_issue(e, *next, NT, next_col_base, 0, lane, rv); next_issued = true; }
        }
        for (int c = 0; c < bw; c += 32) {
            const int cw = min(32, bw - c);                            // 16 or 32 columns
            uint32_t v[32];
            if (cw == 32) {
                tmem_ld32(taddr0 + cb + c, v);
            } else {
                uint32_t lo[16];
                tmem_ld16(taddr0 + cb + c, lo);
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = lo[i];
            }
            tmem_ld_wait();
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (h * 16 >= cw) break;
                const int cc = cb + c + h * 16;                        // column within the N tile
                const int col = col_base + cc;                         // global output column
                float f[16];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float4 b = ld_shared_f4(bias_s + (cc + 4 * i) * 4);
                    f[4 * i + 0] = __uint_as_float(v[h * 16 + 4 * i + 0]) + b.x;
                    f[4 * i + 1] = __uint_as_float(v[h * 16 + 4 * i + 1]) + b.y;
                    f[4 * i + 2] = __uint_as_float(v[h * 16 + 4 * i + 2]) + b.z;
                    f[4 * i + 3] = __uint_as_float(v[h * 16 + 4 * i + 3]) + b.w;
                }
                if (e.mask != nullptr && row.valid) {
                    const uint4* m4 = reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(e.mask) +
                                                                     row.opix * e.ld_mask + col);
#pragma unroll
                    for (int i = 0; i < 2; ++i) {
                        const uint4 m = __ldg(m4 + i);
                        const uint32_t w[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            f[8 * i + 2 * j + 0] *= op16::pos_lo(w[j]) ? 1.f : e.lrelu_alpha;
                            f[8 * i + 2 * j + 1] *= op16::pos_hi(w[j]) ? 1.f : e.lrelu_alpha;
                        }
                    }
                }
                if (any_res) {
#pragma unroll
                    for (int i = 0; i < 2; ++i) {
                        const uint4 m = ld_shared_v4(stA + lane * kEpiPitch + (c + h * 16) * 2 + i * 16);
                        const uint32_t w[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            { const float2 u2 = op16::unpack<kF16>(w[j]); f[8 * i + 2 * j + 0] += u2.x; f[8 * i + 2 * j + 1] += u2.y; }
                        }
                    }
                }
                if (kSplit && e.res_lo != nullptr && row.has_res) {              // low half of a split-precision residual (this lane's row)
                    const uint4* r4 = reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(e.res_lo) +
                                                                     row.rpix * e.ld_res + col);
#pragma unroll
                    for (int i = 0; i < 2; ++i) {
                        const uint4 m = __ldg(r4 + i);
                        const uint32_t w[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            { const float2 u2 = op16::unpack<kF16>(w[j]); f[8 * i + 2 * j + 0] += u2.x; f[8 * i + 2 * j + 1] += u2.y; }
                        }
                    }
                }
                if (col < e.tanh_cols) {
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (col + i < e.tanh_cols) f[i] = tanhf(f[i]);
                }
                if (e.out_f32 != nullptr && col < e.f32_cols && row.valid) {
                    float* o = e.out_f32 + row.opix * e.ld_f32 + col;
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (col + i < e.f32_cols) o[i] = f[i];
                }
                if (col + 16 > e.bf16_zero_from) {
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (col + i >= e.bf16_zero_from) f[i] = 0.f;
                }
                const uint32_t so = lane * kEpiPitch + (c + h * 16) * 2;
                if (kSplit && e.out_raw_lo != nullptr && row.valid) {            // v - bf16(v), written straight from registers
                    uint32_t lo[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const uint32_t hi = op16::pack<kF16>(f[2 * i], f[2 * i + 1]);
                        lo[i] = op16::pack<kF16>(f[2 * i] - op16::lo<kF16>(hi), f[2 * i + 1] - op16::hi<kF16>(hi));
                    }
                    uint4* o = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(e.out_raw_lo) + row.opix * e.ldo + col);
                    o[0] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                    o[1] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
                }
                if (kSplit && e.out_act_lo != nullptr && row.valid) {
                    uint32_t lo[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float a0 = f[2 * i] > 0.f ? f[2 * i] : e.lrelu_alpha * f[2 * i];
                        const float a1 = f[2 * i + 1] > 0.f ? f[2 * i + 1] : e.lrelu_alpha * f[2 * i + 1];
                        const uint32_t hi = op16::pack<kF16>(a0, a1);
                        lo[i] = op16::pack<kF16>(a0 - op16::lo<kF16>(hi), a1 - op16::hi<kF16>(hi));
                    }
                    uint4* o = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(e.out_act_lo) + row.opix * e.ldo + col);
                    o[0] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                    o[1] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
                }
                if (e.out_raw != nullptr) {
                    st_shared_v4(stA + so, op16::pack<kF16>(f[0], f[1]), op16::pack<kF16>(f[2], f[3]), op16::pack<kF16>(f[4], f[5]), op16::pack<kF16>(f[6], f[7]));
                    st_shared_v4(stA + so + 16, op16::pack<kF16>(f[8], f[9]), op16::pack<kF16>(f[10], f[11]), op16::pack<kF16>(f[12], f[13]), op16::pack<kF16>(f[14], f[15]));
                }
                if (e.out_act != nullptr) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) f[i] = f[i] > 0.f ? f[i] : e.lrelu_alpha * f[i];
                    st_shared_v4(stB + so, op16::pack<kF16>(f[0], f[1]), op16::pack<kF16>(f[2], f[3]), op16::pack<kF16>(f[4], f[5]), op16::pack<kF16>(f[6], f[7]));
                    st_shared_v4(stB + so + 16, op16::pack<kF16>(f[8], f[9]), op16::pack<kF16>(f[10], f[11]), op16::pack<kF16>(f[12], f[13]), op16::pack<kF16>(f[14], f[15]));
                }
            }
        }
        if (e.out_raw != nullptr || e.out_act != nullptr) {
            __syncwarp();
            __nv_bfloat16* oraw = static_cast<__nv_bfloat16*>(e.out_raw);
            __nv_bfloat16* oact = static_cast<__nv_bfloat16*>(e.out_act);
            const int sh = (segs == 8) ? 3 : (segs == 4) ? 2 : (segs == 2) ? 1 : -1;
            const uint32_t vmask = __ballot_sync(0xffffffffu, row.valid);
            const uint32_t my_op = (uint32_t)row.opix;
            for (int idx = lane; idx < total; idx += 32) {
                const int r = sh >= 0 ? (idx >> sh) : (idx / segs);
                const int sg = idx - r * segs;
                const uint32_t op = __shfl_sync(0xffffffffu, my_op, r);
                if ((vmask >> r) & 1u) {
                    const size_t off = (size_t)op * e.ldo + col_base + cb;
                    const uint32_t sa = r * kEpiPitch + sg * 16;
                    if (oraw != nullptr) *(reinterpret_cast<uint4*>(oraw + off) + sg) = ld_shared_v4(stA + sa);
                    if (oact != nullptr) *(reinterpret_cast<uint4*>(oact + off) + sg) = ld_shared_v4(stB + sa);
                }
            }
        }
        __syncwarp();                                                  // staging blocks are reused by the next column block
    }
    return next_issued;
}

// Copies bias[col_base, col_base + NT) into this warp's bias area (call when col_base changes).
__device__ __forceinline__ void epilogue_load_bias(const EpilogueArgs& e, uint32_t bias_s, int NT, int col_base, int lane) {
    for (int i = lane; i < NT; i += 32) {
        const float b = __ldg(e.bias + col_base + i);
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(bias_s + i * 4), "f"(b) : "memory");
    }
    __syncwarp();
}

}  // namespace ptx
}  // namespace apvae
