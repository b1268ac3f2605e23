#include "planner.h"
#include "tuning.h"
#include "op16.h"

#include <stdlib.h>
#include <string.h>
#include <algorithm>

namespace apvae {

namespace {
const apvae_tuning kDefaultTuning = {};
thread_local const apvae_tuning* t_tuning = nullptr;
}  // namespace
const apvae_tuning& tuning() { return t_tuning ? *t_tuning : kDefaultTuning; }
TuningScope::TuningScope(const apvae_tuning* t) : prev_(t_tuning), copy_(t ? *t : kDefaultTuning) { t_tuning = &copy_; }
TuningScope::~TuningScope() { t_tuning = prev_; }


void same_padding(int in, int k, int s, int* out, int* lo, int* hi) {
    const int o = (in + s - 1) / s;
    const int total = std::max((o - 1) * s + k - in, 0);
    *out = o;
    *lo = total / 2;
    *hi = total - total / 2;
}

uint16_t f32_to_bf16(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7F800000u) == 0x7F800000u && (u & 0x007FFFFFu)) return (uint16_t)((u >> 16) | 0x40);
    u += 0x7FFFu + ((u >> 16) & 1u);   // round to nearest even
    return (uint16_t)(u >> 16);
}

float bf16_to_f32(uint16_t h) {
    uint32_t u = (uint32_t)h << 16;
    float f;
    memcpy(&f, &u, 4);
    return f;
}

static int floor_div2(int t) { return (t >= 0) ? t / 2 : -((-t + 1) / 2); }

namespace {

struct Tap {
    int di, dj;       // kernel tap
    int dx, dy, ph;   // source box shift and row phase
    int cbase;        // channel base in the innermost dim (column phase * Cp for s2d)
};

// Taps of one term for output phase (py, px) (phase only meaningful for a
// stride-2 transposed conv).
std::vector<Tap> enumerate_taps(const TermSpec& t, int py, int px) {
    std::vector<Tap> taps;
    int o, lo_h, lo_w, hi;
    if (!t.transpose && t.stride == 1) {
        same_padding(t.in_h, t.k, 1, &o, &lo_h, &hi);
        same_padding(t.in_w, t.k, 1, &o, &lo_w, &hi);
        for (int di = 0; di < t.k; ++di)
            for (int dj = 0; dj < t.k; ++dj)
                taps.push_back({di, dj, dj - lo_w, di - lo_h, 0, 0});
    } else if (!t.transpose) {   // stride-2 conv through the space-to-depth view
        same_padding(t.in_h, t.k, 2, &o, &lo_h, &hi);
        same_padding(t.in_w, t.k, 2, &o, &lo_w, &hi);
        for (int di = 0; di < t.k; ++di)
            for (int dj = 0; dj < t.k; ++dj) {
                const int ty = di - lo_h, tx = dj - lo_w;
                const int qy = floor_div2(ty), qx = floor_div2(tx);
                taps.push_back({di, dj, qx, qy, ty - 2 * qy, (tx - 2 * qx) * t.in_cp});
            }
    } else if (t.stride == 1) {  // transposed conv, stride 1: i = Y - di + pad_lo
        same_padding(t.in_h, t.k, 1, &o, &lo_h, &hi);
        same_padding(t.in_w, t.k, 1, &o, &lo_w, &hi);
        for (int di = 0; di < t.k; ++di)
            for (int dj = 0; dj < t.k; ++dj)
                taps.push_back({di, dj, lo_w - dj, lo_h - di, 0, 0});
    } else {                     // transposed conv, stride 2: Y = 2 i + di - pad_lo
        same_padding(t.in_h * 2, t.k, 2, &o, &lo_h, &hi);
        same_padding(t.in_w * 2, t.k, 2, &o, &lo_w, &hi);
        for (int di = 0; di < t.k; ++di) {
            if ((py + lo_h - di) & 1) continue;
            for (int dj = 0; dj < t.k; ++dj) {
                if ((px + lo_w - dj) & 1) continue;
                taps.push_back({di, dj, floor_div2(px + lo_w - dj), floor_div2(py + lo_h - di), 0, 0});
            }
        }
    }
    return taps;
}

}  // namespace

bool plan_op(const std::vector<TermSpec>& terms, int capacity_n, OpPlan* plan, std::string* err) {
    auto fail = [&](const char* m) { if (err) *err = m; return false; };
    if (terms.empty() || terms.size() > (size_t)kMaxSrc) return fail("1..3 terms expected");
    OpPlan& P = *plan;
    P = OpPlan();
    P.capacity_n = capacity_n;
    const TermSpec& t0 = terms[0];
    const int cp = t0.in_cp;
    P.kc = (cp % 64 == 0) ? 64 : (cp % 32 == 0) ? 32 : 16;
    P.cout = t0.cout;
    P.npad = pad16(t0.cout);
    bool phased = false;
    for (const TermSpec& t : terms) {
        if (t.stride != 1 && t.stride != 2) return fail("stride must be 1 or 2");
        if (t.in_cp % 16 != 0 || t.in_cp < t.cin) return fail("in_cp must be a multiple of 16 and >= cin");
        if (t.in_cp != cp) return fail("terms of one op must share the padded channel count");
        if (t.cout != t0.cout) return fail("terms of one op must share cout");
        if (t.src < 0 || t.src >= kMaxSrc) return fail("bad src index");
        if (t.transpose && t.stride == 2) phased = true;
        if (!t.transpose && t.stride == 2 && ((t.in_h | t.in_w) & 1)) return fail("stride-2 conv needs even input size");
        int oh, ow;
        if (t.transpose) { oh = t.in_h * t.stride; ow = t.in_w * t.stride; }
        else { oh = (t.in_h + t.stride - 1) / t.stride; ow = (t.in_w + t.stride - 1) / t.stride; }
        if (&t == &t0) { P.ho = oh; P.wo = ow; }
        else if (oh != P.ho || ow != P.wo) return fail("terms disagree on the output size");
    }
    if (phased)      // sub-pixel phases: every term must be the same stride-2 transposed geometry (split-precision terms are)
        for (const TermSpec& t : terms)
            if (!t.transpose || t.stride != 2 || t.k != t0.k || t.in_h != t0.in_h || t.in_w != t0.in_w)
                return fail("a stride-2 transposed conv can only be fused with terms of the same geometry");
    P.num_phases = phased ? 4 : 1;
    P.os = phased ? 2 : 1;
    P.hg = phased ? t0.in_h : P.ho;
    P.wg = phased ? t0.in_w : P.wo;
    if (P.wg > kTileM) return fail("grid width > 128 unsupported");
    // pixel tile
    P.bw = P.wg;
    {
        const int bh_max = std::max(1, std::min(P.hg, kTileM / P.bw));
        int best = bh_max, best_tiles = (P.hg + bh_max - 1) / bh_max;
        for (int bh = bh_max; bh >= 1; --bh) {
            const int nt = (P.hg + bh - 1) / bh;
            if (nt < best_tiles) { best = bh; best_tiles = nt; }
        }
        P.bh = best;
        P.bn = (P.bh == P.hg) ? std::max(1, kTileM / (P.bw * P.bh)) : 1;
        if (P.bn > 256) P.bn = 256;
    }
    P.rows = P.bw * P.bh * P.bn;
    P.tiles_x = 1;
    P.tiles_y = (P.hg + P.bh - 1) / P.bh;
    // N tiling
    P.nt = P.npad;
    if (P.npad > 256) {
        P.nt = 256;
        while (P.npad % P.nt) P.nt -= 16;
    }
    P.n_tiles = P.npad / P.nt;
    // sources
    P.nsrc = 0;
    for (const TermSpec& t : terms) P.nsrc = std::max(P.nsrc, t.src + 1);
    bool seen[kMaxSrc] = {false, false, false};
    for (const TermSpec& t : terms) {
        SrcGeom& g = P.src[t.src];
        const bool s2d = (!t.transpose && t.stride == 2);
        const uint64_t C = t.in_cp, W = t.in_w, H = t.in_h;
        SrcGeom n;
        if (!s2d) {
            n.dims[0] = C; n.dims[1] = W; n.dims[2] = 1; n.dims[3] = H; n.dims[4] = capacity_n;
            n.strides[0] = C * 2; n.strides[1] = W * C * 2; n.strides[2] = W * C * 2; n.strides[3] = H * W * C * 2;
        } else {
            n.dims[0] = 2 * C; n.dims[1] = W / 2; n.dims[2] = 2; n.dims[3] = H / 2; n.dims[4] = capacity_n;
            n.strides[0] = 2 * C * 2; n.strides[1] = W * C * 2; n.strides[2] = 2 * W * C * 2; n.strides[3] = H * W * C * 2;
        }
        n.box[0] = P.kc; n.box[1] = P.bw; n.box[2] = 1; n.box[3] = P.bh; n.box[4] = P.bn;
        if (seen[t.src] && memcmp(&g, &n, sizeof(n)) != 0) return fail("two terms view one source differently");
        g = n;
        seen[t.src] = true;
    }
    for (int s = 0; s < P.nsrc; ++s)
        if (!seen[s]) return fail("source indices must be dense");
    // steps + weights + bias
    P.bias.assign(P.npad, 0.f);
    for (const TermSpec& t : terms)
        if (t.bias) for (int n = 0; n < t.cout; ++n) P.bias[n] += t.bias[n];
    const int chunks = cp / P.kc;
    P.macs_per_image = 0;
    for (int ph = 0; ph < P.num_phases; ++ph) {
        P.phase_begin[ph] = (int)P.steps.size();
        for (const TermSpec& t : terms) {
            const std::vector<Tap> taps = enumerate_taps(t, ph >> 1, ph & 1);
            const int used_chunks = (t.cin + P.kc - 1) / P.kc;   // skip all-padding chunks
            for (const Tap& tap : taps) {
                for (int c = 0; c < std::min(chunks, used_chunks); ++c) {
                    TapStep st;
                    st.src = (int16_t)t.src; st.dx = (int16_t)tap.dx; st.dy = (int16_t)tap.dy;
                    st.ph = (int16_t)tap.ph; st.c0 = tap.cbase + c * P.kc;
                    P.steps.push_back(st);
                    const size_t base = P.wpack.size();
                    P.wpack.resize(base + (size_t)P.npad * P.kc, 0);
                    for (int n = 0; n < t.cout; ++n)
                        for (int kk = 0; kk < P.kc; ++kk) {
                            const int ci = c * P.kc + kk;
                            if (ci >= t.cin) break;
                            const size_t tap_off = ((size_t)tap.di * t.k + tap.dj);
                            const float w = t.transpose
                                ? t.kernel[(tap_off * t.cout + n) * t.cin + ci]
                                : t.kernel[(tap_off * t.cin + ci) * t.cout + n];
                            P.wpack[base + (size_t)n * P.kc + kk] = op16::host_from_f32(w, t.f16 != 0);
                        }
                }
            }
            P.macs_per_image += (double)taps.size() * t.cin * t.cout * P.hg * P.wg;
        }
    }
    for (int ph = P.num_phases; ph <= kMaxPhases; ++ph) P.phase_begin[ph] = (int)P.steps.size();
    for (int ph = 0; ph < P.num_phases; ++ph)
        if (P.phase_begin[ph + 1] == P.phase_begin[ph]) return fail("a phase has no taps");
    return true;
}


// ---------------------------------------------------------------------------------------------
// Patch (A-stationary) plan.  Source shift of a tap: conv2d (dj - pad_lo, di - pad_lo);
// conv2d_transpose stride 1 (pad_lo - dj, pad_lo - di)  [SURVEY section 9].
bool plan_patch_op(const std::vector<TermSpec>& terms, int capacity_n, int force_macc, int force_cluster,
                   PatchPlan* plan, std::string* err, int phase) {
    auto fail = [&](const char* m) { if (err) *err = m; return false; };
    // up to kPatchMaxMaps terms over at most kPatchMaxSrc sources: conv + fused shortcut (two sources), or the products of
    // a split-precision op (A_hi W_hi, A_lo W_hi[, A_hi W_lo]: sources hi and lo, one activation map per term)
    if (terms.empty() || terms.size() > (size_t)kPatchMaxMaps) return fail("patch: 1..4 terms");
    PatchPlan& P = *plan;
    P = PatchPlan();
    P.capacity_n = capacity_n;
    const TermSpec& t0 = terms[0];
    // stride-2 conv2d: the grid is the OUTPUT grid; the input is read through its space-to-depth view
    // (row phase = one tensor map each, column phase folded into the channel coordinate)
    bool s2c = (phase < 0);
    for (const TermSpec& t : terms) s2c = s2c && !t.transpose && t.stride == 2;
    if (s2c && !switch_on(tuning().patch_s2, true)) return fail("patch: stride-2 convs disabled (tuning.patch_s2 = -1)");
    if (s2c && ((t0.in_w | t0.in_h) & 1)) return fail("patch: stride-2 conv needs an even input size");
    const int Wreal = s2c ? t0.in_w / 2 : t0.in_w, H = s2c ? t0.in_h / 2 : t0.in_h, cp = t0.in_cp;
    // tile rows are a power of two wide.  A grid up to 1/8 narrower (28, 14, 7: arch M) runs as tiles of the next power of
    // two: the activation / residual / mask boxes are zero-filled and the output boxes clipped by the TMA unit at the image's
    // real width and height, so the memory layout stays the dense [N, H, Wreal, C] and only tile rows are wasted
    int W = 8;
    while (W < Wreal) W <<= 1;
    const bool virt = (W != Wreal);
    if (W > 128 || Wreal * 8 < W * 7) return fail("patch: width must be 8, 16, 32, 64 or 128, or at most 1/8 below one of them");
    if (virt && !switch_on(tuning().patch_virtual, true)) return fail("patch: widths below a power of two disabled (tuning.patch_virtual = -1)");
    // 8-wide images: two images side by side per tile row (PatchGemmParams::fold); needs whole pairs in the buffers
    const int fold = (W == 8) ? 2 : 1;
    if (fold > 1 && (capacity_n & 1)) return fail("patch: 8-wide images need an even capacity");
    if (fold > 1 && !switch_on(tuning().patch_fold, true)) return fail("patch: 8-wide images disabled (tuning.patch_fold = -1)");
    const int Wt = W * fold;
    for (size_t i = 0; i < terms.size(); ++i) {
        const TermSpec& t = terms[i];
        if (phase < 0 && !s2c && t.stride != 1) return fail("patch: every term stride 1, or every term a stride-2 conv2d");
        // sub-pixel phases: stride-2 transposed terms of ONE geometry (one term, or the hi / lo products of a split-precision op)
        if (phase >= 0 && (!t.transpose || t.stride != 2 || t.k != t0.k || t.cin != t0.cin)) return fail("patch: phases are for stride-2 transposed terms of one geometry");
        if (!(t.k & 1)) return fail("patch: odd kernel sizes only");
        if (t.in_w != t0.in_w || t.in_h != t0.in_h) return fail("patch: terms disagree on the grid");
        if (t.in_cp != cp || cp % 16 || cp < t.cin) return fail("patch: bad channel padding");
        if (t.cout != t0.cout) return fail("patch: terms disagree on cout");
        if (t.src < 0 || t.src >= kPatchMaxSrc) return fail("patch: at most two sources");
    }
    P.w = W; P.w_real = Wreal; P.h = H; P.in_cp = cp; P.fold = fold; P.s2d = s2c ? 1 : 0;
    P.kc = (cp % 64 == 0) ? 64 : (cp % 32 == 0) ? 32 : 16;
    // phase 4 + py: both column phases of output rows 2y + py in ONE op.  Its N is [px = 0 | px = 1] x cout, i.e. the two
    // adjacent output pixels (2x, 2x + 1) of input pixel x as one contiguous 2 * cout row (no channel padding allowed)
    const bool merged = phase >= 4;
    if (merged && (phase > 5 || t0.cout % 16 || 2 * t0.cout > 256)) return fail("patch: merged column phases need cout % 16 == 0 and 2 * cout <= 256");
    P.cout = merged ? 2 * t0.cout : t0.cout;
    P.npad = pad16(P.cout);
    if (P.npad > 256) return fail("patch: cout > 256");
    P.nsrc = 0;
    for (const TermSpec& t : terms) P.nsrc = std::max(P.nsrc, t.src + 1);
    const int rows128 = kTileM / Wt;
    if (!virt && H < rows128) return fail("patch: image smaller than one 128-pixel tile");
    // a virtual-width grid may also end inside its last row tile: up to 1/7 more rows than the image has (28 rows as two
    // 16-row tiles of four accumulators instead of seven 4-row tiles of one: the per-tile costs and the patch halo are paid
    // twice instead of seven times per image)
    auto rows_for = [&](int rt) { return virt ? round_up(H, rt) : H; };
    int kmax = 1;
    for (const TermSpec& t : terms) kmax = std::max(kmax, s2c ? (t.k + 1) / 2 : t.k);     // vertical taps that share one patch
    int ktot = 1;                                                                       // taps per output (pair-mode rule)
    for (const TermSpec& t : terms) ktot = std::max(ktot, t.k * t.k);
    // a phase of a stride-2 transposed conv: taps = {dy} x {dx}, consecutive offsets on the input grid
    std::vector<Tap> ph_taps;
    int ph_dy0 = 0, ph_nv = 0, ph_dx0 = 0, ph_nh = 0;
    std::vector<Tap> ph_taps1;            // merged: the taps of column phase px = 1 (ph_taps holds px = 0)
    if (phase >= 0) {
        ph_taps = enumerate_taps(t0, merged ? phase - 4 : phase >> 1, merged ? 0 : phase & 1);
        if (merged) ph_taps1 = enumerate_taps(t0, phase - 4, 1);
        if (ph_taps.empty() || (merged && ph_taps1.empty())) return fail("patch: phase has no taps");
        int dy1 = -1000, dx1 = -1000; ph_dy0 = 1000; ph_dx0 = 1000;
        for (const std::vector<Tap>* v : {&ph_taps, &ph_taps1})
            for (const Tap& tp : *v) {
                ph_dy0 = std::min(ph_dy0, tp.dy); dy1 = std::max(dy1, tp.dy);
                ph_dx0 = std::min(ph_dx0, tp.dx); dx1 = std::max(dx1, tp.dx);
            }
        ph_nv = dy1 - ph_dy0 + 1; ph_nh = dx1 - ph_dx0 + 1;
        // every (dy, dx) of the bounding box must be used by at least one of the phases, and each phase's taps are distinct
        for (int dy = ph_dy0; dy <= dy1; ++dy)
            for (int dx = ph_dx0; dx <= dx1; ++dx) {
                int hits = 0;
                for (const std::vector<Tap>* v : {&ph_taps, &ph_taps1})
                    for (const Tap& tp : *v) hits += (tp.dy == dy && tp.dx == dx);
                if (hits < 1 || hits > (merged ? 2 : 1)) return fail("patch: phase taps are not a product set");
            }
        kmax = ph_nv;
        P.phase = phase; P.out_scale = 2;
    }
    uint32_t budget = 227u * 1024u - kPatchFixedSmem;
    P.tap_bytes = (uint32_t)P.npad * P.kc * 2;
    // weight item = `tpi` vertical taps behind one barrier: up to 32 KB, fewer when the rings would not fit
    int tpi_want = std::max(1, std::min(kmax, (int)(32768u / P.tap_bytes)));
    // CTA pairs (cta_group::2): tuning.patch_2cta = +1 / -1 overrides the default; tuning.patch_2cta_min_k moves the threshold.
    // Measured on B200 (1000 samples, TFLOP/s single -> pair): 96->96 5x5 @32^2 1121 -> 1436, 192->192 3x3 @16^2
    // 1073 -> 1214, 128->128 3x3 @16^2 967 -> 1170, 64->64 3x3 @32^2 697 -> 786, 96->96 3x3 @32^2 778 -> 808; the
    // 16- and 32-channel layers (K < 500 per output) sit on the A-operand fetch floor either way and the whole step
    // was best with the threshold at 500 (sweep 1 / 100 / 250 / 500 / 1500 in one call: profiles/r1_notes.md).
    int two_min_k = 500;
    if (tuning().patch_2cta_min_k > 0) two_min_k = tuning().patch_2cta_min_k;
    int two = ((long long)(phase >= 0 ? ph_nv * ph_nh : ktot) * cp >= two_min_k) ? 1 : 0;
    if (tuning().patch_2cta != 0) two = tuning().patch_2cta > 0;
    if (force_cluster && force_cluster != 2) two = 0;
    P.two_cta = two;
    // in a pair each CTA holds half of every tap tile: a 24 KB tap (192 wide, kc = 64) that would travel alone takes a
    // second tap along (24 KB per CTA and item): the MMA warp then crosses a weight barrier every 8 MMAs instead of every 4
    // (the same effect that made the 96-wide layers need all their taps in one item).  Measured in one call (per-op table, 7400 samples) 192->192 3x3
    // @16^2 1.12 -> 0.97 / 1.23 -> 1.11 ms, @8^2 0.31 -> 0.23 / 0.28 -> 0.25 ms; the 256-wide layers (16 KB per CTA and tap)
    // lose 3 % with two taps and keep one (profiles/r2_notes.md).
    if (two && tpi_want == 1 && kmax >= 2 && P.tap_bytes <= 24576u) tpi_want = 2;
    { const int v = tuning().patch_tpi; if (v >= 1 && v <= kmax) tpi_want = v; }
    // Two epilogue warpgroups when a tile costs the epilogue more than the MMAs.  Measured (clock stamps of CTA 0, 96->96 3x3
    // @32^2): a 64-column block takes one epilogue warp ~2.8 k cycles, a 16-column one ~1.0 k (~400 + 37 per column); an MMA
    // of M = 128 per CTA takes max(N / 2, 32) cycles at best (the 4 KB A-operand read from shared memory).  The wider
    // staging costs 39 KB of ring space, so it is only taken when the tile is clearly epilogue-bound.
    {
        const int nb_acc = (P.npad + 63) / 64;
        const int last_cols = P.npad - (nb_acc - 1) * 64;
        long long taps = 0;                                    // (tap, 16-channel slice) MMAs per accumulator and tile
        if (phase >= 0) taps = (long long)ph_nv * ph_nh * ((t0.cin + P.kc - 1) / P.kc) * (P.kc / 16) * (long long)terms.size();
        else for (const TermSpec& t : terms) taps += (long long)(s2c ? t.k * t.k : t.k * t.k) * ((t.cin + P.kc - 1) / P.kc) * (P.kc / 16);
        const double mma_cyc = (double)taps * std::max(P.npad / 2, 32);
        const double epi_cyc = (nb_acc - 1) * 2800.0 + (400.0 + 37.0 * last_cols);
        const bool want = epi_cyc > 1.25 * mma_cyc;
        const int sw = tuning().patch_epi2;
        P.epi2 = (sw != 0 ? sw > 0 : want) ? 1 : 0;
        if (P.epi2) budget = 227u * 1024u - kPatchFixedSmem2;
    }
    auto b_slot_for = [&](int tpi) { return ((uint32_t)tpi * (P.tap_bytes / (two ? 2 : 1)) + 1023u) & ~1023u; };
    auto a_slot_for = [&](int macc) { return ((uint32_t)(macc * rows128 + kmax - 1) * Wt * P.kc * 2 + 1023u) & ~1023u; };
    // accumulator count, taps per item and ring depths for a given ring budget
    struct Rings { int macc = 0, tpi = 1, a_stages = 0, b_stages = 0; };
    int a_cap = 4;
    { const int v = tuning().patch_a_cap; if (v >= 2 && v <= 12) a_cap = v; }
    auto size_rings = [&](uint32_t bud) {
        Rings r;
        for (int cand : {4, 2, 1}) {
            if (force_macc && cand != force_macc) continue;
            if (cand * P.npad > 256) continue;
            const int rt = cand * rows128;
            const int Hv = rows_for(rt);
            if (rt > Hv || Hv % rt || (Hv - H) * 7 > H) continue;
            int tpi = tpi_want;
            while (tpi > 1 && 2 * a_slot_for(cand) + 2u * b_slot_for(tpi) > bud) --tpi;
            if (2 * a_slot_for(cand) + 2u * b_slot_for(tpi) > bud) continue;
            r.macc = cand;
            r.tpi = tpi;
            break;
        }
        if (!r.macc) return r;
        const uint32_t as = a_slot_for(r.macc), bs = b_slot_for(r.tpi);
        // ring depths: weights first (k+2 tiles keep one dx group ahead), then patches, then more weights
        r.b_stages = std::min(12, (kmax + r.tpi - 1) / r.tpi + 2);
        while (r.b_stages > 2 && 2 * as + r.b_stages * bs > bud) --r.b_stages;
        r.a_stages = 2;
        while (r.a_stages < a_cap && (r.a_stages + 1) * as + r.b_stages * bs <= bud) ++r.a_stages;
        while (r.b_stages < 12 && r.a_stages * as + (r.b_stages + 1) * bs <= bud) ++r.b_stages;
        return r;
    };
    Rings R = size_rings(budget);
    if (P.epi2 && tuning().patch_epi2 == 0) {
        // the wider staging must not cost the op what it was short of: same accumulator count, and a patch ring as deep as
        // without it unless the patches are small (kc <= 32) or the op is a 1x1 (three groups per tile).  Measured layer by
        // layer (tools/r2_epi2_layers.sh): 64->64 3x3 @32^2 +16 % with macc 4 -> 2, 64->128 stride 2 +20 % with 3 -> 2 patch stages,
        // against -7 .. -22 % for the 16- / 32-channel layers, -13 .. -17 % for the 1x1 shortcuts, -6 % for 96->96 3x3.
        const Rings R1 = size_rings(227u * 1024u - kPatchFixedSmem);
        const bool ok = R.macc && R.macc == R1.macc && (R.a_stages >= R1.a_stages || P.kc <= 32 || ktot == 1);
        if (!ok) { P.epi2 = 0; budget = 227u * 1024u - kPatchFixedSmem; R = R1; }
    }
    if (!R.macc) return fail("patch: no accumulator count fits");
    const int macc = R.macc;
    P.tpi = R.tpi;
    P.macc = macc;
    P.rows_tile = macc * rows128;
    P.tiles_y = (H + P.rows_tile - 1) / P.rows_tile;
    P.a_slot_bytes = a_slot_for(macc);
    P.b_slot_bytes = b_slot_for(P.tpi);
    P.a_stages = R.a_stages;
    P.b_stages = R.b_stages;
    { const int v = tuning().patch_a_stages; if (v >= 1 && v <= P.a_stages) P.a_stages = v; }
    { const int v = tuning().patch_b_stages; if (v >= 1 && v <= P.b_stages) P.b_stages = v; }
    // cluster: weight-stream multicast across 2 / 4 CTAs is available (force_cluster) but measured slower
    // than independent CTAs on B200 for every layer of the reference models (the weights live in L2 anyway
    // and the cluster-wide slot release adds a round trip), so the default is 1.
    int c = P.two_cta ? 2 : 1;
    if (force_cluster) {
        if (force_cluster != 1 && force_cluster != 2 && force_cluster != 4) return fail("patch: cluster must be 1, 2 or 4");
        if ((P.npad / force_cluster) % 8) return fail("patch: cout slice per cluster CTA must be a multiple of 8");
        c = force_cluster;
    }
    P.cluster = c;
    // groups + packed weights + bias
    P.bias.assign(P.npad, 0.f);
    for (const TermSpec& t : terms)
        if (t.bias) for (int n = 0; n < t.cout; ++n) { P.bias[n] += t.bias[n]; if (merged) P.bias[t.cout + n] += t.bias[n]; }
    P.macs_per_image = 0;
    // one weight tile [npad][kc] per item; fill_tap writes kernel tap (di, dj) into rows row0 .. row0 + cout - 1 of the newest
    auto new_item = [&]() { P.wpack.resize(P.wpack.size() + (size_t)P.npad * P.kc, 0); ++P.num_items; };
    auto fill_tap = [&](const TermSpec& t, int c0, int di, int dj, int row0) {
        const size_t base = P.wpack.size() - (size_t)P.npad * P.kc;
        const size_t tap_off = (size_t)di * t.k + dj;
        for (int n = 0; n < t.cout; ++n)
            for (int kk = 0; kk < P.kc; ++kk) {
                const int ci = c0 * P.kc + kk;
                if (ci >= t.cin) break;
                const float w = t.transpose ? t.kernel[(tap_off * t.cout + n) * t.cin + ci]
                                            : t.kernel[(tap_off * t.cin + ci) * t.cout + n];
                P.wpack[base + (size_t)(row0 + n) * P.kc + kk] = op16::host_from_f32(w, t.f16 != 0);
            }
    };
    auto pack_tap = [&](const TermSpec& t, int c0, int di, int dj) { new_item(); fill_tap(t, c0, di, dj, 0); };
    for (const TermSpec& t : terms) {
        const int used_chunks = (t.cin + P.kc - 1) / P.kc;
        if (s2c) {
            // taps by (row phase, column phase, dx); the vertical taps of one (row phase) are consecutive row shifts
            const std::vector<Tap> taps = enumerate_taps(t, 0, 0);
            int map_of[2] = {-1, -1}, dy0[2] = {0, 0}, nv[2] = {0, 0};
            for (int ph = 0; ph < 2; ++ph) {
                int lo = 1000, hi2 = -1000;
                for (const Tap& tp : taps) if (tp.ph == ph) { lo = std::min(lo, tp.dy); hi2 = std::max(hi2, tp.dy); }
                if (hi2 < lo) continue;
                if (P.nmaps >= kPatchMaxMaps) return fail("patch: too many activation maps");
                map_of[ph] = P.nmaps; dy0[ph] = lo; nv[ph] = hi2 - lo + 1;
                P.map_src[P.nmaps] = t.src; P.map_ph[P.nmaps] = ph; P.prow[P.nmaps] = P.rows_tile + nv[ph] - 1;
                ++P.nmaps;
            }
            for (int c0 = 0; c0 < used_chunks; ++c0)
                for (int ph = 0; ph < 2; ++ph) {
                    if (map_of[ph] < 0) continue;
                    for (int cph = 0; cph < 2; ++cph)
                        for (int dx = -8; dx <= 8; ++dx) {
                            std::vector<const Tap*> col(nv[ph], nullptr);
                            int found = 0;
                            for (const Tap& tp : taps)
                                if (tp.ph == ph && tp.cbase == cph * cp && tp.dx == dx) { col[tp.dy - dy0[ph]] = &tp; ++found; }
                            if (!found) continue;
                            if (found != nv[ph]) return fail("patch: stride-2 taps are not a product set");
                            PatchGroup gr;
                            gr.src = (int16_t)map_of[ph]; gr.dx = (int16_t)dx; gr.py0 = (int16_t)dy0[ph]; gr.k = (int16_t)nv[ph];
                            gr.c0 = cph * cp + c0 * P.kc; gr.prow = P.prow[map_of[ph]];
                            P.groups.push_back(gr);
                            for (int j = 0; j < nv[ph]; ++j) pack_tap(t, c0, col[j]->di, col[j]->dj);
                        }
                }
            P.macs_per_image += (double)t.k * t.k * t.cin * t.cout * H * Wreal;
            continue;
        }
        int o, lo_h, lo_w, hi;
        same_padding(H, t.k, 1, &o, &lo_h, &hi);
        same_padding(Wreal, t.k, 1, &o, &lo_w, &hi);
        const int nv = phase >= 0 ? ph_nv : t.k, nh = phase >= 0 ? ph_nh : t.k;
        const int dy0 = phase >= 0 ? ph_dy0 : -lo_h, dx0 = phase >= 0 ? ph_dx0 : -lo_w;
        if (P.nmaps >= kPatchMaxMaps) return fail("patch: too many activation maps");
        P.map_src[P.nmaps] = t.src; P.map_ph[P.nmaps] = 0;
        P.prow[P.nmaps] = P.rows_tile + nv - 1;
        const int map = P.nmaps++;
        for (int c0 = 0; c0 < used_chunks; ++c0) {
            for (int g = 0; g < nh; ++g) {             // dx ascending
                const int dx = dx0 + g;
                PatchGroup gr;
                gr.src = (int16_t)map; gr.dx = (int16_t)dx; gr.py0 = (int16_t)dy0; gr.k = (int16_t)nv;
                gr.c0 = c0 * P.kc; gr.prow = P.prow[map];
                P.groups.push_back(gr);
                for (int j = 0; j < nv; ++j) {         // dy ascending
                    const int dy = dy0 + j;
                    int di = -1, dj = -1;
                    if (merged) {                      // rows [0, cout): column phase 0; [cout, 2 cout): column phase 1
                        new_item();
                        for (const Tap& tp : ph_taps) if (tp.dy == dy && tp.dx == dx) fill_tap(t, c0, tp.di, tp.dj, 0);
                        for (const Tap& tp : ph_taps1) if (tp.dy == dy && tp.dx == dx) fill_tap(t, c0, tp.di, tp.dj, t.cout);
                        continue;
                    }
                    if (phase >= 0) {
                        for (const Tap& tp : ph_taps)
                            if (tp.dy == dy && tp.dx == dx) { di = tp.di; dj = tp.dj; }
                        if (di < 0) return fail("patch: phase tap missing");
                    } else {
                        dj = t.transpose ? lo_w - dx : dx + lo_w;
                        di = t.transpose ? lo_h - dy : dy + lo_h;
                    }
                    pack_tap(t, c0, di, dj);
                }
            }
        }
        P.macs_per_image += (merged ? (double)(ph_taps.size() + ph_taps1.size()) : (double)nv * nh) * t.cin * t.cout * H * Wreal;
    }
    if ((int)P.groups.size() > kPatchMaxGroups) return fail("patch: too many groups");
    return true;
}

}  // namespace apvae
