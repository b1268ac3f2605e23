#include "model.h"
#include "tuning.h"

#include <cuda_runtime.h>
#include <stdlib.h>
#include <string.h>
#include "kernels.h"
#include "op16.h"

namespace apvae {

Model::~Model() {
    q_prog.clear(); g_prog.clear(); d_prog.clear(); gd_prog.clear(); gg_prog.clear();
    for (float* p : {mz, mzp, mg, mgp, me, mep, mp, mpp, meps, mu})
        if (p) cudaFree(p);
    for (void* p : allocs) cudaFree(p);
    for (float* p : {eps_buf, log_px, log_q, log_pn, d_g, sq_err})
        if (p) cudaFree(p);
}

long long Model::z_stride() const {
    if (cfg.arch == APVAE_ARCH_A) return (long long)g_base_h() * g_base_w() * pad16(zc());
    return pad16(cfg.z_dim);
}
int Model::z_group() const { return cfg.arch == APVAE_ARCH_A ? zc() : cfg.z_dim; }
int Model::z_cpad() const { return cfg.arch == APVAE_ARCH_A ? pad16(zc()) : pad16(cfg.z_dim); }
float Model::w_xi() const { return cfg.pull_back_energy_weight * (1.f / (1.f + expf(-xi_raw))); }

Tensor Model::alloc_act(int n, int h, int w, int c) {
    Tensor t;
    t.n = n; t.h = h; t.w = w; t.c = c; t.cp = pad16(c);
    const size_t bytes = t.elems() * 2;
    if (cudaMalloc(&t.ptr, bytes) != cudaSuccess) { t.ptr = nullptr; err = "cudaMalloc failed (activation arena)"; return t; }
    cudaMemset(t.ptr, 0, bytes);
    allocs.push_back(t.ptr);
    return t;
}

float* Model::alloc_f32(size_t n) {
    float* p = nullptr;
    if (cudaMalloc(&p, n * sizeof(float)) != cudaSuccess) { err = "cudaMalloc failed (fp32 buffer)"; return nullptr; }
    cudaMemset(p, 0, n * sizeof(float));
    allocs.push_back(p);
    return p;
}

const HostArr* Model::W_(const std::string& name, std::initializer_list<int64_t> shape) {
    auto it = weights.find(name);
    if (it == weights.end()) { err = "missing variable '" + name + "'"; return nullptr; }
    if (it->second.shape != std::vector<int64_t>(shape)) { err = "variable '" + name + "' has the wrong shape"; return nullptr; }
    return &it->second;
}

bool Model::add_op(Program& prog, const std::string& name, const std::vector<TermSpec>& terms_in,
                   const std::vector<const void*>& srcs, const Epilogue& epi_in, int cap, double* macs) {
    std::string perr;
    // every op of a handle packs its kernels and writes its 16-bit outputs in the handle's operand format
    std::vector<TermSpec> terms = terms_in;
    for (TermSpec& t : terms) t.f16 = f16();
    Epilogue epi = epi_in;
    epi.f16 = f16();
    std::unique_ptr<DevOp> op(new DevOp());
    op->name = name;
    // <= 4 -> <= 4 channels, stride 1 (the first conv on the image channels and its adjoint): CUDA cores (tinyconv.cu);
    // tuning.tiny_conv = -1 keeps the tensor-core op (tuning / A-B only)
    {
        const bool lo_parts = epi.out_raw_lo || epi.out_act_lo || epi.res_lo;
        // a split-precision op (push_terms): A_hi W_hi + A_lo W_hi [+ A_hi W_lo] over sources (hi, lo) — here the pair is summed
        // on load and the fp32 kernel is W_hi (+ W_lo); every output and the residual must come as complete pairs
        bool split_tiny = lo_parts && (terms.size() == 2 || terms.size() == 3) && srcs.size() == 2 && !epi.mask &&
                          (!epi.out_raw == !epi.out_raw_lo) && (!epi.out_act == !epi.out_act_lo) &&
                          (epi.res_mode == RES_NONE ? !epi.res_lo : epi.res_lo != nullptr);
        for (size_t i = 0; split_tiny && i < terms.size(); ++i)
            split_tiny = terms[i].src == (i == 1 ? 1 : 0) && terms[i].k == terms[0].k && terms[i].stride == terms[0].stride &&
                         terms[i].transpose == terms[0].transpose && terms[i].cin == terms[0].cin && terms[i].cout == terms[0].cout &&
                         (i == 0 || !terms[i].bias) && (i != 1 || terms[i].kernel == terms[0].kernel);
        if (switch_on(tuning().tiny_conv, true) && (terms.size() == 1 ? !lo_parts : split_tiny) && terms[0].stride == 1 && terms[0].cin <= kTinyMaxC && terms[0].cout <= kTinyMaxC &&
            terms[0].k <= kTinyMaxK && (terms[0].k & 1) && epi.tanh_cols == 0 && epi.bf16_zero_from >= terms[0].cout &&
            terms[0].in_cp % 8 == 0 && (epi.ldo % 8 == 0 || !(epi.out_raw || epi.out_act)) &&
            (!epi.out_f32 || epi.f32_cols == terms[0].cout) && (epi.res_mode == RES_NONE || epi.res)) {
            const TermSpec& t = terms[0];
            op->is_tiny = true;
            op->capacity_n = cap;
            TinyConvParams& tc = op->tiny;
            memset(&tc, 0, sizeof(tc));
            tc.in = srcs[0]; tc.H = t.in_h; tc.W = t.in_w; tc.cp_in = t.in_cp; tc.cin = t.cin; tc.cout = t.cout; tc.k = t.k;
            const int lo = (t.k - 1) / 2;                          // SAME padding, stride 1, odd k
            for (int di = 0; di < t.k; ++di)
                for (int dj = 0; dj < t.k; ++dj) {
                    const int tp = tc.ntaps++;
                    tc.dy[tp] = (int8_t)(t.transpose ? lo - di : di - lo);
                    tc.dx[tp] = (int8_t)(t.transpose ? lo - dj : dj - lo);
                    const size_t tap_off = (size_t)di * t.k + dj;
                    for (int ci = 0; ci < t.cin; ++ci)
                        for (int co = 0; co < t.cout; ++co) {
                            const size_t wi = t.transpose ? (tap_off * t.cout + co) * t.cin + ci : (tap_off * t.cin + ci) * t.cout + co;
                            tc.w[(tp * kTinyMaxC + ci) * kTinyMaxC + co] = t.kernel[wi] + (terms.size() == 3 ? terms[2].kernel[wi] : 0.f);
                        }
                }
            if (t.bias) for (int co = 0; co < t.cout; ++co) tc.bias[co] = t.bias[co];
            tc.out_raw = epi.out_raw; tc.out_act = epi.out_act; tc.ldo = epi.ldo;
            tc.out_f32 = epi.out_f32; tc.ld_f32 = epi.ld_f32;
            tc.res = epi.res; tc.res_mode = epi.res ? epi.res_mode : RES_NONE; tc.ld_res = epi.ld_res;
            tc.mask = epi.mask; tc.ld_mask = epi.ld_mask;
            if (split_tiny) { tc.in_lo = srcs[1]; tc.res_lo = epi.res_lo; tc.out_raw_lo = epi.out_raw_lo; tc.out_act_lo = epi.out_act_lo; }
            tc.alpha = 0.2f;
            tc.f16 = f16();
            op->macs_per_image = (double)t.k * t.k * t.cin * t.cout * t.in_h * t.in_w;
            if (macs) *macs += op->macs_per_image;
            prog.push_back(std::move(op));
            return true;
        }
    }
    // stride-1 convs on wide grids: the A-stationary kernel; everything else: the tap-GEMM kernel
    const bool patch_on = switch_on(tuning().patch, true);
    PatchPlan pplan;
    const int force_macc = tuning().patch_macc, force_cluster = tuning().patch_cluster;
    const int patch_max_n = tuning().patch_max_n > 0 ? tuning().patch_max_n : 256;    // widest output the A-stationary kernel takes
    const bool split_epi = epi.out_raw_lo || epi.out_act_lo || epi.res_lo;     // split precision: low-half outputs / residual (generic epilogue)
    // stride-2 transposed conv: four sub-pixel phases, each an A-stationary stride-1 op on the input grid
    bool all_s2t = !terms.empty();
    for (const TermSpec& t : terms) all_s2t = all_s2t && t.transpose && t.stride == 2 && t.k == terms[0].k && t.cin == terms[0].cin;
    const bool split_on_patch = switch_on(tuning().patch_split, true);
    if (patch_on && switch_on(tuning().patch_phases, true) && (split_epi ? split_on_patch : terms.size() == 1) && all_s2t) {
        std::vector<std::unique_ptr<DevOp>> ops;
        double m = 0;
        bool ok = true;
        // two ops, one per output row phase, each writing both column phases as one 2 * cout-wide row (needs unpadded
        // output rows and 2 * cout <= 256); else one op per sub-pixel phase.  APVAE_PATCH_MERGE=0: always four.
        if (switch_on(tuning().patch_merge, true)) {
            for (int py = 0; py < 2 && ok; ++py) {
                PatchPlan pp;
                std::string e2;
                std::unique_ptr<DevOp> o(new DevOp());
                o->name = name + "/rows" + std::to_string(py);
                ok = plan_patch_op(terms, cap, force_macc, force_cluster, &pp, &e2, 4 + py) && pp.npad <= patch_max_n &&
                     devop_create_patch(dev, pp, srcs.data(), epi, o.get(), &e2);
                if (ok) { m += pp.macs_per_image; ops.push_back(std::move(o)); }
            }
            if (ok) {
                if (macs) *macs += m / split_terms();
                for (auto& o : ops) prog.push_back(std::move(o));
                return true;
            }
            cudaGetLastError();
            ops.clear(); m = 0; ok = true;
        }
        for (int ph = 0; ph < 4 && ok; ++ph) {
            PatchPlan pp;
            std::string e2;
            std::unique_ptr<DevOp> o(new DevOp());
            o->name = name + "/phase" + std::to_string(ph);
            ok = plan_patch_op(terms, cap, force_macc, force_cluster, &pp, &e2, ph) && pp.npad <= patch_max_n &&
                 devop_create_patch(dev, pp, srcs.data(), epi, o.get(), &e2);
            if (ok) { m += pp.macs_per_image; ops.push_back(std::move(o)); }
        }
        if (ok) {
            if (macs) *macs += m / split_terms();
            for (auto& o : ops) prog.push_back(std::move(o));
            return true;
        }
        cudaGetLastError();
    }
    // (a split-precision op fits when it has two terms, hi and lo activations on single kernels — precise = 2; its low
    // halves leave through the generic epilogue)
    if (patch_on && (!split_epi || split_on_patch) &&
        plan_patch_op(terms, cap, force_macc, force_cluster, &pplan, &perr) && pplan.npad <= patch_max_n) {
        if ((int)srcs.size() != pplan.nsrc) return fail("plan " + name + ": source count mismatch");
        const bool made = devop_create_patch(dev, pplan, srcs.data(), epi, op.get(), &perr);
        if (!made && pplan.fold == 1) return fail(perr);
        if (made) {
            if (macs) *macs += pplan.macs_per_image / split_terms();
            prog.push_back(std::move(op));
            return true;
        }
        // a folded (8-wide) op the A-stationary kernel cannot take (half-resolution residual): the tap-GEMM kernel runs it
        cudaGetLastError();
        op.reset(new DevOp());
        op->name = name;
    }
    OpPlan plan;
    if (!plan_op(terms, cap, &plan, &perr)) return fail("plan " + name + ": " + perr);
    if ((int)srcs.size() != plan.nsrc) return fail("plan " + name + ": source count mismatch");
    if (!devop_create(dev, plan, srcs.data(), epi, op.get(), &perr)) return fail(perr);
    if (macs) *macs += plan.macs_per_image / split_terms();   // split precision: 3 (or 2) tensor-core terms per algorithmic product
    prog.push_back(std::move(op));
    return true;
}

static TermSpec term(int src, bool transpose, int k, int stride, int cin, int cout, const Tensor& in,
                     const HostArr* kernel, const HostArr* bias) {
    TermSpec t;
    t.src = src; t.transpose = transpose ? 1 : 0; t.k = k; t.stride = stride; t.cin = cin; t.cout = cout;
    t.in_h = in.h; t.in_w = in.w; t.in_cp = in.cp;
    t.kernel = kernel->v.data();
    t.bias = bias ? bias->v.data() : nullptr;
    return t;
}

bool Model::push_terms(std::vector<TermSpec>& terms, std::vector<const void*>& srcs, bool split, const Tensor& hi,
                       const Tensor& lo, bool transpose, int k, int stride, int cin, int cout, const std::string& wname,
                       const HostArr* kernel, const HostArr* bias) {
    const int s0 = (int)srcs.size();
    if (!split) {
        terms.push_back(term(s0, transpose, k, stride, cin, cout, hi, kernel, bias));
        srcs.push_back(hi.ptr);
        return true;
    }
    if (!lo.ptr) return fail("split-precision op " + wname + " has no low-half input");
    HostArr& whi = weights[wname + "#hi"];
    HostArr& wlo = weights[wname + "#lo"];
    if (whi.v.empty()) {
        whi.shape = kernel->shape; wlo.shape = kernel->shape;
        whi.v.resize(kernel->v.size()); wlo.v.resize(kernel->v.size());
        for (size_t i = 0; i < kernel->v.size(); ++i) {
            whi.v[i] = op16::host_round(kernel->v[i], f16() != 0);
            wlo.v[i] = op16::host_round(kernel->v[i] - whi.v[i], f16() != 0);
        }
    }
    terms.push_back(term(s0, transpose, k, stride, cin, cout, hi, &whi, bias));
    terms.push_back(term(s0 + 1, transpose, k, stride, cin, cout, lo, &whi, nullptr));
    // precise = 2 ("activation split"): kernels stay single 16-bit values (A_hi W + A_lo W).  Their rounding is a fixed,
    // smooth perturbation of the energy that cancels in MALA's E(z) - E(z'); the per-element rounding of the ACTIVATIONS,
    // which differs between z and z' and is what turns accept / reject into noise, is removed at 2x instead of 3x the MMAs.
    if (cfg.precise != 2) terms.push_back(term(s0, transpose, k, stride, cin, cout, hi, &wlo, nullptr));
    srcs.push_back(hi.ptr);
    srcs.push_back(lo.ptr);
    return true;
}

// resnet_conv2d_block (resnet.py:272-364, resize_at_exit=True):
//   op1: r = lrelu(conv_k,1(lrelu(x)) + b0)                       [Cin -> Cin]
//   op2: out = conv_k,s(r) + b1 + (shortcut: conv_1,s(x) + bs | x) [Cin -> Cout]
// The shortcut is a second source term of op2 (bf16 mode) or its own 1x1 op whose output is op2's residual
// (split-precision mode, where op2 already has three terms).
bool Model::conv_block(Program& prog, const std::string& root, const BlockSpec& b, const Pair& in, bool need_raw,
                       bool need_act, int cap, Pair* out, double* macs, std::vector<BlockTape>* tape) {
    const std::string pre = root + "/" + b.scope + "/";
    const bool sp = split_mode;
    const int cin = in.c(), cout = b.cout, k = b.k, s = b.stride, H = in.h(), Wd = in.w();
    const bool use_sc = (s != 1) || (cin != cout);
    const HostArr *k0 = W_(pre + "residual/conv_0/kernel", {k, k, cin, cin}), *b0 = W_(pre + "residual/conv_0/bias", {cin});
    const HostArr *k1 = W_(pre + "residual/conv_1/kernel", {k, k, cin, cout}), *b1 = W_(pre + "residual/conv_1/bias", {cout});
    if (!k0 || !b0 || !k1 || !b1) return false;
    Tensor r = alloc_act(cap, H, Wd, cin), r_lo;
    if (!r.ptr) return false;
    if (sp) { r_lo = alloc_act(cap, H, Wd, cin); if (!r_lo.ptr) return false; }
    // <= 3 input channels (D_psi/level_0: 3 -> 3 -> 16 at 32 x 32 in the CIFAR models, 3 -> 3 -> 32 with a stride-1 / stride-2
    // exit at 32 x 32 / 64 x 64 in arch B): the whole block is one HBM-bound CUDA-core kernel (block0.cu) with fp32 weights,
    // also for split-precision tensors; tuning.small_block = -1 keeps the tensor-core ops (tuning / A-B only)
    if (switch_on(tuning().small_block, true) && (s == 1 || s == 2) && k == kSmallBlockMaxK && cin <= kSmallBlockMaxCin &&
        (cout == 16 || cout == 32) && use_sc && H <= kSmallBlockWideMaxHW && Wd <= kSmallBlockWideMaxHW && !(H % s) && !(Wd % s)) {
        const HostArr *ks = W_(pre + "shortcut/kernel", {1, 1, cin, cout}), *bs = W_(pre + "shortcut/bias", {cout});
        if (!ks || !bs) return false;
        const int Ho = H / s, Wo = Wd / s;
        Pair o;
        if (need_raw) { o.raw = alloc_act(cap, Ho, Wo, cout); if (!o.raw.ptr) return false; }
        if (need_act) { o.act = alloc_act(cap, Ho, Wo, cout); if (!o.act.ptr) return false; }
        if (sp && need_raw) { o.raw_lo = alloc_act(cap, Ho, Wo, cout); if (!o.raw_lo.ptr) return false; }
        if (sp && need_act) { o.act_lo = alloc_act(cap, Ho, Wo, cout); if (!o.act_lo.ptr) return false; }
        if (sp && !in.raw_lo.ptr) return fail("split-precision block " + pre + " has no low-half input");
        std::unique_ptr<DevOp> op(new DevOp());
        op->name = pre + "block";
        op->is_small_block = true;
        op->capacity_n = cap;
        op->macs_per_image = (double)k * k * cin * cin * H * Wd + ((double)k * k * cin * cout + (double)cin * cout) * Ho * Wo;
        SmallBlockParams& sb = op->sblock;
        memset(&sb, 0, sizeof(sb));
        sb.in_raw = in.raw.ptr; sb.r_act = r.ptr; sb.out_raw = o.raw.ptr; sb.out_act = o.act.ptr;
        if (sp) { sb.in_raw_lo = in.raw_lo.ptr; sb.r_act_lo = r_lo.ptr; sb.out_raw_lo = o.raw_lo.ptr; sb.out_act_lo = o.act_lo.ptr; }
        sb.H = H; sb.W = Wd; sb.cp_in = in.raw.cp; sb.cin = cin; sb.cout = cout; sb.ldo = pad16(cout); sb.k = k; sb.stride = s;
        sb.alpha = 0.2f;
        sb.f16 = f16();
        // precise = 2 keeps the kernels at their 16-bit values (the model whose energy the sampler walks, model.h)
        auto wv = [&](float v) { return cfg.precise == 2 && sp ? op16::host_round(v, f16() != 0) : v; };
        for (int c = 0; c < cin; ++c) sb.b0[c] = b0->v[c];
        for (int c = 0; c < cout; ++c) sb.b1[c] = b1->v[c] + bs->v[c];
        // kernel template is instantiated on cin (weights packed [tap][cin][cout] with that cin)
        for (int tp = 0; tp < k * k; ++tp)
            for (int ci = 0; ci < cin; ++ci) {
                for (int co = 0; co < cin; ++co) sb.w0[(tp * cin + ci) * cin + co] = wv(k0->v[((size_t)tp * cin + ci) * cin + co]);
                for (int co = 0; co < cout; ++co) sb.w1[(tp * cin + ci) * cout + co] = wv(k1->v[((size_t)tp * cin + ci) * cout + co]);
            }
        for (int ci = 0; ci < cin; ++ci)
            for (int co = 0; co < cout; ++co) sb.ws[ci * cout + co] = wv(ks->v[(size_t)ci * cout + co]);
        if (macs) *macs += op->macs_per_image;
        prog.push_back(std::move(op));
        if (tape) {
            BlockTape t; t.scope = pre; t.deconv = false; t.use_shortcut = use_sc; t.k = k; t.stride = s;
            t.cin = cin; t.cout = cout; t.in = in; t.r_act = r; t.out = o; t.fused = true;
            tape->push_back(t);
        }
        *out = o;
        return true;
    }
    {
        Epilogue e; e.out_act = r.ptr; e.out_act_lo = r_lo.ptr; e.ldo = r.cp;
        std::vector<TermSpec> terms; std::vector<const void*> srcs;
        if (!push_terms(terms, srcs, sp, in.act, in.act_lo, false, k, 1, cin, cin, pre + "residual/conv_0/kernel", k0, b0)) return false;
        if (!add_op(prog, pre + "conv_0", terms, srcs, e, cap, macs)) return false;
    }
    const int Ho = (H + s - 1) / s, Wo = (Wd + s - 1) / s;
    Pair o;
    if (need_raw) { o.raw = alloc_act(cap, Ho, Wo, cout); if (!o.raw.ptr) return false; }
    if (need_act) { o.act = alloc_act(cap, Ho, Wo, cout); if (!o.act.ptr) return false; }
    if (sp && need_raw) { o.raw_lo = alloc_act(cap, Ho, Wo, cout); if (!o.raw_lo.ptr) return false; }
    if (sp && need_act) { o.act_lo = alloc_act(cap, Ho, Wo, cout); if (!o.act_lo.ptr) return false; }
    Epilogue e;
    e.out_raw = o.raw.ptr; e.out_act = o.act.ptr; e.out_raw_lo = o.raw_lo.ptr; e.out_act_lo = o.act_lo.ptr; e.ldo = pad16(cout);
    std::vector<TermSpec> terms;
    std::vector<const void*> srcs;
    if (!push_terms(terms, srcs, sp, r, r_lo, false, k, s, cin, cout, pre + "residual/conv_1/kernel", k1, b1)) return false;
    if (use_sc) {
        const HostArr *ks = W_(pre + "shortcut/kernel", {1, 1, cin, cout}), *bs = W_(pre + "shortcut/bias", {cout});
        if (!ks || !bs) return false;
        if (!sp) {
            terms.push_back(term(1, false, 1, s, cin, cout, in.raw, ks, bs));
            srcs.push_back(in.raw.ptr);
        } else {
            Tensor sc = alloc_act(cap, Ho, Wo, cout), sc_lo = alloc_act(cap, Ho, Wo, cout);
            if (!sc.ptr || !sc_lo.ptr) return false;
            Epilogue es; es.out_raw = sc.ptr; es.out_raw_lo = sc_lo.ptr; es.ldo = sc.cp;
            std::vector<TermSpec> ts; std::vector<const void*> ss;
            if (!push_terms(ts, ss, true, in.raw, in.raw_lo, false, 1, s, cin, cout, pre + "shortcut/kernel", ks, bs)) return false;
            if (!add_op(prog, pre + "shortcut", ts, ss, es, cap, macs)) return false;
            e.res = sc.ptr; e.res_lo = sc_lo.ptr; e.res_mode = RES_SAME; e.ld_res = sc.cp;
        }
    } else {
        e.res = in.raw.ptr; e.res_lo = in.raw_lo.ptr; e.res_mode = RES_SAME; e.ld_res = in.raw.cp;
    }
    if (!add_op(prog, pre + "conv_1", terms, srcs, e, cap, macs)) return false;
    if (tape) {
        BlockTape t; t.scope = pre; t.deconv = false; t.use_shortcut = use_sc; t.k = k; t.stride = s;
        t.cin = cin; t.cout = cout; t.in = in; t.r_act = r; t.out = o;
        tape->push_back(t);
    }
    *out = o;
    return true;
}

// resnet_deconv2d_block (resnet.py:369-485, resize_at_exit=False):
//   op1: r = lrelu(deconv_k,s(lrelu(x)) + b0)                      [Cin -> Cout]
//   op2: out = deconv_k,1(r) + b1 + shortcut                       [Cout -> Cout]
// shortcut = deconv_1,s(x) + bs: fused as a second term when s == 1 and the padded channel
// counts agree; otherwise computed by its own 1x1 op at the INPUT resolution and added in
// op2's epilogue (for s == 2 only the even output pixels receive x W, conv2d_.py:464-471
// semantics of a 1x1 stride-2 transpose; bs goes into op2's bias).
bool Model::deconv_block(Program& prog, const std::string& root, const BlockSpec& b, const Pair& in, bool need_raw,
                         bool need_act, int cap, Pair* out, double* macs, std::vector<BlockTape>* tape) {
    const std::string pre = root + "/" + b.scope + "/";
    const bool sp = split_mode;
    const int cin = in.c(), cout = b.cout, k = b.k, s = b.stride, H = in.h(), Wd = in.w();
    const bool use_sc = (s != 1) || (cin != cout);
    const HostArr *k0 = W_(pre + "residual/conv_0/kernel", {k, k, cout, cin}), *b0 = W_(pre + "residual/conv_0/bias", {cout});
    const HostArr *k1 = W_(pre + "residual/conv_1/kernel", {k, k, cout, cout}), *b1 = W_(pre + "residual/conv_1/bias", {cout});
    if (!k0 || !b0 || !k1 || !b1) return false;
    const int Ho = H * s, Wo = Wd * s;
    Tensor r = alloc_act(cap, Ho, Wo, cout), r_lo;
    if (!r.ptr) return false;
    if (sp) { r_lo = alloc_act(cap, Ho, Wo, cout); if (!r_lo.ptr) return false; }
    {
        Epilogue e; e.out_act = r.ptr; e.out_act_lo = r_lo.ptr; e.ldo = r.cp;
        std::vector<TermSpec> terms; std::vector<const void*> srcs;
        if (!push_terms(terms, srcs, sp, in.act, in.act_lo, true, k, s, cin, cout, pre + "residual/conv_0/kernel", k0, b0)) return false;
        if (!add_op(prog, pre + "conv_0", terms, srcs, e, cap, macs)) return false;
    }
    Pair o;
    if (need_raw) { o.raw = alloc_act(cap, Ho, Wo, cout); if (!o.raw.ptr) return false; }
    if (need_act) { o.act = alloc_act(cap, Ho, Wo, cout); if (!o.act.ptr) return false; }
    if (sp && need_raw) { o.raw_lo = alloc_act(cap, Ho, Wo, cout); if (!o.raw_lo.ptr) return false; }
    if (sp && need_act) { o.act_lo = alloc_act(cap, Ho, Wo, cout); if (!o.act_lo.ptr) return false; }
    Epilogue e;
    e.out_raw = o.raw.ptr; e.out_act = o.act.ptr; e.out_raw_lo = o.raw_lo.ptr; e.out_act_lo = o.act_lo.ptr; e.ldo = pad16(cout);
    HostArr bias_sum = *b1;
    std::vector<TermSpec> terms;
    std::vector<const void*> srcs;
    if (!push_terms(terms, srcs, sp, r, r_lo, true, k, 1, cout, cout, pre + "residual/conv_1/kernel", k1, &bias_sum)) return false;
    if (use_sc) {
        const HostArr *ks = W_(pre + "shortcut/kernel", {1, 1, cout, cin}), *bs = W_(pre + "shortcut/bias", {cout});
        if (!ks || !bs) return false;
        if (!sp && s == 1 && in.raw.cp == r.cp) {
            terms.push_back(term(1, true, 1, 1, cin, cout, in.raw, ks, bs));
            srcs.push_back(in.raw.ptr);
        } else {
            Tensor sc = alloc_act(cap, H, Wd, cout), sc_lo;
            if (!sc.ptr) return false;
            if (sp) { sc_lo = alloc_act(cap, H, Wd, cout); if (!sc_lo.ptr) return false; }
            Epilogue es; es.out_raw = sc.ptr; es.out_raw_lo = sc_lo.ptr; es.ldo = sc.cp;
            std::vector<TermSpec> ts; std::vector<const void*> ss;
            if (!push_terms(ts, ss, sp, in.raw, in.raw_lo, true, 1, 1, cin, cout, pre + "shortcut/kernel", ks, nullptr)) return false;
            if (!add_op(prog, pre + "shortcut", ts, ss, es, cap, macs)) return false;
            for (int i = 0; i < cout; ++i) bias_sum.v[i] += bs->v[i];
            terms[0].bias = bias_sum.v.data();
            e.res = sc.ptr; e.res_lo = sc_lo.ptr; e.res_mode = (s == 1) ? RES_SAME : RES_EVEN; e.ld_res = sc.cp;
        }
    } else {
        e.res = in.raw.ptr; e.res_lo = in.raw_lo.ptr; e.res_mode = RES_SAME; e.ld_res = in.raw.cp;
    }
    if (!add_op(prog, pre + "conv_1", terms, srcs, e, cap, macs)) return false;
    if (tape) {
        BlockTape t; t.scope = pre; t.deconv = true; t.use_shortcut = use_sc; t.k = k; t.stride = s;
        t.cin = cin; t.cout = cout; t.in = in; t.r_act = r; t.out = o;
        tape->push_back(t);
    }
    *out = o;
    return true;
}

// spt.layers.dense (dense_.py:142-147) over a flattened NHWC tensor: a 1x1 conv on a 1x1 image
// with Cin = in_features.  `in_flat` must have cp == c so the flatten order matches reshape_tail.
bool Model::dense_op(Program& prog, const std::string& name, const Tensor& in_flat, int in_features, int out_features,
                     const Epilogue& epi, int cap, double* macs, const Tensor* in_flat_lo) {
    const HostArr *kk = W_(name + "/kernel", {in_features, out_features}), *bb = W_(name + "/bias", {out_features});
    if (!kk || !bb) return false;
    Tensor v = in_flat, vlo;
    v.h = 1; v.w = 1; v.c = in_features; v.cp = pad16(in_features);
    if ((size_t)v.cp * 1 != (size_t)in_flat.h * in_flat.w * in_flat.cp)
        return fail("dense " + name + ": flattened input has channel padding");
    const bool sp = split_mode && in_flat_lo && in_flat_lo->ptr;
    if (sp) { vlo = v; vlo.ptr = in_flat_lo->ptr; }
    std::vector<TermSpec> terms; std::vector<const void*> srcs;
    if (!push_terms(terms, srcs, sp, v, vlo, false, 1, 1, in_features, out_features, name + "/kernel", kk, bb)) return false;
    return add_op(prog, name, terms, srcs, epi, cap, macs);
}

bool Model::run_program(Program& prog, int n, void* stream) {
    for (auto& op : prog) {
        const int e = devop_launch(dev, *op, n, stream);
        if (e != 0) return fail("launch " + op->name + ": " + cudaGetErrorString((cudaError_t)e));
        ++launches;
    }
    return true;
}

}  // namespace apvae
