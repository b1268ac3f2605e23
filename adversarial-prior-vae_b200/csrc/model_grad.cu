// Input-gradient programs: grad_z of E(z) = w xi D_psi(G_theta(z)) + |z|^2 / 2, which is what
// tf.gradients(energy, [z]) returns in get_sgld_proposal (experiment/dmm/dmm_cifar10.py:241-247),
// and the MALA loop built on it (EnergyDistribution.sample, :179-239).
//
// The adjoint of every layer is again a planned op on the same two tcgen05 kernels:
//   conv2d(K[kh,kw,Ci,Co], stride s)^T   = deconv2d with the SAME array read as [kh,kw,Cout'=Ci,Cin'=Co]
//   deconv2d(K[kh,kw,Co,Ci], stride s)^T = conv2d   with the SAME array read as [kh,kw,Cin'=Co,Cout'=Ci]
// (conv2d_transpose(SAME) is the exact adjoint of the SAME forward conv, conv2d_.py:464-471), biases drop
// out, and leaky_relu'(v) = (v > 0 ? 1 : 0.2) is applied in the epilogue from the saved activation
// (`mask`: sign(leaky_relu(v)) == sign(v)).  Pre-activation block (resnet.py:243-265):
//   out = shortcut(x) + conv_1(act(r)),  r = conv_0(act(x)) + b0
//   g_r = conv_1^T(g_out) * act'(r)
//   g_x = conv_0^T(g_r) * act'(x) + shortcut^T(g_out)
// D_psi's two dense layers have no activation between them (dmm_cifar10.py:559-561), so
// dD/d(flatten) = W64 w_last is a constant vector, uploaded once per sample slot.
#include <cuda_runtime.h>
#include <algorithm>
#include <cmath>
#include <cstring>
#include "kernels.h"
#include "model.h"
#include "op16.h"
#include "tuning.h"

namespace apvae {

static inline int pad16g(int c) { return (c + 15) & ~15; }

static TermSpec gterm(bool transpose, int k, int stride, int cin, int cout, const Tensor& in, const HostArr* kernel) {
    TermSpec t;
    t.src = 0; t.transpose = transpose ? 1 : 0; t.k = k; t.stride = stride; t.cin = cin; t.cout = cout;
    t.in_h = in.h; t.in_w = in.w; t.in_cp = in.cp;
    t.kernel = kernel->v.data();
    t.bias = nullptr;
    return t;
}

// Backward of a resnet_conv2d_block (D_psi).  g_out: gradient w.r.t. the block output [cap, Ho, Wo, cout].
bool Model::grad_conv_block(const BlockTape& t, const Tensor& g_out, float* out_f32, int f32_ld, Tensor* g_x_out) {
    const int cap = cfg.max_chunk, k = t.k, s = t.stride, cin = t.cin, cout = t.cout;
    const int H = t.in.h(), Wd = t.in.w();
    const HostArr* k0 = W_(t.scope + "residual/conv_0/kernel", {k, k, cin, cin});
    const HostArr* k1 = W_(t.scope + "residual/conv_1/kernel", {k, k, cin, cout});
    if (!k0 || !k1) return false;
    // the block whose forward is one CUDA-core kernel (<= 3 input channels): its adjoint as one kernel too (block0.cu)
    if (t.fused && t.use_shortcut && switch_on(tuning().small_block, true) && k == kSmallBlockMaxK && cin <= kSmallBlockMaxCin &&
        (cout == 16 || cout == 32) && g_out.cp % 8 == 0 && !(Wd & 1) && !(H % s)) {
        const HostArr* ks = W_(t.scope + "shortcut/kernel", {1, 1, cin, cout});
        if (!ks) return false;
        Tensor g_x = alloc_act(cap, H, Wd, cin);
        if (!g_x.ptr) return false;
        std::unique_ptr<DevOp> op(new DevOp());
        op->name = t.scope + "block^T";
        op->is_small_block_grad = true;
        op->capacity_n = cap;
        op->macs_per_image = (double)k * k * cin * cin * H * Wd + ((double)k * k * cin * cout + (double)cin * cout) * (H / s) * (Wd / s);
        SmallBlockGradParams& sg = op->sbgrad;
        memset(&sg, 0, sizeof(sg));
        sg.g_out = g_out.ptr; sg.ldg = g_out.cp;
        sg.r_act = t.r_act.ptr;
        sg.x_in = t.in.act.ptr ? t.in.act.ptr : t.in.raw.ptr;
        sg.gx_f32 = out_f32; sg.ld_f32 = f32_ld;
        sg.gx_raw = out_f32 ? nullptr : g_x.ptr; sg.ldo = g_x.cp;      // the first block's gradient is only wanted in fp32
        sg.H = H; sg.W = Wd; sg.cp_in = t.r_act.cp; sg.cin = cin; sg.cout = cout; sg.stride = s;
        sg.alpha = 0.2f; sg.f16 = f16();
        if ((t.in.act.ptr ? t.in.act.cp : t.in.raw.cp) != t.r_act.cp) return fail("fused block adjoint: input and r disagree on the channel padding");
        for (int tp = 0; tp < k * k; ++tp)
            for (int ci = 0; ci < cin; ++ci) {
                for (int co = 0; co < cin; ++co) sg.w0[(tp * cin + ci) * cin + co] = k0->v[((size_t)tp * cin + ci) * cin + co];
                for (int co = 0; co < cout; ++co) sg.w1[(tp * cin + ci) * cout + co] = k1->v[((size_t)tp * cin + ci) * cout + co];
            }
        for (int ci = 0; ci < cin; ++ci)
            for (int co = 0; co < cout; ++co) sg.ws[ci * cout + co] = ks->v[(size_t)ci * cout + co];
        macs_grad += op->macs_per_image;
        gd_prog.push_back(std::move(op));
        *g_x_out = g_x;
        return true;
    }
    Tensor g_r = alloc_act(cap, H, Wd, cin);
    if (!g_r.ptr) return false;
    {   // g_r = conv_1^T(g_out) * act'(r)
        Epilogue e; e.out_raw = g_r.ptr; e.ldo = g_r.cp; e.mask = t.r_act.ptr; e.ld_mask = t.r_act.cp;
        if (!add_op(gd_prog, t.scope + "conv_1^T", {gterm(true, k, s, cout, cin, g_out, k1)}, {g_out.ptr}, e, cap, &macs_grad)) return false;
    }
    Epilogue e;
    // sign(leaky_relu(x)) == sign(x): the raw block input serves as the mask too (D_psi's first block may have no act copy)
    e.mask = t.in.act.ptr ? t.in.act.ptr : t.in.raw.ptr; e.ld_mask = t.in.act.ptr ? t.in.act.cp : t.in.raw.cp;
    if (t.use_shortcut) {   // shortcut^T: 1x1 transpose at the OUTPUT resolution; for s == 2 it lands on even input pixels
        const HostArr* ks = W_(t.scope + "shortcut/kernel", {1, 1, cin, cout});
        if (!ks) return false;
        Tensor g_sc = alloc_act(cap, g_out.h, g_out.w, cin);
        if (!g_sc.ptr) return false;
        Epilogue es; es.out_raw = g_sc.ptr; es.ldo = g_sc.cp;
        if (!add_op(gd_prog, t.scope + "shortcut^T", {gterm(true, 1, 1, cout, cin, g_out, ks)}, {g_out.ptr}, es, cap, &macs_grad)) return false;
        e.res = g_sc.ptr; e.ld_res = g_sc.cp; e.res_mode = (s == 1) ? RES_SAME : RES_EVEN;
    } else {
        e.res = g_out.ptr; e.ld_res = g_out.cp; e.res_mode = RES_SAME;
    }
    Tensor g_x = alloc_act(cap, H, Wd, cin);
    if (!g_x.ptr) return false;
    e.out_raw = g_x.ptr; e.ldo = g_x.cp;
    if (out_f32) { e.out_f32 = out_f32; e.ld_f32 = f32_ld; e.f32_cols = cin; }
    if (!add_op(gd_prog, t.scope + "conv_0^T", {gterm(true, k, 1, cin, cin, g_r, k0)}, {g_r.ptr}, e, cap, &macs_grad)) return false;
    *g_x_out = g_x;
    return true;
}

// Backward of a resnet_deconv2d_block (G_theta).  g_out: [cap, H*s, W*s, cout].
bool Model::grad_deconv_block(const BlockTape& t, const Tensor& g_out, float* out_f32, int f32_ld, Tensor* g_x_out) {
    const int cap = cfg.max_chunk, k = t.k, s = t.stride, cin = t.cin, cout = t.cout;
    const int H = t.in.h(), Wd = t.in.w();
    const HostArr* k0 = W_(t.scope + "residual/conv_0/kernel", {k, k, cout, cin});
    const HostArr* k1 = W_(t.scope + "residual/conv_1/kernel", {k, k, cout, cout});
    if (!k0 || !k1) return false;
    Tensor g_r = alloc_act(cap, g_out.h, g_out.w, cout);
    if (!g_r.ptr) return false;
    {   // g_r = deconv_1^T(g_out) * act'(r): a stride-1 conv
        Epilogue e; e.out_raw = g_r.ptr; e.ldo = g_r.cp; e.mask = t.r_act.ptr; e.ld_mask = t.r_act.cp;
        if (!add_op(gg_prog, t.scope + "conv_1^T", {gterm(false, k, 1, cout, cout, g_out, k1)}, {g_out.ptr}, e, cap, &macs_grad)) return false;
    }
    Epilogue e;
    e.mask = t.in.act.ptr; e.ld_mask = t.in.act.cp;
    if (t.use_shortcut) {   // shortcut^T = 1x1 conv with stride s of g_out (input resolution)
        const HostArr* ks = W_(t.scope + "shortcut/kernel", {1, 1, cout, cin});
        if (!ks) return false;
        Tensor g_sc = alloc_act(cap, H, Wd, cin);
        if (!g_sc.ptr) return false;
        Epilogue es; es.out_raw = g_sc.ptr; es.ldo = g_sc.cp;
        if (!add_op(gg_prog, t.scope + "shortcut^T", {gterm(false, 1, s, cout, cin, g_out, ks)}, {g_out.ptr}, es, cap, &macs_grad)) return false;
        e.res = g_sc.ptr; e.ld_res = g_sc.cp; e.res_mode = RES_SAME;
    } else {
        e.res = g_out.ptr; e.ld_res = g_out.cp; e.res_mode = RES_SAME;
    }
    Tensor g_x = alloc_act(cap, H, Wd, cin);
    if (!g_x.ptr) return false;
    e.out_raw = g_x.ptr; e.ldo = g_x.cp;
    if (out_f32) { e.out_f32 = out_f32; e.ld_f32 = f32_ld; e.f32_cols = cin; }
    if (!add_op(gg_prog, t.scope + "conv_0^T", {gterm(false, k, s, cout, cin, g_r, k0)}, {g_r.ptr}, e, cap, &macs_grad)) return false;
    *g_x_out = g_x;
    return true;
}

bool Model::build_grad() {
    const int cap = cfg.max_chunk;
    if (d_tape.empty() || g_tape.empty()) return fail("gradient programs need the forward tapes");
    // ---- seed: dD/d(flatten) = W64 w_last  (dense(64) then dense(1), no activation in between)
    const int feat = d_flat.h * d_flat.w * d_flat.c;
    const HostArr* k64 = W_("D_psi/level_-2/kernel", {feat, 64});
    const HostArr* kl = W_("D_psi/level_-1/kernel", {64, 1});
    if (!k64 || !kl) return false;
    std::vector<uint16_t> seed((size_t)feat);
    for (int i = 0; i < feat; ++i) {
        double a = 0;
        for (int j = 0; j < 64; ++j) a += (double)k64->v[(size_t)i * 64 + j] * kl->v[j];
        seed[i] = op16::host_from_f32((float)a, f16() != 0);
    }
    d_seed = alloc_act(cap, d_flat.h, d_flat.w, d_flat.c);
    if (!d_seed.ptr) return false;
    {
        std::vector<uint16_t> rep((size_t)cap * feat);
        for (int n = 0; n < cap; ++n) memcpy(rep.data() + (size_t)n * feat, seed.data(), (size_t)feat * 2);
        if (cudaMemcpy(d_seed.ptr, rep.data(), rep.size() * 2, cudaMemcpyHostToDevice) != cudaSuccess) return fail("upload of the gradient seed failed");
    }
    gx_f32 = alloc_f32((size_t)cap * HW() * C());
    gz_f32 = alloc_f32((size_t)cap * cfg.z_dim);
    g_pre = alloc_act(cap, H(), W(), C());
    if (!gx_f32 || !gz_f32 || !g_pre.ptr) return false;
    // ---- D_psi backward: blocks in reverse; the first block's input gradient is d D / d x_mean (fp32)
    Tensor g = d_seed;
    for (int i = (int)d_tape.size() - 1; i >= 0; --i) {
        Tensor gx;
        if (!grad_conv_block(d_tape[i], g, i == 0 ? gx_f32 : nullptr, C(), &gx)) return false;
        g = gx;
    }
    // ---- x_mean = tanh(head): elementwise between the two programs (run_energy_grad), then the 1x1 head^T
    const std::string mean_name = cfg.arch == APVAE_ARCH_M ? "G_theta/feature_map_mean_to_pixel" : "G_theta/x_mean";
    const int cg = g_head_in.c;
    const HostArr* km = W_(mean_name + "/kernel", {1, 1, cg, C()});
    if (!km) return false;
    Tensor g_h = alloc_act(cap, H(), W(), cg);
    if (!g_h.ptr) return false;
    // C <= 4 -> cg channels: tanh' and this 1x1 transpose are one HBM-streaming CUDA-core kernel (as a GEMM it pads K from
    // 3 to 16 and runs at 4 TFLOP/s); tuning.tiny_conv = -1 keeps the elementwise pass + the tensor-core op
    if (switch_on(tuning().tiny_conv, true) && C() <= kHeadGradMaxC && cg <= kHeadGradMaxCg && g_h.cp % 8 == 0) {
        head_grad_fused = true;
        memset(&head_grad, 0, sizeof(head_grad));
        head_grad.g = gx_f32; head_grad.head = head_f32; head_grad.ld = head_ld();
        head_grad.C = C(); head_grad.cg = cg; head_grad.cp_out = g_h.cp; head_grad.f16 = f16();
        head_grad.out = g_h.ptr;
        for (int co = 0; co < cg; ++co)
            for (int c = 0; c < C(); ++c) head_grad.w[co * kHeadGradMaxC + c] = km->v[(size_t)co * C() + c];
        macs_grad += (double)cg * C() * HW();
    } else {
        Epilogue e; e.out_raw = g_h.ptr; e.ldo = g_h.cp;
        if (!add_op(gg_prog, mean_name + "^T", {gterm(true, 1, 1, C(), cg, g_pre, km)}, {g_pre.ptr}, e, cap, &macs_grad)) return false;
    }
    // ---- G_theta backward
    g = g_h;
    for (int i = (int)g_tape.size() - 1; i >= 0; --i) {
        const bool to_z = (i == 0 && cfg.arch == APVAE_ARCH_A);      // block 0's input IS z reshaped to [8,8,zc]
        Tensor gx;
        if (!grad_deconv_block(g_tape[i], g, to_z ? gz_f32 : nullptr, zc(), &gx)) return false;
        g = gx;
    }
    if (cfg.arch != APVAE_ARCH_A) {      // dense(z, 64*8*8)^T: g_z = g_h0 W^T
        const int feat0 = 64 * g_base_h() * g_base_w();
        const HostArr* kd = W_("G_theta/level_0/kernel", {cfg.z_dim, feat0});
        if (!kd) return false;
        HostArr& kt = weights["G_theta/level_0/kernel^T"];
        kt.shape = {feat0, cfg.z_dim};
        kt.v.resize((size_t)feat0 * cfg.z_dim);
        for (int a = 0; a < cfg.z_dim; ++a)
            for (int b = 0; b < feat0; ++b) kt.v[(size_t)b * cfg.z_dim + a] = kd->v[(size_t)a * feat0 + b];
        if (g.cp != g.c) return fail("G_theta first block input has channel padding");
        Tensor v = g;
        v.h = 1; v.w = 1; v.c = feat0; v.cp = pad16g(feat0);
        Epilogue e; e.out_f32 = gz_f32; e.ld_f32 = cfg.z_dim; e.f32_cols = cfg.z_dim;
        TermSpec t = gterm(false, 1, 1, feat0, cfg.z_dim, v, &kt);
        if (!add_op(gg_prog, "G_theta/level_0^T", {t}, {v.ptr}, e, cap, &macs_grad)) return false;
    }
    return true;
}

#define GCHECK(call, what)                                                                        \
    do {                                                                                          \
        const int _e = (call);                                                                    \
        ++launches;                                                                               \
        if (_e != 0) return fail(std::string(what) + ": " + cudaGetErrorString((cudaError_t)_e)); \
    } while (0)

bool Model::energy_grad_chunk(const float* z_dev, int n, float* energy, float* grad, float* pure, void* stream) {
    if (!cfg.enable_grad || gd_prog.empty()) return fail("this handle was created without enable_grad");
    if (n > cfg.max_chunk) return fail("energy_grad_chunk: n exceeds max_chunk");
    // forward (keeps every block's activations: they are the masks of the backward ops)
    GCHECK(launch_sample_z(nullptr, nullptr, z_dev, n, 0, 1, cfg.z_dim, 0.f, 0, nullptr, g_in.raw.ptr, g_in.act.ptr,
                           z_stride(), z_group(), z_cpad(), 0.2f, nullptr, nullptr, f16(), stream, g_in.raw_lo.ptr, g_in.act_lo.ptr),
           "stage_latents");
    if (!run_decoder_chunk(n, stream)) return false;
    float* d_out = pure ? pure : z_tmp;                 // z_tmp: [max_chunk, z_dim] scratch, large enough for n floats
    if (!run_critic_chunk(n, d_out, stream)) return false;
    if (grad) {
        if (!run_program(gd_prog, n, stream)) return false;
        if (head_grad_fused) {
            head_grad.pixels = (long long)n * HW();
            GCHECK(launch_tanh_backward_head(head_grad, stream), "tanh_backward_head");
        } else {
            GCHECK(launch_tanh_backward_pack(gx_f32, head_f32, head_ld(), (long long)n * HW(), C(), g_pre.cp, g_pre.ptr, f16(), stream),
                   "tanh_backward");
        }
        if (!run_program(gg_prog, n, stream)) return false;
    }
    if (energy || grad)
        GCHECK(launch_energy_combine(d_out, gz_f32, z_dev, n, cfg.z_dim, w_xi(), energy, grad, stream), "energy_combine");
    return true;
}

bool Model::critic_grad_of_images(const float* x_dev, int N, float* d_out, float* grad_out, float* sqnorm_out, void* stream) {
    if (!cfg.enable_grad || gd_prog.empty()) return fail("this handle was created without enable_grad");
    const size_t per = (size_t)HW() * C();
    for (int s0 = 0; s0 < N; s0 += cfg.max_chunk) {
        const int n = std::min(cfg.max_chunk, N - s0);
        GCHECK(launch_pack_image(x_dev + (size_t)s0 * per, (long long)n * HW(), C(), d_in.raw.cp, d_in.raw.ptr, d_in.act.ptr,
                                 0.2f, f16(), stream, d_in.raw_lo.ptr, d_in.act_lo.ptr), "pack_image");
        if (!run_critic_chunk(n, d_out ? d_out + s0 : z_tmp, stream)) return false;
        if (!run_program(gd_prog, n, stream)) return false;
        if (grad_out && cudaMemcpyAsync(grad_out + (size_t)s0 * per, gx_f32, (size_t)n * per * sizeof(float),
                                        cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(stream)) != cudaSuccess)
            return fail("copy of the image gradient failed");
        if (sqnorm_out) GCHECK(launch_row_sqnorm(gx_f32, (long long)per, n, sqnorm_out + s0, stream), "row_sqnorm");
    }
    return true;
}

bool Model::mala(float* z_io, int n, int steps, float alpha, const float* eps, const float* u, uint64_t seed,
                 float* hist_e, float* hist_pure, float* hist_ratio, float* hist_z, void* stream) {
    const int d = cfg.z_dim;
    if ((size_t)n > mala_cap) {
        for (float** p : {&mz, &mzp, &mg, &mgp, &me, &mep, &mp, &mpp, &meps, &mu}) { if (*p) cudaFree(*p); *p = nullptr; }
        mala_cap = 0;
        for (float** p : {&mzp, &mg, &mgp, &meps})
            if (cudaMalloc(p, (size_t)n * d * sizeof(float)) != cudaSuccess) return fail("cudaMalloc (MALA state)");
        for (float** p : {&me, &mep, &mp, &mpp, &mu})
            if (cudaMalloc(p, (size_t)n * sizeof(float)) != cudaSuccess) return fail("cudaMalloc (MALA state)");
        mala_cap = (size_t)n;
    }
    auto eval_all = [&](const float* z, float* e, float* g, float* p) {
        for (int s0 = 0; s0 < n; s0 += cfg.max_chunk) {
            const int m = std::min(cfg.max_chunk, n - s0);
            if (!energy_grad_chunk(z + (size_t)s0 * d, m, e + s0, g + (size_t)s0 * d, p + s0, stream)) return false;
        }
        return true;
    };
    // energy / gradient at the start; afterwards the accepted state's values are carried over (the
    // reference recomputes them every step, :197-199; the values are the same)
    if (!eval_all(z_io, me, mg, mp)) return false;
    for (int i = 0; i < steps; ++i) {
        const float* eps_i = eps ? eps + (size_t)i * n * d : meps;
        const float* u_i = u ? u + (size_t)i * n : mu;
        // history_z[i] = z BEFORE step i's accept (dmm_cifar10.py:200,235)
        if (hist_z && cudaMemcpyAsync(hist_z + (size_t)i * n * d, z_io, (size_t)n * d * sizeof(float), cudaMemcpyDeviceToDevice,
                                      static_cast<cudaStream_t>(stream)) != cudaSuccess)
            return fail("copy of history_z failed");
        if (!eps) GCHECK(launch_fill_normal(meps, (long long)n * d, seed, (unsigned long long)i * 2 * n * d, 0, stream), "fill_normal");
        if (!u) GCHECK(launch_fill_uniform(mu, n, seed ^ 0x9e3779b97f4a7c15ull, (unsigned long long)i * n, stream), "fill_uniform");
        GCHECK(launch_mala_propose(z_io, mg, eps_i, n, d, alpha, mzp, stream), "mala_propose");
        if (!eval_all(mzp, mep, mgp, mpp)) return false;
        GCHECK(launch_mala_accept(z_io, me, mg, mp, mzp, mep, mgp, mpp, u_i, n, d, alpha,
                                  hist_e ? hist_e + (size_t)i * n : nullptr, hist_pure ? hist_pure + (size_t)i * n : nullptr,
                                  hist_ratio ? hist_ratio + (size_t)i * n : nullptr, stream), "mala_accept");
    }
    return true;
}

}  // namespace apvae
