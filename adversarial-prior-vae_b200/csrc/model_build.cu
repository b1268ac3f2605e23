// Program construction and forward drivers for q_net / G_theta / D_psi.
#include <cuda_runtime.h>
#include <string.h>
#include <math.h>
#include <string>
#include <vector>
#include "kernels.h"
#include "model.h"
#include "tuning.h"

namespace apvae {

#define KCHECK(call, what)                                                                        \
    do {                                                                                          \
        const int _e = (call);                                                                    \
        ++launches;                                                                               \
        if (_e != 0) return fail(std::string(what) + ": " + cudaGetErrorString((cudaError_t)_e)); \
    } while (0)

bool Model::build_q() {
    const int cap = cfg.max_batch;
    q_in.raw = alloc_act(cap, H(), W(), C());
    q_in.act = alloc_act(cap, H(), W(), C());
    q_mu = alloc_f32((size_t)cap * cfg.z_dim);
    q_logstd = alloc_f32((size_t)cap * cfg.z_dim);
    if (!q_in.raw.ptr || !q_in.act.ptr || !q_mu || !q_logstd) return false;
    if (cfg.nf_layers > 0) {
        // u_hat = u + (m(w.u) - w.u) w / |w|^2, m(a) = -1 + softplus(a)  (planar_nf.py:104-106); the log-determinant
        // needs only w . u_hat per layer: det = 1 + (1 - tanh^2(w.x + b)) (w . u_hat)  (:127-131)
        const int L = cfg.nf_layers, d = cfg.z_dim;
        std::vector<float> hw((size_t)L * d), hu((size_t)L * d), hb(L), hwu(L);
        for (int l = 0; l < L; ++l) {
            const std::string pre = L == 1 ? std::string("posterior_flow/") : "posterior_flow/_" + std::to_string(l) + "/";
            const HostArr& w = weights[pre + "w"];
            const HostArr& u = weights[pre + "u"];
            const HostArr& b = weights[pre + "b"];
            double wu = 0, ww = 0;
            for (int j = 0; j < d; ++j) { wu += (double)w.v[j] * u.v[j]; ww += (double)w.v[j] * w.v[j]; }
            const double sp = wu > 30 ? wu : log1p(exp(wu));
            const double coef = (-1.0 + sp - wu) / ww;
            double wuh = 0;
            for (int j = 0; j < d; ++j) {
                const double uh = u.v[j] + coef * w.v[j];
                hw[(size_t)l * d + j] = w.v[j];
                hu[(size_t)l * d + j] = (float)uh;
                wuh += (double)w.v[j] * uh;
            }
            hb[l] = b.v[0];
            hwu[l] = (float)wuh;
        }
        flow_w = alloc_f32(hw.size()); flow_uhat = alloc_f32(hu.size()); flow_b = alloc_f32(L); flow_wu = alloc_f32(L);
        if (!flow_w || !flow_uhat || !flow_b || !flow_wu) return false;
        cudaMemcpy(flow_w, hw.data(), hw.size() * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(flow_uhat, hu.data(), hu.size() * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(flow_b, hb.data(), L * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(flow_wu, hwu.data(), L * 4, cudaMemcpyHostToDevice);
    }
    const std::vector<BlockSpec> blocks = q_blocks(cfg);
    Pair h = q_in;
    for (size_t i = 0; i < blocks.size(); ++i) {
        const bool last = (i + 1 == blocks.size());
        // arch A heads are resnet blocks (need raw + act); B/M heads are dense on the raw flatten
        const bool need_act = !last || cfg.arch == APVAE_ARCH_A;
        Pair o;
        if (!conv_block(q_prog, "q_net", blocks[i], h, true, need_act, cap, &o, &macs_q, nullptr)) return false;
        h = o;
    }
    if (cfg.arch == APVAE_ARCH_A) {          // dmm_cifar10.py:440-446
        const int zcn = zc();
        float* outs[2] = {q_mu, q_logstd};
        const char* names[2] = {"z_mean", "z_logstd"};
        for (int t = 0; t < 2; ++t) {
            // a resnet_conv2d_block whose conv_1 output goes straight to fp32 [B, 8*8*zc] = [B, z_dim]
            const std::string pre = std::string("q_net/") + names[t] + "/";
            const int cin = h.c();
            const HostArr *k0 = W_(pre + "residual/conv_0/kernel", {3, 3, cin, cin}), *b0 = W_(pre + "residual/conv_0/bias", {cin});
            const HostArr *k1 = W_(pre + "residual/conv_1/kernel", {3, 3, cin, zcn}), *b1 = W_(pre + "residual/conv_1/bias", {zcn});
            if (!k0 || !b0 || !k1 || !b1) return false;
            Tensor r = alloc_act(cap, h.h(), h.w(), cin);
            if (!r.ptr) return false;
            Epilogue e0; e0.out_act = r.ptr; e0.ldo = r.cp;
            TermSpec t0; t0.src = 0; t0.transpose = 0; t0.k = 3; t0.stride = 1; t0.cin = cin; t0.cout = cin;
            t0.in_h = h.h(); t0.in_w = h.w(); t0.in_cp = h.act.cp; t0.kernel = k0->v.data(); t0.bias = b0->v.data();
            if (!add_op(q_prog, pre + "conv_0", {t0}, {h.act.ptr}, e0, cap, &macs_q)) return false;
            Epilogue e1; e1.out_f32 = outs[t]; e1.ld_f32 = zcn; e1.f32_cols = zcn;
            TermSpec t1 = t0; t1.cout = zcn; t1.in_cp = r.cp; t1.kernel = k1->v.data(); t1.bias = b1->v.data();
            std::vector<TermSpec> terms = {t1};
            std::vector<const void*> srcs = {r.ptr};
            if (cin != zcn) {
                const HostArr *ks = W_(pre + "shortcut/kernel", {1, 1, cin, zcn}), *bs = W_(pre + "shortcut/bias", {zcn});
                if (!ks || !bs) return false;
                TermSpec ts = t1; ts.src = 1; ts.k = 1; ts.in_cp = h.raw.cp; ts.kernel = ks->v.data(); ts.bias = bs->v.data();
                terms.push_back(ts);
                srcs.push_back(h.raw.ptr);
            } else {
                e1.res = h.raw.ptr; e1.res_mode = RES_SAME; e1.ld_res = h.raw.cp;
            }
            if (!add_op(q_prog, pre + "conv_1", terms, srcs, e1, cap, &macs_q)) return false;
        }
    } else {                                  // dmm_for_sampling.py:441-446
        const int feat = h.h() * h.w() * h.c();
        if (h.raw.cp != h.raw.c) return fail("q_net flatten has channel padding");
        Epilogue e; e.out_f32 = q_mu; e.ld_f32 = cfg.z_dim; e.f32_cols = cfg.z_dim;
        if (!dense_op(q_prog, "q_net/z_mean", h.raw, feat, cfg.z_dim, e, cap, &macs_q)) return false;
        e.out_f32 = q_logstd;
        if (!dense_op(q_prog, "q_net/z_logstd", h.raw, feat, cfg.z_dim, e, cap, &macs_q)) return false;
    }
    return true;
}

bool Model::build_g() {
    const int cap = cfg.max_chunk;
    const int gh = g_base_h(), gw = g_base_w();
    std::vector<BlockTape>* tape = cfg.enable_grad ? &g_tape : nullptr;
    const bool sp = cfg.precise != 0;
    split_mode = sp;
    Pair h;
    if (cfg.arch == APVAE_ARCH_A) {           // dmm_cifar10.py:478-480: z reshaped to [8,8,zc]
        g_in.raw = alloc_act(cap, gh, gw, zc());
        g_in.act = alloc_act(cap, gh, gw, zc());
        if (!g_in.raw.ptr || !g_in.act.ptr) return false;
        if (sp) {
            g_in.raw_lo = alloc_act(cap, gh, gw, zc());
            g_in.act_lo = alloc_act(cap, gh, gw, zc());
            if (!g_in.raw_lo.ptr || !g_in.act_lo.ptr) return false;
        }
        h = g_in;
    } else {                                  // dmm_for_sampling.py:486-492: dense(z, 64*8*8) -> [8,8,64]
        g_in.raw = alloc_act(cap, 1, 1, cfg.z_dim);
        if (!g_in.raw.ptr) return false;
        if (sp) { g_in.raw_lo = alloc_act(cap, 1, 1, cfg.z_dim); if (!g_in.raw_lo.ptr) return false; }
        const int feat = 64 * gh * gw;
        h.raw = alloc_act(cap, gh, gw, 64);
        h.act = alloc_act(cap, gh, gw, 64);
        if (!h.raw.ptr || !h.act.ptr) return false;
        if (sp) {
            h.raw_lo = alloc_act(cap, gh, gw, 64);
            h.act_lo = alloc_act(cap, gh, gw, 64);
            if (!h.raw_lo.ptr || !h.act_lo.ptr) return false;
        }
        Epilogue e; e.out_raw = h.raw.ptr; e.out_act = h.act.ptr; e.out_raw_lo = h.raw_lo.ptr; e.out_act_lo = h.act_lo.ptr; e.ldo = feat;
        if (!dense_op(g_prog, "G_theta/level_0", g_in.raw, cfg.z_dim, feat, e, cap, &macs_g, &g_in.raw_lo)) return false;
    }
    const std::vector<BlockSpec> blocks = g_blocks(cfg);
    for (size_t i = 0; i < blocks.size(); ++i) {
        const bool last = (i + 1 == blocks.size());
        Pair o;
        if (!deconv_block(g_prog, "G_theta", blocks[i], h, true, !last, cap, &o, &macs_g, tape)) return false;
        h = o;
    }
    g_head_in = h.raw;
    // heads (dmm_cifar10.py:489-496): one 1x1 GEMM for [x_mean | x_logstd]; tanh on the x_mean columns.
    // The bf16 copy of x_mean (raw + leaky_relu) is D_psi's input.
    const int cg = h.c(), Cx = C();
    d_in.raw = alloc_act(cap, H(), W(), Cx);
    d_in.act = alloc_act(cap, H(), W(), Cx);
    head_f32 = alloc_f32((size_t)cap * HW() * head_ld());
    if (!d_in.raw.ptr || !d_in.act.ptr || !head_f32) return false;
    if (sp) {
        d_in.raw_lo = alloc_act(cap, H(), W(), Cx);
        d_in.act_lo = alloc_act(cap, H(), W(), Cx);
        if (!d_in.raw_lo.ptr || !d_in.act_lo.ptr) return false;
    }
    HostArr kcat, bcat;
    const bool has_std = head_ld() == 2 * Cx;
    const std::string mean_name = cfg.arch == APVAE_ARCH_M ? "G_theta/feature_map_mean_to_pixel" : "G_theta/x_mean";
    const HostArr *km = W_(mean_name + "/kernel", {1, 1, cg, Cx}), *bm = W_(mean_name + "/bias", {Cx});
    if (!km || !bm) return false;
    const int nout = head_ld();
    kcat.v.assign((size_t)cg * nout, 0.f);
    bcat.v.assign(nout, 0.f);
    const HostArr *ksd = nullptr, *bsd = nullptr;
    if (has_std) {
        ksd = W_("G_theta/x_logstd/kernel", {1, 1, cg, Cx});
        bsd = W_("G_theta/x_logstd/bias", {Cx});
        if (!ksd || !bsd) return false;
    }
    for (int ci = 0; ci < cg; ++ci)
        for (int c = 0; c < Cx; ++c) {
            kcat.v[(size_t)ci * nout + c] = km->v[(size_t)ci * Cx + c];
            if (has_std) kcat.v[(size_t)ci * nout + Cx + c] = ksd->v[(size_t)ci * Cx + c];
        }
    for (int c = 0; c < Cx; ++c) { bcat.v[c] = bm->v[c]; if (has_std) bcat.v[Cx + c] = bsd->v[c]; }
    HostArr& kc = weights["G_theta/heads/kernel"];      // [1,1,cg,nout]: x_mean | x_logstd side by side
    kc.shape = {1, 1, cg, nout};
    kc.v = kcat.v;
    Epilogue e;
    e.out_raw = d_in.raw.ptr; e.out_act = d_in.act.ptr; e.out_raw_lo = d_in.raw_lo.ptr; e.out_act_lo = d_in.act_lo.ptr;
    e.ldo = d_in.raw.cp; e.bf16_zero_from = Cx;
    e.out_f32 = head_f32; e.ld_f32 = nout; e.f32_cols = nout; e.tanh_cols = Cx;
    // nout <= 8 outputs per pixel: an HBM-bound stream, run on the CUDA cores (heads.cu) with fp32 weights;
    // tuning.heads_gemm = +1 keeps the 1x1 tensor-core op (tuning / A-B only)
    if (!switch_on(tuning().heads_gemm, false) && nout <= 8 && ((cg + 7) / 8) * 8 * nout <= kHeadsMaxW && d_in.raw.cp % 8 == 0) {
        std::unique_ptr<DevOp> op(new DevOp());
        op->name = "G_theta/heads";
        op->is_heads = true;
        op->capacity_n = cap;
        op->macs_per_image = (double)cg * nout * HW();
        HeadsParams& hp = op->heads;
        memset(&hp, 0, sizeof(hp));
        hp.in = h.raw.ptr; hp.in_lo = sp ? h.raw_lo.ptr : nullptr;
        hp.cp = h.raw.cp; hp.cin = cg; hp.nout = nout; hp.tanh_cols = Cx;
        hp.pixels_per_image = HW();
        hp.out_f32 = head_f32; hp.ld_f32 = nout;
        hp.raw = d_in.raw.ptr; hp.act = d_in.act.ptr; hp.raw_lo = d_in.raw_lo.ptr; hp.act_lo = d_in.act_lo.ptr;
        hp.ldo = d_in.raw.cp;
        hp.alpha = 0.2f;
        hp.f16 = f16();
        for (int c = 0; c < nout; ++c) hp.bias[c] = bcat.v[c];
        for (int ci = 0; ci < cg; ++ci)
            for (int c = 0; c < nout; ++c) hp.w[ci * nout + c] = kcat.v[(size_t)ci * nout + c];
        macs_g += op->macs_per_image;
        heads_op = op.get();
        lik_part = alloc_f32((size_t)cap * heads_tiles_per_image(HW()) * 2);
        if (!lik_part) return false;
        g_prog.push_back(std::move(op));
        split_mode = false;
        return true;
    }
    std::vector<TermSpec> terms; std::vector<const void*> srcs;
    if (!push_terms(terms, srcs, sp, h.raw, h.raw_lo, false, 1, 1, cg, nout, "G_theta/heads/kernel", &kc, &bcat)) return false;
    const bool ok = add_op(g_prog, "G_theta/heads", terms, srcs, e, cap, &macs_g);
    split_mode = false;
    return ok;
}

// G_omega (dmm_cifar10.py:503-532): dense(z_omega[128]) -> [oh, ow, 256] -> six resnet deconv blocks -> 1x1 conv to
// z_channel at [2 oh, 2 ow]; flattened, that is gan_z = the initial_z of the MALA chains (:623-635,1061-1066).
bool Model::build_omega() {
    const int cap = cfg.max_chunk;
    if (cfg.arch == APVAE_ARCH_M) {
        // dmm_bernouli_fashion.py:526-543: h = z; six times h = leaky_relu(h W_eff + b) with the weight-normalised kernel
        // W_eff = scale * kernel / sqrt(max(sum_in kernel^2, 1e-12)) (weight_norm_.py:76-97, tf.nn.l2_normalize over the
        // input axis), folded here once at load time.  The last layer's activation runs on the fp32 output.
        const int din = cfg.z_dim;
        om_in = alloc_act(cap, 1, 1, din);
        om_out = alloc_f32((size_t)cap * cfg.z_dim);
        om_noise = alloc_f32((size_t)cap * din);
        if (!om_in.ptr || !om_out || !om_noise) return false;
        Tensor h = om_in;
        for (int i = 0; i < 6; ++i) {
            const std::string pre = std::string("G_omega/dense") + (i ? "_" + std::to_string(i) : "");
            const int fin = i == 0 ? din : 40, fout = i == 5 ? cfg.z_dim : 40;
            const HostArr *k = W_(pre + "/kernel", {fin, fout}), *sc = W_(pre + "/weight_norm/scale", {fin, fout}),
                          *b = W_(pre + "/bias", {fout});
            if (!k || !sc || !b) return false;
            HostArr& we = weights[pre + "#wn/kernel"];
            we.shape = {fin, fout};
            we.v.resize((size_t)fin * fout);
            for (int o = 0; o < fout; ++o) {
                double ss = 0;
                for (int a = 0; a < fin; ++a) ss += (double)k->v[(size_t)a * fout + o] * k->v[(size_t)a * fout + o];
                const double inv = 1.0 / sqrt(ss > 1e-12 ? ss : 1e-12);
                for (int a = 0; a < fin; ++a) we.v[(size_t)a * fout + o] = (float)(k->v[(size_t)a * fout + o] * inv * sc->v[(size_t)a * fout + o]);
            }
            weights[pre + "#wn/bias"] = *b;
            Epilogue e;
            Tensor o;
            if (i < 5) {
                o = alloc_act(cap, 1, 1, fout);
                if (!o.ptr) return false;
                e.out_act = o.ptr; e.ldo = o.cp;
            } else {
                e.out_f32 = om_out; e.ld_f32 = cfg.z_dim; e.f32_cols = cfg.z_dim;
            }
            if (!dense_op(om_prog, pre + "#wn", h, fin, fout, e, cap, &macs_omega)) return false;
            h = o;
        }
        return true;
    }
    const int oh = omega_base_h(cfg), ow = omega_base_w(cfg);
    const int feat = 256 * oh * ow;
    om_in = alloc_act(cap, 1, 1, 128);
    om_out = alloc_f32((size_t)cap * cfg.z_dim);
    om_noise = alloc_f32((size_t)cap * 128);
    Pair h;
    h.raw = alloc_act(cap, oh, ow, 256);
    h.act = alloc_act(cap, oh, ow, 256);
    if (!om_in.ptr || !om_out || !om_noise || !h.raw.ptr || !h.act.ptr) return false;
    {
        Epilogue e; e.out_raw = h.raw.ptr; e.out_act = h.act.ptr; e.ldo = feat;
        if (!dense_op(om_prog, "G_omega/level_0", om_in, 128, feat, e, cap, &macs_omega)) return false;
    }
    const std::vector<BlockSpec> blocks = omega_blocks(cfg);
    for (size_t i = 0; i < blocks.size(); ++i) {
        const bool last = (i + 1 == blocks.size());
        Pair o;
        if (!deconv_block(om_prog, "G_omega", blocks[i], h, true, !last, cap, &o, &macs_omega, nullptr)) return false;
        h = o;
    }
    const int co = h.c(), zch = cfg.z_dim / (4 * oh * ow);
    const HostArr* k = W_("G_omega/feature_map_mean_to_pixel/kernel", {1, 1, co, zch});
    const HostArr* b = W_("G_omega/feature_map_mean_to_pixel/bias", {zch});
    if (!k || !b) return false;
    Epilogue e;
    e.out_f32 = om_out; e.ld_f32 = zch; e.f32_cols = zch;     // [n][2oh][2ow][zch] == reshape_tail(x_mean, 3, (-1,))
    std::vector<TermSpec> terms; std::vector<const void*> srcs;
    if (!push_terms(terms, srcs, false, h.raw, h.raw_lo, false, 1, 1, co, zch, "G_omega/feature_map_mean_to_pixel/kernel", k, b))
        return false;
    return add_op(om_prog, "G_omega/feature_map_mean_to_pixel", terms, srcs, e, cap, &macs_omega);
}

bool Model::latent_gan_chunk(const float* z_omega_dev, int n, float* z0_dev, void* stream) {
    if (om_prog.empty()) return fail("latent_gan: create the handle with latent_gan = 1");
    const int din = omega_in_dim();
    KCHECK(launch_sample_z(nullptr, nullptr, z_omega_dev, n, 0, 1, din, 0.f, 0, nullptr, om_in.ptr, nullptr, om_in.cp, din,
                           om_in.cp, 0.2f, nullptr, nullptr, f16(), stream), "pack z_omega");
    if (!run_program(om_prog, n, stream)) return false;
    if (cfg.arch == APVAE_ARCH_M)      // the last dense layer's leaky_relu (dmm_bernouli_fashion.py:530-541)
        KCHECK(launch_leaky_relu_f32(om_out, (long long)n * cfg.z_dim, 0.2f, stream), "leaky_relu");
    const cudaError_t ce = cudaMemcpyAsync(z0_dev, om_out, (size_t)n * cfg.z_dim * sizeof(float), cudaMemcpyDeviceToDevice,
                                           static_cast<cudaStream_t>(stream));
    if (ce != cudaSuccess) return fail(std::string("latent_gan copy: ") + cudaGetErrorString(ce));
    return true;
}

bool Model::build_d() {
    const int cap = cfg.max_chunk;
    std::vector<BlockTape>* tape = cfg.enable_grad ? &d_tape : nullptr;
    const std::vector<BlockSpec> blocks = d_blocks(cfg);
    split_mode = cfg.precise != 0;
    Pair h = d_in;
    for (size_t i = 0; i < blocks.size(); ++i) {
        const bool last = (i + 1 == blocks.size());
        Pair o;
        if (!conv_block(d_prog, "D_psi", blocks[i], h, true, !last, cap, &o, &macs_d, tape)) return false;
        h = o;
    }
    d_flat = h.raw;
    if (h.raw.cp != h.raw.c) return fail("D_psi flatten has channel padding");
    const int feat = h.h() * h.w() * h.c();
    d64 = alloc_f32((size_t)cap * 64);
    if (!d64) return false;
    Epilogue e; e.out_f32 = d64; e.ld_f32 = 64; e.f32_cols = 64;
    if (!dense_op(d_prog, "D_psi/level_-2", h.raw, feat, 64, e, cap, &macs_d, &h.raw_lo)) return false;   // dmm_cifar10.py:559
    split_mode = false;
    const HostArr *kl = W_("D_psi/level_-1/kernel", {64, 1}), *bl = W_("D_psi/level_-1/bias", {1});
    if (!kl || !bl) return false;
    d_last_w = alloc_f32(64);
    d_last_b = alloc_f32(1);
    if (!d_last_w || !d_last_b) return false;
    cudaMemcpy(d_last_w, kl->v.data(), 64 * sizeof(float), cudaMemcpyHostToDevice);
    cudaMemcpy(d_last_b, bl->v.data(), sizeof(float), cudaMemcpyHostToDevice);
    macs_d += 64;
    return true;
}

bool Model::build() {
    if (built) return fail("weights already loaded (create a new handle to reload)");
    if (cudaSetDevice(dev.device) != cudaSuccess) return fail("cudaSetDevice failed");
    if (!build_q() || !build_g() || !build_d()) return false;
    // D_psi's first block as the fused CUDA-core block reads only the raw x_mean (it applies the activation itself):
    // the heads kernel then skips the activated copy (32 B per pixel)
    if (heads_op && !d_prog.empty() && d_prog[0]->is_small_block) {
        heads_op->heads.act = nullptr;
        heads_op->heads.act_lo = nullptr;
        if (!d_tape.empty()) d_tape[0].in.act = Tensor();      // the adjoint's leaky-relu' mask then reads the raw input (same signs)
    }
    prior4 = alloc_f32(4);
    if (!prior4) return false;
    if (cfg.latent_gan && !build_omega()) return false;
    if (cfg.enable_grad && !build_grad()) return false;
    z_tmp = alloc_f32((size_t)cfg.max_chunk * cfg.z_dim);
    if (!z_tmp) return false;
    if (cudaDeviceSynchronize() != cudaSuccess) return fail("device error while uploading weights");
    built = true;
    return true;
}

bool Model::ensure_sample_buffers(size_t samples) {
    if (samples <= sample_cap) return true;
    for (float** p : {&eps_buf, &log_px, &log_q, &log_pn, &d_g, &sq_err}) {
        if (*p) cudaFree(*p);
        *p = nullptr;
    }
    sample_cap = 0;
    if (cudaMalloc(&eps_buf, samples * cfg.z_dim * sizeof(float)) != cudaSuccess) return fail("cudaMalloc eps");
    for (float** p : {&log_px, &log_q, &log_pn, &d_g, &sq_err})
        if (cudaMalloc(p, samples * sizeof(float)) != cudaSuccess) return fail("cudaMalloc per-sample stats");
    sample_cap = samples;
    return true;
}


bool Model::run_encoder(const float* x_dev, int B, void* stream) {
    if (B > cfg.max_batch) return fail("batch exceeds max_batch");
    KCHECK(launch_pack_image(x_dev, (long long)B * HW(), C(), q_in.raw.cp, q_in.raw.ptr, q_in.act.ptr, 0.2f, f16(), stream),
           "pack_image");
    if (!run_program(q_prog, B, stream)) return false;
    // maybe_clip_value(z_logstd, min_val=min_logstd_of_q)  (dmm_cifar10.py:449)
    KCHECK(launch_clip_min(q_logstd, (long long)B * cfg.z_dim, cfg.min_logstd_of_q, stream), "clip_min");
    return true;
}

bool Model::run_decoder_chunk(int n, void* stream, bool want_f32, const LikArgs* lik) {
    const LikelihoodSpec spec = script_likelihood(cfg.bernoulli, cfg.epsilon);
    if (heads_op) {                               // per-call arguments of the fused heads kernel
        HeadsParams& hp = heads_op->heads;
        hp.out_f32 = want_f32 ? head_f32 : nullptr;
        hp.lik_part = lik ? lik_part : nullptr;
        if (lik) {
            hp.lik_x = lik->x; hp.lik_s0 = lik->s0; hp.lik_B = lik->B;
            hp.lik_has_std = head_ld() == 2 * C() ? 1 : 0;
            hp.lik_uniform = cfg.uniform_scale ? 1 : 0;
            hp.lik_beta = beta;
            hp.lik = spec;
        }
    }
    if (!run_program(g_prog, n, stream)) return false;
    if (!lik) return true;
    if (heads_op) {
        KCHECK(launch_sum_tiles(lik_part, heads_tiles_per_image(HW()), n, lik->log_px_out, lik->sq_err_out, stream), "sum_tiles");
    } else {                                      // heads ran as a GEMM op: score the fp32 heads in a separate pass
        KCHECK(launch_pixel_logprob(head_f32, head_ld(), lik->x, n, lik->s0, lik->B, HW(), C(), spec, cfg.uniform_scale, beta,
                                    lik->log_px_out, lik->sq_err_out, stream), "pixel_logprob");
    }
    return true;
}

bool Model::run_critic_chunk(int n, float* d_out, void* stream) {
    if (!run_program(d_prog, n, stream)) return false;
    KCHECK(launch_rowdot(d64, 64, d_last_w, d_last_b, 64, n, d_out, stream), "rowdot");   // dmm_cifar10.py:561-562
    return true;
}

bool Model::critic_of_images(const float* x_dev, int N, float* d_out, void* stream) {
    const size_t per = (size_t)HW() * C();
    for (int s0 = 0; s0 < N; s0 += cfg.max_chunk) {
        const int n = std::min(cfg.max_chunk, N - s0);
        KCHECK(launch_pack_image(x_dev + (size_t)s0 * per, (long long)n * HW(), C(), d_in.raw.cp, d_in.raw.ptr,
                                 d_in.act.ptr, 0.2f, f16(), stream, d_in.raw_lo.ptr, d_in.act_lo.ptr), "pack_image");
        if (!run_critic_chunk(n, d_out + s0, stream)) return false;
    }
    return true;
}

}  // namespace apvae
