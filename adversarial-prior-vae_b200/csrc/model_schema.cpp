// Architecture tables + TF variable schema (host only, no CUDA).
#include "model.h"

namespace apvae {

std::vector<BlockSpec> q_blocks(const apvae_config& c) {
    if (c.arch == APVAE_ARCH_A)          // dmm_cifar10.py:432-438
        return {{"level_0", 96, 5, 1}, {"level_1", 96, 3, 1}, {"level_2", 192, 3, 2}, {"level_4", 192, 3, 1},
                {"level_5", 192, 3, 2}, {"level_7", 192, 3, 1}, {"level_9", 192, 3, 1}};
    if (c.arch == APVAE_ARCH_B) {        // dmm_for_sampling.py:433-438
        const int s0 = c.celeba ? 2 : 1;
        return {{"level_0", 32, 5, s0}, {"level_1", 32, 3, 1}, {"level_2", 64, 3, 2}, {"level_4", 64, 3, 1},
                {"level_5", 128, 3, 2}};
    }
    return {{"level_0", 16, 3, 1}, {"level_2", 32, 3, 1}, {"level_3", 32, 3, 1}, {"level_4", 32, 3, 2},
            {"level_6", 64, 3, 2}};     // dmm_bernouli_fashion.py:444-448
}

std::vector<BlockSpec> g_blocks(const apvae_config& c) {
    if (c.arch == APVAE_ARCH_A)          // dmm_cifar10.py:481-488
        return {{"level_0", 192, 3, 1}, {"level_1", 192, 3, 1}, {"level_2", 192, 3, 1}, {"level_3", 192, 3, 2},
                {"level_5", 192, 3, 1}, {"level_6", 192, 3, 1}, {"level_8", 96, 3, 2}, {"level_9", 96, 5, 1}};
    if (c.arch == APVAE_ARCH_B) {        // dmm_for_sampling.py:497-502
        const int s9 = c.celeba ? 2 : 1;
        return {{"level_3", 64, 3, 2}, {"level_5", 64, 3, 1}, {"level_6", 32, 3, 2}, {"level_8", 32, 3, 1},
                {"level_9", 32, 5, s9}};
    }
    return {{"level_3", 64, 3, 2}, {"level_5", 32, 3, 2}, {"level_6", 32, 3, 1}, {"level_7", 32, 3, 1},
            {"level_8", 16, 3, 1}};     // dmm_bernouli_fashion.py:510-514
}

std::vector<BlockSpec> d_blocks(const apvae_config& c) {
    if (c.arch == APVAE_ARCH_A)          // dmm_cifar10.py:551-556
        return {{"level_0", 16, 3, 1}, {"level_1", 32, 3, 1}, {"level_2", 64, 3, 1}, {"level_3", 128, 3, 2},
                {"level_4", 256, 3, 2}, {"level_5", 256, 3, 2}};
    if (c.arch == APVAE_ARCH_B) {        // dmm_for_sampling.py:565-570
        const int s0 = c.celeba ? 2 : 1;
        return {{"level_0", 32, 3, s0}, {"level_1", 32, 3, 1}, {"level_3", 64, 3, 2}, {"level_4", 64, 3, 1},
                {"level_5", 64, 3, 2}};
    }
    return {{"level_0", 16, 3, 1}, {"level_2", 32, 3, 1}, {"level_3", 32, 3, 1}, {"level_4", 32, 3, 2},
            {"level_6", 64, 3, 2}};     // dmm_bernouli_fashion.py:561-565
}

std::vector<BlockSpec> omega_blocks(const apvae_config&) {      // dmm_cifar10.py:523-528 (same in dmm_for_sampling.py)
    return {{"level_1", 256, 3, 1}, {"level_2", 256, 3, 1}, {"level_3", 128, 3, 2}, {"level_4", 64, 3, 1},
            {"level_5", 32, 3, 1}, {"level_6", 16, 3, 1}};
}
int omega_base_h(const apvae_config& c) { return c.arch == APVAE_ARCH_A ? c.x_h / 8 : 4; }   // :516-521 / for_sampling :529-535
int omega_base_w(const apvae_config& c) { return c.arch == APVAE_ARCH_A ? c.x_w / 8 : 4; }

bool validate_config(const apvae_config& c, std::string* err) {
    auto fail = [&](const char* m) { if (err) *err = m; return false; };
    if (c.arch < 0 || c.arch > 2) return fail("arch must be 0 (A), 1 (B) or 2 (M)");
    if (c.x_h <= 0 || c.x_w <= 0 || c.x_c <= 0 || c.x_c > 16) return fail("bad x_shape");
    if (c.x_h % 4 || c.x_w % 4) return fail("x_shape must be divisible by 4");
    if (c.z_dim <= 0) return fail("z_dim must be positive");
    if (c.arch == APVAE_ARCH_A) {
        const int zc = 16 * c.z_dim / c.x_h / c.x_w;   // dmm_cifar10.py:440
        if (zc < 1 || zc * (c.x_h / 4) * (c.x_w / 4) != c.z_dim)
            return fail("arch A needs z_dim = zc * (H/4) * (W/4) with zc = 16*z_dim/(H*W) >= 1");
    }
    if (c.arch == APVAE_ARCH_B) {
        const int want = c.celeba ? 64 : 32;
        if (c.x_h != want || c.x_w != want) return fail("arch B: x_shape must be 64x64 (celebA) or 32x32");
    }
    if (c.max_batch <= 0 || c.max_chunk <= 0) return fail("max_batch / max_chunk must be positive");
    if (c.latent_gan) {
        const int cells = 4 * omega_base_h(c) * omega_base_w(c);
        if (c.arch != APVAE_ARCH_M && (c.z_dim % cells || c.z_dim / cells < 1))
            return fail("latent_gan: z_dim must be a multiple of the G_omega output grid");
    }
    if (c.nf_layers < 0 || c.nf_layers > 256) return fail("nf_layers must be in [0, 256]");
    if (c.nf_layers > 0 && c.z_dim > 512) return fail("posterior flow: z_dim <= 512");
    return true;
}

static int block_schema(std::vector<WeightInfo>& out, const std::string& root, const std::vector<BlockSpec>& blocks,
                        int cin, bool deconv) {
    for (const BlockSpec& b : blocks) {
        const std::string pre = root + "/" + b.scope + "/";
        const int64_t k = b.k, co = b.cout, ci = cin;
        if (deconv) {   // resnet.py:219-223 (resize_at_exit=False); deconv kernels [kh,kw,Cout,Cin] (conv2d_.py:339)
            out.push_back({pre + "residual/conv_0/kernel", {k, k, co, ci}});
            out.push_back({pre + "residual/conv_0/bias", {co}});
            out.push_back({pre + "residual/conv_1/kernel", {k, k, co, co}});
            out.push_back({pre + "residual/conv_1/bias", {co}});
            if (b.stride != 1 || ci != co) {
                out.push_back({pre + "shortcut/kernel", {1, 1, co, ci}});
                out.push_back({pre + "shortcut/bias", {co}});
            }
        } else {        // resnet.py:214-218 (resize_at_exit=True)
            out.push_back({pre + "residual/conv_0/kernel", {k, k, ci, ci}});
            out.push_back({pre + "residual/conv_0/bias", {ci}});
            out.push_back({pre + "residual/conv_1/kernel", {k, k, ci, co}});
            out.push_back({pre + "residual/conv_1/bias", {co}});
            if (b.stride != 1 || ci != co) {
                out.push_back({pre + "shortcut/kernel", {1, 1, ci, co}});
                out.push_back({pre + "shortcut/bias", {co}});
            }
        }
        cin = b.cout;
    }
    return cin;
}

std::vector<WeightInfo> weight_schema(const apvae_config& c) {
    std::vector<WeightInfo> out;
    const int64_t d = c.z_dim;
    // q_net
    int cq = block_schema(out, "q_net", q_blocks(c), c.x_c, false);
    if (c.arch == APVAE_ARCH_A) {
        const int zc = 16 * c.z_dim / c.x_h / c.x_w;
        block_schema(out, "q_net", {{"z_mean", zc, 3, 1}}, cq, false);
        block_schema(out, "q_net", {{"z_logstd", zc, 3, 1}}, cq, false);
    } else {
        int h = c.x_h, w = c.x_w;
        for (const BlockSpec& b : q_blocks(c)) { h = (h + b.stride - 1) / b.stride; w = (w + b.stride - 1) / b.stride; }
        for (const char* head : {"z_mean", "z_logstd"}) {
            out.push_back({std::string("q_net/") + head + "/kernel", {(int64_t)h * w * cq, d}});
            out.push_back({std::string("q_net/") + head + "/bias", {d}});
        }
    }
    // G_theta
    int cg;
    if (c.arch == APVAE_ARCH_A) {
        cg = 16 * c.z_dim / c.x_h / c.x_w;
    } else {
        const int gh = c.arch == APVAE_ARCH_B ? 8 : c.x_h / 4, gw = c.arch == APVAE_ARCH_B ? 8 : c.x_w / 4;
        cg = 64;
        out.push_back({"G_theta/level_0/kernel", {d, (int64_t)64 * gh * gw}});
        out.push_back({"G_theta/level_0/bias", {(int64_t)64 * gh * gw}});
    }
    cg = block_schema(out, "G_theta", g_blocks(c), cg, true);
    if (c.arch == APVAE_ARCH_M) {
        out.push_back({"G_theta/feature_map_mean_to_pixel/kernel", {1, 1, cg, c.x_c}});
        out.push_back({"G_theta/feature_map_mean_to_pixel/bias", {c.x_c}});
    } else {
        for (const char* head : {"x_mean", "x_logstd"}) {
            out.push_back({std::string("G_theta/") + head + "/kernel", {1, 1, cg, c.x_c}});
            out.push_back({std::string("G_theta/") + head + "/bias", {c.x_c}});
        }
    }
    // D_psi
    int cd = block_schema(out, "D_psi", d_blocks(c), c.x_c, false);
    int h = c.x_h, w = c.x_w;
    for (const BlockSpec& b : d_blocks(c)) { h = (h + b.stride - 1) / b.stride; w = (w + b.stride - 1) / b.stride; }
    out.push_back({"D_psi/level_-2/kernel", {(int64_t)h * w * cd, 64}});
    out.push_back({"D_psi/level_-2/bias", {64}});
    out.push_back({"D_psi/level_-1/kernel", {64, 1}});
    out.push_back({"D_psi/level_-1/bias", {1}});
    // G_omega (dmm_cifar10.py:503-532, dmm_for_sampling.py:519-546): dense 128 -> 256*oh*ow, six deconv blocks, 1x1 conv
    // arch M (dmm_bernouli_fashion.py:526-543): six weight-normalised dense layers z_dim -> 40 x5 -> z_dim; variables per
    // layer (tf default scope names dense, dense_1, ...): kernel, weight_norm/scale (kernel-shaped, weight_norm_.py:84-92), bias
    if (c.latent_gan && c.arch == APVAE_ARCH_M) {
        for (int i = 0; i < 6; ++i) {
            const std::string pre = std::string("G_omega/dense") + (i ? "_" + std::to_string(i) : "") + "/";
            const int64_t fin = i == 0 ? c.z_dim : 40, fout = i == 5 ? c.z_dim : 40;
            out.push_back({pre + "kernel", {fin, fout}});
            out.push_back({pre + "weight_norm/scale", {fin, fout}});
            out.push_back({pre + "bias", {fout}});
        }
    } else if (c.latent_gan) {
        const int oh = omega_base_h(c), ow = omega_base_w(c);
        out.push_back({"G_omega/level_0/kernel", {128, (int64_t)256 * oh * ow}});
        out.push_back({"G_omega/level_0/bias", {(int64_t)256 * oh * ow}});
        const int co = block_schema(out, "G_omega", omega_blocks(c), 256, true);
        const int64_t zch = c.z_dim / (4 * oh * ow);
        out.push_back({"G_omega/feature_map_mean_to_pixel/kernel", {1, 1, co, zch}});
        out.push_back({"G_omega/feature_map_mean_to_pixel/bias", {zch}});
    }
    // posterior flow: planar_normalizing_flows(config.nf_layers, name='posterior_flow') (dmm_cifar10.py:850-851;
    // variable scopes planar_nf.py:208-221, variables w [1,d], b [1], u [1,d] :79-103)
    for (int i = 0; i < c.nf_layers; ++i) {
        const std::string pre = c.nf_layers == 1 ? std::string("posterior_flow/") : "posterior_flow/_" + std::to_string(i) + "/";
        out.push_back({pre + "w", {1, d}});
        out.push_back({pre + "b", {1}});
        out.push_back({pre + "u", {1, d}});
    }
    return out;
}

}  // namespace apvae
