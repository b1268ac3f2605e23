"""Layer arithmetic of the hot path, restated in numpy (test oracle).

Reference: tfsnippet/layers/convolutional/conv2d_.py, resnet.py, utils.py,
tfsnippet/layers/core/dense_.py, and TensorFlow 1.x op semantics (SAME padding,
conv2d_transpose as the exact adjoint, leaky_relu alpha=0.2).
"""
import numpy as np

__all__ = [
    'same_padding', 'leaky_relu', 'leaky_relu_grad', 'conv2d', 'deconv2d',
    'dense', 'conv2d_input_grad', 'deconv2d_input_grad',
    'resnet_conv2d_block', 'resnet_deconv2d_block', 'set_conv_backend', 'get_conv_backend',
]

# Which library evaluates the k > 1 convolutions.  'numpy': sliding-window + tensordot (the default: plain numpy,
# any dtype).  'torch': torch-CPU conv2d / conv_transpose2d (oneDNN) on the SAME explicitly padded arithmetic — the
# closest available stand-in for TensorFlow-CPU's Eigen / MKL-DNN kernels, used by bench.py's CPU baseline so that the
# GPU / CPU ratio is not flattered by a slow numpy conv.  Both are cross-checked in tests/test_oracle_cpu.py.
_BACKEND = 'numpy'


def set_conv_backend(name):
    global _BACKEND
    if name not in ('numpy', 'torch'):
        raise ValueError("backend must be 'numpy' or 'torch'")
    old, _BACKEND = _BACKEND, name
    return old


def get_conv_backend():
    return _BACKEND


def _torch_conv(x4, kernel, s, pads):
    """SAME cross-correlation of conv2d_.py:204-211 through torch: explicit asymmetric zero padding, then a VALID conv."""
    import torch
    import torch.nn.functional as F
    (plo_h, phi_h), (plo_w, phi_w) = pads
    xt = torch.from_numpy(np.ascontiguousarray(x4)).permute(0, 3, 1, 2)
    xt = F.pad(xt, (plo_w, phi_w, plo_h, phi_h))
    wt = torch.from_numpy(np.ascontiguousarray(kernel)).permute(3, 2, 0, 1)
    return F.conv2d(xt, wt, stride=s).permute(0, 2, 3, 1).numpy()


def _torch_deconv_full(x4, kernel, s):
    """sum_{di,dj} x[i, j] K[di, dj]^T scattered to (i s + di, j s + dj): torch's conv_transpose2d without padding."""
    import torch
    import torch.nn.functional as F
    xt = torch.from_numpy(np.ascontiguousarray(x4)).permute(0, 3, 1, 2)
    wt = torch.from_numpy(np.ascontiguousarray(kernel)).permute(3, 2, 0, 1)      # [Cin, Cout, kh, kw]
    return F.conv_transpose2d(xt, wt, stride=s).permute(0, 2, 3, 1).numpy()


def same_padding(in_size, k, s):
    """TF 'SAME' rule: out = ceil(in/s); pad_total = max((out-1)*s + k - in, 0);
    pad_lo = pad_total // 2 (tf.nn.conv2d, called at
    conv2d_.py:204-211)."""
    out = -(-in_size // s)
    total = max((out - 1) * s + k - in_size, 0)
    lo = total // 2
    return out, lo, total - lo


def leaky_relu(v, alpha=0.2):
    """tf.nn.leaky_relu default alpha (scripts' activation_fn,
    dmm_cifar10.py:428,475,546)."""
    return np.where(v > 0, v, alpha * v)


def leaky_relu_grad(v, alpha=0.2):
    return np.where(v > 0, 1.0, alpha).astype(v.dtype)


def _flatten4(x):
    lead = x.shape[:-3]
    return x.reshape((-1,) + x.shape[-3:]), lead


def conv2d(x, kernel, bias=None, strides=1, activation_fn=None):
    """spt.layers.conv2d, channels_last, padding='same' (conv2d_.py:23-238).

    kernel [kh, kw, Cin, Cout].  A 1x1 kernel takes the strided-slice + matmul
    fast path of conv2d_.py:178-189; anything else is tf.nn.conv2d(SAME, NHWC)
    (:204-211) = zero-padded cross-correlation.  Bias is added before the
    activation (:214-227).  Leading dims beyond 4-D are flattened (:193)."""
    kh, kw, cin, cout = kernel.shape
    x4, lead = _flatten4(x)
    n, h, w, c = x4.shape
    assert c == cin, (x.shape, kernel.shape)
    s = int(strides)
    if kh == 1 and kw == 1:
        xs = x4[:, ::s, ::s, :]
        y = xs @ kernel.reshape(cin, cout)
    else:
        ho, plo_h, phi_h = same_padding(h, kh, s)
        wo, plo_w, phi_w = same_padding(w, kw, s)
        if _BACKEND == 'torch':
            y = _torch_conv(x4, np.asarray(kernel, dtype=x4.dtype), s, ((plo_h, phi_h), (plo_w, phi_w)))[:, :ho, :wo]
        else:
            xp = np.pad(x4, ((0, 0), (plo_h, phi_h), (plo_w, phi_w), (0, 0)))
            # windows: [n, Hp-kh+1, Wp-kw+1, c, kh, kw] -> stride subsample
            win = np.lib.stride_tricks.sliding_window_view(xp, (kh, kw), axis=(1, 2))
            win = win[:, ::s, ::s][:, :ho, :wo]
            y = np.tensordot(win, kernel, axes=([4, 5, 3], [0, 1, 2]))
    if bias is not None:
        y = y + bias
    if activation_fn is not None:
        y = activation_fn(y)
    return y.reshape(lead + y.shape[1:])


def deconv2d(x, kernel, bias=None, strides=1, activation_fn=None):
    """spt.layers.deconv2d, channels_last, padding='same' (conv2d_.py:243-501).

    kernel [kh, kw, Cout, Cin] (:339); output size in*s
    (layers/convolutional/utils.py:78-96); tf.nn.conv2d_transpose(SAME)
    (:464-471) = exact adjoint of the SAME forward conv that maps the output
    grid back to the input grid.  No 1x1 fast path."""
    kh, kw, cout, cin = kernel.shape
    x4, lead = _flatten4(x)
    n, h, w, c = x4.shape
    assert c == cin, (x.shape, kernel.shape)
    s = int(strides)
    oh, ow = h * s, w * s
    _, plo_h, _ = same_padding(oh, kh, s)
    _, plo_w, _ = same_padding(ow, kw, s)
    fh = max((h - 1) * s + kh, plo_h + oh)
    fw = max((w - 1) * s + kw, plo_w + ow)
    full = np.zeros((n, fh, fw, cout), dtype=x4.dtype)
    if _BACKEND == 'torch' and (kh > 1 or kw > 1):
        part = _torch_deconv_full(x4, np.asarray(kernel, dtype=x4.dtype), s)
        full[:, :part.shape[1], :part.shape[2], :] = part
    else:
        for di in range(kh):
            for dj in range(kw):
                full[:, di:di + (h - 1) * s + 1:s, dj:dj + (w - 1) * s + 1:s, :] += \
                    x4 @ kernel[di, dj].T
    y = full[:, plo_h:plo_h + oh, plo_w:plo_w + ow, :]
    if bias is not None:
        y = y + bias
    if activation_fn is not None:
        y = activation_fn(y)
    return y.reshape(lead + y.shape[1:])


def dense(x, kernel, bias=None, activation_fn=None):
    """spt.layers.dense (layers/core/dense_.py:14-171): flatten to 2-D,
    x @ W[in,out] + b, optional activation."""
    y = x @ kernel
    if bias is not None:
        y = y + bias
    if activation_fn is not None:
        y = activation_fn(y)
    return y


def conv2d_input_grad(gy, kernel, strides, in_hw):
    """d/dx of conv2d(x, kernel, strides) contracted with gy (what
    tf.gradients(energy, z) needs at dmm_cifar10.py:246): the transposed conv
    with the same kernel.  For the 1x1 strided-slice path the gradient is
    scattered to the picked pixels only."""
    kh, kw, cin, cout = kernel.shape
    s = int(strides)
    g4, lead = _flatten4(gy)
    if kh == 1 and kw == 1:
        gx = np.zeros((g4.shape[0], in_hw[0], in_hw[1], cin), dtype=gy.dtype)
        gx[:, ::s, ::s, :] = g4 @ kernel.reshape(cin, cout).T
    else:
        assert g4.shape[1] * s == in_hw[0] and g4.shape[2] * s == in_hw[1], \
            'oracle dgrad supports sizes divisible by the stride'
        gx = deconv2d(g4, kernel, None, s)
    return gx.reshape(lead + gx.shape[1:])


def deconv2d_input_grad(gy, kernel, strides):
    """d/dx of deconv2d(x, kernel, strides) contracted with gy = the SAME
    forward conv with the same [kh,kw,Cout,Cin] array read as a conv kernel
    [kh,kw,Cin_conv=Cout,Cout_conv=Cin]."""
    kh, kw, cout, cin = kernel.shape
    s = int(strides)
    g4, lead = _flatten4(gy)
    n, h, w, _ = g4.shape
    ho, plo_h, phi_h = same_padding(h, kh, s)
    wo, plo_w, phi_w = same_padding(w, kw, s)
    gp = np.pad(g4, ((0, 0), (plo_h, phi_h), (plo_w, phi_w), (0, 0)))
    win = np.lib.stride_tricks.sliding_window_view(gp, (kh, kw), axis=(1, 2))
    win = win[:, ::s, ::s][:, :ho, :wo]
    gx = np.tensordot(win, kernel, axes=([4, 5, 3], [0, 1, 2]))
    return gx.reshape(lead + gx.shape[1:])


def _block(x, params, scope, out_channels, kernel_size, strides, conv_fn,
           resize_at_exit, tape=None):
    """resnet_general_block with normalizer_fn=None, dropout_fn=None,
    gated=False, activation_fn=leaky_relu (resnet.py:145-267):

      shortcut = conv1x1_s(input, use_bias=True) iff strides != 1 or Cin != Cout
                 (on the RAW input, :180-182,225-243)
      residual = conv_1(act(conv_0(act(input))))            (:246-256)
      output   = shortcut + residual                         (:265)

    resize_at_exit=True  (conv block,  :281): conv_0 Cin->Cin s1, conv_1 Cin->Cout s.
    resize_at_exit=False (deconv block,:379): conv_0 Cin->Cout s, conv_1 Cout->Cout s1.
    Every conv has a bias because normalizer_fn is None (:185-186)."""
    in_channels = x.shape[-1]
    s = int(strides)
    use_shortcut = (s != 1) or (in_channels != out_channels)
    p = lambda name: params[scope + '/' + name]
    if use_shortcut:
        shortcut = conv_fn(x, p('shortcut/kernel'), p('shortcut/bias'), s)
    else:
        shortcut = x
    a0 = leaky_relu(x)
    s0, s1 = (1, s) if resize_at_exit else (s, 1)
    r0 = conv_fn(a0, p('residual/conv_0/kernel'), p('residual/conv_0/bias'), s0)
    a1 = leaky_relu(r0)
    r1 = conv_fn(a1, p('residual/conv_1/kernel'), p('residual/conv_1/bias'), s1)
    if tape is not None:
        tape.append(dict(scope=scope, x=x, r0=r0, s0=s0, s1=s1, s=s,
                         use_shortcut=use_shortcut, deconv=conv_fn is deconv2d))
    return shortcut + r1


def resnet_conv2d_block(x, params, scope, out_channels, kernel_size=3,
                        strides=1, tape=None):
    """spt.layers.resnet_conv2d_block (resnet.py:272-364), shortcut k=1."""
    return _block(x, params, scope, out_channels, kernel_size, strides, conv2d,
                  True, tape)


def resnet_deconv2d_block(x, params, scope, out_channels, kernel_size=3,
                          strides=1, tape=None):
    """spt.layers.resnet_deconv2d_block (resnet.py:369-485), shortcut k=1."""
    return _block(x, params, scope, out_channels, kernel_size, strides,
                  deconv2d, False, tape)


def block_input_grad(g_out, params, rec):
    """Input gradient of one resnet block from its tape record (manual
    reverse-mode of `_block`; used for grad_z E in the MALA oracle)."""
    scope = rec['scope']
    p = lambda name: params[scope + '/' + name]
    x, r0 = rec['x'], rec['r0']
    hw = x.shape[-3:-1]
    if rec['deconv']:
        g_a1 = deconv2d_input_grad(g_out, p('residual/conv_1/kernel'), rec['s1'])
        g_r0 = g_a1 * leaky_relu_grad(r0)
        g_a0 = deconv2d_input_grad(g_r0, p('residual/conv_0/kernel'), rec['s0'])
        g_x = g_a0 * leaky_relu_grad(x)
        if rec['use_shortcut']:
            g_x = g_x + deconv2d_input_grad(g_out, p('shortcut/kernel'), rec['s'])
        else:
            g_x = g_x + g_out
    else:
        g_a1 = conv2d_input_grad(g_out, p('residual/conv_1/kernel'), rec['s1'],
                                 r0.shape[-3:-1])
        g_r0 = g_a1 * leaky_relu_grad(r0)
        g_a0 = conv2d_input_grad(g_r0, p('residual/conv_0/kernel'), rec['s0'], hw)
        g_x = g_a0 * leaky_relu_grad(x)
        if rec['use_shortcut']:
            g_x = g_x + conv2d_input_grad(g_out, p('shortcut/kernel'), rec['s'], hw)
        else:
            g_x = g_x + g_out
    return g_x


__all__.append('block_input_grad')
