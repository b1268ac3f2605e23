"""GPU parity of the tap-GEMM kernel against the oracle, through the C ABI
(apvae_plan_create / apvae_plan_run).  Modelled on the reference's
tests/layers/convolutional/test_conv2d.py (wrapper == raw op + bias + act) and
tests/layers/core/test_dense.py (dense == np.dot + b), with the numpy oracle in
place of TensorFlow.  Inputs and kernels are bf16-exact, so the only difference
is fp32 accumulation order: tolerance 2e-4 relative to the output scale."""
import numpy as np
import pytest

import oracle as O
from helpers import bf16_round, exact16, pad_channels

pytestmark = pytest.mark.gpu

DT = 'fp16'     # the 16-bit operand format under test (set per test by the fixture below)


@pytest.fixture(autouse=True, params=['fp16', 'bf16'])
def operand_format(request, monkeypatch):
    """Every layer test runs in both operand formats of the library (op16.h): fp16 (default) and bf16.  The inputs are
    bf16-exact numbers of moderate magnitude, hence exact in fp16 as well."""
    import sys
    monkeypatch.setattr(sys.modules[__name__], 'DT', request.param)
    yield request.param


def _torch():
    import torch
    return torch


def _td():
    import torch
    return torch.float16 if DT == "fp16" else torch.bfloat16


def _check(got, want, tol=2e-4):
    scale = max(1.0, float(np.abs(want).max()))
    err = float(np.abs(got - want).max())
    assert err <= tol * scale, 'max err %g (scale %g)' % (err, scale)


LAYER_CASES = [
    # transpose, k, stride, N, H, W, cin, cout
    (False, 3, 1, 5, 8, 8, 64, 64),        # KC=64
    (False, 3, 1, 3, 32, 32, 96, 96),      # KC=32, N=96
    (False, 3, 1, 3, 16, 16, 192, 192),    # N=192
    (False, 5, 1, 2, 32, 32, 3, 96),       # KC=16, Cin=3
    (False, 3, 2, 4, 32, 32, 64, 128),     # stride-2 through space-to-depth view
    (False, 3, 2, 9, 8, 8, 256, 256),      # N=256, 4x4 output, 8 images per tile
    (False, 5, 2, 2, 64, 64, 3, 32),       # celebA first block
    (False, 1, 2, 3, 16, 16, 128, 256),    # 1x1 stride-2 shortcut
    (False, 3, 1, 3, 28, 28, 16, 16),      # 112-row tiles
    (False, 3, 2, 3, 14, 14, 32, 64),      # 7x7 output
    (True, 3, 1, 5, 8, 8, 192, 192),
    (True, 5, 1, 2, 32, 32, 96, 96),
    (True, 3, 2, 5, 8, 8, 192, 192),       # 4 sub-pixel phases
    (True, 3, 2, 3, 16, 16, 192, 96),
    (True, 5, 2, 2, 32, 32, 32, 32),
    (True, 1, 1, 130, 8, 8, 2, 192),       # ragged: 130 images, 2 per tile
    (True, 3, 2, 3, 7, 7, 64, 64),         # arch M
]


@pytest.mark.parametrize('transpose,k,stride,N,H,W,cin,cout', LAYER_CASES)
def test_conv_layers(transpose, k, stride, N, H, W, cin, cout):
    torch = _torch()
    import apvae
    rs = np.random.RandomState(k + 10 * stride + cin + cout)
    x = exact16(rs.normal(size=(N, H, W, cin)))
    shp = (k, k, cout, cin) if transpose else (k, k, cin, cout)
    kern = exact16(rs.normal(size=shp) / np.sqrt(k * k * cin))
    bias = rs.normal(size=(cout,)).astype(np.float32)
    fn = apvae.deconv2d if transpose else apvae.conv2d
    got = fn(torch.tensor(x, device='cuda'), kern, bias, stride, dtype=DT).cpu().numpy()
    ofn = O.deconv2d if transpose else O.conv2d
    want = ofn(x.astype(np.float64), kern.astype(np.float64), bias.astype(np.float64), stride)
    assert got.shape == want.shape
    _check(got, want)


@pytest.mark.parametrize('n,fin,fout', [(7, 4096, 64), (300, 128, 4096), (5, 40, 3136), (64, 8192, 128)])
def test_dense(n, fin, fout):
    torch = _torch()
    import apvae
    rs = np.random.RandomState(fin + fout)
    x = exact16(rs.normal(size=(n, fin)))
    k = exact16(rs.normal(size=(fin, fout)) / np.sqrt(fin))
    b = rs.normal(size=fout).astype(np.float32)
    got = apvae.dense(torch.tensor(x, device='cuda'), k, b, dtype=DT).cpu().numpy()
    _check(got, x.astype(np.float64) @ k.astype(np.float64) + b)


def test_tanh_activation():
    torch = _torch()
    import apvae
    rs = np.random.RandomState(0)
    x = exact16(rs.normal(size=(2, 32, 32, 96)))
    k = exact16(rs.normal(size=(1, 1, 96, 3)) * 0.2)
    b = rs.normal(size=3).astype(np.float32)
    got = apvae.conv2d(torch.tensor(x, device='cuda'), k, b, 1, 'tanh', dtype=DT).cpu().numpy()
    _check(got, O.conv2d(x.astype(np.float64), k.astype(np.float64), b.astype(np.float64), 1, np.tanh))


def test_epilogue_residual_mask_and_bf16_outputs():
    """v = (acc + bias) * lrelu'(mask) + res ; out_raw = bf16(v), out_act = bf16(lrelu(v));
    columns >= bf16_zero_from stored as 0; only n_valid images written."""
    torch = _torch()
    from apvae.plan import Plan, pad16
    from apvae import _lib
    rs = np.random.RandomState(5)
    cap, n_valid, H, cin, cout = 6, 5, 8, 32, 48
    x = exact16(rs.normal(size=(cap, H, H, cin)))
    kern = exact16(rs.normal(size=(3, 3, cin, cout)) * 0.1)
    bias = rs.normal(size=cout).astype(np.float32)
    res = exact16(rs.normal(size=(cap, H, H, cout)))
    msk = exact16(rs.normal(size=(cap, H, H, cout)))
    plan = Plan([dict(src=0, in_h=H, in_w=H, kernel=kern, bias=bias)], cap, dtype=DT)
    dev = 'cuda'
    t = lambda a: torch.tensor(a, device=dev).to(_td())
    out_raw = torch.full((cap, H, H, cout), 7.0, dtype=_td(), device=dev)
    out_act = torch.full((cap, H, H, cout), 7.0, dtype=_td(), device=dev)
    out_f32 = torch.full((cap, H, H, cout), 7.0, dtype=torch.float32, device=dev)
    plan.run([t(pad_channels(x, pad16(cin)))], n_valid, out_raw=out_raw, out_act=out_act, out_f32=out_f32,
             f32_cols=cout, res=t(res), res_mode=_lib.RES_SAME, mask=t(msk), bf16_zero_from=40)
    conv = O.conv2d(x.astype(np.float64), kern.astype(np.float64), bias.astype(np.float64), 1)
    want = conv * np.where(msk > 0, 1.0, 0.2) + res
    got = out_f32.cpu().numpy()
    _check(got[:n_valid], want[:n_valid])
    assert np.all(got[n_valid:] == 7.0), 'rows beyond n_valid must not be written'
    raw = out_raw.float().cpu().numpy()
    act = out_act.float().cpu().numpy()
    w40 = want[:n_valid].copy(); w40[..., 40:] = 0
    _check(raw[:n_valid], w40, tol=1e-2)
    _check(act[:n_valid], np.where(w40 > 0, w40, 0.2 * w40), tol=1e-2)
    assert np.all(raw[n_valid:] == 7.0)


def test_epilogue_even_pixel_residual():
    """RES_EVEN: the 1x1 stride-2 transposed shortcut adds x W only at even output pixels."""
    torch = _torch()
    from apvae.plan import Plan, pad16
    from apvae import _lib
    rs = np.random.RandomState(9)
    n, H, c = 3, 8, 64
    r = exact16(rs.normal(size=(n, 2 * H, 2 * H, c)))
    kern = exact16(rs.normal(size=(3, 3, c, c)) * 0.05)
    sc_low = exact16(rs.normal(size=(n, H, H, c)))
    plan = Plan([dict(src=0, transpose=True, in_h=2 * H, in_w=2 * H, kernel=kern)], n, dtype=DT)
    t = lambda a: torch.tensor(a, device='cuda').to(_td())
    out = torch.zeros((n, 2 * H, 2 * H, c), dtype=torch.float32, device='cuda')
    plan.run([t(r)], n, out_f32=out, f32_cols=c, res=t(sc_low), res_mode=_lib.RES_EVEN)
    want = O.deconv2d(r.astype(np.float64), kern.astype(np.float64), None, 1)
    want[:, ::2, ::2, :] += sc_low
    _check(out.cpu().numpy(), want)


def test_fused_shortcut_gpu():
    torch = _torch()
    from apvae.plan import Plan
    rs = np.random.RandomState(7)
    n, H, cin, cout = 4, 16, 128, 256
    r = exact16(rs.normal(size=(n, H, H, cin)))
    x = exact16(rs.normal(size=(n, H, H, cin)))
    k1 = exact16(rs.normal(size=(3, 3, cin, cout)) * 0.03)
    ks = exact16(rs.normal(size=(1, 1, cin, cout)) * 0.1)
    b1, bs = rs.normal(size=cout).astype(np.float32), rs.normal(size=cout).astype(np.float32)
    plan = Plan([dict(src=0, stride=2, in_h=H, in_w=H, kernel=k1, bias=b1),
                 dict(src=1, stride=2, in_h=H, in_w=H, kernel=ks, bias=bs)], n, dtype=DT)
    t = lambda a: torch.tensor(a, device='cuda').to(_td())
    out = torch.zeros((n, H // 2, H // 2, cout), dtype=torch.float32, device='cuda')
    plan.run([t(r), t(x)], n, out_f32=out, f32_cols=cout)
    want = O.conv2d(r.astype(np.float64), k1.astype(np.float64), b1.astype(np.float64), 2) + \
        O.conv2d(x.astype(np.float64), ks.astype(np.float64), bs.astype(np.float64), 2)
    _check(out.cpu().numpy(), want)


PATCH_GPU_CASES = [
    # transpose, k, N, H, W, cin, cout
    (False, 3, 5, 16, 16, 192, 192),
    (True, 3, 5, 16, 16, 192, 192),
    (True, 5, 3, 32, 32, 96, 96),
    (False, 3, 7, 32, 32, 16, 16),
    (False, 3, 3, 32, 32, 64, 64),
    (True, 3, 2, 64, 64, 32, 32),
    (False, 3, 9, 16, 16, 128, 128),
    (False, 3, 6, 8, 8, 192, 192),         # 8-wide images run as pairs (fold): even capacity
    (True, 3, 4, 8, 8, 64, 64),
    (True, 1, 130, 8, 8, 2, 192),
    (False, 3, 5, 28, 28, 32, 32),         # arch M: virtual widths (28 in 32, 14 in 16, 7 in 8 as pairs), fp32 output
    (True, 3, 4, 14, 14, 64, 64),          # through the generic epilogue, which skips the pixels past the image
    (False, 3, 6, 7, 7, 64, 64),
    (False, 5, 3, 14, 28, 16, 48),
]

VIRTUAL_W = {7: 8, 14: 16, 28: 32}         # grid width -> tile-row width of the A-stationary kernel (planner.cpp)


@pytest.mark.parametrize('cluster', [1, 2, 4])
@pytest.mark.parametrize('macc', [1, 2, 4])
@pytest.mark.parametrize('transpose,k,N,H,W,cin,cout', PATCH_GPU_CASES)
def test_patch_kernel(transpose, k, N, H, W, cin, cout, macc, cluster):
    """A-stationary kernel (patchgemm.cu) for every accumulator count / cluster size it supports."""
    torch = _torch()
    import apvae
    from apvae.plan import Plan, pad16
    Wv = VIRTUAL_W.get(W, W)
    Wt = 16 if Wv == 8 else Wv                 # 8-wide images: two side by side per tile row
    rt = macc * 128 // Wt                      # rows per tile
    Hv = -(-H // rt) * rt if Wv != W else H    # virtual widths: the last row tile may hang over the image by up to H / 7 rows
    if macc * pad16(cout) > 256 or Hv % rt or (Hv - H) * 7 > H or (pad16(cout) // cluster) % 8:
        pytest.skip('combination not supported by the planner')
    tuning = dict(patch=1)
    tuning['patch_macc'] = macc
    tuning['patch_cluster'] = cluster
    rs = np.random.RandomState(k + cin + cout + N)
    x = exact16(rs.normal(size=(N, H, W, cin)))
    shp = (k, k, cout, cin) if transpose else (k, k, cin, cout)
    kern = exact16(rs.normal(size=shp) / np.sqrt(k * k * cin))
    bias = rs.normal(size=(cout,)).astype(np.float32)
    plan = Plan([dict(src=0, transpose=transpose, in_h=H, in_w=W, kernel=kern, bias=bias)], N, dtype=DT, tuning=tuning)
    if plan.info.patch != 1:
        pytest.skip('forced accumulator count does not fit the shared-memory rings')
    assert plan.info.patch_macc == macc and plan.info.patch_cluster == cluster
    src = torch.tensor(pad_channels(x, pad16(cin)), device='cuda').to(_td())
    out = torch.zeros((N, H, W, cout), dtype=torch.float32, device='cuda')
    plan.run([src], N, out_f32=out, f32_cols=cout)
    ofn = O.deconv2d if transpose else O.conv2d
    want = ofn(x.astype(np.float64), kern.astype(np.float64), bias.astype(np.float64), 1)
    _check(out.cpu().numpy(), want)


def test_patch_kernel_matches_tap_kernel_bitwise_inputs():
    """Both kernels on the same op (fused 1x1 shortcut + residual-free epilogue, bf16 outputs)."""
    torch = _torch()
    from apvae.plan import Plan
    rs = np.random.RandomState(21)
    n, H, cin, cout = 6, 32, 32, 64
    r = exact16(rs.normal(size=(n, H, H, cin)))
    x = exact16(rs.normal(size=(n, H, H, cin)))
    k1 = exact16(rs.normal(size=(3, 3, cin, cout)) * 0.06)
    ks = exact16(rs.normal(size=(1, 1, cin, cout)) * 0.2)
    b1, bs = rs.normal(size=cout).astype(np.float32), rs.normal(size=cout).astype(np.float32)
    t = lambda a: torch.tensor(a, device='cuda').to(_td())
    outs = {}
    for mode in ('0', '1'):
        tuning = dict(patch=(1 if int(mode) else -1))
        plan = Plan([dict(src=0, in_h=H, in_w=H, kernel=k1, bias=b1), dict(src=1, in_h=H, in_w=H, kernel=ks, bias=bs)], n, dtype=DT, tuning=tuning)
        assert plan.info.patch == int(mode)
        o32 = torch.zeros((n, H, H, cout), dtype=torch.float32, device='cuda')
        oact = torch.zeros((n, H, H, cout), dtype=_td(), device='cuda')
        plan.run([t(r), t(x)], n - 1, out_f32=o32, f32_cols=cout, out_act=oact)
        outs[mode] = (o32.cpu().numpy(), oact.float().cpu().numpy())
    want = O.conv2d(r.astype(np.float64), k1.astype(np.float64), b1.astype(np.float64), 1) + \
        O.conv2d(x.astype(np.float64), ks.astype(np.float64), bs.astype(np.float64), 1)
    _check(outs['1'][0][:n - 1], want[:n - 1])
    _check(outs['1'][0], outs['0'][0], tol=1e-5)
    assert np.all(outs['1'][0][n - 1] == 0) and np.all(outs['1'][1][n - 1] == 0)
    _check(outs['1'][1], outs['0'][1], tol=1e-2)


TMA_EPI_CASES = [
    # H, W, cin, cout, res ('none' | 'same' | 'even'), n_valid, cap
    (16, 16, 192, 192, 'same', 5, 6),
    (16, 16, 192, 192, 'even', 3, 3),
    (32, 32, 96, 96, 'same', 3, 4),
    (32, 32, 96, 96, 'even', 3, 3),
    (32, 32, 64, 64, 'none', 2, 2),
    (64, 64, 32, 64, 'even', 2, 2),
    (16, 16, 128, 256, 'same', 4, 4),
    (32, 32, 16, 16, 'same', 3, 3),        # narrower than one 64-column box
    (32, 32, 32, 48, 'none', 2, 3),
    (128, 128, 16, 64, 'even', 1, 1),
    (8, 8, 192, 192, 'same', 5, 6),        # 8-wide images run as pairs: 5-D epilogue maps
    (8, 8, 192, 192, 'none', 4, 4),
    (8, 8, 2, 192, 'none', 3, 4),
    (8, 8, 256, 256, 'same', 7, 8),
    (8, 8, 64, 48, 'same', 2, 2),
    (8, 8, 64, 64, 'even', 3, 4),          # half-resolution residual: falls back to the tap-GEMM kernel
    (16, 8, 32, 32, 'same', 3, 4),         # 8 wide, 16 high: two row tiles per pair
    (28, 28, 32, 32, 'same', 3, 4),        # arch M (virtual widths): residual through its 4-D view, stores clipped by the TMA unit
    (28, 28, 16, 32, 'none', 2, 2),
    (28, 28, 32, 16, 'even', 2, 3),        # G_theta/level_5/conv_1: half-resolution residual from the 14 x 14 grid
    (14, 14, 64, 64, 'same', 3, 3),        # one 16-row tile of two accumulators (macc = 1: two tiles, the second 6 rows)
    (14, 14, 64, 64, 'even', 2, 3),        # G_theta/level_3/conv_1: residual from the 7 x 7 grid
    (7, 7, 64, 64, 'same', 3, 4),          # image pairs, row 7 and column 7 of every tile hang over
    (7, 7, 64, 48, 'none', 4, 4),
    (14, 28, 32, 96, 'same', 2, 2),
]


@pytest.mark.parametrize('epi2', [-1, 1])
@pytest.mark.parametrize('two_cta', [0, 1])
@pytest.mark.parametrize('macc', [0, 1])
@pytest.mark.parametrize('H,W,cin,cout,res,n_valid,cap', TMA_EPI_CASES)
def test_patch_kernel_tma_epilogue(H, W, cin, cout, res, n_valid, cap, macc, two_cta, epi2):
    """16-bit-only epilogue of the A-stationary kernel (bulk tensor loads of the residual, bulk tensor stores of
    out_raw / out_act), with one and with two epilogue warpgroups: v = acc + bias (+ res on every / on even pixels);
    raw = round16(v), act = round16(lrelu(v));
    images >= n_valid stay untouched (a folded 8-wide op may also write the second image of its last pair)."""
    torch = _torch()
    from apvae import _lib
    from apvae.plan import Plan, pad16
    tuning = dict(patch=1)
    tuning['patch_2cta'] = (1 if int(two_cta) else -1)     # cta_group::2 pairs: one 256-row MMA per two tiles
    tuning['patch_epi2'] = epi2                            # one / two epilogue warpgroups (256 / 384 threads)
    if macc:
        tuning['patch_macc'] = macc
    rs = np.random.RandomState(H + cin + cout + len(res))
    x = exact16(rs.normal(size=(cap, H, W, cin)))
    kern = exact16(rs.normal(size=(3, 3, cin, cout)) / np.sqrt(9 * cin))
    bias = rs.normal(size=(cout,)).astype(np.float32)
    plan = Plan([dict(src=0, in_h=H, in_w=W, kernel=kern, bias=bias)], cap, dtype=DT, tuning=tuning)
    if plan.info.patch != 1:
        pytest.skip('the rings do not fit next to the wider epilogue staging')
    assert plan.info.patch_epi2 == int(epi2 > 0)
    npad = pad16(cout)
    t = lambda a: torch.tensor(a, device='cuda').to(_td())
    out_raw = torch.full((cap, H, W, npad), 7.0, dtype=_td(), device='cuda')
    out_act = torch.full((cap, H, W, npad), 7.0, dtype=_td(), device='cuda')
    want = O.conv2d(x.astype(np.float64), kern.astype(np.float64), bias.astype(np.float64), 1)
    kw = {}
    if res == 'same':
        r = exact16(rs.normal(size=(cap, H, W, npad)))
        want = want + r[..., :cout]
        kw = dict(res=t(r), res_mode=_lib.RES_SAME)
    elif res == 'even':
        r = exact16(rs.normal(size=(cap, H // 2, W // 2, npad)))
        want[:, ::2, ::2, :] += r[..., :cout]
        kw = dict(res=t(r), res_mode=_lib.RES_EVEN)
    plan.run([t(pad_channels(x, pad16(cin)))], n_valid, out_raw=out_raw, out_act=out_act, **kw)
    torch.cuda.synchronize()
    raw = out_raw.float().cpu().numpy()
    act = out_act.float().cpu().numpy()
    scale = max(1.0, float(np.abs(want).max()))
    assert np.abs(raw[:n_valid, ..., :cout] - want[:n_valid]).max() <= 2 ** -7 * scale       # bf16 rounding of the output
    assert np.abs(act[:n_valid, ..., :cout] - O.leaky_relu(want[:n_valid])).max() <= 2 ** -7 * scale
    assert np.all(raw[:n_valid, ..., cout:] == 0) and np.all(act[:n_valid, ..., cout:] == 0)   # padding columns stay zero
    keep = n_valid + (n_valid & 1 if plan.info.patch_fold == 2 else 0)
    assert np.all(raw[keep:] == 7.0) and np.all(act[keep:] == 7.0)


SPLIT_EPI_CASES = [
    # transpose, k, stride, H, W (input), cin, cout, res, outs, n_valid, cap
    (False, 3, 1, 16, 16, 64, 64, 'same', 'raw+act', 3, 4),     # arch B block output with identity shortcut
    (False, 3, 1, 28, 28, 32, 32, 'same', 'raw+act', 2, 3),     # arch M widths: 4-D residual views, clipped stores
    (True, 3, 1, 14, 14, 64, 64, 'even', 'raw+act', 2, 2),
    (False, 3, 1, 7, 7, 64, 64, 'same', 'raw+act', 3, 4),
    (False, 3, 2, 28, 28, 32, 32, 'none', 'act', 2, 2),
    (True, 3, 2, 7, 7, 64, 64, 'none', 'act', 4, 4),
    (False, 3, 1, 32, 32, 32, 32, 'none', 'act', 2, 3),         # conv_0 of a block
    (True, 5, 1, 64, 64, 32, 32, 'same', 'raw', 1, 2),          # arch B celebA G_theta/level_9/conv_1
    (True, 3, 1, 32, 32, 32, 32, 'even', 'raw+act', 2, 2),      # stride-2 block: half-resolution shortcut on even pixels
    (False, 3, 1, 8, 8, 64, 48, 'same', 'raw+act', 3, 4),       # 8-wide images as pairs, 48 columns
    (False, 3, 1, 8, 8, 64, 64, 'none', 'act', 4, 4),
    (False, 3, 2, 32, 32, 32, 64, 'none', 'act', 2, 2),         # stride-2 conv through the space-to-depth maps
    (True, 3, 2, 16, 16, 64, 32, 'none', 'act', 3, 4),          # stride-2 transposed conv: merged row-phase ops
    (True, 5, 2, 32, 32, 32, 32, 'none', 'act', 2, 2),          # arch B celebA G_theta/level_9/conv_0
    (True, 3, 2, 8, 8, 64, 64, 'none', 'act', 4, 4),            # 8-wide input grid, phases on image pairs
]


@pytest.mark.parametrize('nterms', [2, 3])
@pytest.mark.parametrize('transpose,k,stride,H,W,cin,cout,res,outs,n_valid,cap', SPLIT_EPI_CASES)
def test_split_precision_ops_on_the_patch_kernel(transpose, k, stride, H, W, cin, cout, res, outs, n_valid, cap, nterms):
    """A split-precision op (apvae_config.precise): activations as hi + lo 16-bit pairs, kernels single (two terms,
    precise = 2) or hi + lo as well (three terms, precise = 1), outputs and residual as hi + lo pairs — on the
    A-stationary kernel with its split TMA epilogue (two store passes, hi then lo).  hi + lo of every output must
    reproduce the float64 result to fp32-accumulation accuracy: 2^-17 of the scale (the pair itself carries 22 significant
    bits; what is left is the fp32 accumulation of k^2 * cin products, measured ~1e-5 at scale 7)."""
    torch = _torch()
    from apvae import _lib
    from apvae.plan import Plan, pad16
    from helpers import op16_round
    rs = np.random.RandomState(H + cin + cout + k + nterms)
    a = rs.normal(size=(cap, H, W, cin)).astype(np.float32)
    a_hi = op16_round(a, DT); a_lo = op16_round(a - a_hi, DT)
    shp = (k, k, cout, cin) if transpose else (k, k, cin, cout)
    w = (rs.normal(size=shp) / np.sqrt(k * k * cin)).astype(np.float32)
    w_hi = op16_round(w, DT); w_lo = op16_round(w - w_hi, DT)
    bias = rs.normal(size=(cout,)).astype(np.float32)
    geo = dict(transpose=transpose, stride=stride, in_h=H, in_w=W)
    terms = [dict(src=0, kernel=w_hi, bias=bias, **geo), dict(src=1, kernel=w_hi, **geo)]
    if nterms == 3:
        terms.append(dict(src=0, kernel=w_lo, **geo))
    plan = Plan(terms, cap, dtype=DT, tuning=dict(patch=1))
    Ho, Wo = plan.info.ho, plan.info.wo
    npad = pad16(cout)
    t = lambda x: torch.tensor(np.ascontiguousarray(x), device='cuda').to(_td())
    ofn = O.deconv2d if transpose else O.conv2d
    act64 = a_hi.astype(np.float64) + a_lo
    want = ofn(act64, w_hi.astype(np.float64), bias.astype(np.float64), stride)
    if nterms == 3:
        want = want + ofn(a_hi.astype(np.float64), w_lo.astype(np.float64), None, stride)
    kw = {}
    if res != 'none':
        rshape = (cap, Ho, Wo, npad) if res == 'same' else (cap, Ho // 2, Wo // 2, npad)
        r = rs.normal(size=rshape).astype(np.float32)
        r[..., cout:] = 0
        r_hi = op16_round(r, DT); r_lo = op16_round(r - r_hi, DT)
        r64 = r_hi.astype(np.float64) + r_lo
        if res == 'same':
            want = want + r64[..., :cout]
        else:
            want[:, ::2, ::2, :] += r64[..., :cout]
        kw = dict(res=t(r_hi), res_lo=t(r_lo), res_mode=_lib.RES_SAME if res == 'same' else _lib.RES_EVEN)
    bufs = {}
    for name in ('raw', 'act'):
        if name in outs:
            bufs['out_' + name] = torch.full((cap, Ho, Wo, npad), 7.0, dtype=_td(), device='cuda')
            bufs['out_' + name + '_lo'] = torch.full((cap, Ho, Wo, npad), 7.0, dtype=_td(), device='cuda')
    plan.run([t(pad_channels(a_hi, pad16(cin))), t(pad_channels(a_lo, pad16(cin)))], n_valid, **bufs, **kw)
    torch.cuda.synchronize()
    scale = max(1.0, float(np.abs(want).max()))
    for name, ref in (('raw', want), ('act', O.leaky_relu(want))):
        if name not in outs:
            continue
        hi = bufs['out_' + name].float().cpu().numpy().astype(np.float64)
        lo = bufs['out_' + name + '_lo'].float().cpu().numpy().astype(np.float64)
        tol = (2.0 ** -17 if DT == 'fp16' else 2.0 ** -14) * scale
        assert np.abs((hi + lo)[:n_valid, ..., :cout] - ref[:n_valid]).max() <= tol, (name, np.abs((hi + lo)[:n_valid, ..., :cout] - ref[:n_valid]).max(), tol)
        assert np.abs(hi[:n_valid, ..., :cout] - ref[:n_valid]).max() <= 2.0 ** -7 * scale
        assert np.all(hi[:n_valid, ..., cout:] == 0) and np.all(lo[:n_valid, ..., cout:] == 0)
        keep = n_valid + (n_valid & 1 if plan.info.patch_fold == 2 or W in (7, 8) else 0)
        assert np.all(hi[keep:] == 7.0) and np.all(lo[keep:] == 7.0)


PHASE_CASES = [
    # k, H, W, cin, cout, n_valid, cap
    (3, 16, 16, 192, 96, 3, 4),        # arch A G_theta/level_8/conv_0
    (3, 16, 16, 64, 64, 5, 5),
    (5, 32, 32, 32, 32, 2, 2),         # arch B celebA
    (3, 32, 32, 64, 32, 2, 3),
    (5, 16, 16, 16, 48, 3, 3),
    (3, 64, 64, 16, 16, 1, 1),
    (3, 8, 8, 192, 192, 3, 4),         # arch A G_theta/level_3/conv_0: 8-wide inputs run as pairs
    (3, 8, 8, 64, 32, 4, 4),
    (5, 8, 8, 32, 32, 2, 2),
    (3, 7, 7, 64, 64, 3, 4),           # arch M G_theta/level_3/conv_0: 7 -> 14, pairs of 7-wide images in 8-wide tile rows
    (3, 14, 14, 64, 32, 2, 3),         # G_theta/level_5/conv_0: 14 -> 28
]


@pytest.mark.parametrize('merge', ['1', '0'])
@pytest.mark.parametrize('outs', ['act', 'raw+act'])
@pytest.mark.parametrize('k,H,W,cin,cout,n_valid,cap', PHASE_CASES)
def test_stride2_deconv_as_four_patch_phases(k, H, W, cin, cout, n_valid, cap, outs, merge):
    """conv2d_transpose(stride 2, SAME) (conv2d_.py:464-471) as four sub-pixel phase ops of the A-stationary kernel:
    phase (py, px) is a stride-1 correlation on the input grid written to output pixels (2y + py, 2x + px) by strided
    bulk tensor stores.  merge=1: two ops, one per output row phase, whose N = [px 0 | px 1] x cout is the pair of
    adjacent output pixels (2x, 2x + 1) written as one contiguous row (when 2 * cout <= 256)."""
    torch = _torch()
    from apvae.plan import Plan, pad16
    tuning = dict(patch=1)
    tuning['patch_merge'] = (1 if int(merge) else -1)
    rs = np.random.RandomState(k + H + cin + cout)
    x = exact16(rs.normal(size=(cap, H, W, cin)))
    kern = exact16(rs.normal(size=(k, k, cout, cin)) / np.sqrt(k * k * cin / 4.0))
    bias = rs.normal(size=(cout,)).astype(np.float32)
    plan = Plan([dict(src=0, transpose=True, stride=2, in_h=H, in_w=W, kernel=kern, bias=bias)], cap, dtype=DT, tuning=tuning)
    npad = pad16(cout)
    t = lambda a: torch.tensor(a, device='cuda').to(_td())
    out_act = torch.full((cap, 2 * H, 2 * W, npad), 7.0, dtype=_td(), device='cuda')
    out_raw = torch.full((cap, 2 * H, 2 * W, npad), 7.0, dtype=_td(), device='cuda') if outs != 'act' else None
    plan.run([t(pad_channels(x, pad16(cin)))], n_valid, out_raw=out_raw, out_act=out_act)
    torch.cuda.synchronize()
    want = O.deconv2d(x.astype(np.float64), kern.astype(np.float64), bias.astype(np.float64), 2)
    scale = max(1.0, float(np.abs(want).max()))
    act = out_act.float().cpu().numpy()
    assert np.abs(act[:n_valid, ..., :cout] - O.leaky_relu(want[:n_valid])).max() <= 2 ** -7 * scale
    keep = n_valid + (n_valid & 1 if W in (7, 8) else 0)      # pairs of 8-wide images: the last pair is written whole
    assert np.all(act[keep:] == 7.0)
    if out_raw is not None:
        raw = out_raw.float().cpu().numpy()
        assert np.abs(raw[:n_valid, ..., :cout] - want[:n_valid]).max() <= 2 ** -7 * scale
        assert np.all(raw[keep:] == 7.0)


S2_CASES = [
    # k, H, W (input), cin, cout, fused 1x1 stride-2 shortcut, n_valid, cap
    (3, 32, 32, 64, 128, True, 3, 4),      # arch A D_psi/level_3/conv_1
    (3, 16, 16, 128, 256, True, 5, 6),     # D_psi/level_4/conv_1: 8-wide output, image pairs
    (3, 32, 32, 96, 192, True, 2, 2),      # q_net/level_2/conv_1
    (5, 64, 64, 3, 32, False, 2, 3),       # celebA first block
    (3, 64, 64, 32, 64, True, 1, 2),
    (1, 32, 32, 16, 16, False, 3, 3),
    (3, 28, 28, 32, 32, True, 3, 4),       # arch M level_4/conv_1: 28 -> 14 (14 in 16-wide tile rows)
    (3, 14, 14, 32, 64, True, 3, 4),       # level_6/conv_1: 14 -> 7 (image pairs)
]


@pytest.mark.parametrize('two_cta', [0, 1])
@pytest.mark.parametrize('k,H,W,cin,cout,shortcut,n_valid,cap', S2_CASES)
def test_stride2_conv_on_patch_kernel(k, H, W, cin, cout, shortcut, n_valid, cap, two_cta):
    """conv2d(stride 2, SAME) (conv2d_.py:204-211) on the A-stationary kernel through the space-to-depth view of the
    input (one tensor map per row phase), optionally with the block's 1x1 stride-2 shortcut as a second term
    (resnet.py:234-241); bf16 raw + act outputs through the TMA epilogue."""
    torch = _torch()
    from apvae.plan import Plan, pad16
    tuning = dict(patch=1)
    tuning['patch_2cta'] = (1 if int(two_cta) else -1)
    rs = np.random.RandomState(k + H + cin + cout)
    t = lambda a: torch.tensor(a, device='cuda').to(_td())
    x = exact16(rs.normal(size=(cap, H, W, cin)))
    kern = exact16(rs.normal(size=(k, k, cin, cout)) / np.sqrt(k * k * cin))
    bias = rs.normal(size=(cout,)).astype(np.float32)
    terms = [dict(src=0, stride=2, in_h=H, in_w=W, kernel=kern, bias=bias)]
    srcs = [t(pad_channels(x, pad16(cin)))]
    want = O.conv2d(x.astype(np.float64), kern.astype(np.float64), bias.astype(np.float64), 2)
    if shortcut:
        x2 = exact16(rs.normal(size=(cap, H, W, cin)))
        ks = exact16(rs.normal(size=(1, 1, cin, cout)) / np.sqrt(cin))
        bs = rs.normal(size=(cout,)).astype(np.float32)
        terms.append(dict(src=1, stride=2, in_h=H, in_w=W, kernel=ks, bias=bs))
        srcs.append(t(pad_channels(x2, pad16(cin))))
        want = want + O.conv2d(x2.astype(np.float64), ks.astype(np.float64), bs.astype(np.float64), 2)
    plan = Plan(terms, cap, dtype=DT, tuning=tuning)
    assert plan.info.patch == 1 and plan.info.patch_s2d == 1
    npad = pad16(cout)
    out_raw = torch.full((cap, H // 2, W // 2, npad), 7.0, dtype=_td(), device='cuda')
    out_act = torch.full((cap, H // 2, W // 2, npad), 7.0, dtype=_td(), device='cuda')
    plan.run(srcs, n_valid, out_raw=out_raw, out_act=out_act)
    torch.cuda.synchronize()
    raw = out_raw.float().cpu().numpy()
    act = out_act.float().cpu().numpy()
    scale = max(1.0, float(np.abs(want).max()))
    assert np.abs(raw[:n_valid, ..., :cout] - want[:n_valid]).max() <= 2 ** -7 * scale
    assert np.abs(act[:n_valid, ..., :cout] - O.leaky_relu(want[:n_valid])).max() <= 2 ** -7 * scale
    keep = n_valid + (n_valid & 1 if plan.info.patch_fold == 2 else 0)
    assert np.all(raw[keep:] == 7.0) and np.all(act[keep:] == 7.0)
    # fp32 output through the generic epilogue of the same kernel
    o32 = torch.zeros((cap, H // 2, W // 2, cout), dtype=torch.float32, device='cuda')
    plan.run(srcs, n_valid, out_f32=o32, f32_cols=cout)
    _check(o32.cpu().numpy()[:n_valid], want[:n_valid])


MASK_CASES = [
    # H, W, cin, cout, res ('none' | 'same' | 'even'), n_valid, cap
    (32, 32, 96, 96, 'none', 3, 4),
    (32, 32, 96, 96, 'same', 2, 2),
    (16, 16, 192, 192, 'even', 3, 3),
    (32, 32, 32, 16, 'same', 3, 3),
    (64, 64, 32, 32, 'none', 1, 2),
    (8, 8, 192, 192, 'same', 5, 6),        # image pairs
    (8, 8, 64, 48, 'none', 4, 4),
    (28, 28, 32, 32, 'same', 2, 3),        # arch M widths: mask and residual through 4-D views
    (28, 28, 32, 16, 'none', 3, 3),
    (14, 14, 64, 64, 'even', 2, 2),
    (7, 7, 64, 64, 'same', 3, 4),
]


@pytest.mark.parametrize('epi2', [-1, 1])
@pytest.mark.parametrize('two_cta', [0, 1])
@pytest.mark.parametrize('H,W,cin,cout,res,n_valid,cap', MASK_CASES)
def test_patch_kernel_tma_epilogue_mask(H, W, cin, cout, res, n_valid, cap, two_cta, epi2):
    """Adjoint ops (model_grad.cu): v = (acc + bias) * leaky_relu'(mask) (+ res), raw bf16 output only, through the TMA
    epilogue of the A-stationary kernel: the mask (the forward activation, dmm_cifar10.py:241-247 via tf.gradients of
    tf.nn.leaky_relu) is fetched one block ahead into the unused act staging box."""
    torch = _torch()
    from apvae import _lib
    from apvae.plan import Plan, pad16
    tuning = dict(patch=1)
    tuning['patch_2cta'] = (1 if int(two_cta) else -1)
    tuning['patch_epi2'] = epi2                            # one / two epilogue warpgroups
    rs = np.random.RandomState(H + cin + cout + len(res))
    x = exact16(rs.normal(size=(cap, H, W, cin)))
    kern = exact16(rs.normal(size=(3, 3, cin, cout)) / np.sqrt(9 * cin))
    bias = rs.normal(size=(cout,)).astype(np.float32)
    plan = Plan([dict(src=0, in_h=H, in_w=W, kernel=kern, bias=bias)], cap, dtype=DT, tuning=tuning)
    if plan.info.patch != 1:
        pytest.skip('the rings do not fit next to the wider epilogue staging')
    npad = pad16(cout)
    t = lambda a: torch.tensor(a, device='cuda').to(_td())
    mask = exact16(rs.normal(size=(cap, H, W, npad)))
    mask[rs.uniform(size=mask.shape) < 0.1] = 0.0                 # lrelu'(0) = alpha
    want = O.conv2d(x.astype(np.float64), kern.astype(np.float64), bias.astype(np.float64), 1)
    want = want * np.where(mask[..., :cout] > 0, 1.0, 0.2)
    kw = {}
    if res == 'same':
        r = exact16(rs.normal(size=(cap, H, W, npad)))
        want = want + r[..., :cout]
        kw = dict(res=t(r), res_mode=_lib.RES_SAME)
    elif res == 'even':
        r = exact16(rs.normal(size=(cap, H // 2, W // 2, npad)))
        want[:, ::2, ::2, :] += r[..., :cout]
        kw = dict(res=t(r), res_mode=_lib.RES_EVEN)
    out_raw = torch.full((cap, H, W, npad), 7.0, dtype=_td(), device='cuda')
    plan.run([t(pad_channels(x, pad16(cin)))], n_valid, out_raw=out_raw, mask=t(mask), **kw)
    torch.cuda.synchronize()
    raw = out_raw.float().cpu().numpy()
    scale = max(1.0, float(np.abs(want).max()))
    assert np.abs(raw[:n_valid, ..., :cout] - want[:n_valid]).max() <= 2 ** -7 * scale
    keep = n_valid + (n_valid & 1 if plan.info.patch_fold == 2 else 0)
    assert np.all(raw[keep:] == 7.0)
    # same op through the generic epilogue (tuning.fast_mask = -1): same values up to the 16-bit rounding of the output
    plan = Plan([dict(src=0, in_h=H, in_w=W, kernel=kern, bias=bias)], cap, dtype=DT, tuning=dict(tuning, fast_mask=-1))
    out2 = torch.full((cap, H, W, npad), 7.0, dtype=_td(), device='cuda')
    plan.run([t(pad_channels(x, pad16(cin)))], n_valid, out_raw=out2, mask=t(mask), **kw)
    torch.cuda.synchronize()
    assert np.abs(out2.float().cpu().numpy()[:n_valid, ..., :cout] - want[:n_valid]).max() <= 2 ** -7 * scale


PHASE_MASK_CASES = [
    # k, H, W (input), cin, cout, n_valid, cap
    (3, 16, 16, 128, 64, 3, 4),        # merged column phases (2 * cout <= 256, unpadded rows)
    (5, 32, 32, 32, 3, 2, 2),          # arch B D_psi/level_0/conv_1^T: 3 output channels, four phases
    (3, 8, 8, 192, 192, 3, 4),         # image pairs, four phases
    (3, 32, 32, 64, 48, 2, 3),
    (3, 14, 14, 32, 32, 2, 3),         # arch M level_4/conv_1^T: 14 -> 28
    (3, 7, 7, 64, 32, 3, 4),           # level_6/conv_1^T: 7 -> 14, image pairs
]


@pytest.mark.parametrize('k,H,W,cin,cout,n_valid,cap', PHASE_MASK_CASES)
def test_stride2_deconv_phases_with_mask(k, H, W, cin, cout, n_valid, cap):
    """The adjoint of a stride-2 conv (model_grad.cu: conv_1^T of a D_psi block) = conv2d_transpose(stride 2) whose
    epilogue multiplies by leaky_relu'(forward activation): sub-pixel phase ops whose mask is the same strided view as
    their output, fetched by 4-D bulk tensor loads."""
    torch = _torch()
    from apvae.plan import Plan, pad16
    tuning = dict(patch=1)
    rs = np.random.RandomState(k + H + cin + cout)
    x = exact16(rs.normal(size=(cap, H, W, cin)))
    kern = exact16(rs.normal(size=(k, k, cout, cin)) / np.sqrt(k * k * cin / 4.0))
    plan = Plan([dict(src=0, transpose=True, stride=2, in_h=H, in_w=W, kernel=kern, bias=None)], cap, dtype=DT, tuning=tuning)
    npad = pad16(cout)
    t = lambda a: torch.tensor(a, device='cuda').to(_td())
    mask = exact16(rs.normal(size=(cap, 2 * H, 2 * W, npad)))
    out_raw = torch.full((cap, 2 * H, 2 * W, npad), 7.0, dtype=_td(), device='cuda')
    plan.run([t(pad_channels(x, pad16(cin)))], n_valid, out_raw=out_raw, mask=t(mask))
    torch.cuda.synchronize()
    want = O.deconv2d(x.astype(np.float64), kern.astype(np.float64), np.zeros(cout), 2) * np.where(mask[..., :cout] > 0, 1.0, 0.2)
    scale = max(1.0, float(np.abs(want).max()))
    raw = out_raw.float().cpu().numpy()
    assert np.abs(raw[:n_valid, ..., :cout] - want[:n_valid]).max() <= 2 ** -7 * scale
    keep = n_valid + (n_valid & 1 if W in (7, 8) else 0)
    assert np.all(raw[keep:] == 7.0)
