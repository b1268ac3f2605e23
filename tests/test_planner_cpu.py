"""Host planner vs the oracle: the step table + TMA view geometry + packed
weights must reproduce conv2d / deconv2d / dense for every (k, stride) the
scripts use.  Runs without a GPU (planning is host-only)."""
import os

import numpy as np
import pytest

import oracle as O
from apvae.plan import Plan, pad16
from helpers import bf16_round, exact16, emulate_plan, pad_channels

CASES = [
    # transpose, k, stride, H, W, cin, cout
    (False, 3, 1, 8, 8, 16, 16),
    (False, 5, 1, 8, 8, 3, 24),
    (False, 3, 2, 8, 8, 32, 16),
    (False, 5, 2, 8, 8, 3, 8),
    (False, 1, 1, 4, 4, 64, 6),
    (False, 1, 2, 8, 8, 16, 32),
    (False, 3, 1, 7, 7, 20, 16),
    (False, 3, 2, 14, 14, 16, 16),
    (True, 3, 1, 8, 8, 16, 16),
    (True, 5, 1, 4, 4, 96, 16),
    (True, 3, 2, 4, 4, 16, 32),
    (True, 5, 2, 4, 4, 8, 16),
    (True, 1, 1, 8, 8, 2, 48),
    (True, 3, 2, 7, 7, 64, 16),
]


@pytest.mark.parametrize('transpose,k,stride,H,W,cin,cout', CASES)
def test_plan_matches_oracle(transpose, k, stride, H, W, cin, cout):
    rs = np.random.RandomState(k * 100 + stride * 10 + cin)
    n = 3
    x = exact16(rs.normal(size=(n, H, W, cin)))
    shp = (k, k, cout, cin) if transpose else (k, k, cin, cout)
    kern = exact16(rs.normal(size=shp) * 0.2)
    bias = rs.normal(size=(cout,)).astype(np.float32)
    plan = Plan([dict(src=0, transpose=transpose, stride=stride, in_h=H, in_w=W, kernel=kern, bias=bias)], n)
    got = emulate_plan(plan, [pad_channels(x, pad16(cin))], n)[..., :cout]
    fn = O.deconv2d if transpose else O.conv2d
    want = fn(x.astype(np.float64), kern.astype(np.float64), bias.astype(np.float64), stride)
    assert got.shape == want.shape
    np.testing.assert_allclose(got, want, rtol=1e-9, atol=1e-9)
    assert plan.info.macs_per_image > 0


def test_fused_shortcut_terms():
    """conv_1 (3x3, s2) + 1x1 s2 shortcut on another source accumulate in one op
    (resnet.py:225-265)."""
    rs = np.random.RandomState(7)
    n, H, cin, cout = 2, 8, 32, 64
    r = exact16(rs.normal(size=(n, H, H, cin)))
    x = exact16(rs.normal(size=(n, H, H, cin)))
    k1 = exact16(rs.normal(size=(3, 3, cin, cout)) * 0.1)
    ks = exact16(rs.normal(size=(1, 1, cin, cout)) * 0.1)
    b1, bs = rs.normal(size=cout).astype(np.float32), rs.normal(size=cout).astype(np.float32)
    plan = Plan([dict(src=0, stride=2, in_h=H, in_w=H, kernel=k1, bias=b1),
                 dict(src=1, stride=2, in_h=H, in_w=H, kernel=ks, bias=bs)], n)
    got = emulate_plan(plan, [r, x], n)[..., :cout]
    want = O.conv2d(r.astype(np.float64), k1.astype(np.float64), b1.astype(np.float64), 2) + \
        O.conv2d(x.astype(np.float64), ks.astype(np.float64), bs.astype(np.float64), 2)
    # the two biases are summed in float32 by the planner
    np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-6)


def test_dense_as_1x1():
    """dense == np.dot(x, k) + b (reference tests/layers/core/test_dense.py)."""
    rs = np.random.RandomState(3)
    n, fin, fout = 5, 4096, 64
    x = exact16(rs.normal(size=(n, 1, 1, fin)))
    k = exact16(rs.normal(size=(fin, fout)) * 0.05)
    b = rs.normal(size=fout).astype(np.float32)
    plan = Plan([dict(src=0, in_h=1, in_w=1, kernel=k.reshape(1, 1, fin, fout), bias=b)], n)
    assert plan.info.bn == 128 and plan.info.kc == 64
    got = emulate_plan(plan, [x], n)[:, 0, 0, :fout]
    np.testing.assert_allclose(got, x[:, 0, 0].astype(np.float64) @ k.astype(np.float64) + b, rtol=1e-9, atol=1e-9)


def test_wide_n_is_tiled():
    k = np.zeros((1, 1, 16, 3136), dtype=np.float32)
    plan = Plan([dict(src=0, in_h=1, in_w=1, kernel=k)], 4)
    assert plan.info.nt == 224 and plan.info.n_tiles == 14


def test_unsupported_shapes_raise():
    from apvae import ApvaeError
    with pytest.raises(ApvaeError):
        Plan([dict(src=0, stride=3, in_h=9, in_w=9, kernel=np.zeros((3, 3, 16, 16), np.float32))], 1)
    with pytest.raises(ApvaeError):     # stride-2 conv on an odd grid
        Plan([dict(src=0, stride=2, in_h=7, in_w=7, kernel=np.zeros((3, 3, 16, 16), np.float32))], 1)


PATCH_CASES = [
    # transpose, k, H, W, cin, cout
    (False, 3, 16, 16, 16, 16),
    (False, 5, 32, 32, 3, 24),
    (True, 3, 16, 16, 32, 16),
    (True, 5, 32, 32, 20, 16),
    (False, 1, 16, 16, 64, 48),
    (False, 3, 8, 16, 16, 200),      # one 128-pixel tile per image, wide N
    (True, 3, 64, 64, 16, 32),       # celebA width
    (False, 3, 8, 8, 32, 48),        # 8-wide images: two per tile row (fold)
    (True, 3, 8, 8, 16, 200),
    (False, 1, 16, 8, 16, 16),       # 8 wide, two row tiles per pair
    (False, 3, 28, 28, 16, 32),      # arch M: 28 wide in 32-wide tile rows (two 16-row tiles, the second hangs over by 4 rows)
    (True, 3, 28, 28, 32, 16),
    (False, 3, 14, 14, 32, 32),      # 14 in 16: one 16-row tile of two accumulators, rows 14 and 15 hang over
    (True, 3, 7, 7, 64, 64),         # 7 in 8, as image pairs: one 8-row tile per pair
    (False, 5, 14, 28, 16, 16),      # non-square, 5x5
]


@pytest.mark.parametrize('transpose,k,H,W,cin,cout', PATCH_CASES)
@pytest.mark.parametrize('macc', [0, 1])
def test_patch_plan_matches_oracle(transpose, k, H, W, cin, cout, macc):
    """The A-stationary plan (patchgemm.h): groups, packed weights and patch geometry reproduce
    conv2d / deconv2d (stride 1)."""
    from helpers import emulate_patch_plan
    tuning = dict(patch=1)
    tuning['patch_macc'] = macc
    rs = np.random.RandomState(k * 100 + cin + cout)
    n = 2
    x = exact16(rs.normal(size=(n, H, W, cin)))
    shp = (k, k, cout, cin) if transpose else (k, k, cin, cout)
    kern = exact16(rs.normal(size=shp) * 0.2)
    bias = rs.normal(size=(cout,)).astype(np.float32)
    plan = Plan([dict(src=0, transpose=transpose, stride=1, in_h=H, in_w=W, kernel=kern, bias=bias)], n, tuning=tuning)
    assert plan.info.patch == 1
    assert plan.info.patch_macc * plan.info.npad <= 256
    assert plan.info.patch_fold == (2 if W in (7, 8) else 1)
    assert plan.info.patch_wv == {7: 8, 14: 16, 28: 32}.get(W, W)
    got = emulate_patch_plan(plan, [pad_channels(x, pad16(cin))], n)[..., :cout]
    fn = O.deconv2d if transpose else O.conv2d
    want = fn(x.astype(np.float64), kern.astype(np.float64), bias.astype(np.float64), 1)
    np.testing.assert_allclose(got, want, rtol=1e-9, atol=1e-9)


S2_PATCH_CASES = [
    # k, H, W (input), cin, cout, with a fused 1x1 stride-2 shortcut term
    (3, 32, 32, 64, 128, True),      # arch A D_psi/level_3/conv_1
    (3, 16, 16, 32, 48, True),       # 8-wide output: image pairs
    (5, 64, 64, 3, 32, False),       # celebA first block
    (1, 32, 32, 16, 16, False),
    (3, 32, 64, 16, 24, False),      # non-square
    (3, 28, 28, 32, 32, True),       # arch M level_4: 28 -> 14 (14 in 16-wide tile rows)
    (3, 14, 14, 32, 64, True),       # arch M level_6: 14 -> 7 (7 in 8, image pairs)
]


@pytest.mark.parametrize('k,H,W,cin,cout,shortcut', S2_PATCH_CASES)
def test_patch_plan_stride2_conv_matches_oracle(k, H, W, cin, cout, shortcut):
    """Stride-2 conv2d on the A-stationary kernel: the input is read through its space-to-depth view, one tensor map
    per row phase, column phase folded into the channel coordinate; vertical taps of one row phase share a patch."""
    from helpers import emulate_patch_plan
    tuning = dict(patch=1)
    rs = np.random.RandomState(k * 100 + cin + cout)
    n = 2
    x = exact16(rs.normal(size=(n, H, W, cin)))
    kern = exact16(rs.normal(size=(k, k, cin, cout)) * 0.2)
    bias = rs.normal(size=(cout,)).astype(np.float32)
    terms = [dict(src=0, stride=2, in_h=H, in_w=W, kernel=kern, bias=bias)]
    srcs = [pad_channels(x, pad16(cin))]
    want = O.conv2d(x.astype(np.float64), kern.astype(np.float64), bias.astype(np.float64), 2)
    if shortcut:
        x2 = exact16(rs.normal(size=(n, H, W, cin)))
        ks = exact16(rs.normal(size=(1, 1, cin, cout)) * 0.2)
        bs = rs.normal(size=(cout,)).astype(np.float32)
        terms.append(dict(src=1, stride=2, in_h=H, in_w=W, kernel=ks, bias=bs))
        srcs.append(pad_channels(x2, pad16(cin)))
        want = want + O.conv2d(x2.astype(np.float64), ks.astype(np.float64), bs.astype(np.float64), 2)
    plan = Plan(terms, n, tuning=tuning)
    assert plan.info.patch == 1 and plan.info.patch_s2d == 1
    assert plan.info.patch_fold == (2 if W in (14, 16) else 1)
    assert plan.info.patch_nmaps == (1 if k == 1 else 2) + (1 if shortcut else 0)
    got = emulate_patch_plan(plan, srcs, n)[..., :cout]
    assert got.shape == want.shape
    np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-6)


PHASE_PLAN_CASES = [
    # k, H, W (input), cin, cout
    (3, 16, 16, 32, 48),       # 2 * cout <= 256, cout % 16 == 0: row-phase plans exist
    (5, 16, 16, 16, 16),
    (3, 8, 8, 32, 32),         # 8-wide input: image pairs
    (3, 16, 16, 16, 3),        # cout % 16 != 0: four sub-pixel phases only
    (3, 16, 32, 16, 192),      # 2 * cout > 256: four sub-pixel phases only
    (3, 7, 7, 64, 64),         # arch M G_theta/level_3: 7 -> 14 (7 in 8, image pairs)
    (3, 14, 14, 64, 32),       # arch M G_theta/level_5: 14 -> 28
]


@pytest.mark.parametrize('k,H,W,cin,cout', PHASE_PLAN_CASES)
def test_stride2_deconv_phase_plans_match_oracle(k, H, W, cin, cout):
    """conv2d_transpose(stride 2, SAME) (conv2d_.py:464-471) assembled from its sub-pixel phase plans: four plans (py, px),
    or two row-phase plans whose N is [px 0 | px 1] x cout — each emulated exactly as the kernel walks it."""
    from helpers import emulate_patch_plan
    tuning = dict(patch=1)
    rs = np.random.RandomState(k * 100 + cin + cout)
    n = 2
    x = exact16(rs.normal(size=(n, H, W, cin)))
    kern = exact16(rs.normal(size=(k, k, cout, cin)) * 0.2)
    bias = rs.normal(size=(cout,)).astype(np.float32)
    plan = Plan([dict(src=0, transpose=True, stride=2, in_h=H, in_w=W, kernel=kern, bias=bias)], n, tuning=tuning)
    want = O.deconv2d(x.astype(np.float64), kern.astype(np.float64), bias.astype(np.float64), 2)
    src = [pad_channels(x, pad16(cin))]
    got = np.zeros_like(want)
    for ph in range(4):
        plan.select_patch(ph)
        assert plan.info.patch_phase == ph and plan.info.patch_out_scale == 2
        got += emulate_patch_plan(plan, src, n)
    np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-6)
    merged_ok = cout % 16 == 0 and 2 * cout <= 256
    if merged_ok:
        got = np.zeros_like(want)
        for py in range(2):
            plan.select_patch(4 + py)
            assert plan.info.patch_npad == 2 * cout
            got += emulate_patch_plan(plan, src, n)
        np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-6)
    else:
        with pytest.raises(Exception):
            plan.select_patch(4)


def test_patch_plan_fused_shortcut_and_eligibility():
    from helpers import emulate_patch_plan
    tuning = dict(patch=1)
    rs = np.random.RandomState(11)
    n, H, cin, cout = 2, 16, 32, 64
    r = exact16(rs.normal(size=(n, H, H, cin)))
    x = exact16(rs.normal(size=(n, H, H, cin)))
    k1 = exact16(rs.normal(size=(3, 3, cin, cout)) * 0.1)
    ks = exact16(rs.normal(size=(1, 1, cin, cout)) * 0.1)
    b1, bs = rs.normal(size=cout).astype(np.float32), rs.normal(size=cout).astype(np.float32)
    plan = Plan([dict(src=0, in_h=H, in_w=H, kernel=k1, bias=b1), dict(src=1, in_h=H, in_w=H, kernel=ks, bias=bs)], n, tuning=tuning)
    assert plan.info.patch == 1 and plan.info.patch_prow[1] == plan.info.patch_rows_tile
    got = emulate_patch_plan(plan, [r, x], n)[..., :cout]
    want = O.conv2d(r.astype(np.float64), k1.astype(np.float64), b1.astype(np.float64), 1) + \
        O.conv2d(x.astype(np.float64), ks.astype(np.float64), bs.astype(np.float64), 1)
    np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-6)
    # not eligible: stride 2 onto a 4x4 grid, 8x8 grids with an odd capacity (they run as image pairs), widths more than
    # 1/8 below a power of two, and the arch-M widths with tuning.patch_virtual = -1
    assert Plan([dict(src=0, kernel=np.zeros((3, 3, 16, 16), np.float32), in_h=8, in_w=8)], 4).info.patch_fold == 2
    assert Plan([dict(src=0, kernel=np.zeros((3, 3, 16, 16), np.float32), in_h=28, in_w=28)], 1, tuning=tuning).info.patch_wv == 32
    for kw in (dict(stride=2, in_h=8, in_w=8), dict(stride=2, in_h=16, in_w=16), dict(in_h=8, in_w=8), dict(in_h=24, in_w=24),
               dict(in_h=7, in_w=7), dict(in_h=28, in_w=28, virt=-1)):
        tn = dict(tuning)
        if kw.pop('virt', 0):
            tn['patch_virtual'] = -1
        p = Plan([dict(src=0, kernel=np.zeros((3, 3, 16, 16), np.float32), **kw)], 1, tuning=tn)
        assert p.info.patch == 0
    tuning['patch'] = -1
    p = Plan([dict(src=0, in_h=16, in_w=16, kernel=np.zeros((3, 3, 16, 16), np.float32))], 1, tuning=tuning)
    assert p.info.patch == 0


def test_patch_plan_takes_the_products_of_a_split_precision_op():
    """A split-precision op is two or three terms over TWO sources (hi and lo activations): A_hi W_hi + A_lo W_hi
    [+ A_hi W_lo] (model.cu push_terms).  The A-stationary plan gives every term its own activation map (up to four),
    also for the sub-pixel phases of a stride-2 transposed conv."""
    from helpers import emulate_patch_plan, fp16_round
    tuning = dict(patch=1)
    rs = np.random.RandomState(17)
    n, H, cin, cout = 2, 16, 32, 48
    a = rs.normal(size=(n, H, H, cin)).astype(np.float32)
    a_hi = fp16_round(a)
    a_lo = fp16_round(a - a_hi)
    w = (rs.normal(size=(3, 3, cin, cout)) * 0.1).astype(np.float32)
    w_hi = fp16_round(w)
    w_lo = fp16_round(w - w_hi)
    b = rs.normal(size=cout).astype(np.float32)
    for nterms in (2, 3):
        terms = [dict(src=0, in_h=H, in_w=H, kernel=w_hi, bias=b), dict(src=1, in_h=H, in_w=H, kernel=w_hi)]
        if nterms == 3:
            terms.append(dict(src=0, in_h=H, in_w=H, kernel=w_lo))
        plan = Plan(terms, n, tuning=tuning)
        assert plan.info.patch == 1 and plan.info.patch_nmaps == nterms and plan.info.nsrc == 2
        assert list(plan.info.patch_map_src)[:nterms] == [0, 1, 0][:nterms]
        got = emulate_patch_plan(plan, [a_hi, a_lo], n)[..., :cout]
        act = (a_hi.astype(np.float64) + a_lo)
        want = O.conv2d(act, w_hi.astype(np.float64), b.astype(np.float64), 1)
        if nterms == 3:
            want = want + O.conv2d(a_hi.astype(np.float64), w_lo.astype(np.float64), None, 1)
        np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-6)
    # stride-2 transposed conv: the phase plans take both products as well
    wt = (rs.normal(size=(3, 3, cout, cin)) * 0.1).astype(np.float32)
    wt_hi = fp16_round(wt)
    plan = Plan([dict(src=0, transpose=True, stride=2, in_h=H, in_w=H, kernel=wt_hi, bias=b),
                 dict(src=1, transpose=True, stride=2, in_h=H, in_w=H, kernel=wt_hi)], n, tuning=tuning)
    want = O.deconv2d(a_hi.astype(np.float64) + a_lo, wt_hi.astype(np.float64), b.astype(np.float64), 2)
    got = np.zeros_like(want)
    for ph in range(4):
        plan.select_patch(ph)
        assert plan.info.patch_nmaps == 2
        got += emulate_patch_plan(plan, [a_hi, a_lo], n)[..., :cout]
    np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-6)


def test_python_tuning_struct_matches_the_header():
    """apvae/_lib.py's Tuning fields, in order, are the int fields of `struct apvae_tuning` in include/apvae.h."""
    import re
    from apvae import _lib
    text = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'include', 'apvae.h')).read()
    body = re.search(r'typedef struct apvae_tuning \{(.*?)\} apvae_tuning;', text, re.S).group(1)
    body = re.sub(r'/\*.*?\*/', '', body, flags=re.S)
    fields = []
    for decl in re.findall(r'int\s+([^;]+);', body):
        fields += [f.strip() for f in decl.split(',')]
    assert tuple(fields) == tuple(_lib.TUNING_FIELDS)
    assert _lib.Tuning.make(dict(patch=-1, patch_split=1)).patch == -1
    with pytest.raises(AttributeError):
        _lib.Tuning.make(dict(no_such_knob=1))


def test_planner_picks_the_second_epilogue_warpgroup_for_epilogue_bound_tiles():
    """PatchPlan::epi2 (planner.cpp): two epilogue warpgroups when the estimated epilogue cycles of a tile exceed its MMA
    cycles, unless the wider staging would cost the op an accumulator or (large patches) a patch stage; tuning.patch_epi2
    forces it either way.  Also the weight-item rule of the 192-wide pair layers: two taps per item (four 24 KB weight slots
    and three patch stages instead of six one-tap slots and four)."""
    def plan(transpose, k, H, W, cin, cout, stride=1, tuning=None, which=None):
        shp = (k, k, cout, cin) if transpose else (k, k, cin, cout)
        p = Plan([dict(src=0, transpose=transpose, stride=stride, in_h=H, in_w=W, kernel=np.zeros(shp, np.float32))], 7400, tuning=tuning)
        if which is not None:
            p.select_patch(which)
        return p.info
    # epilogue-bound: narrow layers, 1x1 shortcuts, the 96-wide 3x3
    for args in ((False, 3, 32, 32, 16, 16), (False, 3, 32, 32, 16, 32), (False, 3, 32, 32, 32, 32), (True, 1, 16, 16, 192, 96),
                 (True, 3, 32, 32, 96, 96), (True, 3, 8, 8, 2, 192)):
        assert plan(*args).patch_epi2 == 1, args
        assert plan(*args, tuning=dict(patch_epi2=-1)).patch_epi2 == 0
    # MMA-bound, or not worth an accumulator / a patch stage: one warpgroup
    for args in ((True, 3, 16, 16, 192, 192), (True, 5, 32, 32, 96, 96), (False, 3, 32, 32, 64, 64), (False, 3, 32, 32, 64, 128, 2),
                 (False, 3, 16, 16, 128, 128), (True, 5, 64, 64, 32, 32)):
        assert plan(*args).patch_epi2 == 0, args
    assert plan(False, 3, 32, 32, 64, 64).patch_macc == 4 and plan(False, 3, 32, 32, 64, 64, tuning=dict(patch_epi2=1)).patch_macc == 2
    # 192-wide pair layers: four two-tap weight slots
    i = plan(True, 3, 16, 16, 192, 192)
    assert (i.patch_cluster, i.patch_a_stages, i.patch_b_stages) == (2, 3, 4)
    j = plan(True, 3, 16, 16, 192, 192, tuning=dict(patch_tpi=1))
    assert (j.patch_a_stages, j.patch_b_stages) == (4, 6)


def test_every_arch_m_conv_plans_onto_the_patch_kernel():
    """dmm_bernouli_fashion.py:444-448, 510-514, 561-565: every conv / deconv of arch M's q_net, G_theta and D_psi
    (28 / 14 / 7-wide grids, 16 - 64 channels) has an A-stationary plan through the virtual widths (even capacity: the
    7-wide grids run as image pairs); stride-2 transposed convs as merged row-phase plans."""
    z = lambda *s: np.zeros(s, np.float32)
    convs = [
        # transpose, k, stride, H (= W), cin, cout, with a 1x1 shortcut term of the same stride
        (False, 3, 1, 28, 16, 16, False), (False, 3, 1, 28, 16, 32, True),        # level_2
        (False, 3, 1, 28, 32, 32, False),                                         # level_3, level_4/conv_0
        (False, 3, 2, 28, 32, 32, True),                                          # level_4/conv_1: 28 -> 14
        (False, 3, 1, 14, 32, 32, False), (False, 3, 2, 14, 32, 64, True),        # level_6: 14 -> 7
        (True, 3, 1, 14, 64, 64, False),                                          # G level_3/conv_1
        (True, 3, 1, 28, 32, 32, False),                                          # G level_5/conv_1, level_6, level_7
        (True, 3, 1, 28, 32, 16, False), (True, 3, 1, 28, 16, 16, False),         # G level_8
        (True, 1, 1, 28, 32, 16, False),                                          # G level_8/shortcut
        (True, 1, 1, 7, 64, 64, False), (True, 1, 1, 14, 64, 32, False),          # half-resolution shortcuts of level_3 / level_5
    ]
    for tr, k, s, H, cin, cout, sc in convs:
        shp = (k, k, cout, cin) if tr else (k, k, cin, cout)
        terms = [dict(src=0, transpose=tr, stride=s, in_h=H, in_w=H, kernel=z(*shp))]
        if sc:
            terms.append(dict(src=1, stride=s, in_h=H, in_w=H, kernel=z(1, 1, cin, cout)))
        i = Plan(terms, 64).info
        Wg = H // s
        assert i.patch == 1 and i.patch_wv == {7: 8, 14: 16, 28: 32}[Wg], (tr, k, s, H, cin, cout)
        assert i.patch_fold == (2 if Wg == 7 else 1)
        assert i.patch_rows_tile * i.patch_wv * i.patch_fold == 128 * i.patch_macc
    for H, cin, cout in ((7, 64, 64), (14, 64, 32)):                              # G level_3 / level_5 conv_0: stride-2 transposed
        plan = Plan([dict(src=0, transpose=True, stride=2, in_h=H, in_w=H, kernel=z(3, 3, cout, cin))], 64)
        plan.select_patch(4)
        assert plan.info.patch_phase == 4 and plan.info.patch_npad == 2 * cout and plan.info.patch_wv == {7: 8, 14: 16}[H]
    # an odd capacity keeps the 7-wide grids on the tap-GEMM kernel (pairs need whole pairs in the buffers)
    assert Plan([dict(src=0, in_h=7, in_w=7, kernel=z(3, 3, 64, 64))], 5).info.patch == 0
