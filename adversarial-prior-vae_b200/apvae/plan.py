"""Layer-level interface: spt.layers.conv2d / deconv2d / dense on the tap-GEMM kernel.

Mirrors tfsnippet/layers/convolutional/conv2d_.py:23-238 (conv2d), :243-501
(deconv2d) and layers/core/dense_.py:14-171 (dense) for channels_last,
padding='same', square kernels and strides 1 or 2 — the only forms the
experiment scripts use.  Planning is host-only (works without a GPU); running
needs a B200.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import lib, check


def pad16(c):
    return (c + 15) // 16 * 16


class Plan(object):
    """A planned op: sum of conv terms sharing one output grid."""

    def __init__(self, terms, capacity_n, dtype='fp16', tuning=None):
        """dtype: the 16-bit operand format ('fp16' default, 'bf16'): kernels are packed in it, sources and 16-bit
        outputs hold it.  tuning: dict of apvae_tuning overrides (kernel variant selection; None = defaults)."""
        if dtype not in ('fp16', 'bf16'):
            raise ValueError("dtype must be 'fp16' or 'bf16'")
        self.dtype = dtype
        self._keep = []
        arr = (_lib.ConvTerm * len(terms))()
        for i, t in enumerate(terms):
            k = np.ascontiguousarray(t['kernel'], dtype=np.float32)
            self._keep.append(k)
            arr[i].src = t.get('src', i)
            arr[i].transpose = int(bool(t.get('transpose', False)))
            arr[i].k = int(k.shape[0])
            arr[i].stride = int(t.get('stride', 1))
            if arr[i].transpose:
                arr[i].cout, arr[i].cin = int(k.shape[2]), int(k.shape[3])
            else:
                arr[i].cin, arr[i].cout = int(k.shape[2]), int(k.shape[3])
            arr[i].in_h, arr[i].in_w = int(t['in_h']), int(t['in_w'])
            arr[i].in_cp = int(t.get('in_cp', pad16(arr[i].cin)))
            arr[i].kernel = k.ctypes.data_as(C.POINTER(C.c_float))
            if t.get('bias') is not None:
                b = np.ascontiguousarray(t['bias'], dtype=np.float32)
                self._keep.append(b)
                arr[i].bias = b.ctypes.data_as(C.POINTER(C.c_float))
        self._h = C.c_void_p()
        self._tuning = _lib.Tuning.make(tuning)
        check(lib.apvae_plan_create(arr, len(terms), int(capacity_n), int(dtype == 'bf16'), C.byref(self._tuning),
                                    C.byref(self._h)))
        self.info = _lib.PlanInfo()
        check(lib.apvae_plan_get_info(self._h, C.byref(self.info)))

    def __del__(self):
        if getattr(self, '_h', None):
            lib.apvae_plan_destroy(self._h)
            self._h = None

    @property
    def steps(self):
        n = self.info.total_steps
        out = np.zeros((n, 5), dtype=np.int32)
        check(lib.apvae_plan_copy_steps(self._h, out.ctypes.data_as(C.POINTER(C.c_int32))))
        return out

    @property
    def weights(self):
        i = self.info
        out = np.zeros((i.total_steps, i.npad, i.kc), dtype=np.uint16)
        check(lib.apvae_plan_copy_weights(self._h, out.ctypes.data_as(C.POINTER(C.c_uint16))))
        return out

    @property
    def patch_groups(self):
        out = np.zeros((self.info.patch_groups, 6), dtype=np.int32)
        check(lib.apvae_plan_copy_patch_groups(self._h, out.ctypes.data_as(C.POINTER(C.c_int32))))
        return out

    def select_patch(self, which):
        """-1: the op's own patch plan; 0..3: sub-pixel phase plans of a stride-2 transposed conv; 4..5: its row-phase
        plans with both column phases merged into N.  Refreshes `info`."""
        check(lib.apvae_plan_select_patch(self._h, int(which)))
        check(lib.apvae_plan_get_info(self._h, C.byref(self.info)))
        return self

    @property
    def patch_bias(self):
        out = np.zeros((self.info.patch_npad,), dtype=np.float32)
        check(lib.apvae_plan_copy_patch_bias(self._h, out.ctypes.data_as(C.POINTER(C.c_float))))
        return out

    @property
    def patch_weights(self):
        i = self.info
        out = np.zeros((i.patch_items, i.patch_npad or i.npad, i.kc), dtype=np.uint16)
        check(lib.apvae_plan_copy_patch_weights(self._h, out.ctypes.data_as(C.POINTER(C.c_uint16))))
        return out

    @property
    def bias(self):
        out = np.zeros((self.info.npad,), dtype=np.float32)
        check(lib.apvae_plan_copy_bias(self._h, out.ctypes.data_as(C.POINTER(C.c_float))))
        return out

    def run(self, srcs, n_valid, out_raw=None, out_act=None, out_f32=None, f32_cols=0,
            res=None, res_mode=_lib.RES_NONE, mask=None, tanh_cols=0, bf16_zero_from=0,
            device=0, bench_reps=0, out_raw_lo=None, out_act_lo=None, res_lo=None):
        """srcs: bf16 CUDA tensors [capacity_n, H, W, Cp]; outputs are CUDA tensors
        whose last dim is the row pitch."""
        ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
        e = _lib.Epilogue()
        e.out_raw_dev, e.out_act_dev = ptr(out_raw), ptr(out_act)
        ref = out_raw if out_raw is not None else out_act
        e.ldo = int(ref.shape[-1]) if ref is not None else 0
        e.bf16_zero_from = int(bf16_zero_from)
        e.out_f32_dev = ptr(out_f32)
        e.ld_f32 = int(out_f32.shape[-1]) if out_f32 is not None else 0
        e.f32_cols = int(f32_cols)
        e.res_dev, e.res_mode = ptr(res), int(res_mode if res is not None else 0)
        e.ld_res = int(res.shape[-1]) if res is not None else 0
        e.mask_dev = ptr(mask)
        e.ld_mask = int(mask.shape[-1]) if mask is not None else 0
        e.tanh_cols = int(tanh_cols)
        e.out_raw_lo_dev, e.out_act_lo_dev, e.res_lo_dev = ptr(out_raw_lo), ptr(out_act_lo), ptr(res_lo)   # split precision
        arr = (C.c_void_p * len(srcs))(*[t.data_ptr() for t in srcs])
        if bench_reps > 0:
            ms = C.c_float()
            check(lib.apvae_plan_bench(self._h, int(device), arr, int(n_valid), C.byref(e), int(bench_reps), C.byref(ms)))
            return ms.value
        check(lib.apvae_plan_run(self._h, int(device), arr, int(n_valid), C.byref(e)))


def torch_dtype(dtype):
    import torch
    return torch.float16 if dtype == 'fp16' else torch.bfloat16


def _to_nhwc_16(x, cp, dtype='fp16'):
    """float32 CUDA [N,H,W,C] -> 16-bit operand tensor [N,H,W,cp] (zero padded). torch is plumbing only."""
    import torch
    n, h, w, c = x.shape
    out = torch.zeros((n, h, w, cp), dtype=torch_dtype(dtype), device=x.device)
    out[..., :c] = x.to(out.dtype)
    return out


def _run_single(x, kernel, bias, strides, transpose, activation_fn, dtype='fp16'):
    import torch
    lead = tuple(x.shape[:-3])
    x4 = x.reshape((-1,) + tuple(x.shape[-3:])).contiguous()
    n, h, w, c = x4.shape
    kernel = np.asarray(kernel, dtype=np.float32)
    cout = kernel.shape[2] if transpose else kernel.shape[3]
    plan = Plan([dict(src=0, transpose=transpose, stride=strides, in_h=h, in_w=w, kernel=kernel,
                      bias=bias)], n, dtype=dtype)
    src = _to_nhwc_16(x4, pad16(c), dtype)
    out = torch.zeros((n, plan.info.ho, plan.info.wo, cout), dtype=torch.float32, device=x.device)
    if activation_fn not in (None, 'tanh'):
        raise ValueError('activation_fn must be None or "tanh"')
    plan.run([src], n, out_f32=out, f32_cols=cout,
             tanh_cols=cout if activation_fn == 'tanh' else 0, device=x.device.index or 0)
    return out.reshape(lead + tuple(out.shape[1:]))


def conv2d(x, kernel, bias=None, strides=1, activation_fn=None, dtype='fp16'):
    """spt.layers.conv2d(channels_last=True, padding='same'); x float32 CUDA [...,H,W,C],
    kernel [kh,kw,Cin,Cout] (conv2d_.py:23-238).  16-bit operands (`dtype`), fp32 accumulate."""
    return _run_single(x, kernel, bias, strides, False, activation_fn, dtype)


def deconv2d(x, kernel, bias=None, strides=1, activation_fn=None, dtype='fp16'):
    """spt.layers.deconv2d; kernel [kh,kw,Cout,Cin] (conv2d_.py:243-501)."""
    return _run_single(x, kernel, bias, strides, True, activation_fn, dtype)


def dense(x, kernel, bias=None, dtype='fp16'):
    """spt.layers.dense (dense_.py:14-171): x [..., in] @ kernel [in, out] + bias."""
    lead = tuple(x.shape[:-1])
    k = np.asarray(kernel, dtype=np.float32)
    x4 = x.reshape((-1, 1, 1, x.shape[-1]))
    y = _run_single(x4, k.reshape((1, 1) + k.shape), bias, 1, False, None, dtype)
    return y.reshape(lead + (k.shape[1],))
