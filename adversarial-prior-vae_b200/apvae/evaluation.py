"""Evaluation driver: the host loop of spt.Evaluator.run (tfsnippet/trainer/evaluator.py:193-259)
plus the log-Z loop (dmm_cifar10.py:1215-1227), sharded over ranks.

Reference: per mini-batch one session.run of the metric scalars, then np.average weighted by
batch size (evaluator.py:17-31,249-254).  Here every batch is ONE fused C call
(apvae_is_nll) + ONE accumulation kernel; per-shard sums stay on the device and ranks are
combined with a single sum-allreduce (NCCL over NVLink on GPUs).  With the per-image mean as
batch metric and batch weight = batch size, the weighted mean of batch means equals the mean over
images, which is what the summed accumulators give.
"""
import math
import time

import numpy as np

from . import _lib
from ._lib import lib, check
from .model import _ptr, _stream

# metric name -> (accumulator slot, sign)   names as in dmm_cifar10.py:1131-1136
METRIC_SLOTS = {
    'test_nll': (_lib.ROW_LL, -1.0),
    'test_lb': (_lib.ROW_LB, 1.0),
    'adv_test_nll': (_lib.ROW_ADV_LL, -1.0),
    'adv_test_lb': (_lib.ROW_ADV_LB, 1.0),
    'test_recon': (_lib.ROW_RECON, 1.0),
    'reconstruct_energy': (_lib.ROW_DG, 1.0),
    'real_energy': (_lib.ROW_REAL_E, 1.0),
}
_SUM_SLOTS = [0, 1, 2, 3, 4, 5, 8, _lib.ACCUM_WEIGHT, _lib.ACCUM_BATCHES]


def shard_range(n, rank, world):
    """Contiguous ceil(n/world) slicing — the rule of the reference's (unused) tower helper,
    tfsnippet/examples/utils/multi_gpu.py:239-249.  Returns (start, stop)."""
    per = -(-n // world)
    lo = min(n, rank * per)
    return lo, min(n, lo + per)


def combine_accumulators(accum, group=None):
    """All ranks' accumulator vectors (float64 torch tensor [ACCUM_SLOTS], any device) -> global one.
    ONE sum-allreduce carries every additive slot; the log-Z (max, sum-exp) pair is made additive
    first by rescaling to a max that travels in a tiny max-allreduce only when log-Z batches exist."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return accum
    out = accum.clone()
    gmax = out[_lib.ROW_LZ_MAX:_lib.ROW_LZ_MAX + 1].clone()
    if float(out[_lib.ACCUM_BATCHES]) == 0:
        gmax.fill_(-float('inf'))
    dist.all_reduce(gmax, op=dist.ReduceOp.MAX, group=group)
    if math.isfinite(float(gmax)):
        scale = torch.exp(out[_lib.ROW_LZ_MAX] - gmax[0]) if float(out[_lib.ACCUM_BATCHES]) > 0 else out.new_zeros(())
        out[_lib.ROW_LZ_SUM] = out[_lib.ROW_LZ_SUM] * scale
    else:
        out[_lib.ROW_LZ_SUM] = 0
    out[_lib.ROW_LZ_MAX] = 0
    dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)       # THE allreduce of the metric sums
    out[_lib.ROW_LZ_MAX] = gmax[0]
    return out


def metrics_from_accumulator(accum, len_train=None):
    """float64 accumulator -> dict of dataset metrics (names of dmm_cifar10.py:1131-1136) and,
    if log-Z batches were accumulated, 'log_Z' = logsumexp_b(lz_b) - log(#batches) (+ log len_train)."""
    a = np.asarray(accum.detach().cpu().numpy() if hasattr(accum, 'detach') else accum, dtype=np.float64)
    w = a[_lib.ACCUM_WEIGHT]
    out = {}
    if w > 0:
        for name, (slot, sign) in METRIC_SLOTS.items():
            out[name] = sign * a[slot] / w
    nb = a[_lib.ACCUM_BATCHES]
    if nb > 0 and a[_lib.ROW_LZ_SUM] > 0:
        out['log_Z'] = a[_lib.ROW_LZ_MAX] + math.log(a[_lib.ROW_LZ_SUM]) - math.log(nb)
        if len_train:
            out['log_Z'] += math.log(len_train)
    return out


class Evaluator(object):
    """spt.Evaluator-shaped driver of the IS-NLL test graph.

    Evaluator(model, n_z=K, batch_size=B).run(x[, x_origin]) evaluates THIS rank's shard of the
    N images (contiguous ceil(N/R) slicing), combines ranks with one allreduce and returns the
    metric dict (every rank gets the same values).  `x` may be a CUDA tensor (resident data) or a
    host array / pinned tensor (copied batch by batch inside the loop, like feed_dict)."""

    def __init__(self, model, n_z=None, batch_size=None, seed=0, group=None, time_metric_name='eval_time'):
        self.model = model
        self.n_z = int(n_z if n_z is not None else model.config.test_n_qz)
        self.batch_size = int(batch_size if batch_size is not None else model.config.test_batch_size)
        if self.batch_size > model.config.max_batch:
            raise ValueError('batch_size exceeds the model capacity max_batch')
        self.seed = int(seed)
        self.group = group
        self.time_metric_name = time_metric_name
        self.last_accumulator = None

    def _rank_world(self):
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(self.group), dist.get_world_size(self.group)
        return 0, 1

    def run_shard(self, x, x_origin=None, eps=None, accum=None, first_index=0):
        """Accumulate the metrics of images x [n,H,W,C] (already this rank's shard) into `accum`."""
        import torch
        m = self.model
        dev = m.device
        if accum is None:
            accum = torch.zeros(_lib.ACCUM_SLOTS, dtype=torch.float64, device=dev)
        n = int(x.shape[0])
        pi = torch.empty((_lib.PER_IMAGE_ROWS, self.batch_size), dtype=torch.float32, device=dev)
        for b0 in range(0, n, self.batch_size):
            b1 = min(n, b0 + self.batch_size)
            B = b1 - b0
            xb = m._images(x[b0:b1] if isinstance(x, torch.Tensor) else np.asarray(x[b0:b1], dtype=np.float32))
            xo = None
            if x_origin is not None:
                xo = m._images(x_origin[b0:b1] if isinstance(x_origin, torch.Tensor)
                               else np.asarray(x_origin[b0:b1], dtype=np.float32))
            eb = m._f32(eps[:, b0:b1]) if eps is not None else None
            view = pi.reshape(-1)[:_lib.PER_IMAGE_ROWS * B]
            check(lib.apvae_is_nll(m._h, _ptr(xb), _ptr(xo), B, self.n_z, _ptr(eb),
                                   self.seed + first_index + b0, _ptr(view), None, _stream()), m._h)
            check(lib.apvae_accumulate_metrics(m._h, _ptr(view), B, self.n_z, _ptr(accum), _stream()), m._h)
        return accum

    def run(self, x, x_origin=None, eps=None):
        import torch
        rank, world = self._rank_world()
        lo, hi = shard_range(int(x.shape[0]), rank, world)
        t0 = time.time()
        accum = self.run_shard(x[lo:hi], None if x_origin is None else x_origin[lo:hi],
                               None if eps is None else eps[:, lo:hi], first_index=lo)
        accum = combine_accumulators(accum, self.group)
        self.last_accumulator = accum
        out = metrics_from_accumulator(accum, self.model.config.len_train)
        torch.cuda.synchronize(self.model.device)
        out[self.time_metric_name] = time.time() - t0
        return out

    def estimate_log_Z(self, x_train, x_origin=None, eps=None, set_model=True):
        """The log-Z loop (dmm_cifar10.py:1215-1227): another_log_Z_compute_op over the train set,
        logsumexp(list) - log(#batches); stores it like get_log_Z().set(...)."""
        out = self.run(x_train, x_origin, eps)
        if set_model and 'log_Z' in out:
            m = self.model
            m.set_scalars(m.xi_raw, m.beta, out['log_Z'])
        return out.get('log_Z')
