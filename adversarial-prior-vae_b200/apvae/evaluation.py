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

# metric name -> (accumulator slot, sign)   the eleven names of dmm_cifar10.py:1131-1136
METRIC_SLOTS = {
    'test_nll': (_lib.ROW_LL, -1.0),
    'test_lb': (_lib.ROW_LB, 1.0),
    'adv_test_nll': (_lib.ROW_ADV_LL, -1.0),
    'adv_test_lb': (_lib.ROW_ADV_LB, 1.0),
    'test_recon': (_lib.ROW_RECON, 1.0),
    'reconstruct_energy': (_lib.ROW_DG, 1.0),
    'real_energy': (_lib.ROW_REAL_E, 1.0),
    'test_mse': (_lib.ACCUM_MSE, 1.0),
}
# the test_pn_net branch (:883,915-927): batch scalars, size-weighted like every Evaluator metric
PRIOR_METRIC_SLOTS = {
    'pn_energy': _lib.ACCUM_PN_ENERGY,
    'pd_energy': _lib.ACCUM_PD_ENERGY,
    'kl_adv_and_gaussian': _lib.ACCUM_KL_ADV,
    'log_Z_compute_op': _lib.ACCUM_LOG_Z_OP,       # not in the Evaluator's dict in the reference; reported alongside
}


def shard_range(n, rank, world):
    """Contiguous ceil(n/world) slicing — the rule of the reference's (unused) tower helper,
    tfsnippet/examples/utils/multi_gpu.py:239-249.  Returns (start, stop)."""
    per = -(-n // world)
    lo = min(n, rank * per)
    return lo, min(n, lo + per)


def combine_accumulators(accum, group=None):
    """All ranks' accumulator vectors (float64 torch tensor [ACCUM_SLOTS], any device) -> the global one, on every
    rank, with ONE collective and no host synchronisation: an all-gather of the R x ACCUM_SLOTS doubles (NCCL over
    NVLink on GPUs), then every rank sums the additive slots and merges the log-Z (max, sum-exp) pairs by rescaling to
    the common max, all as device tensor ops."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return accum
    world = dist.get_world_size(group)
    mine = accum.contiguous()
    gathered = torch.empty((world, mine.numel()), dtype=mine.dtype, device=mine.device)
    try:
        dist.all_gather_into_tensor(gathered, mine, group=group)        # THE collective of an evaluation
    except (RuntimeError, NotImplementedError):                         # backends without the flat variant
        parts = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine, group=group)
        gathered = torch.stack(parts)
    return merge_accumulators(gathered)


def merge_accumulators(gathered):
    """[R, ACCUM_SLOTS] per-rank accumulators -> one [ACCUM_SLOTS] (pure tensor arithmetic, no host sync)."""
    import torch
    out = gathered.sum(dim=0)
    has = gathered[:, _lib.ACCUM_BATCHES] > 0                           # ranks that saw log-Z batches
    ninf = torch.full_like(gathered[:, 0], -float('inf'))
    gmax = torch.where(has, gathered[:, _lib.ROW_LZ_MAX], ninf).max()
    terms = torch.where(has, gathered[:, _lib.ROW_LZ_SUM] * torch.exp(gathered[:, _lib.ROW_LZ_MAX] - gmax),
                        torch.zeros_like(ninf))
    any_has = has.any()
    out[_lib.ROW_LZ_MAX] = torch.where(any_has, gmax, torch.zeros_like(gmax))
    out[_lib.ROW_LZ_SUM] = torch.where(any_has, terms.sum(), torch.zeros_like(gmax))
    return out


def metrics_from_accumulator(accum, len_train=None):
    """float64 accumulator -> dict of dataset metrics (the eleven names of dmm_cifar10.py:1131-1136; the prior-branch
    ones when that branch was accumulated) and, if log-Z batches were accumulated, 'log_Z' = logsumexp_b(lz_b) -
    log(#batches) (+ log len_train).  test_mse: the reference reshapes the per-image minima to (-1, test_x_samples)
    and averages twice (:893-895), which is the plain mean over images."""
    a = np.asarray(accum.detach().cpu().numpy() if hasattr(accum, 'detach') else accum, dtype=np.float64)
    w = a[_lib.ACCUM_WEIGHT]
    out = {}
    if w > 0:
        for name, (slot, sign) in METRIC_SLOTS.items():
            out[name] = sign * a[slot] / w
    pw = a[_lib.ACCUM_PRIOR_WEIGHT]
    if pw > 0:
        for name, slot in PRIOR_METRIC_SLOTS.items():
            out[name] = a[slot] / pw
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

    def __init__(self, model, n_z=None, batch_size=None, seed=0, group=None, time_metric_name='eval_time',
                 prior_branch=True, n_pz=None):
        """prior_branch: evaluate the test_pn_net branch (config.test_n_pz prior draws through G_theta and D_psi per
        mini-batch -> pn_energy, pd_energy, kl_adv_and_gaussian) with every batch, as the reference's graph does
        (dmm_cifar10.py:883,915-927); False leaves those three metrics out (throughput runs that say so)."""
        self.model = model
        self.prior_branch = bool(prior_branch)
        self.n_pz = int(n_pz if n_pz is not None else model.config.test_n_pz)
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

    def run_shard(self, x, x_origin=None, eps=None, accum=None, first_index=0, eps_prior=None):
        """Accumulate the metrics of images x [n,H,W,C] (already this rank's shard) into `accum`.
        eps_prior [n_batches, n_pz, z_dim]: fed-in prior draws of the test_pn_net branch (parity runs)."""
        import torch
        m = self.model
        dev = m.device
        if accum is None:
            accum = torch.zeros(_lib.ACCUM_SLOTS, dtype=torch.float64, device=dev)
        n = int(x.shape[0])
        pi = torch.empty((_lib.PER_IMAGE_ROWS, self.batch_size), dtype=torch.float32, device=dev)
        prior4 = torch.empty(4, dtype=torch.float32, device=dev) if self.prior_branch else None
        for ib, b0 in enumerate(range(0, n, self.batch_size)):
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
            if self.prior_branch:
                m.prior_branch(self.n_pz, eps=None if eps_prior is None else eps_prior[ib],
                               seed=self.seed + first_index + b0, out=prior4)
            check(lib.apvae_accumulate_metrics(m._h, _ptr(view), _ptr(prior4), B, self.n_z, _ptr(accum), _stream()), m._h)
        return accum

    def run(self, x, x_origin=None, eps=None, eps_prior=None):
        import torch
        rank, world = self._rank_world()
        lo, hi = shard_range(int(x.shape[0]), rank, world)
        t0 = time.time()
        accum = self.run_shard(x[lo:hi], None if x_origin is None else x_origin[lo:hi],
                               None if eps is None else eps[:, lo:hi], first_index=lo, eps_prior=eps_prior)
        accum = combine_accumulators(accum, self.group)
        self.last_accumulator = accum
        out = metrics_from_accumulator(accum, self.model.config.len_train)
        torch.cuda.synchronize(self.model.device)
        out[self.time_metric_name] = time.time() - t0
        return out

    def estimate_log_Z(self, x_train, x_origin=None, eps=None, set_model=True):
        """The log-Z loop (dmm_cifar10.py:1215-1227): another_log_Z_compute_op over the train set,
        logsumexp(list) - log(#batches); stores it like get_log_Z().set(...)."""
        keep, self.prior_branch = self.prior_branch, False        # the log-Z pass reads only the chain's energies
        try:
            out = self.run(x_train, x_origin, eps)
        finally:
            self.prior_branch = keep
        if set_model and 'log_Z' in out:
            m = self.model
            m.set_scalars(m.xi_raw, m.beta, out['log_Z'])
        return out.get('log_Z')
