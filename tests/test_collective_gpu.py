"""apvae_allreduce_metrics / apvae_merge_metrics: the one collective of a sharded evaluation on the C ABI.

* the device-side merge against the host-side definition (sum of the additive slots, log-sum-exp merge of the log-Z
  (max, sum-exp) pair over the ranks that saw log-Z batches) — one GPU;
* the NCCL path end to end on two GPUs (skipped on a one-GPU box): each rank builds its own ncclComm_t through ctypes
  (ncclGetUniqueId on rank 0, shared through a file; ncclCommInitRank), fills a different accumulator, calls
  apvae_allreduce_metrics through the C ABI and must end with the merged vector — the same one on both ranks.
The gloo / world_size-2 CPU coverage of the Python combine is tests/test_evaluation_cpu.py::test_two_rank_combine_over_gloo.
"""
import ctypes as C
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _p(t):
    return C.c_void_p(t.data_ptr())


def _accumulators(R, seed=0, logz_ranks=None):
    from apvae import _lib
    rs = np.random.RandomState(seed)
    a = rs.normal(size=(R, _lib.ACCUM_SLOTS)) * 100.0
    a[:, _lib.ACCUM_WEIGHT] = rs.randint(1, 50, size=R)
    a[:, _lib.ACCUM_BATCHES] = rs.randint(1, 9, size=R)
    a[:, _lib.ROW_LZ_MAX] = rs.normal(size=R) * 40.0 - 300.0
    a[:, _lib.ROW_LZ_SUM] = rs.uniform(1.0, 5.0, size=R)
    if logz_ranks is not None:                                 # ranks outside the list saw no log-Z batch
        for r in range(R):
            if r not in logz_ranks:
                a[r, _lib.ACCUM_BATCHES] = 0.0
                a[r, _lib.ROW_LZ_MAX] = 0.0
                a[r, _lib.ROW_LZ_SUM] = 0.0
    return a


def _merge_host(a):
    from apvae import _lib
    out = a.sum(axis=0)
    has = a[:, _lib.ACCUM_BATCHES] > 0
    if has.any():
        gmax = a[has, _lib.ROW_LZ_MAX].max()
        out[_lib.ROW_LZ_MAX] = gmax
        out[_lib.ROW_LZ_SUM] = (a[has, _lib.ROW_LZ_SUM] * np.exp(a[has, _lib.ROW_LZ_MAX] - gmax)).sum()
    else:
        out[_lib.ROW_LZ_MAX] = out[_lib.ROW_LZ_SUM] = 0.0
    return out


@pytest.mark.parametrize('R,logz', [(1, None), (2, None), (8, None), (8, [1, 6]), (4, [])])
def test_merge_metrics_matches_the_host_definition(R, logz):
    import torch
    from apvae import _lib
    from apvae._lib import lib, check
    from apvae.evaluation import merge_accumulators
    a = _accumulators(R, seed=R, logz_ranks=logz)
    g = torch.tensor(a, dtype=torch.float64, device='cuda')
    out = torch.full((_lib.ACCUM_SLOTS,), 7.0, dtype=torch.float64, device='cuda')
    check(lib.apvae_merge_metrics(_p(g), R, _p(out), None))
    torch.cuda.synchronize()
    want = _merge_host(a)
    np.testing.assert_allclose(out.cpu().numpy(), want, rtol=1e-12, atol=1e-12)
    # and the torch.distributed path's merge (the Python host side) is the same arithmetic
    np.testing.assert_allclose(merge_accumulators(g).cpu().numpy(), want, rtol=1e-12, atol=1e-12)


def _nccl():
    import torch  # noqa: F401  (maps torch's bundled libnccl.so.2 first, so the soname resolves to the same copy)
    return C.CDLL('libnccl.so.2')


class _UniqueId(C.Structure):
    _fields_ = [('internal', C.c_byte * 128)]


def _rank_main(rank, world, tmpdir):
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import conftest  # noqa: F401  (sys.path of the repo and the package)
    import time
    import torch
    from apvae import _lib
    from apvae._lib import lib, check
    torch.cuda.set_device(rank)
    nccl = _nccl()
    uid = _UniqueId()
    idf = os.path.join(tmpdir, 'nccl_id')
    if rank == 0:
        assert nccl.ncclGetUniqueId(C.byref(uid)) == 0
        with open(idf + '.tmp', 'wb') as f:
            f.write(bytes(uid.internal))
        os.rename(idf + '.tmp', idf)
    else:
        for _ in range(600):
            if os.path.exists(idf):
                break
            time.sleep(0.1)
        C.memmove(C.byref(uid), open(idf, 'rb').read(), 128)
    comm = C.c_void_p()
    nccl.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, _UniqueId, C.c_int]
    assert nccl.ncclCommInitRank(C.byref(comm), world, uid, rank) == 0
    a = _accumulators(world, seed=42)
    mine = torch.tensor(a[rank], dtype=torch.float64, device='cuda')
    gather = torch.empty(world * _lib.ACCUM_SLOTS, dtype=torch.float64, device='cuda')
    for _ in range(2):                                          # twice: the second call merges the merged vector again
        mine.copy_(torch.tensor(a[rank], dtype=torch.float64))
        check(lib.apvae_allreduce_metrics(comm, world, _p(mine), _p(gather), None))
        torch.cuda.synchronize()
    np.save(os.path.join(tmpdir, 'out%d.npy' % rank), mine.cpu().numpy())
    nccl.ncclCommDestroy.argtypes = [C.c_void_p]
    nccl.ncclCommDestroy(comm)


def test_allreduce_metrics_over_nccl_two_ranks(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs (run with gpurun --gpus 2)')
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_rank_main, args=(world, str(tmp_path)), nprocs=world, join=True)
    want = _merge_host(_accumulators(world, seed=42))
    for r in range(world):
        got = np.load(str(tmp_path / ('out%d.npy' % r)))
        np.testing.assert_allclose(got, want, rtol=1e-12, atol=1e-12)


def test_allreduce_metrics_single_rank_is_a_no_op_and_bad_arguments_are_loud():
    import torch
    from apvae import _lib
    from apvae._lib import lib, last_error
    a = torch.arange(_lib.ACCUM_SLOTS, dtype=torch.float64, device='cuda')
    assert lib.apvae_allreduce_metrics(None, 1, _p(a), None, None) == _lib.OK
    assert a.cpu().tolist() == list(range(_lib.ACCUM_SLOTS))
    assert lib.apvae_allreduce_metrics(None, 2, _p(a), None, None) == _lib.ERR_INVALID
    assert 'allreduce_metrics' in last_error(None)
