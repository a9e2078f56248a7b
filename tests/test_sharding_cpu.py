"""Multi-process row sharding on CPU: gloo backend, world_size 2 (and 3, uneven blocks).

The per-rank compute is the CPU oracle (allowed in tests): the test checks the
partition + all-gather logic of gpsig_b200/sharding.py, i.e. that the row blocks
computed independently by the ranks assemble into the single-process Gram bit for bit.
"""
import importlib.util
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _sharding():
    # loaded by path: importing the gpsig_b200 package would require the CUDA library
    spec = importlib.util.spec_from_file_location('gpsig_sharding', os.path.join(ROOT, 'gpsig_b200', 'sharding.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_sym_row_ranges_balance_the_upper_triangle():
    sh = _sharding()
    for n in (8, 2000, 50001):
        for w in (1, 2, 4, 8):
            seen = np.zeros(n, dtype=int)
            work = []
            for r in range(w):
                tot = 0
                for b, e in sh.sym_row_ranges(n, w, r):
                    seen[b:e] += 1
                    tot += sum(n - i for i in range(b, e)) if n < 5000 else (e - b) * (2 * n - b - e + 1) // 2
                work.append(tot)
            assert (seen == 1).all()
            assert max(work) - min(work) <= 2 * n + 4          # equal shares up to one row per block


def test_row_ranges_partition():
    sh = _sharding()
    for n in (0, 1, 7, 8, 2000, 50001):
        for w in (1, 2, 3, 8):
            r = sh.all_row_ranges(n, w)
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            sizes = [e - b for b, e in r]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sh.row_range(10, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, return_levels, q):
    import sys
    sys.path.insert(0, ROOT)
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    sh = _sharding()
    from oracle.trunc_np import SignatureKernelOracle
    from tests.util import brownian
    L, d, m = 12, 3, 3
    X, Y = brownian(n, L, d, 0), brownian(5, L, d, 1)
    o = SignatureKernelOracle(L * d, d, m, base='rbf')

    class K:  # same call signature as gpsig_b200.kernels.SignatureKernel.K(..., rows=)
        @staticmethod
        def K(X_, X2_, return_levels=False, rows=None):
            b, e = rows if rows is not None else (0, n)
            full = o.K(X.reshape(n, -1)[b:e], Y.reshape(5, -1), return_levels=return_levels) if e > b else \
                np.zeros(((m + 1, 0, 5) if return_levels else (0, 5)))
            return torch.from_numpy(np.ascontiguousarray(full))

    Yt = torch.from_numpy(Y.reshape(5, -1))
    out = sh.sharded_gram(K, torch.from_numpy(X.reshape(n, -1)), Yt, return_levels=return_levels, gather=True)
    ref = o.K(X.reshape(n, -1), Y.reshape(5, -1), return_levels=return_levels)
    ok = out.shape == ref.shape and bool(np.array_equal(out.numpy(), ref))
    # tiled (collective overlapped with the next tile's compute): same matrix
    out2 = sh.sharded_gram(K, torch.from_numpy(X.reshape(n, -1)), Yt, return_levels=return_levels, gather=True, tiles=2)
    ok = ok and bool(np.array_equal(out2.numpy(), ref))

    class KS:  # symmetric: honours upper_only by POISONING the entries below the diagonal of its rows
        @staticmethod
        def K(X_, X2_, return_levels=False, rows=None, upper_only=False):
            assert X2_ is None
            b, e = rows if rows is not None else (0, n)
            full = o.K(X.reshape(n, -1), return_levels=return_levels)
            blk = np.array(full[..., b:e, :])
            if upper_only:
                for i in range(b, e):
                    blk[..., i - b, :i] = np.nan
            return torch.from_numpy(np.ascontiguousarray(blk))

    def mirror(full):
        a = full.numpy()
        iu = np.triu_indices(n, 1)
        a[..., iu[1], iu[0]] = a[..., iu[0], iu[1]]
        return full

    refs = o.K(X.reshape(n, -1), return_levels=return_levels)
    refs = np.triu(refs) + np.swapaxes(np.triu(refs, 1), -1, -2)
    for t in (1, 2):
        outs = sh.sharded_gram(KS, torch.from_numpy(X.reshape(n, -1)), None, return_levels=return_levels, gather=True,
                               tiles=t, mirror=mirror)
        ok = ok and outs.shape == refs.shape and bool(np.array_equal(outs.numpy(), refs))
    q.put((rank, ok))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n,return_levels", [(2, 8, False), (2, 7, True), (3, 8, False)])
def test_sharded_gram_gloo(world, n, return_levels):
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, return_levels, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(r for r, _ in res) == list(range(world))
    assert all(ok for _, ok in res)
