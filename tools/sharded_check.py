"""Multi-GPU check + timing of gpsig_b200.sharding.sharded_gram (run under torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tools/sharded_check.py [N]

Checks the all-gathered K(X, X2) and the mirrored symmetric K(X, X) against the single-GPU result and prints the
device time (max over ranks) of: rows only (no collective), gather after compute, tiled gather overlapped with compute.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpsig_b200 import kernels, sharding  # noqa: E402
from tests.util import brownian  # noqa: E402


def main():
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    dist.init_process_group('nccl', device_id=dev)
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
    L, d, m = 200, 8, 5
    X = torch.as_tensor(brownian(N, L, d, 0).reshape(N, -1), dtype=torch.float64, device=dev)
    X2 = torch.as_tensor(brownian(N, L, d, 1).reshape(N, -1), dtype=torch.float64, device=dev)
    k = kernels.SignatureRBF(L * d, d, m)
    res = {"N": N, "world": world}

    def timed(fn):
        fn()
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); out = fn(); e1.record(); e1.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return out, float(t.item())

    _, res["rows_only_ms"] = timed(lambda: sharding.sharded_gram(k, X, X2, gather=False))
    full1, res["gather_ms"] = timed(lambda: sharding.sharded_gram(k, X, X2))
    full4, res["gather_tiled4_ms"] = timed(lambda: sharding.sharded_gram(k, X, X2, tiles=4))
    sym, res["sym_gather_ms"] = timed(lambda: sharding.sharded_gram(k, X, None, tiles=2))
    if N <= 4000:                                    # single-GPU reference of the same matrices
        ref = k.K(X, X2)
        res["gram_equal"] = bool(torch.equal(full1, ref)) and bool(torch.equal(full4, ref))
        refs = k.K(X)
        res["sym_max_rel_diff"] = float((sym - refs).abs().max() / refs.abs().max())
        res["sym_exactly_symmetric"] = bool(torch.equal(sym, sym.T))
    res["pairs_per_s_rows_only"] = N * N / (res["rows_only_ms"] * 1e-3)
    res["pairs_per_s_gathered"] = N * N / (res["gather_tiled4_ms"] * 1e-3)
    res["delivered_pairs_per_s_sym"] = N * N / (res["sym_gather_ms"] * 1e-3)
    if rank == 0:
        print(json.dumps(res))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
