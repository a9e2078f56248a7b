"""BASELINE config c5 at its REAL size: SignatureRBF m=5, normalised K(X, X), N = 50 000, L = 200, d = 8, float64,
row-sharded over the GPUs of one box with the NCCL exchange of the finished rows (run under torchrun):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 \
        tools/c5_full.py [N]

gpsig_b200.sharding.sharded_gram with X2 = None: only the entries on and above the diagonal are computed (mirrored row
blocks, every rank the same number of pairs), the full [N, N] matrix (20 GB) is assembled on every rank and mirrored.
Device time = max over ranks (CUDA events).  Parity: a seeded sub-block against the CPU oracle (C port), the exact
symmetry of a band, and bit-equality of a block with a single-rank compute of the same rows.
"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpsig_b200 import kernels, sharding  # noqa: E402
from tests.util import brownian  # noqa: E402


def main():
    rank, world, local = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('LOCAL_RANK', 0))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
    L, d, m = 200, 8, 5
    Xn = brownian(N, L, d, 0)
    X = torch.as_tensor(Xn.reshape(N, -1), dtype=torch.float64, device=dev)
    k = kernels.SignatureRBF(L * d, d, m)
    full = torch.empty((N, N), dtype=torch.float64, device=dev)
    # warm-up on a small problem (NCCL communicator, module load)
    sharding.sharded_gram(k, X[:512 * world].contiguous(), None)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    K = sharding.sharded_gram(k, X, None, out=full)
    e1.record()
    e1.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    res = {"workload": "c5: SignatureRBF m=5 normalised K(X, X), N=%d, L=200, d=8, float64, %d GPU(s)" % (N, world),
           "ms": ms, "delivered_pairs_per_s": N * N / (ms * 1e-3), "computed_pairs_per_s": (N * (N + 1) / 2) / (ms * 1e-3),
           "matrix_bytes": N * N * 8, "collective": "one broadcast per owner and mirrored row block (NCCL), %d bytes received per rank" % (N * N * 8 * (world - 1) // max(world, 1))}
    if rank == 0:
        # parity of a seeded sub-block with the CPU oracle: rows near the start, columns near the end (another rank's rows)
        from oracle.trunc_np import SignatureKernelOracle
        o = SignatureKernelOracle(L * d, d, m, base='rbf')
        ri = np.arange(0, 12)
        ci = np.arange(N - 12, N)
        ref = o.K(Xn[ri].reshape(len(ri), -1), Xn[ci].reshape(len(ci), -1))
        got = K[ri[0]:ri[-1] + 1, ci[0]:ci[-1] + 1].cpu().numpy()
        res["oracle_subblock_max_rel_err"] = float(np.max(np.abs(got - ref)) / np.max(np.abs(ref)))
        band = K[N // 2 - 256:N // 2 + 256, N // 2 - 256:N // 2 + 256]
        res["band_exactly_symmetric"] = bool(torch.equal(band, band.T))
        res["corner_symmetric"] = bool(torch.equal(K[:64, N - 64:], K[N - 64:, :64].T))
        alone = k.K(X[N - 64:].contiguous(), X[:256].contiguous())        # single-rank compute of rows the last rank owns
        res["block_max_rel_diff_to_single_rank"] = float((K[N - 64:, :256] - alone).abs().max() / alone.abs().max())
        print(json.dumps(res))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
