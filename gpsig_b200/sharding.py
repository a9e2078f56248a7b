"""Row-block sharding of the Gram across the GPUs of one box.

Every ordered pair (x_i, y_j) is independent, so rank r of G computes the rows
``row_range(N, G, r)`` of the ``[N, N2]`` Gram with no data-path collective;
``X2`` (and the normalisation diagonals) are replicated.  Only when the FULL
matrix is needed (Kff / Kuu for a Cholesky) the blocks are all-gathered
(``all_gather_rows``: NCCL over NVLink on GPUs, gloo in the CPU tests).

The compute is injected (``compute_rows(begin, end) -> tensor``) so that this
module has no dependency on the CUDA library: the multi-process logic is
tested on CPU with gloo, world_size 2 (tests/test_sharding_cpu.py).
"""
import torch
import torch.distributed as dist


def row_range(n_rows, world_size, rank):
    """Contiguous balanced partition: the first ``n_rows % world_size`` ranks get one extra row."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad world_size/rank %r/%r" % (world_size, rank))
    base, rem = divmod(int(n_rows), world_size)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def all_row_ranges(n_rows, world_size):
    return [row_range(n_rows, world_size, r) for r in range(world_size)]


def sharded_rows(compute_rows, n_rows, group=None):
    """This rank's row block ``compute_rows(begin, end)`` and its range."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    begin, end = row_range(n_rows, world, rank)
    return compute_rows(begin, end), (begin, end)


def all_gather_rows(block, n_rows, group=None, row_dim=0):
    """Assemble the full matrix from the ranks' row blocks (uneven blocks allowed).

    ``block``: this rank's rows along ``row_dim`` (e.g. ``[R, N2]`` or ``[m+1, R, N2]`` with row_dim=1).
    """
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return block
    world = dist.get_world_size(group)
    ranges = all_row_ranges(n_rows, world)
    blk = block.movedim(row_dim, 0).contiguous()
    maxr = max(e - b for b, e in ranges)
    if all(e - b == maxr for b, e in ranges):
        out = blk.new_empty((world * maxr,) + tuple(blk.shape[1:]))
        dist.all_gather_into_tensor(out, blk, group=group)
    else:
        pad = blk.new_zeros((maxr,) + tuple(blk.shape[1:]))
        pad[:blk.shape[0]] = blk
        bufs = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(bufs, pad, group=group)
        out = torch.cat([bufs[r][:e - b] for r, (b, e) in enumerate(ranges)], dim=0)
    return out.movedim(0, row_dim)


def sharded_gram(kernel, X, X2=None, return_levels=False, gather=True, group=None):
    """``kernel.K(X, X2)`` with the rows split over the ranks; ``gather=True`` returns the full matrix on every rank."""
    n = X.shape[0]
    blk, _ = sharded_rows(lambda b, e: kernel.K(X, X2, return_levels=return_levels, rows=(b, e)), n, group)
    if not gather:
        return blk
    return all_gather_rows(blk, n, group=group, row_dim=1 if return_levels else 0)
