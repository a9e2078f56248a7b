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


def sym_row_ranges(n_rows, world_size, rank):
    """K(X, X) with only the entries i <= j computed: row i costs N - i pairs, so rank r takes blocks r and
    2G-1-r of a partition into 2G blocks -- every rank gets the same number of pairs."""
    blocks = all_row_ranges(n_rows, 2 * world_size)
    return [blocks[rank], blocks[2 * world_size - 1 - rank]]


def _gather_into(full, blk, ranges, group, handles):
    """All-gather one row block per rank (``ranges[r]`` = rows of rank r, uneven allowed) into ``full`` (rows on
    dim 0).  Asynchronous: appends (work, finish) to ``handles``; ``finish()`` places the pieces."""
    world = len(ranges)
    maxr = max(e - b for b, e in ranges)
    if maxr == 0:
        return
    pad = blk
    if blk.shape[0] != maxr:
        pad = blk.new_zeros((maxr,) + tuple(blk.shape[1:]))
        pad[:blk.shape[0]] = blk
    buf = blk.new_empty((world * maxr,) + tuple(blk.shape[1:]))
    work = dist.all_gather_into_tensor(buf, pad.contiguous(), group=group, async_op=True)

    def finish():
        for r, (b, e) in enumerate(ranges):
            full[b:e] = buf[r * maxr:r * maxr + (e - b)]
    handles.append((work, finish))


def sharded_gram(kernel, X, X2=None, return_levels=False, gather=True, group=None, tiles=1, mirror=None):
    """``kernel.K(X, X2)`` with the rows split over the ranks; ``gather=True`` returns the full matrix on every rank.

    * ``tiles > 1``: the rank's rows are computed in ``tiles`` pieces and the all-gather of a finished piece
      (NCCL, its own stream) runs while the next piece is computed -- the collective hides behind the kernel.
    * ``X2 is None`` and ``gather``: only the entries on and above the diagonal are computed
      (``kernel.K(..., upper_only=True)`` on the mirrored row blocks of ``sym_row_ranges``), the assembled matrix
      is completed by ``mirror`` (default ``gpsig_b200.ops.mirror_lower``): half the pairs, every rank the same share.
    """
    n = X.shape[0]
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world == 1:
        return kernel.K(X, X2, return_levels=return_levels)
    if not gather:
        b, e = row_range(n, world, rank)
        return kernel.K(X, X2, return_levels=return_levels, rows=(b, e))
    sym = X2 is None
    if sym:
        per_rank = [sym_row_ranges(n, world, r) for r in range(world)]           # two blocks per rank
    else:
        per_rank = [[row_range(n, world, r)] for r in range(world)]
    # split every block into `tiles` pieces (same piece count on every rank)
    pieces = []                                                                    # [piece][rank] -> (b, e)
    for k in range(len(per_rank[0])):
        for t in range(max(1, int(tiles))):
            pieces.append([tuple(bb + x for x in row_range(ee - bb, max(1, int(tiles)), t)) for bb, ee in (pr[k] for pr in per_rank)])
    full = None
    handles = []
    for ranges in pieces:
        b, e = ranges[rank]
        kw = dict(upper_only=True) if sym else {}
        blk = kernel.K(X, X2, return_levels=return_levels, rows=(b, e), **kw)
        blk = blk.movedim(1, 0) if return_levels else blk                          # rows first
        if full is None:
            full = blk.new_empty((n,) + tuple(blk.shape[1:]))
        _gather_into(full, blk, ranges, group, handles)
    for work, finish in handles:
        work.wait()
        finish()
    full = full.movedim(0, 1).contiguous() if return_levels else full
    if sym:
        if mirror is None:
            from . import ops
            mirror = ops.mirror_lower
        full = mirror(full)
    return full
