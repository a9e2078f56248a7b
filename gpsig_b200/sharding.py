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


def _global_rank(group, r):
    return r if group is None else dist.get_global_rank(group, r)


def _gather_piece(full, ranges, rank, group, handles, inplace_ok):
    """Make every rank's rows ``ranges[r]`` of ``full`` (rows on dim 0; this rank's rows already computed IN PLACE)
    visible on all ranks.  No staging buffer and no copy: equal blocks in rank order are one in-place all-gather
    (``ncclAllGather`` with the send buffer inside the receive buffer), anything else (uneven blocks, the mirrored
    second block of the symmetric sweep, tiles) is one broadcast per owner straight into its row slice.
    Asynchronous: appends the work handles to ``handles``."""
    world = len(ranges)
    sizes = [e - b for b, e in ranges]
    if max(sizes) == 0:
        return
    contiguous = all(ranges[r][1] == ranges[r + 1][0] for r in range(world - 1))
    if inplace_ok and contiguous and min(sizes) == max(sizes):
        b0, e1 = ranges[0][0], ranges[-1][1]
        b, e = ranges[rank]
        handles.append(dist.all_gather_into_tensor(full[b0:e1], full[b:e], group=group, async_op=True))
        return
    for r, (b, e) in enumerate(ranges):
        if e > b:
            handles.append(dist.broadcast(full[b:e], src=_global_rank(group, r), group=group, async_op=True))


def _call_K(kernel, X, X2, rows, out, return_levels, **kw):
    """kernel.K on a row block, written straight into ``out`` when the kernel supports it."""
    if out is not None and getattr(kernel, 'supports_out', False):
        kernel.K(X, X2, return_levels=return_levels, rows=rows, out=out, **kw)
        return
    blk = kernel.K(X, X2, return_levels=return_levels, rows=rows, **kw)
    if out is not None:
        out.copy_(blk.movedim(1, 0) if return_levels else blk)


def sharded_gram(kernel, X, X2=None, return_levels=False, gather=True, group=None, tiles=1, mirror=None, out=None):
    """``kernel.K(X, X2)`` with the rows split over the ranks; ``gather=True`` returns the full matrix on every rank.

    Who needs the full matrix: the Cholesky of Kuu / Kff (``inducing_variables.py:51-66`` of the reference).
    * Every rank computes its rows directly into its slice of the full matrix and the slices are exchanged in
      place (see ``_gather_piece``): one ``[N, N2]`` allocation per rank, no staging buffers, no copies.
    * ``tiles > 1``: the rank's rows are computed in ``tiles`` pieces and the exchange of a finished piece (NCCL, its
      own stream) runs while the next piece is computed.  Default 1: at the Gram shapes of BASELINE.json the
      exchange is < 1 % of a step and splitting the kernel launch costs more than the overlap hides (measured on
      8 B200: 247 vs 235 ms at N=4000, profiles/r01n_sharded_check_8gpu.txt).
    * ``X2 is None`` and ``gather``: only the entries on and above the diagonal are computed
      (``kernel.K(..., upper_only=True)`` on the mirrored row blocks of ``sym_row_ranges``), the assembled matrix
      is completed by ``mirror`` (default ``gpsig_b200.ops.mirror_lower``): half the pairs, every rank the same share.
    * ``out``: optional preallocated result (``[N, N2]``; rows-first ``[N, m+1, N2]`` scratch is internal for levels).
    """
    n = X.shape[0]
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world == 1:
        if out is not None and not return_levels and getattr(kernel, 'supports_out', False):
            return kernel.K(X, X2, out=out)
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
    nt = max(1, int(tiles))
    pieces = []                                                                    # [piece][rank] -> (b, e)
    for k in range(len(per_rank[0])):
        for t in range(nt):
            pieces.append([tuple(bb + x for x in row_range(ee - bb, nt, t)) for bb, ee in (pr[k] for pr in per_rank)])
    # in-place all-gather: NCCL only (gloo's all_gather_into_tensor stages the input; broadcasts are used there)
    inplace_ok = dist.get_backend(group) == 'nccl'
    handles = []
    kw = dict(upper_only=True) if sym else {}
    n2 = n if sym else X2.shape[0]
    full = None
    if not return_levels:                       # result shape and dtype are known: allocate once, compute in place
        full = out if out is not None else X.new_empty((n, n2))
        if tuple(full.shape) != (n, n2) or not full.is_contiguous():
            raise ValueError("out must be a contiguous [%d, %d] tensor" % (n, n2))
    for ranges in pieces:
        b, e = ranges[rank]
        if full is None:                        # levels: rows-first scratch [N, m+1, N2], shaped after the first block
            blk0 = kernel.K(X, X2, return_levels=True, rows=(b, e), **kw).movedim(1, 0)
            full = blk0.new_empty((n,) + tuple(blk0.shape[1:]))
            full[b:e] = blk0
            del blk0
        elif e > b:
            _call_K(kernel, X, X2, (b, e), full[b:e], return_levels, **kw)
        _gather_piece(full, ranges, rank, group, handles, inplace_ok)
    for work in handles:
        work.wait()
    full = full.movedim(0, 1).contiguous() if return_levels else full
    if sym:
        if mirror is None:
            from . import ops
            mirror = ops.mirror_lower
        full = mirror(full)
    return full
