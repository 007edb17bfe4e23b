"""Row-tile sharding of the pixel domain across the GPUs of one box.

Every pixel's gather, fit and evaluation is independent (analysis.py:221-232, 350-363),
so the ROI is cut into contiguous row slabs, one per rank, with NO collective on the hot
path.  Pixel coordinates stay global (row0 + local row) because p = (x, y, 0) enters the
light vector.  The only communication is the final gather of the small per-pixel model
(PTM coefficients: 48 B/pixel) to one rank; the int32 tables stay sharded.

One process per GPU, launched by torchrun; torch.distributed (NCCL on GPUs, gloo in the
CPU tests) is plumbing only.
"""
from __future__ import annotations


def row_partition(rows: int, world_size: int):
    """[(row0, n_rows)] per rank: rows // world_size each, the remainder spread one row at
    a time over the LAST ranks (so rank 0, which also gathers, never has the extra row)."""
    if rows < 0 or world_size <= 0:
        raise ValueError("rows >= 0 and world_size > 0 required")
    base, rem = divmod(rows, world_size)
    out, r0 = [], 0
    for r in range(world_size):
        n = base + (1 if r >= world_size - rem else 0)
        out.append((r0, n))
        r0 += n
    return out


def gather_planar(local, rows_total: int, cols: int, dst: int = 0, group=None, method: str = "p2p"):
    """Gather per-rank planar maps (K, rows_r * cols) into (K, rows_total, cols) on `dst`.

    method "p2p": the point-to-point form below.  "allgather": when every rank holds the same number of
    rows, one `all_gather_into_tensor` (NCCL's ring / NVLS all-gather over NVLink, every rank receives all
    slabs) and a device-side permute on `dst` -- more bytes moved than needed, but a bandwidth-optimal
    collective instead of K * (world - 1) messages converging on one GPU.  Measured on B200
    (tools/bench_gather.py, 805 MB map): 8 ranks p2p 1.47 ms (478 GB/s into rank 0) / allgather 1.49 ms;
    4 ranks 1.20 / 1.41 ms -- no gain, so "p2p" (which also handles uneven slabs) is the default.

    `local` is this rank's tensor for its row_partition() slab (any K, any dtype, CPU or
    CUDA).  One batch of point-to-point operations (`batch_isend_irecv`: a single NCCL group
    on GPUs, so all sources stream to `dst` concurrently over NVLink): every rank sends its K
    planes, `dst` receives each plane slab straight into its place in the assembled tensor
    (a plane's row slab is contiguous there, so there is no staging copy and uneven slabs need
    no padding).  Returns the assembled tensor on `dst` and None elsewhere.
    """
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    parts = row_partition(rows_total, world)
    k = local.shape[0]
    r0, nr = parts[rank]
    if local.shape != (k, nr * cols):
        raise ValueError("rank %d: expected local shape %s, got %s" % (rank, (k, nr * cols), tuple(local.shape)))
    local = local.contiguous()
    equal = all(sn == parts[0][1] for _, sn in parts) and nr > 0
    if method == "allgather":
        if not equal:
            raise ValueError("allgather needs equal row slabs")
        stage = torch.empty((world, k, nr * cols), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(stage.view(-1), local.view(-1), group=group)
        if rank != dst:
            return None
        # (world, K, nr, cols) -> (K, world * nr, cols)
        return stage.view(world, k, nr, cols).permute(1, 0, 2, 3).reshape(k, rows_total, cols).contiguous()
    ops = []
    full = None
    if rank != dst:
        if nr > 0:
            ops = [dist.P2POp(dist.isend, local[i], dst, group) for i in range(k)]
    else:
        full = torch.empty((k, rows_total, cols), dtype=local.dtype, device=local.device)
        for src, (s0, sn) in enumerate(parts):
            if sn == 0:
                continue
            if src == dst:
                full[:, s0:s0 + sn, :] = local.view(k, sn, cols)
            else:
                ops += [dist.P2POp(dist.irecv, full[i, s0:s0 + sn, :].view(-1), src, group) for i in range(k)]
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return full


def fit_ptm_slab(frames_slab, cams, row0: int):
    """This rank's share of the PTM fit: frames_slab uint8 (N, rows_r, cols) holding ROI rows
    [row0, row0 + rows_r), cams float64 (N, 3).  Returns the device InterpolationResult."""
    from . import analysis as A

    stack = A.FrameStack(frames_slab, cams, row0=row0)
    coef, bad = A._fit_ptm_from_stack(stack)
    A._raise_if_bad(bad, "PTM")
    return A.InterpolationResult("ptm", stack.rows, stack.cols, A.light_axis(), coef=coef, row0=row0)
