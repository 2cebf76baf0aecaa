"""Sample-pass data parallelism over the GPUs of one box (north_star: "each GPU tracing a
disjoint spp slice seeded by global sample index, float4 accumulation buffers summed with one NCCL reduce
over NVLink per progressive frame").

The reference has no multi-GPU path at all (single device 0, utilfun.cpp:73,95); this is new.  The scene is
replicated, rank r renders global sample indices [begin_r, begin_r + count_r) into its own float4 buffer, and
ONE collective per progressive frame sums the buffers onto rank 0 (torch.distributed: NCCL on GPUs, gloo in
the CPU tests).  The per-frame clamp of the legacy contract (tracer.cu:390) is not linear, so only the
native unclamped float4 sum contract can be reduced; clamping happens once, in resolve, on the root.
"""


def sample_slice(total, world, rank):
    """Contiguous block partition of `total` samples over `world` ranks -> (offset, count).
    Counts differ by at most one; offsets are increasing in rank; the slices tile [0,total)."""
    if world < 1 or not (0 <= rank < world) or total < 0:
        raise ValueError("bad partition (total=%r world=%r rank=%r)" % (total, world, rank))
    base, rem = divmod(total, world)
    count = base + (1 if rank < rem else 0)
    offset = rank * base + min(rank, rem)
    return offset, count


def render_progressive_frame(render_fn, frame_accum, total_accum, frame_begin, frame_spp,
                             dist=None, group=None, root=0, all_ranks=False):
    """One progressive frame of `frame_spp` samples starting at global index `frame_begin`.

    frame_accum : per-rank float tensor (W*H*4), zeroed here, filled by render_fn(sample_begin, sample_count)
                  which must ADD that slice into it, then summed over ranks with ONE collective
                  (reduce to `root`, or all_reduce when every rank needs the image).
    total_accum : the running float4 sum over all frames; updated on the root (on every rank with all_ranks).
                  May be None.  Keeping the frame buffer separate from the running sum is what makes the
                  reduce safe to repeat frame after frame (reducing a running sum would double count).
    Returns (offset, count) of the slice this rank rendered."""
    world = dist.get_world_size(group) if dist is not None else 1
    rank = dist.get_rank(group) if dist is not None else 0
    off, cnt = sample_slice(frame_spp, world, rank)
    frame_accum.zero_()
    if cnt:
        render_fn(frame_begin + off, cnt)
    if dist is not None and world > 1:
        if all_ranks:
            dist.all_reduce(frame_accum, op=dist.ReduceOp.SUM, group=group)
        else:
            dist.reduce(frame_accum, dst=root, op=dist.ReduceOp.SUM, group=group)
    if total_accum is not None and (all_ranks or rank == root):
        total_accum.add_(frame_accum)
    return off, cnt
