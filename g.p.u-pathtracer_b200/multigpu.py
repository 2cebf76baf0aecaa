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


class PeerFrames:
    """Fused alternative to render_progressive_frame's NCCL reduce (SURVEY 8-e "later" row): every rank renders into a
    library-owned, IPC-exported frame buffer; the root reads all peers' buffers over NVLink inside ONE kernel that
    sums the ranks in rank order, adds the frame into the running sum and resolves it (gpt_reduce_resolve).  The only
    collective left per frame is a one-element all_reduce used as a stream-ordered barrier ("the peers' renders are
    done"); double buffering makes the next frame's clear safe while the root may still be reading.  With more than one
    rank the root runs that kernel on a SIDE stream behind an event, so its own next render overlaps the
    reduce; read the outputs after wait_reduced().

    ctx: gpt_b200.Context; dist: an initialised torch.distributed (NCCL) or None for a single rank."""

    def __init__(self, ctx, width, height, dist=None, device=None, root=0):
        import torch
        self.ctx, self.dist, self.root = ctx, dist, root
        self.world = dist.get_world_size() if dist is not None else 1
        self.rank = dist.get_rank() if dist is not None else 0
        self.width, self.height = int(width), int(height)
        self.frame_index = 0
        self._side = None                  # root: the stream the fused reduce runs on, so that the next frame's render overlaps it
        self._reduced = [None, None]       # root: event per frame buffer, recorded behind its reduce
        handle = ctx.frame_alloc(width, height, 2)
        self._flag = torch.zeros(1, dtype=torch.int32, device=device)
        if self.world > 1:
            mine = torch.tensor(list(handle), dtype=torch.uint8, device=device)
            gathered = [torch.empty_like(mine) for _ in range(self.world)]
            dist.all_gather(gathered, mine)
            ctx.frame_open_peers([bytes(g.cpu().tolist()) for g in gathered], self.rank)
            dist.barrier()

    def frame(self, render_fn, frame_begin, frame_spp, total_accum=None, out_rgba8=None, out_rgb_f32=None, stream=None):
        """render_fn(accum_ptr, sample_begin, sample_count) must ADD the slice into the float4 buffer at accum_ptr on
        `stream`.  Returns this rank's (offset, count).

        `stream`: a torch.cuda.Stream (or None = torch's current stream).  Clear, render, barrier and reduce must all be ordered
        on ONE stream: the barrier is a torch.distributed collective, which torch issues on ITS current stream, so the whole
        body runs under torch.cuda.stream(stream).  A raw cudaStream_t handle is accepted only if it is torch's current stream."""
        import torch
        if stream is None or isinstance(stream, int):
            cur = torch.cuda.current_stream()
            if isinstance(stream, int) and stream != cur.cuda_stream:
                raise ValueError("PeerFrames.frame: pass a torch.cuda.Stream (a raw handle must be torch's current stream): the barrier "
                                 "collective runs on torch's current stream and would not order the peers' renders before the reduce")
            stream = cur
        buf = self.frame_index & 1
        self.frame_index += 1
        off, cnt = sample_slice(frame_spp, self.world, self.rank)
        with torch.cuda.stream(stream):
            raw = stream.cuda_stream
            self.ctx.frame_clear(buf, raw)
            if cnt:
                render_fn(self.ctx.frame_ptr(buf), frame_begin + off, cnt)
            if self.world > 1:
                # The root's reduce of the PREVIOUS frame runs on the side stream; the peers clear that frame's buffer again once
                # they are past THIS barrier, so the root must not enter it before that reduce has finished.  Issued here -- behind
                # this frame's render -- the wait costs nothing unless the reduce is slower than a render.
                if self.rank == self.root and self._reduced[buf ^ 1] is not None:
                    stream.wait_event(self._reduced[buf ^ 1])
                self.dist.all_reduce(self._flag)               # stream-ordered barrier: every rank's render precedes it
            if self.rank == self.root:
                if self.world > 1:
                    if self._side is None:
                        self._side = torch.cuda.Stream()
                    ready = torch.cuda.Event()
                    ready.record(stream)
                    self._side.wait_event(ready)
                    self.ctx.reduce_resolve(buf, total_accum, out_rgba8, out_rgb_f32, self._side.cuda_stream)
                    done = torch.cuda.Event()
                    done.record(self._side)
                    self._reduced[buf] = done
                else:
                    self.ctx.reduce_resolve(buf, total_accum, out_rgba8, out_rgb_f32, raw)
        return off, cnt

    def wait_reduced(self, stream=None):
        """Root: make `stream` (default: torch's current stream) wait for the reduces issued so far -- call before reading
        total_accum / out_rgba8 / out_rgb_f32 on it.  A no-op for a single rank (the reduce ran on the caller's stream)."""
        import torch
        stream = stream if stream is not None else torch.cuda.current_stream()
        for ev in self._reduced:
            if ev is not None:
                stream.wait_event(ev)

    def close(self):
        if self.world > 1:
            import torch
            torch.cuda.synchronize()
            self.dist.barrier()                            # nobody unmaps while the root may still be reading
        self.ctx.frame_free()
