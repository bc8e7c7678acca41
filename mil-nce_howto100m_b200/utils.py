"""Drop-in for the reference's ``utils.py::AllGather`` (utils.py:8-24) and the
fused gather -> loss module for the train-step call site
(main_distributed.py:233-238, train.py:217-221).

``AllGather.apply(tensor, args)`` keeps the reference's contract - forward
returns the rank-ordered concatenation, backward returns only the local slice of
the incoming gradient with NO reduction (utils.py:20-24) - but gathers with one
``all_gather_into_tensor`` straight into the output buffer (no list of temporaries,
no ``torch.cat`` copy).

``GatheredMILNCELoss`` goes further: every rank computes only
the row block and the column block of the similarity matrix that its own samples
need, so the loss is sharded instead of being recomputed in full on every rank as
in the reference.  It returns the same global-mean loss on every rank and the same
local gradients ``AllGather.backward`` would deliver.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def _world(args=None):
    if args is not None and hasattr(args, "world_size"):
        return int(args.world_size), int(args.rank)
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(), dist.get_rank()
    return 1, 0


class AllGather(torch.autograd.Function):
    """An autograd function that performs allgather on a tensor (reference utils.py:8-24)."""

    @staticmethod
    def forward(ctx, tensor, args):
        world, rank = _world(args)
        ctx.rank = rank
        ctx.batch_size = tensor.shape[0]
        t = tensor.contiguous()
        if world == 1:
            return t.clone()
        out = torch.empty((world * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(out, t)
        return out

    @staticmethod
    def backward(ctx, grad_output):
        return (
            grad_output[ctx.batch_size * ctx.rank: ctx.batch_size * (ctx.rank + 1)],
            None,
        )


# ---------------------------------------------------------------------------------------------
# the fused gather -> loss -> local-slice module
# ---------------------------------------------------------------------------------------------
def _gather_embeddings(v, t, world, group):
    """All-gather of the video AND text shards as ONE collective call (reference utils.py:13-17 x 2:
    two list all_gathers and two cats).  NCCL runs the pair as a single grouped kernel; a backend
    without coalescing support (gloo on CPU) gets two plain calls."""
    V = torch.empty((world * v.shape[0], v.shape[1]), dtype=v.dtype, device=v.device)
    T = torch.empty((world * t.shape[0], t.shape[1]), dtype=t.dtype, device=t.device)
    pg = group if group is not None else dist.group.WORLD
    if dist.get_backend(pg) == "nccl":
        with dist._coalescing_manager(group=pg, device=v.device, async_ops=True) as cm:
            dist.all_gather_into_tensor(V, v, group=pg)
            dist.all_gather_into_tensor(T, t, group=pg)
        cm.wait()
        return V, T
    h1 = dist.all_gather_into_tensor(V, v, group=group, async_op=True)
    h2 = dist.all_gather_into_tensor(T, t, group=group, async_op=True)
    h1.wait()
    h2.wait()
    return V, T


class PeerRing:
    """This rank's slot in peer-mapped ("symmetric") memory plus the peers' views of theirs.

    The rows scheme can then skip the NCCL all-gather of the embeddings: every rank copies its shard into
    its slot (milnce_peer_publish), the ranks meet at a device-side barrier, and the staging kernel reads
    all shards straight from the peers' slots over NVLink (milnce_forward_shard_peers).  One slot, ONE
    barrier before it is read ("every slot is complete").  No barrier is needed before a rank overwrites its
    slot in the next step: it gets there only after passing that step's statistics exchange (a barrier, or an
    all-gather), which every rank enters only after its staging kernel - the last reader of the slots - has
    finished.  (Barriers are kernels on the current stream, so the sequence can be captured in a CUDA graph.)

    The same allocation holds this rank's partial text gradient: backward writes it there, a third barrier
    follows, and every rank sums its own block of all partials from the peers' memory
    (milnce_peer_reduce_dtext) - the reduce-scatter without a collective kernel competing for SMs with the
    dvideo GEMM.  The next step's two forward barriers keep it from being overwritten while a peer still reads."""

    def __init__(self, b, K, D, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        from . import _lib
        lib = _lib.load()
        self.slot_bytes = int(lib.milnce_peer_slot_bytes(b, K, D))
        pg = group if group is not None else dist.group.WORLD
        world = dist.get_world_size(pg)
        # one allocation: [slot | (B*K, D) bfloat16 partial text gradient of this rank | its statistics pack]
        grad_bytes = world * b * K * D * 2
        pack_off = self.slot_bytes + (grad_bytes + 255) // 256 * 256
        pack_bytes = (2 * b + world * b) * 4
        self.buf = symm_mem.empty(pack_off + pack_bytes, dtype=torch.uint8, device=device)
        self.hdl = symm_mem.rendezvous(self.buf, pg)
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.grad_ptrs = [p + self.slot_bytes for p in self.ptrs]
        self.grad = self.buf[self.slot_bytes:self.slot_bytes + grad_bytes].view(torch.bfloat16)
        self.pack_ptrs = [p + pack_off for p in self.ptrs]
        self.pack = self.buf[pack_off:pack_off + pack_bytes].view(torch.float32)
        self.rank = int(self.hdl.rank)
        self.shape = (b, K, D)
        self.side = torch.cuda.Stream(device=device)     # barrier + gradient reduction run here, beside dvideo
        self.combine_scratch = torch.zeros(int(lib.milnce_shard_combine_scratch_bytes(world, b)), dtype=torch.uint8,
                                           device=device)

    def next_slot(self):
        return self.ptrs[self.rank], list(self.ptrs)

    def barrier(self):
        self.hdl.barrier(channel=0)          # every rank has published


def _mark(marks, name, dev):
    """Phase stamp for the timeline tool (profiles/multi_timeline.py); a no-op in normal use."""
    if marks is not None:
        ev = torch.cuda.Event(enable_timing=True)
        ev.record(torch.cuda.current_stream(dev))
        marks.append((name, ev))


class _GatheredMILNCE(torch.autograd.Function):
    """Row block + column block per rank ("rowcol"): 6/W GEMM units, no gradient collective."""

    @staticmethod
    def forward(ctx, video_local, text_local, world, rank, group, backend, opts):
        b = video_local.shape[0]
        v = video_local.detach().contiguous()
        t = text_local.detach().contiguous()
        marks = opts.get("marks")
        _mark(marks, "start", v.device)
        if world > 1:
            V, T = _gather_embeddings(v, t, world, group)
        else:
            V, T = v, t
        _mark(marks, "gather_embeddings", v.device)
        i0 = rank * b
        loss_part, den, nom, ws = backend.forward(V, T, i0, b)
        _mark(marks, "forward_kernels", v.device)
        if world > 1:
            # one small all-gather: [den of my samples | my share of the loss]; every rank needs den
            # of every sample for its column block, and the loss is the sum of the shares
            pack = torch.cat([den[i0:i0 + b], loss_part.reshape(1).to(den.dtype)])
            allp = torch.empty(world * (b + 1), dtype=pack.dtype, device=pack.device)
            dist.all_gather_into_tensor(allp, pack, group=group)
            allp = allp.view(world, b + 1)
            den = allp[:, :b].reshape(world * b).contiguous()
            loss = allp[:, b].sum()
        else:
            loss = loss_part[0]
        _mark(marks, "stats_exchange", v.device)
        ctx.save_for_backward(V, T, den, nom)
        ctx.ws = ws
        ctx.meta = (i0, b, backend, v.dtype, t.dtype, marks)
        return loss

    @staticmethod
    def backward(ctx, grad_output):
        V, T, den, nom = ctx.saved_tensors
        i0, b, backend, vdt, tdt, marks = ctx.meta
        _mark(marks, "backward_start", V.device)
        g = grad_output.detach().reshape(1).to(torch.float32).contiguous()
        dv, dt = backend.backward(V, T, den, nom, g, i0, b, ctx.ws)
        ctx.ws = None
        _mark(marks, "backward_kernels", V.device)
        # exactly the slice AllGather.backward returns: no gradient collective (utils.py:20-24)
        return dv.to(vdt), dt.to(tdt), None, None, None, None, None


class _GatheredMILNCERows(torch.autograd.Function):
    """Rows-only sharding: each rank computes only its rows of the similarity matrix.

    Exchanges: one grouped all-gather of the embeddings (as in the reference), one all-gather of
    2*b + B floats per rank (row statistic, positives, partial column statistic) and one
    reduce-scatter of the (B*K, D) partial text gradient in backward (bfloat16 or float32 on the
    wire).  Per-rank GEMM work is 4/W of the single-GPU step (6/W for the row+column block scheme)."""

    @staticmethod
    def forward(ctx, video_local, text_local, world, rank, group, backend, opts):
        b = video_local.shape[0]
        v = video_local.detach().contiguous()
        t = text_local.detach().contiguous()
        marks = opts.get("marks")
        _mark(marks, "start", v.device)
        B = world * b
        i0 = rank * b
        ring = opts.get("ring")
        if ring is not None:
            # no embedding collective: publish the shard, meet, stage all shards from the peers' memory
            pack, ws = backend.forward_shard_peers(ring, v, t, world, rank,
                                                   mark=(lambda name: _mark(marks, name, v.device)) if marks is not None else None)
            V, T = v, t                      # (only their shapes are used from here on)
            _mark(marks, "stage+forward_kernels", v.device)
            if opts.get("reduce") != "nccl":
                # the statistics stay where they are: barrier, then every rank reads all W packs
                ring.barrier()
                _mark(marks, "stats_exchange", v.device)
                den, nom_all, loss = backend.shard_combine_peers(ring, world)
                _mark(marks, "combine", v.device)
                pack = None
        else:
            V, T = _gather_embeddings(v, t, world, group)
            _mark(marks, "gather_embeddings", v.device)
            # the kernels write [row_lse of my samples | nom of my samples | my column partials] in place
            pack, ws = backend.forward_shard(V, T, i0, b)
            _mark(marks, "forward_kernels", v.device)
        if pack is not None:
            allp = torch.empty(world * pack.numel(), dtype=pack.dtype, device=pack.device)
            dist.all_gather_into_tensor(allp, pack, group=group)
            _mark(marks, "stats_exchange", v.device)
            den, nom_all, loss = backend.shard_combine(allp.view(world, 2 * b + B), world, b)
            _mark(marks, "combine", v.device)
        ctx.save_for_backward(V, T, den, nom_all)
        ctx.ws = ws
        ctx.meta = (i0, b, backend, v.dtype, t.dtype, group, t.shape[0], marks, opts.get("partial_dtype"),
                    (B, t.shape[0] // b, v.shape[1]), ring, opts.get("reduce"))
        return loss

    @staticmethod
    def backward(ctx, grad_output):
        V, T, den, nom = ctx.saved_tensors
        i0, b, backend, vdt, tdt, group, bk_local, marks, pdt, dims, ring, reduce = ctx.meta
        _mark(marks, "backward_start", V.device)
        g = grad_output.detach().reshape(1).to(torch.float32).contiguous()
        # default: partials in the embeddings' own 16-bit dtype - the result is rounded to that dtype anyway, and
        # a float32 payload (67 MB at B=8192) made the reduce-scatter the longest phase of the step
        # (profiles/r2_multi_timeline_*.txt); float32 on request (partial_dtype=torch.float32)
        # (NCCL adds 16-bit partials in 16 bits: beyond two ranks the rounding of every ring step accumulates -
        # 1.5e-2 on dtext at eight ranks, harness/config5.py - so the collective then carries float32; the peer
        # path sums bf16 partials in float32 and keeps the 16-bit payload at any rank count)
        if pdt is None:
            two_byte = tdt == torch.bfloat16 and ((ring is not None and reduce != "nccl") or dims[0] // b <= 2)
            pdt = torch.bfloat16 if two_byte else torch.float32
        # the text-gradient partial first, so that its reduce-scatter runs while dvideo is computed
        if ring is not None and pdt == torch.bfloat16 and reduce != "nccl":
            world, rank = dims[0] // b, i0 // b
            push = reduce == "push" and (b * dims[1]) % 128 == 0
            if push:
                # the dtext GEMM's epilogue stores every owner's rows into that rank's memory
                backend.backward_shard_push(ring, den, nom, g, world, rank, ctx.ws)
            else:
                # partial into my own peer-readable buffer; every rank later reads its block of all partials
                backend.backward_shard(V, T, den, nom, g, i0, b, ctx.ws, parts=1, partial_dtype=pdt, dims=dims,
                                       partial_out=ring.grad)
            _mark(marks, "gradient_tiles+dtext_partial", V.device)
            # barrier + reduction on a side stream, next to the dvideo GEMM (whose split-K units leave SMs free)
            main = torch.cuda.current_stream(V.device)
            dt = torch.empty(dims[1] * b, dims[2], dtype=tdt, device=V.device)
            ring.side.wait_stream(main)
            with torch.cuda.stream(ring.side):
                ring.barrier()
                backend.peer_reduce_dtext(ring, world, rank, tdt, pushed=push, out=dt)
            dv, _ = backend.backward_shard(V, T, den, nom, g, i0, b, ctx.ws, parts=2, grad_dtype=vdt, dims=dims)
            ctx.ws = None
            _mark(marks, "dvideo", V.device)
            main.wait_stream(ring.side)
            _mark(marks, "peer_reduce_dtext", V.device)
            return dv, dt, None, None, None, None, None
        _, dt_partial = backend.backward_shard(V, T, den, nom, g, i0, b, ctx.ws, parts=1, partial_dtype=pdt, dims=dims)
        _mark(marks, "gradient_tiles+dtext_partial", V.device)
        dt = torch.empty((bk_local, dt_partial.shape[1]), dtype=dt_partial.dtype, device=dt_partial.device)
        work = dist.reduce_scatter_tensor(dt, dt_partial, op=dist.ReduceOp.SUM, group=group, async_op=True)
        dv, _ = backend.backward_shard(V, T, den, nom, g, i0, b, ctx.ws, parts=2, grad_dtype=vdt, dims=dims)
        ctx.ws = None
        _mark(marks, "dvideo", V.device)
        work.wait()
        _mark(marks, "reduce_scatter_done", V.device)
        return dv, dt.to(tdt), None, None, None, None, None


class GatheredMILNCELoss(torch.nn.Module):
    """allgather(video), allgather(text), MILNCELoss - fused and sharded.

    Replaces the three lines ``video_embd = allgather(video_embd, args); text_embd =
    allgather(text_embd, args); loss = loss_fun(video_embd, text_embd)``
    (main_distributed.py:235-237).  Inputs are the LOCAL (b, D) / (b*K, D)
    embeddings of this rank.
    """

    def __init__(self, args=None, group=None, scheme="auto", partial_dtype=None, gather="auto", reduce="auto",
                 _backend=None):
        """scheme: "rows" (each rank computes only its rows of S, reduce-scatter of the text
        gradient; bf16 embeddings), "rowcol" (row block + column block per rank, no gradient
        collective - the reference's communication pattern), or "auto" (rows when available).
        partial_dtype: dtype of the rows scheme's text-gradient partials on the wire (default: the
        embeddings' dtype, i.e. bfloat16; torch.float32 doubles the reduce-scatter payload).
        gather: how the rows scheme obtains the other ranks' embeddings - "nccl" (one grouped all-gather),
        "peer" (no collective: shards are read from the peers' memory, see PeerRing; needs CUDA bf16
        embeddings, a local batch that is a multiple of 64, D a multiple of 8, at most 16 ranks), or
        "auto" (peer when those conditions hold and symmetric memory can be set up, else nccl).
        reduce: the gradient collective of the rows scheme on the peer path - "pull" / "auto" (bfloat16 partials
        stay in the rank's peer-readable buffer and every rank sums its block of all W partials, beside the
        dvideo GEMM), "push" (the text-gradient GEMM's epilogue stores each owner's rows into that rank's
        memory; measured slower: 16-byte stores over NVLink stall the epilogue), or "nccl" (reduce-scatter; also
        keeps the statistics all-gather on NCCL).
        _backend: private test seam (the CPU tests drive the orchestration with a stand-in)."""
        super().__init__()
        if reduce not in ("auto", "nccl", "pull", "push"):
            raise ValueError("reduce must be 'auto', 'nccl', 'pull' or 'push'")
        self.reduce = reduce
        if scheme not in ("auto", "rows", "rowcol"):
            raise ValueError("scheme must be 'auto', 'rows' or 'rowcol'")
        if gather not in ("auto", "nccl", "peer"):
            raise ValueError("gather must be 'auto', 'nccl' or 'peer'")
        self.gather = gather
        self._ring = None
        self._ring_failed = False
        self.last_gather = None
        self.args = args
        self.group = group
        self._backend = _backend
        self.scheme = scheme
        self.partial_dtype = partial_dtype
        self.last_scheme = None
        self._marks = None   # set to a list by profiles/multi_timeline.py to collect phase stamps

    def forward(self, video_embd, text_embd):
        from .loss import CudaBackend, _validate
        backend = self._backend
        if backend is None:
            _validate(video_embd, text_embd)
            backend = CudaBackend
        world, rank = _world(self.args)
        use_rows = False
        if world > 1 and self.scheme != "rowcol":
            ok = backend.supports_rows_scheme(video_embd, rank * video_embd.shape[0])
            if self.scheme == "rows" and not ok:
                raise RuntimeError("scheme='rows' needs bfloat16 embeddings and a local batch that is a "
                                   "multiple of 8")
            use_rows = ok
        self.last_scheme = "rows" if use_rows else ("rowcol" if world > 1 else "single")
        fn = _GatheredMILNCERows if use_rows else _GatheredMILNCE
        ring = self._peer_ring(video_embd, text_embd, world) if (use_rows and self._backend is None) else None
        self.last_gather = "peer" if ring is not None else ("nccl" if world > 1 else None)
        opts = {"marks": self._marks, "partial_dtype": self.partial_dtype, "ring": ring, "reduce": self.reduce}
        return fn.apply(video_embd, text_embd, world, rank, self.group, backend, opts)


def _peer_ring(self, video_embd, text_embd, world):
    """The PeerRing for these shapes, or None when the peer path does not apply (see `gather`)."""
    if self.gather == "nccl" or self._ring_failed:
        return None
    b, D = video_embd.shape
    K = text_embd.shape[0] // max(b, 1)
    ok = (video_embd.is_cuda and video_embd.dtype == torch.bfloat16 and b % 64 == 0 and D % 8 == 0 and world <= 16
          and dist.get_backend(self.group if self.group is not None else dist.group.WORLD) == "nccl")
    if not ok:
        if self.gather == "peer":
            raise RuntimeError("gather='peer' needs CUDA bfloat16 embeddings, a local batch that is a multiple of 64, "
                               "D a multiple of 8, at most 16 ranks and the NCCL backend")
        return None
    if self._ring is None or self._ring.shape != (b, K, D):
        try:
            self._ring = PeerRing(b, K, D, video_embd.device, self.group)
        except Exception:
            if self.gather == "peer":
                raise
            self._ring_failed = True        # symmetric memory is not available here: stay on NCCL
            return None
    return self._ring


GatheredMILNCELoss._peer_ring = _peer_ring


class GraphedGatheredMILNCEStep:
    """One sharded loss step - embedding all-gather, sharded forward, backward, gradient exchange -
    for fixed local shapes, captured ONCE as a CUDA graph (NCCL collectives included) and replayed.

    From two ranks on, a step is a few hundred microseconds of GPU work behind some forty host-side
    operations (launches, allocations, NCCL calls, autograd), and the eager module is bound by the
    host.  Replaying removes that time from every step.  Same contract as GraphedMILNCEStep:

        step = GraphedGatheredMILNCEStep(b, K, D)        # every rank, same arguments; captures
        loss, dvideo, dtext = step(video_local, text_local)
        ...
        step.close()                                      # before destroy_process_group()

    Outputs are the graph's static tensors (overwritten by the next replay): the global loss and
    d loss / d(local embeddings), i.e. what AllGather.backward (reference utils.py:20-24) hands back.
    All ranks must construct, replay and close in step with each other (the graph holds collectives).
    """

    def __init__(self, b, K, D, dtype=torch.bfloat16, device=None, group=None, scheme="auto",
                 partial_dtype=None, gather="auto", reduce="auto", warmup=3):
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.video = torch.zeros(b, D, dtype=dtype, device=dev).requires_grad_(True)
        self.text = torch.zeros(b * K, D, dtype=dtype, device=dev).requires_grad_(True)
        self.module = GatheredMILNCELoss(group=group, scheme=scheme, partial_dtype=partial_dtype, gather=gather,
                                         reduce=reduce)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):       # warm-up off the capture: communicators, library, allocator
            for _ in range(max(1, warmup)):
                self.video.grad = None
                self.text.grad = None
                self.module(self.video, self.text).backward()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.video.grad = None
        self.text.grad = None
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss = self.module(self.video, self.text)
            self.loss.backward()
        self.dvideo = self.video.grad
        self.dtext = self.text.grad

    def replay(self):
        self.graph.replay()
        return self.loss, self.dvideo, self.dtext

    def __call__(self, video_embd, text_embd):
        with torch.no_grad():
            self.video.copy_(video_embd, non_blocking=True)
            self.text.copy_(text_embd, non_blocking=True)
        return self.replay()

    def close(self):
        """Release the captured graph (it pins NCCL work); call on every rank before the process
        group is destroyed."""
        torch.cuda.synchronize()
        self.graph = None
