"""Drop-in for the reference's ``loss.py::MILNCELoss`` (reference loss.py:6-18).

Same constructor, same ``forward(video_embd, text_embd)`` -> 0-d tensor, same
gradients; the B x B*K similarity matrix is never materialised in fp32 and the
host->device copy of ``eye(B)`` the reference does per call (loss.py:13) is gone.
All arithmetic runs in libb200head.so (hand-written sm_100a kernels); there is no
PyTorch or CPU fallback.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib


def _dtype_code(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return _lib.F32
    if t.dtype == torch.bfloat16:
        return _lib.BF16
    raise TypeError(f"MILNCELoss supports float32 and bfloat16 embeddings, got {t.dtype}")


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


class CudaBackend:
    """The product backend: calls the C ABI on the tensors' CUDA device."""

    @staticmethod
    def forward(video, text, i0, n):
        """Returns (loss_part[1], den[B], nom[B], workspace) for samples [i0, i0+n)."""
        lib = _lib.load()
        B, D = video.shape
        K = text.shape[0] // B
        code = _dtype_code(video)
        dev = video.device
        nbytes = lib.milnce_workspace_bytes(B, K, D, code, n)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        den = torch.empty(B, dtype=torch.float32, device=dev)
        nom = torch.empty(B, dtype=torch.float32, device=dev)
        loss = torch.empty(1, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            rc = lib.milnce_forward(_ptr(video), _ptr(text), B, K, D, code, i0, n, _ptr(den),
                                    _ptr(nom), _ptr(loss), _ptr(ws), nbytes, _lib.stream_ptr(dev))
        _lib.check(rc, "milnce_forward")
        return loss, den, nom, ws

    @staticmethod
    def backward(video, text, den, nom, grad_loss, i0, n, ws, grad_dtype=torch.float32):
        """Returns (dvideo (n, D), dtext (n*K, D)) in grad_dtype: the rows this range owns."""
        lib = _lib.load()
        B, D = video.shape
        K = text.shape[0] // B
        code = _dtype_code(video)
        dev = video.device
        reuse = 1
        if ws is None:
            nbytes = lib.milnce_workspace_bytes(B, K, D, code, n)
            ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            reuse = 0
        if grad_dtype not in (torch.float32, torch.bfloat16):
            raise TypeError("grad_dtype must be float32 or bfloat16")
        dv = torch.empty(n, D, dtype=grad_dtype, device=dev)
        dt = torch.empty(n * K, D, dtype=grad_dtype, device=dev)
        gcode = 0 if grad_dtype == torch.float32 else 1
        with torch.cuda.device(dev):
            rc = lib.milnce_backward(_ptr(video), _ptr(text), _ptr(den), _ptr(nom), _ptr(grad_loss),
                                     _ptr(dv), _ptr(dt), gcode, B, K, D, code, i0, n, reuse, _ptr(ws),
                                     ws.numel(), _lib.stream_ptr(dev))
        _lib.check(rc, "milnce_backward")
        return dv, dt


    # ---- rows-only sharding (bf16): see include/b200head.h, milnce_forward_shard ----
    @staticmethod
    def supports_rows_scheme(video, i0):
        return video.dtype == torch.bfloat16 and i0 % 8 == 0 and video.shape[1] % 8 == 0

    @staticmethod
    def forward_shard(video, text, i0, n):
        """Returns (pack, ws): pack = [row_lse of the n owned samples | their nom | col_part of all B
        samples], float32, 2n + B entries - the buffer the ranks all-gather (the kernels write it
        in place: no torch.cat in the step)."""
        lib = _lib.load()
        B, D = video.shape
        K = text.shape[0] // B
        code = _dtype_code(video)
        dev = video.device
        nbytes = lib.milnce_workspace_bytes(B, K, D, code, n)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        pack = torch.empty(2 * n + B, dtype=torch.float32, device=dev)
        # the C ABI writes entries [i0, i0+n) of row_lse and nom: hand it bases such that those
        # entries are pack[0:n] and pack[n:2n]
        row_base = ctypes.c_void_p(pack.data_ptr() - 4 * i0)
        nom_base = ctypes.c_void_p(pack.data_ptr() + 4 * n - 4 * i0)
        col_base = ctypes.c_void_p(pack.data_ptr() + 8 * n)
        with torch.cuda.device(dev):
            rc = lib.milnce_forward_shard(_ptr(video), _ptr(text), B, K, D, code, i0, n, row_base,
                                          col_base, nom_base, _ptr(ws), nbytes, _lib.stream_ptr(dev))
        _lib.check(rc, "milnce_forward_shard")
        return pack, ws

    @staticmethod
    def forward_shard_peers(ring, video_local, text_local, world, rank, mark=None):
        """The rows-only forward without an embedding collective (see PeerRing in utils.py and
        include/b200head.h, milnce_forward_shard_peers).  Returns (pack, ws) like forward_shard."""
        lib = _lib.load()
        b, D = video_local.shape
        K = text_local.shape[0] // b
        B = world * b
        i0 = rank * b
        dev = video_local.device
        nbytes = lib.milnce_workspace_bytes(B, K, D, _lib.BF16, b)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        pack = ring.pack        # this rank's statistics, where the peers can read them
        mark = mark or (lambda name: None)
        slot_ptr, peer_ptrs = ring.next_slot()
        with torch.cuda.device(dev):
            st = _lib.stream_ptr(dev)
            rc = lib.milnce_peer_publish(_ptr(video_local), _ptr(text_local), b, K, D, ctypes.c_void_p(slot_ptr),
                                         ctypes.c_void_p(pack.data_ptr() + 4 * b), st)
            _lib.check(rc, "milnce_peer_publish")
            mark("publish")
            ring.barrier()      # every rank's slot is complete
            mark("barrier_published")
            arr = (ctypes.c_void_p * world)(*peer_ptrs)
            rc = lib.milnce_forward_shard_peers(arr, world, B, K, D, i0, b, ctypes.c_void_p(pack.data_ptr() - 4 * i0),
                                                ctypes.c_void_p(pack.data_ptr() + 8 * b), _ptr(ws), nbytes, st)
            _lib.check(rc, "milnce_forward_shard_peers")
        return pack, ws

    @staticmethod
    def backward_shard_push(ring, den, nom, grad_loss, world, rank, ws):
        """Gradient tiles + partial text gradient, each owner's rows written straight into that rank's receive
        area in peer-mapped memory (include/b200head.h, milnce_backward_shard_push)."""
        lib = _lib.load()
        b, K, D = ring.shape
        dev = ring.grad.device
        block = b * K * D * 2        # bytes of one (b*K, D) bf16 block; my block in q's area is block number `rank`
        arr = (ctypes.c_void_p * world)(*[p + rank * block for p in ring.grad_ptrs])
        with torch.cuda.device(dev):
            rc = lib.milnce_backward_shard_push(_ptr(den), _ptr(nom), _ptr(grad_loss), arr, world, world * b, K, D,
                                                rank * b, b, _ptr(ws), ws.numel(), _lib.stream_ptr(dev))
        _lib.check(rc, "milnce_backward_shard_push")

    @staticmethod
    def peer_reduce_dtext(ring, world, rank, out_dtype, pushed=False, out=None):
        """This rank's (b*K, D) text gradient.  pushed=False: the sum of its block of every rank's partial, read
        from the peers' memory; pushed=True: the sum of the W blocks the peers delivered into this rank's own
        receive area (include/b200head.h, milnce_peer_reduce_dtext).  The caller has passed ring.barrier()."""
        lib = _lib.load()
        b, K, D = ring.shape
        dev = ring.grad.device
        dt = out if out is not None else torch.empty(b * K, D, dtype=out_dtype, device=dev)
        if pushed:
            mine = ring.grad_ptrs[rank]
            arr = (ctypes.c_void_p * world)(*[mine + q * b * K * D * 2 for q in range(world)])
            rank = 0
        else:
            arr = (ctypes.c_void_p * world)(*ring.grad_ptrs)
        with torch.cuda.device(dev):
            rc = lib.milnce_peer_reduce_dtext(arr, world, rank, b, K, D, _ptr(dt),
                                              0 if out_dtype == torch.float32 else 1, _lib.stream_ptr(dev))
        _lib.check(rc, "milnce_peer_reduce_dtext")
        return dt

    @staticmethod
    def shard_combine_peers(ring, world):
        """(den[B], nom[B], loss) from the W statistics packs in the ranks' peer-readable memory; the caller has
        passed ring.barrier() after forward_shard_peers."""
        lib = _lib.load()
        b = ring.shape[0]
        dev = ring.pack.device
        B = world * b
        den = torch.empty(B, dtype=torch.float32, device=dev)
        nom = torch.empty(B, dtype=torch.float32, device=dev)
        loss = torch.empty(1, dtype=torch.float32, device=dev)
        arr = (ctypes.c_void_p * world)(*ring.pack_ptrs)
        with torch.cuda.device(dev):
            rc = lib.milnce_shard_combine_peers(arr, world, b, _ptr(den), _ptr(nom), _ptr(loss), _ptr(ring.combine_scratch),
                                                _lib.stream_ptr(dev))
        _lib.check(rc, "milnce_shard_combine_peers")
        return den, nom, loss.reshape(())

    @staticmethod
    def shard_combine(gathered, world, b):
        """(den[B], nom[B], loss) from the all-gathered (W, 2b + B) statistics of forward_shard."""
        lib = _lib.load()
        dev = gathered.device
        B = world * b
        den = torch.empty(B, dtype=torch.float32, device=dev)
        nom = torch.empty(B, dtype=torch.float32, device=dev)
        loss = torch.empty(1, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            rc = lib.milnce_shard_combine(_ptr(gathered), world, b, _ptr(den), _ptr(nom), _ptr(loss),
                                          _lib.stream_ptr(dev))
        _lib.check(rc, "milnce_shard_combine")
        return den, nom, loss.reshape(())

    @staticmethod
    def backward_shard(video, text, den, nom, grad_loss, i0, n, ws, parts=3, grad_dtype=torch.float32,
                       partial_dtype=torch.float32, dims=None, partial_out=None):
        """Returns (dvideo (n, D) in grad_dtype or None, dtext_partial (B*K, D) in partial_dtype or None).

        parts=1: gradient tiles + dtext_partial only; parts=2: dvideo only, from the workspace a
        parts=1 call left; parts=3: both.  dtext_partial is to be reduce-scattered (NCCL, or
        peer_reduce_dtext when `partial_out` is a buffer the peers can read)."""
        lib = _lib.load()
        if dims is not None:        # (B, K, D) given explicitly: video / text are then only the local shards
            B, K, D = dims          # and are not read (the workspace of the forward holds the operands)
            if ws is None:
                raise ValueError("explicit dims need the workspace of the forward")
        else:
            B, D = video.shape
            K = text.shape[0] // B
        code = _dtype_code(video)
        dev = video.device
        reuse = 1
        if ws is None:
            if parts == 2:
                raise ValueError("parts=2 needs the workspace of the parts=1 call")
            nbytes = lib.milnce_workspace_bytes(B, K, D, code, n)
            ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            reuse = 0
        dv = torch.empty(n, D, dtype=grad_dtype, device=dev) if parts & 2 else None
        dtp = None
        if parts & 1:
            if partial_out is not None:
                if partial_out.dtype != partial_dtype or partial_out.numel() != B * K * D:
                    raise ValueError("partial_out must hold (B*K, D) values of partial_dtype")
                dtp = partial_out.view(B * K, D)
            else:
                dtp = torch.empty(B * K, D, dtype=partial_dtype, device=dev)
        gcode = 0 if grad_dtype == torch.float32 else 1
        pcode = 0 if partial_dtype == torch.float32 else 1
        null = ctypes.c_void_p(0)
        with torch.cuda.device(dev):
            rc = lib.milnce_backward_shard(_ptr(video), _ptr(text), _ptr(den), _ptr(nom), _ptr(grad_loss),
                                           _ptr(dv) if dv is not None else null, gcode,
                                           _ptr(dtp) if dtp is not None else null, pcode, parts,
                                           B, K, D, code, i0, n, reuse, _ptr(ws), ws.numel(),
                                           _lib.stream_ptr(dev))
        _lib.check(rc, "milnce_backward_shard")
        return dv, dtp


def _validate(video_embd, text_embd):
    if video_embd.dim() != 2 or text_embd.dim() != 2:
        raise ValueError("video_embd must be (B, D) and text_embd (B*K, D)")
    if video_embd.shape[1] != text_embd.shape[1]:
        raise ValueError("embedding widths differ")
    b = video_embd.shape[0]
    if b == 0 or text_embd.shape[0] == 0 or text_embd.shape[0] % b != 0:
        # the reference's x.view(B, B, -1) (loss.py:12) raises here as well
        raise ValueError("text_embd rows must be a positive multiple of video_embd rows")
    if video_embd.dtype != text_embd.dtype:
        raise TypeError("video_embd and text_embd must share a dtype")
    if not (video_embd.is_cuda and text_embd.is_cuda):
        raise RuntimeError("MILNCELoss runs on CUDA tensors only (no CPU path)")
    if video_embd.device != text_embd.device:
        raise RuntimeError("video_embd and text_embd live on different devices")
    _dtype_code(video_embd)


class _MILNCEFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, video_embd, text_embd):
        _validate(video_embd, text_embd)
        v = video_embd.detach().contiguous()
        t = text_embd.detach().contiguous()
        loss, den, nom, ws = CudaBackend.forward(v, t, 0, v.shape[0])
        ctx.save_for_backward(v, t, den, nom)
        ctx.ws = ws
        return loss[0]

    @staticmethod
    def backward(ctx, grad_output):
        v, t, den, nom = ctx.saved_tensors
        g = grad_output.detach().reshape(1).to(torch.float32).contiguous()
        # gradients come back in the input dtype straight from the GEMM epilogues
        dv, dt = CudaBackend.backward(v, t, den, nom, g, 0, v.shape[0], ctx.ws, grad_dtype=v.dtype)
        ctx.ws = None
        return dv, dt


class MILNCELoss(torch.nn.Module):
    """``MILNCELoss()(video_embd, text_embd)`` - see reference loss.py:6-18.

    video_embd: (B, D); text_embd: (B*K, D) with the K candidates of a sample
    contiguous.  Returns ``mean_i(log sum over row i and column-group i of
    exp(S) - log sum_k exp(S[i, i, k]))`` as a float32 0-d tensor (the reference
    returns float32 for bf16 inputs too, because its mask is float32).
    """

    def __init__(self):
        super().__init__()

    def forward(self, video_embd, text_embd):
        return _MILNCEFunction.apply(video_embd, text_embd)


class GraphedMILNCEStep:
    """Loss and input gradients of :class:`MILNCELoss` for a fixed shape, replayed as ONE CUDA graph.

    A training loop that reads the loss on the host every step (the reference does,
    main_distributed.py:241) is launch-bound around a ~1 ms loss head: the next step's eight kernels
    are only submitted after the host has seen the previous loss.  Capturing forward + backward
    once and replaying it removes that submission time from every step.

        step = GraphedMILNCEStep(B, K, D)            # captures on the current device
        loss, dvideo, dtext = step(video, text)      # copies into the static inputs, replays
        # or write step.video / step.text directly (e.g. an H2D copy) and call step.replay()

    The returned tensors are the graph's static outputs: they are overwritten by the next call.
    Gradients are d loss / d input (grad_output = 1), in the input dtype.
    """

    def __init__(self, B, K, D, dtype=torch.bfloat16, device=None, warmup=3):
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.video = torch.zeros(B, D, dtype=dtype, device=dev).requires_grad_(True)
        self.text = torch.zeros(B * K, D, dtype=dtype, device=dev).requires_grad_(True)
        mod = MILNCELoss()
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):       # warm-up off the capture: builds/loads the library, primes allocators
            for _ in range(max(1, warmup)):
                self.video.grad = None
                self.text.grad = None
                mod(self.video, self.text).backward()
        torch.cuda.current_stream(dev).wait_stream(side)
        self.video.grad = None
        self.text.grad = None
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss = mod(self.video, self.text)
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
