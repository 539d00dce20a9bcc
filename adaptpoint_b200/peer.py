"""Small all-reduce over NVLink peer memory (csrc/peer.cu) for SyncBN statistics.

The reference converts every BatchNorm to torch SyncBatchNorm when world_size > 1
(examples/classification/main.py:27), one small NCCL collective per layer per direction.  `PeerAllreduce` replaces
those with one single-CTA kernel each (P2P stores + flags), replayable inside the step's CUDA graph; gradient
buckets (MBs) stay on NCCL.  One process per GPU, all GPUs of ONE node (CUDA IPC); anything else keeps NCCL.
"""
import ctypes
import os

import torch
import torch.distributed as dist

from . import _cabi

_CTX = {}          # id(process group) -> PeerAllreduce | None (None = tried, unavailable)


class PeerAllreduce:
    def __init__(self, group=None):
        if not (dist.is_available() and dist.is_initialized()):
            raise _cabi.ApbError("PeerAllreduce needs an initialised process group")
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        if self.world > 8:
            raise _cabi.ApbError("PeerAllreduce covers the <= 8 GPUs of one node")
        lib = _cabi.load()
        self.max_elems = lib.apb_peer_max_elems()
        self.device = torch.device("cuda", torch.cuda.current_device())
        mine = ctypes.c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        with torch.cuda.device(self.device):
            status = lib.apb_peer_ws_create(ctypes.byref(mine), handle)
        self._own = mine.value if status == 0 else None
        self._opened = []                       # peer mappings this process opened (closed by close())
        self._last = None                       # (stream handle, event, capturing) of the previous all-reduce
        try:
            handles = [None] * self.world
            dist.all_gather_object(handles, (self.rank, bytes(handle) if status == 0 else None, os.uname().nodename),
                                   group=self.group)
            if any(h[1] is None for h in handles):           # the same verdict on every rank
                raise _cabi.ApbError(f"apb_peer_ws_create failed on a rank (mine: status {status})")
            if len({h[2] for h in handles}) != 1:
                raise _cabi.ApbError("PeerAllreduce: ranks span several nodes")
            self._ptrs = (ctypes.c_void_p * self.world)()
            for r, hb, _ in handles:
                if r == self.rank:
                    self._ptrs[r] = self._own
                    continue
                p = ctypes.c_void_p()
                buf = (ctypes.c_ubyte * 64).from_buffer_copy(hb)
                with torch.cuda.device(self.device):
                    status = lib.apb_peer_ws_open(buf, ctypes.byref(p))
                if status != 0:
                    raise _cabi.ApbError(f"apb_peer_ws_open(rank {r}) failed (status {status})")
                self._ptrs[r] = p.value
                self._opened.append(p.value)
        except Exception:
            self.close()                        # no leaked cudaMalloc / IPC mappings on the failure path
            raise
        self.calls = 0                          # enable() lines the ranks up before the first store can land

    def close(self):
        """Unmap the peers' workspaces and free this rank's own.  Collective in effect: every rank must have stopped
        issuing all-reduces (call after a device synchronize + barrier)."""
        lib = _cabi.load()
        with torch.cuda.device(self.device):
            for p in self._opened:
                lib.apb_peer_ws_close(ctypes.c_void_p(p), 0)
            self._opened = []
            if self._own is not None:
                lib.apb_peer_ws_close(ctypes.c_void_p(self._own), 1)
                self._own = None

    def _serialise(self, cur):
        """The two-parity protocol of csrc/peer.cu has ONE sequence counter per workspace: two all-reduces of a rank
        must never be in flight together.  A call issued on a different stream than the previous one is therefore
        ordered after it with an event (same-stream calls are ordered already)."""
        capturing = torch.cuda.is_current_stream_capturing()
        last = self._last
        if last is not None and last[0] != cur.cuda_stream and last[2] == capturing:
            cur.wait_event(last[1])
        return capturing

    def forget_order(self):
        """After a device-wide synchronize (capture_graph does one) no earlier all-reduce is in flight."""
        self._last = None

    def supports(self, t):
        return (t.is_cuda and t.device == self.device and t.is_contiguous() and t.numel() <= self.max_elems
                and t.dtype in (torch.float32, torch.float64))

    def allreduce_(self, t):
        """In-place sum of `t` over the group on the current stream."""
        lib = _cabi.load()
        if self._own is None:
            raise _cabi.ApbError("PeerAllreduce used after close()")
        with torch.cuda.device(self.device):
            cur = torch.cuda.current_stream(self.device)
            capturing = self._serialise(cur)
            stream = ctypes.c_void_p(cur.cuda_stream)
            trace = _cabi.TRACE
            if trace is not None:          # bench.py's per-launch CUDA-event trace (eager steps only)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(cur)
            status = lib.apb_peer_allreduce(ctypes.c_void_p(t.data_ptr()), t.numel(),
                                            1 if t.dtype == torch.float64 else 0, self._ptrs, self.world, self.rank,
                                            stream)
            if trace is not None:
                e1.record(cur)
                trace.append(("apb_peer_allreduce", (int(t.numel()), self.world), e0, e1))
            ev = torch.cuda.Event()
            ev.record(cur)
            self._last = (cur.cuda_stream, ev, capturing)
        if status != 0:
            raise _cabi.ApbError(f"apb_peer_allreduce failed: {lib.apb_error_string(status).decode()} ({status})")
        _cabi.LAUNCHES += 1
        self.calls += 1
        return t


def enable(group=None):
    """Create (once) the peer context of `group`; returns it, or None when the group is not one node of CUDA ranks
    (gloo/CPU tests, world size 1).  Collective: every rank of the group must call it."""
    if not (dist.is_available() and dist.is_initialized()):
        return None
    g = group if group is not None else dist.group.WORLD
    key = id(g)
    if key in _CTX:
        return _CTX[key]
    ctx = None
    if (dist.get_world_size(g) > 1 and torch.cuda.is_available() and dist.get_backend(g) == "nccl"
            and os.environ.get("APB_PEER_ALLREDUCE", "1") != "0"):
        err = None
        try:
            ctx = PeerAllreduce(g)
        except _cabi.ApbError as e:      # e.g. no peer access between two of the GPUs
            ctx, err = None, e
        # all-or-nothing: a rank that could not map its peers must not leave the others on the peer path; this
        # collective is also the barrier after which every workspace is zeroed and mapped everywhere
        ok = torch.tensor([0 if ctx is None else 1], device=torch.device("cuda", torch.cuda.current_device()))
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=g)
        if int(ok.item()) == 0:
            if err is not None:
                import warnings
                warnings.warn(f"peer all-reduce unavailable ({err}); SyncBN statistics stay on NCCL")
            ctx = None
    _CTX[key] = ctx
    return ctx


def disable(group=None):
    """Close and forget the peer context of `group` (every rank, after a synchronize + barrier)."""
    g = group if group is not None else (dist.group.WORLD if dist.is_initialized() else None)
    ctx = _CTX.pop(id(g), None)
    if ctx is not None:
        ctx.close()


def get(group=None):
    """The peer context `enable` created for `group`, or None."""
    if not _CTX:
        return None
    g = group if group is not None else (dist.group.WORLD if dist.is_initialized() else None)
    return _CTX.get(id(g))
