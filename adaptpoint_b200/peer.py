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
        self.calls = 0                          # enable() lines the ranks up before the first store can land

    def supports(self, t):
        return (t.is_cuda and t.device == self.device and t.is_contiguous() and t.numel() <= self.max_elems
                and t.dtype in (torch.float32, torch.float64))

    def allreduce_(self, t):
        """In-place sum of `t` over the group on the current stream."""
        lib = _cabi.load()
        with torch.cuda.device(self.device):
            stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
            status = lib.apb_peer_allreduce(ctypes.c_void_p(t.data_ptr()), t.numel(),
                                            1 if t.dtype == torch.float64 else 0, self._ptrs, self.world, self.rank,
                                            stream)
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


def get(group=None):
    """The peer context `enable` created for `group`, or None."""
    if not _CTX:
        return None
    g = group if group is not None else (dist.group.WORLD if dist.is_initialized() else None)
    return _CTX.get(id(g))
