"""TEST INFRASTRUCTURE ONLY -- Python front-end of the parity oracle.

* `oracle.cpu`  : numpy wrappers over liboracle.so, the plain-C CPU restatement of the reference's
                  pointnet2_batch algorithms (oracle/pointnet2_oracle.c; every function cites the
                  reference file:line it follows).
* `oracle.refcuda`: torch wrappers over oracle/_ref/libpointnet2_ref.so, the UNMODIFIED reference CUDA
                  kernels compiled from /root/reference (oracle/Makefile).  Only usable on a GPU box.

Nothing under adaptpoint_b200/ imports this package.  Allowed importers: tests/,
__graft_entry__.smoke(), bench.py's cpu_baseline / --impl reference legs.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")
_REF = os.path.join(_HERE, "_ref", "libpointnet2_ref.so")


def build(verbose=False):
    """Compile liboracle.so (gcc) and, when /root/reference is present, _ref/libpointnet2_ref.so (nvcc)."""
    out = subprocess.run(["make", "-C", _HERE, "all"], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + out.stdout + out.stderr)
    if verbose:
        print(out.stdout)


def _ensure_cpu_lib():
    src = os.path.join(_HERE, "pointnet2_oracle.c")
    if (not os.path.exists(_LIB)) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        out = subprocess.run(["make", "-C", _HERE, "liboracle.so"], capture_output=True, text=True)
        if out.returncode != 0:
            raise RuntimeError("oracle build failed:\n" + out.stdout + out.stderr)
    return ctypes.CDLL(_LIB)


_f32 = np.float32
_i32 = np.int32


def _fp(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class _Cpu:
    def __init__(self):
        self._lib = None

    @property
    def lib(self):
        if self._lib is None:
            self._lib = _ensure_cpu_lib()
            self._lib.orc_opt_n_threads.restype = ctypes.c_int
        return self._lib

    def opt_n_threads(self, n):
        return self.lib.orc_opt_n_threads(ctypes.c_int(n))

    def fps(self, xyz, m, temp=None):
        xyz = np.ascontiguousarray(xyz, _f32)
        B, N, _ = xyz.shape
        if temp is None:
            temp = np.full((B, N), 1e10, _f32)
        idx = np.zeros((B, m), _i32)
        self.lib.orc_fps(B, N, m, _fp(xyz), _fp(temp), _fp(idx))
        return idx

    def ball_query(self, radius, nsample, xyz, new_xyz):
        xyz = np.ascontiguousarray(xyz, _f32)
        new_xyz = np.ascontiguousarray(new_xyz, _f32)
        B, N, _ = xyz.shape
        M = new_xyz.shape[1]
        idx = np.zeros((B, M, nsample), _i32)
        self.lib.orc_ball_query(B, N, M, ctypes.c_float(radius), nsample, _fp(new_xyz), _fp(xyz), _fp(idx))
        return idx

    def knn(self, k, xyz, new_xyz):
        xyz = np.ascontiguousarray(xyz, _f32)
        new_xyz = np.ascontiguousarray(new_xyz, _f32)
        B, N, _ = xyz.shape
        M = new_xyz.shape[1]
        idx = np.zeros((B, M, k), _i32)
        d2 = np.zeros((B, M, k), _f32)
        self.lib.orc_knn(B, N, M, k, _fp(new_xyz), _fp(xyz), _fp(idx), _fp(d2))
        return idx, d2

    def group(self, points, idx):
        points = np.ascontiguousarray(points, _f32)
        idx = np.ascontiguousarray(idx, _i32)
        B, C, N = points.shape
        _, M, K = idx.shape
        out = np.zeros((B, C, M, K), _f32)
        self.lib.orc_group(B, C, N, M, K, _fp(points), _fp(idx), _fp(out))
        return out

    def group_grad(self, grad_out, idx, N):
        grad_out = np.ascontiguousarray(grad_out, _f32)
        idx = np.ascontiguousarray(idx, _i32)
        B, C, M, K = grad_out.shape
        gp = np.zeros((B, C, N), _f32)
        self.lib.orc_group_grad(B, C, N, M, K, _fp(grad_out), _fp(idx), _fp(gp))
        return gp

    def gather(self, points, idx):
        return self.group(points, np.ascontiguousarray(idx, _i32)[:, :, None])[..., 0]

    def gather_grad(self, grad_out, idx, N):
        return self.group_grad(np.ascontiguousarray(grad_out, _f32)[..., None], np.ascontiguousarray(idx, _i32)[:, :, None], N)

    def three_nn(self, unknown, known):
        unknown = np.ascontiguousarray(unknown, _f32)
        known = np.ascontiguousarray(known, _f32)
        B, n, _ = unknown.shape
        m = known.shape[1]
        d2 = np.zeros((B, n, 3), _f32)
        idx = np.zeros((B, n, 3), _i32)
        self.lib.orc_three_nn(B, n, m, _fp(unknown), _fp(known), _fp(d2), _fp(idx))
        return d2, idx

    def three_interp(self, points, idx, weight):
        points = np.ascontiguousarray(points, _f32)
        idx = np.ascontiguousarray(idx, _i32)
        weight = np.ascontiguousarray(weight, _f32)
        B, C, m = points.shape
        n = idx.shape[1]
        out = np.zeros((B, C, n), _f32)
        self.lib.orc_three_interp(B, C, m, n, _fp(points), _fp(idx), _fp(weight), _fp(out))
        return out

    def three_interp_grad(self, grad_out, idx, weight, m):
        grad_out = np.ascontiguousarray(grad_out, _f32)
        idx = np.ascontiguousarray(idx, _i32)
        weight = np.ascontiguousarray(weight, _f32)
        B, C, n = grad_out.shape
        gp = np.zeros((B, C, m), _f32)
        self.lib.orc_three_interp_grad(B, C, n, m, _fp(grad_out), _fp(idx), _fp(weight), _fp(gp))
        return gp


cpu = _Cpu()


class _RefCuda:
    """The reference's own CUDA kernels (default stream); torch CUDA tensors in, torch tensors out."""

    def __init__(self):
        self._lib = None

    def available(self):
        return os.path.exists(_REF)

    @property
    def lib(self):
        if self._lib is None:
            if not os.path.exists(_REF):
                raise RuntimeError(f"{_REF} missing: run `make -C oracle ref` where /root/reference exists")
            self._lib = ctypes.CDLL(_REF)
        return self._lib

    @staticmethod
    def _p(t):
        return ctypes.c_void_p(t.data_ptr())

    def _sync(self):
        import torch
        torch.cuda.synchronize()

    def fps(self, xyz, m):
        import torch
        B, N, _ = xyz.shape
        temp = torch.full((B, N), 1e10, dtype=torch.float32, device=xyz.device)
        idx = torch.zeros((B, m), dtype=torch.int32, device=xyz.device)
        self._sync()
        self.lib.ref_fps(B, N, m, self._p(xyz), self._p(temp), self._p(idx))
        self._sync()
        return idx

    def ball_query(self, radius, nsample, xyz, new_xyz):
        import torch
        B, N, _ = xyz.shape
        M = new_xyz.shape[1]
        idx = torch.zeros((B, M, nsample), dtype=torch.int32, device=xyz.device)
        self._sync()
        self.lib.ref_ball_query(B, N, M, ctypes.c_float(radius), nsample, self._p(new_xyz), self._p(xyz), self._p(idx))
        self._sync()
        return idx

    def group(self, points, idx):
        import torch
        B, C, N = points.shape
        _, M, K = idx.shape
        out = torch.zeros((B, C, M, K), dtype=torch.float32, device=points.device)
        self._sync()
        self.lib.ref_group(B, C, N, M, K, self._p(points), self._p(idx), self._p(out))
        self._sync()
        return out

    def group_grad(self, grad_out, idx, N):
        import torch
        B, C, M, K = grad_out.shape
        gp = torch.zeros((B, C, N), dtype=torch.float32, device=grad_out.device)
        self._sync()
        self.lib.ref_group_grad(B, C, N, M, K, self._p(grad_out), self._p(idx), self._p(gp))
        self._sync()
        return gp

    def gather(self, points, idx):
        import torch
        B, C, N = points.shape
        m = idx.shape[1]
        out = torch.zeros((B, C, m), dtype=torch.float32, device=points.device)
        self._sync()
        self.lib.ref_gather(B, C, N, m, self._p(points), self._p(idx), self._p(out))
        self._sync()
        return out

    def gather_grad(self, grad_out, idx, N):
        import torch
        B, C, m = grad_out.shape
        gp = torch.zeros((B, C, N), dtype=torch.float32, device=grad_out.device)
        self._sync()
        self.lib.ref_gather_grad(B, C, N, m, self._p(grad_out), self._p(idx), self._p(gp))
        self._sync()
        return gp

    def three_nn(self, unknown, known):
        import torch
        B, n, _ = unknown.shape
        m = known.shape[1]
        d2 = torch.zeros((B, n, 3), dtype=torch.float32, device=unknown.device)
        idx = torch.zeros((B, n, 3), dtype=torch.int32, device=unknown.device)
        self._sync()
        self.lib.ref_three_nn(B, n, m, self._p(unknown), self._p(known), self._p(d2), self._p(idx))
        self._sync()
        return d2, idx

    def three_interp(self, points, idx, weight):
        import torch
        B, C, m = points.shape
        n = idx.shape[1]
        out = torch.zeros((B, C, n), dtype=torch.float32, device=points.device)
        self._sync()
        self.lib.ref_three_interp(B, C, m, n, self._p(points), self._p(idx), self._p(weight), self._p(out))
        self._sync()
        return out

    def three_interp_grad(self, grad_out, idx, weight, m):
        import torch
        B, C, n = grad_out.shape
        gp = torch.zeros((B, C, m), dtype=torch.float32, device=grad_out.device)
        self._sync()
        self.lib.ref_three_interp_grad(B, C, n, m, self._p(grad_out), self._p(idx), self._p(weight), self._p(gp))
        self._sync()
        return gp


refcuda = _RefCuda()
