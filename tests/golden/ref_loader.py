"""TEST INFRASTRUCTURE ONLY (runs only where /root/reference exists, i.e. in the build container).

Imports the reference's OWN Python modules for the hot path --
  openpoints/models/layers/{activation,norm,conv,group,subsample,upsampling}.py
  openpoints/models/backbone/pointnext.py, openpoints/models/classification/cls_base.py,
  openpoints/models/segmentation/base_seg.py, openpoints/loss/build.py, openpoints/utils/registry.py
  openpoints/models_adaptpoint/{generator_component4_15,point_discriminator}.py, openpoints/online_aug/pointwolf.py
-- on the CPU, WITHOUT importing the `openpoints` package (its __init__ chain needs easydict, multimethod, h5py,
termcolor, ... which are not installed).  The files are executed unmodified from where they lie;
only their environment is stubbed:
  * `openpoints.cpp.pointnet2_cuda` (the compiled extension) -> a shim whose *_wrapper functions fill the
    caller-allocated tensors from oracle/liboracle.so (the C restatement that is itself pinned against the
    compiled reference CUDA kernels, tests/golden/case*.npz);
  * legacy `torch.cuda.FloatTensor/IntTensor(...)` constructors -> CPU tensors; `device='cuda'` -> CPU;
  * `easydict`, `distutils.command.build`, `sklearn.decomposition` -> minimal stand-ins.
Used by make_model_golden.py to produce tests/golden/model_*.npz.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference/openpoints"
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


class _EasyDict(dict):
    def __init__(self, d=None, **kw):
        super().__init__()
        for k, v in dict(d or {}, **kw).items():
            self[k] = v

    def __setitem__(self, k, v):
        if isinstance(v, dict) and not isinstance(v, _EasyDict):
            v = _EasyDict(v)
        super().__setitem__(k, v)

    __setattr__ = __setitem__

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            raise AttributeError(k)

    def pop(self, k, *a):
        return super().pop(k, *a)


def _np(t):
    return t.detach().contiguous().numpy()


class _Pointnet2Shim:
    """pointnet2_batch_cuda's 9 functions (pointnet2_api.cpp:10-24) over the CPU oracle."""

    def __init__(self):
        from oracle import cpu
        self.o = cpu

    def furthest_point_sampling_wrapper(self, B, N, m, xyz, temp, out):
        out.copy_(torch.from_numpy(self.o.fps(_np(xyz), m)))

    def ball_query_wrapper(self, B, N, M, radius, nsample, new_xyz, xyz, idx):
        idx.copy_(torch.from_numpy(self.o.ball_query(radius, nsample, _np(xyz), _np(new_xyz))))

    def group_points_wrapper(self, B, C, N, M, K, points, idx, out):
        out.copy_(torch.from_numpy(self.o.group(_np(points), _np(idx))))

    def group_points_grad_wrapper(self, B, C, N, M, K, grad_out, idx, grad_points):
        grad_points.add_(torch.from_numpy(self.o.group_grad(_np(grad_out), _np(idx), N)))

    def gather_points_wrapper(self, B, C, N, m, points, idx, out):
        out.copy_(torch.from_numpy(self.o.gather(_np(points), _np(idx))))

    def gather_points_grad_wrapper(self, B, C, N, m, grad_out, idx, grad_points):
        grad_points.add_(torch.from_numpy(self.o.gather_grad(_np(grad_out), _np(idx), N)))

    def three_nn_wrapper(self, B, n, m, unknown, known, dist2, idx):
        d2, i = self.o.three_nn(_np(unknown), _np(known))
        dist2.copy_(torch.from_numpy(d2))
        idx.copy_(torch.from_numpy(i))

    def three_interpolate_wrapper(self, B, c, m, n, points, idx, weight, out):
        out.copy_(torch.from_numpy(self.o.three_interp(_np(points), _np(idx), _np(weight))))

    def three_interpolate_grad_wrapper(self, B, c, n, m, grad_out, idx, weight, grad_points):
        grad_points.add_(torch.from_numpy(self.o.three_interp_grad(_np(grad_out), _np(idx), _np(weight), m)))


def _pkg(name):
    m = types.ModuleType(name)
    m.__path__ = []
    sys.modules[name] = m
    return m


def _load(name, relpath):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, relpath))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    parent, _, leaf = name.rpartition(".")
    if parent in sys.modules:
        setattr(sys.modules[parent], leaf, mod)
    return mod


_loaded = None


def load_reference():
    """Returns a namespace with the reference classes/functions (CPU-runnable)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not os.path.isdir(REF):
        raise RuntimeError("/root/reference is not available here")

    # -- environment stubs ------------------------------------------------------------------------------
    ed = types.ModuleType("easydict")
    ed.EasyDict = _EasyDict
    sys.modules["easydict"] = ed
    if "distutils" not in sys.modules:
        _pkg("distutils")
    _pkg("distutils.command")
    b = types.ModuleType("distutils.command.build")
    b.build = object
    sys.modules["distutils.command.build"] = b
    sk = sys.modules.get("sklearn.decomposition")
    if sk is None:
        _pkg("sklearn")
        sk = types.ModuleType("sklearn.decomposition")
        sk.FactorAnalysis = object
        sys.modules["sklearn.decomposition"] = sk

    torch.cuda.FloatTensor = lambda *s, device=None: torch.empty(*s, dtype=torch.float32)
    torch.cuda.IntTensor = lambda *s, device=None: torch.empty(*s, dtype=torch.int32)
    _zeros = torch.zeros

    def zeros(*a, **k):
        if k.get("device") == "cuda":
            k["device"] = "cpu"
        return _zeros(*a, **k)
    torch.zeros = zeros
    torch.Tensor.cuda = lambda self, *a, **k: self
    torch.nn.Module.cuda = lambda self, *a, **k: self

    # -- package skeleton -------------------------------------------------------------------------------
    op = _pkg("openpoints")
    cpp = _pkg("openpoints.cpp")
    shim = _Pointnet2Shim()
    cpp.pointnet2_cuda = shim
    pb = _pkg("openpoints.cpp.pointnet2_batch")
    pb.pointnet2_cuda = shim
    utils = _pkg("openpoints.utils")
    registry = _load("openpoints.utils.registry", "utils/registry.py")
    utils.registry = registry
    for n in ("get_missing_parameters_message", "get_unexpected_parameters_message", "load_checkpoint"):
        setattr(utils, n, lambda *a, **k: "")
    _pkg("openpoints.loss")
    loss_build = _load("openpoints.loss.build", "loss/build.py")
    sys.modules["openpoints.loss"].build_criterion_from_cfg = loss_build.build_criterion_from_cfg

    _pkg("openpoints.models")
    layers = _pkg("openpoints.models.layers")
    act = _load("openpoints.models.layers.activation", "models/layers/activation.py")
    norm = _load("openpoints.models.layers.norm", "models/layers/norm.py")
    conv = _load("openpoints.models.layers.conv", "models/layers/conv.py")
    for n in ("create_convblock1d", "create_convblock2d", "create_linearblock"):
        setattr(layers, n, getattr(conv, n))
    layers.create_act, layers.create_norm = act.create_act, norm.create_norm
    group = _load("openpoints.models.layers.group", "models/layers/group.py")
    sub = _load("openpoints.models.layers.subsample", "models/layers/subsample.py")
    up = _load("openpoints.models.layers.upsampling", "models/layers/upsampling.py")
    for n in ("torch_grouping_operation", "grouping_operation", "gather_operation", "create_grouper",
              "get_aggregation_feautres", "ball_query"):
        setattr(layers, n, getattr(group, n))
    for n in ("random_sample", "furthest_point_sample", "fps"):
        setattr(layers, n, getattr(sub, n))
    for n in ("three_interpolate", "three_nn", "three_interpolation"):
        setattr(layers, n, getattr(up, n))
    # CHANNEL_MAP lives in local_aggregation.py:13-29 (a file whose other imports are outside the path)
    src = open(os.path.join(REF, "models/layers/local_aggregation.py")).read()
    start = src.index("CHANNEL_MAP = {")
    end = src.index("}", start) + 1
    ns = {}
    exec(src[start:end], ns)
    layers.CHANNEL_MAP = ns["CHANNEL_MAP"]

    mbuild = _load("openpoints.models.build", "models/build.py")
    sys.modules["openpoints.models"].build = mbuild
    _pkg("openpoints.models.backbone")
    pointnext = _load("openpoints.models.backbone.pointnext", "models/backbone/pointnext.py")
    _pkg("openpoints.models.classification")
    cls_base = _load("openpoints.models.classification.cls_base", "models/classification/cls_base.py")
    _pkg("openpoints.models.segmentation")
    base_seg = _load("openpoints.models.segmentation.base_seg", "models/segmentation/base_seg.py")

    _pkg("openpoints.models_adaptpoint")
    _load("openpoints.models_adaptpoint.build", "models_adaptpoint/build.py")
    gen = _load("openpoints.models_adaptpoint.generator_component4_15",
                "models_adaptpoint/generator_component4_15.py")
    dis = _load("openpoints.models_adaptpoint.point_discriminator", "models_adaptpoint/point_discriminator.py")
    _pkg("openpoints.online_aug")
    wolf = _load("openpoints.online_aug.pointwolf", "online_aug/pointwolf.py")

    _loaded = types.SimpleNamespace(
        EasyDict=_EasyDict, build_model_from_cfg=mbuild.build_model_from_cfg, pointnext=pointnext,
        cls_base=cls_base, base_seg=base_seg, group=group, subsample=sub, upsampling=up,
        AdaptPoint_Augmentor=gen.AdaptPoint_Augmentor, PointDiscriminator1=dis.PointDiscriminator1,
        PointWOLF_classversion=wolf.PointWOLF_classversion, generator=gen, SmoothCrossEntropy=loss_build.SmoothCrossEntropy)
    return _loaded
