"""Shared bodies of the module/model parity tests against tests/golden/model_*.npz (outputs of the reference's
own Python modules, see tests/golden/make_model_golden.py).  `dev` is 'cpu' (ops patched to the CPU stand-ins by
the caller) or a CUDA device (real kernels through the C-ABI)."""
import copy
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")

import sys
sys.path.insert(0, GOLD)
from fill import fill_state  # noqa: E402
from adaptpoint_b200.synth import make_clouds  # noqa: E402

# fp32 tolerance of north_star (1e-5 relative) is applied to op-level results; whole-network outputs and
# gradients accumulate rounding through ~40 layers of cuDNN/cuBLAS GEMMs in a different summation order than the
# CPU run that produced the goldens, so they are compared at 2e-4 / 2e-3 relative to the tensor's max magnitude.
TOL_OUT = 2e-4
TOL_GRAD = 2e-3


def strict_fp32():
    """cuDNN convolutions default to TF32 on Ampere+; parity is stated for fp32."""
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False


def load(name):
    return np.load(os.path.join(GOLD, f"model_{name}.npz"))


def rel_err(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)


def batch(B, N, seed, kind="shell"):
    pos = make_clouds(B, N, kind, seed)
    h = pos[:, :, 1:2] - pos[:, :, 1:2].amin(dim=1, keepdim=True)
    x = torch.cat([pos, h], -1).transpose(1, 2).contiguous()
    y = torch.randint(0, 15, (B,), generator=torch.Generator().manual_seed(seed + 1))
    return pos, x, y


def train_no_dropout(m):
    m.train()
    for sub in m.modules():
        if isinstance(sub, torch.nn.Dropout):
            sub.eval()


def check_names(names, g):
    assert [n for n, _ in names] == list(g["param_names"]), "state_dict names differ from the reference"


def check_grads(model, g, tol=TOL_GRAD):
    params = dict(model.named_parameters())
    for k in g.files:
        if k.startswith("grad::"):
            got = params[k[6:]].grad.detach().cpu().numpy()
            e = rel_err(got, g[k])
            assert e < tol, f"{k}: rel err {e:.2e}"


def case_cls_s(dev, fused=False):
    from adaptpoint_b200.models import build_model
    g = load("cls_s")
    model = build_model("pointnext-s", fused=fused)
    check_names(fill_state(model, 0), g)
    model.to(dev)
    pos, x, y = batch(int(g["B"]), int(g["N"]), int(g["seed"]))
    model.eval()
    with torch.no_grad():
        logits = model({"pos": pos.to(dev), "x": x.to(dev)})
    tol = 3e-2 if fused else TOL_OUT   # bf16 tensor-core MLP path: north_star tolerance 2e-2 per layer
    e = rel_err(logits.cpu().numpy(), g["logits_eval"])
    assert e < tol, f"eval logits rel err {e:.3e}"
    if fused:
        return
    train_no_dropout(model)
    pos_g = pos.to(dev).requires_grad_(True)
    x_g = x.to(dev).requires_grad_(True)
    logits_tr, loss = model.get_logits_loss({"pos": pos_g, "x": x_g}, y.to(dev))
    loss.backward()
    assert rel_err(logits_tr.detach().cpu().numpy(), g["logits_train"]) < TOL_OUT
    assert abs(loss.item() - float(g["loss"])) < 1e-4 * abs(float(g["loss"]))
    assert rel_err(pos_g.grad.cpu().numpy(), g["grad_pos"]) < TOL_GRAD
    assert rel_err(x_g.grad.cpu().numpy(), g["grad_x"]) < TOL_GRAD
    check_grads(model, g)


def case_seg_b(dev, fused=False):
    from adaptpoint_b200.models import build_model
    g = load("seg_b")
    model = build_model("pointnext-b-seg", fused=fused)
    check_names(fill_state(model, 1), g)
    model.to(dev)
    pos, x, _ = batch(int(g["B"]), int(g["N"]), int(g["seed"]), "ball")
    model.eval()
    with torch.no_grad():
        out = model({"pos": pos.to(dev), "x": x.to(dev)})
    tol = 3e-2 if fused else TOL_OUT
    e = rel_err(out.cpu().numpy(), g["seg_eval"])
    assert e < tol, f"eval seg logits rel err {e:.3e}"
    if fused:
        return
    train_no_dropout(model)
    x_g = x.to(dev).requires_grad_(True)
    out = model({"pos": pos.to(dev), "x": x_g})
    tgt = torch.randint(0, 13, (int(g["B"]), int(g["N"])), generator=torch.Generator().manual_seed(5)).to(dev)
    loss = torch.nn.functional.cross_entropy(out, tgt)
    loss.backward()
    assert rel_err(out.detach().cpu().numpy(), g["seg_train"]) < TOL_OUT
    assert abs(loss.item() - float(g["loss"])) < 1e-4 * abs(float(g["loss"]))
    assert rel_err(x_g.grad.cpu().numpy(), g["grad_x"]) < TOL_GRAD
    check_grads(model, g)


def case_groupers(dev):
    from adaptpoint_b200 import ops
    from adaptpoint_b200.layers import KNNGroup, QueryAndGroup
    g = load("groupers")
    pos, x, _ = batch(2, 600, int(g["seed"]), "ball")
    pos, x = pos.to(dev), x.to(dev)
    fidx = ops.furthest_point_sample(pos, 150).long()
    q = torch.gather(pos, 1, fidx.unsqueeze(-1).expand(-1, -1, 3)).contiguous()
    dp, fj = QueryAndGroup(0.2, 16, normalize_dp=True)(q, pos, x)
    assert rel_err(dp.cpu().numpy(), g["qg_dp"]) < 1e-6
    np.testing.assert_array_equal(fj.cpu().numpy(), g["qg_fj"])
    kdp, kfj = KNNGroup(12, normalize_dp=True)(q, pos, x)
    # the reference's kNN is torch.cdist + topk (tie order unspecified): compare as neighbour SETS per query
    got = np.sort(kfj.cpu().numpy(), axis=-1)
    want = np.sort(g["knn_fj"], axis=-1)
    assert (np.abs(got - want).max(axis=(1, 3)) < 1e-6).mean() > 0.995
    assert rel_err(np.sort(kdp.cpu().numpy(), -1), np.sort(g["knn_dp"], -1)) < 1e-4


def _noise(g, keys, dev):
    return {k: torch.from_numpy(g[k]).to(dev) for k in keys}


def case_generator(dev):
    from adaptpoint_b200.imitator import AdaptPoint_Augmentor
    g = load("generator")
    gen = AdaptPoint_Augmentor(w_num_anchor=4, w_sigma=0.5, w_R_range=10, w_S_range=3, w_T_range=0.25)
    check_names(fill_state(gen, 2), g)
    gen.to(dev).train()
    pos, _, _ = batch(int(g["B"]), int(g["N"]), int(g["seed"]))
    noise = _noise(g, ["gumbel", "dropout", "axis", "proj_axis"], dev)
    _, xyz_new = gen(pos.to(dev), noise=noise)
    got, want = xyz_new.detach().cpu().numpy(), g["xyz_new"]
    # a Gumbel-argmax flip on a near-tie logit would zero / un-zero single points: allow none on CPU, <=0.5% on GPU
    bad = (np.abs(got - want).max(-1) > 1e-4).mean()
    assert bad <= (0.0 if str(dev) == "cpu" else 0.005), f"{bad:.4f} of the points differ"
    (xyz_new * torch.from_numpy(g["w"]).to(dev)).sum().backward()
    if bad == 0.0:
        check_grads(gen, g, tol=5e-3)


def case_discriminator(dev):
    from adaptpoint_b200.imitator import PointDiscriminator1
    g = load("discriminator")
    dis = PointDiscriminator1(num_classes=15, normal_channel=False)
    check_names(fill_state(dis, 3), g)
    dis.to(dev).eval()
    pos, _, _ = batch(2, 512, int(g["seed"]))
    pos_g = pos.to(dev).requires_grad_(True)
    prob = dis(pos_g)
    prob.sum().backward()
    assert rel_err(prob.detach().cpu().numpy(), g["prob"]) < TOL_OUT
    assert rel_err(pos_g.grad.cpu().numpy(), g["grad_pos"]) < TOL_GRAD


def case_pointwolf(dev):
    from adaptpoint_b200.imitator import PointWOLF_classversion
    g = load("pointwolf")
    wolf = PointWOLF_classversion(w_num_anchor=4, w_sigma=0.5, w_R_range=10, w_S_range=3, w_T_range=0.25)
    pos, _, _ = batch(2, 512, int(g["seed"]))
    noise = _noise(g, ["dropout", "axis", "R", "S", "T", "proj_axis"], dev)
    _, out = wolf(pos.to(dev), noise=noise)
    assert rel_err(out.cpu().numpy(), g["out"]) < 2e-5


ALL = [case_cls_s, case_seg_b, case_groupers, case_generator, case_discriminator, case_pointwolf]
