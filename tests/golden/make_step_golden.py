"""Generates tests/golden/model_step.npz: ONE AdaptPoint training iteration -- the body of `train_gan`
(examples/classification/train_autoaug.py:131-207) followed by the body of `train_one_epoch` (:474-506) on the
generated batch -- run through the REFERENCE's own modules (generator_component4_15.AdaptPoint_Augmentor,
point_discriminator.PointDiscriminator1, BaseCls/PointNeXt-S, online_aug.PointWOLF_classversion) and its own
`get_feedback_loss_ver1` (openpoints/function_adaptpoint/ganloss_cls.py:31-65), loaded by ref_loader.py from
/root/reference (CPU, native ops served by the C oracle).  The two loop bodies are restated here statement by
statement because the loop functions themselves need a DataLoader, tqdm, h5py and a TensorBoard writer; every
tensor operation inside them is the reference's.

All noise is recorded so that adaptpoint_b200.train.AdaptPointTrainer can be driven with the same draws:
PointWOLF / imitator draws (CPU generator), Gumbel noise, the resampling subset.  Dropout layers are disabled in
both runs (their draws are device-specific).

Run in the build container:  python tests/golden/make_step_golden.py
"""
import copy
import os
import sys
import types

import numpy as np
import torch
import torch._dynamo  # noqa: F401  (imported by torch.optim; must precede ref_loader's sklearn stub)

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import ref_loader  # noqa: E402
from fill import fill_state  # noqa: E402
from adaptpoint_b200.synth import make_clouds  # noqa: E402

B, N, SEED, S = 4, 2048, 21, 4321
GNAMES = {
    "generator": ["predict_prob_layer.embedding.net.0.weight", "predict_prob_layer.head.prob_head.0.weight",
                  "predict_prob_layer.fuse_masking.0.weight", "predict_prob_layer.extract_feat_list.0.net.0.weight"],
    "discriminator": ["sa1.mlp_convs.1.parametrizations.weight.original", "fc3.parametrizations.weight.original",
                      "prob_head.0.parametrizations.weight.original"],
    "classifier": ["encoder.encoder.0.0.convs.0.0.weight", "encoder.encoder.1.0.convs.1.0.weight",
                   "encoder.encoder.3.0.convs.0.0.weight", "prediction.head.2.0.weight"],
}


def no_dropout(m):
    for sub in m.modules():
        if isinstance(sub, torch.nn.Dropout):
            sub.p = 0.0


def grad_pack(tag, module, out):
    total = 0.0
    for n, p in module.named_parameters():
        if p.grad is not None:
            total += float(p.grad.double().pow(2).sum())
    out[f"gnorm::{tag}"] = total ** 0.5
    sd = dict(module.named_parameters())
    for n in GNAMES[tag]:
        out[f"grad::{tag}::{n}"] = sd[n].grad.detach().numpy().copy()


def main():
    R = ref_loader.load_reference()
    E = R.EasyDict
    ref_loader._pkg("openpoints.function_adaptpoint")
    gl = ref_loader._load("openpoints.function_adaptpoint.ganloss_cls", "function_adaptpoint/ganloss_cls.py")
    from adaptpoint_b200.models import CONFIGS
    from adaptpoint_b200.train import ADAPTPOINT_CFG

    mcfg = copy.deepcopy(CONFIGS["pointnext-s"])
    mcfg.pop("kind")
    mcfg["NAME"] = "BaseCls"
    model = R.build_model_from_cfg(E(mcfg))
    fill_state(model, 0)
    generator = R.AdaptPoint_Augmentor(**ADAPTPOINT_CFG["adaptmodel_gan"])
    fill_state(generator, 2)
    discriminator = R.PointDiscriminator1(**ADAPTPOINT_CFG["adaptmodel_dis"])
    fill_state(discriminator, 3)
    for m in (model, generator, discriminator):
        no_dropout(m)
    ap = ADAPTPOINT_CFG["adaptpoint_params"]
    cfg = E(dict(criterion_args=dict(NAME="SmoothCrossEntropy", label_smoothing=ADAPTPOINT_CFG["label_smoothing"]),
                 adaptpoint_params=ap, epochs=ADAPTPOINT_CFG["epochs"], model=dict(in_channels=4), num_points=1024,
                 feedbackloss_ratio=1, grad_norm_clip=ADAPTPOINT_CFG["grad_norm_clip"]))
    optimizer_G = torch.optim.Adam(generator.parameters(), lr=ap["lr_generator"], betas=(ap["b1"], ap["b2"]))
    optimizer_D = torch.optim.Adam(discriminator.parameters(), lr=ap["lr_discriminator"], betas=(ap["b1"], ap["b2"]))
    optimizer = torch.optim.AdamW(model.parameters(), lr=ADAPTPOINT_CFG["lr"], weight_decay=ADAPTPOINT_CFG["weight_decay"])
    criterion_gan = torch.nn.BCELoss()
    pointwolf = R.PointWOLF_classversion(**ADAPTPOINT_CFG["pointwolf"])
    summary = types.SimpleNamespace(train_iter_num=0)
    writer = types.SimpleNamespace(add_scalar=lambda *a, **k: None)
    epoch = 0

    pos = make_clouds(B, N, "shell", SEED)
    h = pos[:, :, 1:2] - pos[:, :, 1:2].amin(dim=1, keepdim=True)
    loader_x = torch.cat([pos, h], -1).contiguous()                       # data['x'] of the loader: (B, N, 4)
    label = torch.randint(0, 15, (B,), generator=torch.Generator().manual_seed(SEED + 1))
    M = ADAPTPOINT_CFG["adaptmodel_gan"]["w_num_anchor"]

    # ---- record the draws of one iteration, in the reference's order (pointwolf.py:113-131,96;
    # generator_component4_15.py:714 gumbel, :245-247 dropout/axes, :221 projection axis) -------------------------
    def axes(n):
        return ((torch.randint(1, 8, (B, n))[:, :, None] & (1 << torch.arange(3))) > 0).int()
    torch.manual_seed(S)
    wolf = dict(dropout=torch.bernoulli(torch.Tensor(B, M, 3).uniform_(0, 1)), axis=axes(M),
                R=torch.FloatTensor(B, M, 3).uniform_(-10, 10), S=torch.FloatTensor(B, M, 3).uniform_(1., 3),
                T=torch.FloatTensor(B, M, 3).uniform_(-0.25, 0.25), proj_axis=axes(1))
    gumbel = -torch.empty(B, N, 2).exponential_().log()
    gen_noise = dict(dropout=torch.bernoulli(torch.Tensor(B, M, 3).uniform_(0, 1)), axis=axes(M), proj_axis=axes(1))
    np.random.seed(S)
    keep = np.random.choice(1200, 1024, False)

    out = {}
    # ================= train_gan body (train_autoaug.py:131-207) ==================================================
    torch.manual_seed(S)
    generator.train()
    discriminator.train()
    model.eval()
    points = loader_x.clone()
    points_clone = points.clone()
    input_pointcloud = points[:, :, :3].contiguous()
    _, pointcloud_pointwolf = pointwolf(input_pointcloud)
    real_label = torch.full((B, 1), 0.9, requires_grad=True)
    fake_label = torch.full((B, 1), 0.1, requires_grad=True)
    _, gen_imgs = generator(input_pointcloud)
    g_loss_raw = criterion_gan(discriminator(gen_imgs), real_label)
    points[:, :, :3] = gen_imgs
    data_fake = {"pos": points[:, :, :3].contiguous(), "y": label,
                 "x": points[:, :, :cfg.model.in_channels].transpose(1, 2).contiguous()}
    data_real = {"pos": points_clone[:, :, :3].contiguous(), "y": label,
                 "x": points_clone[:, :, :cfg.model.in_channels].transpose(1, 2).contiguous()}
    feedback_loss = gl.get_feedback_loss_ver1(cfg=cfg, model_pointcloud=model, data_real=data_real,
                                              data_fake=data_fake, epoch=epoch, summary=summary, writer=writer)
    g_loss = g_loss_raw + feedback_loss * cfg.feedbackloss_ratio
    optimizer_G.zero_grad()
    g_loss.backward()
    grad_pack("generator", generator, out)
    optimizer_G.step()
    real_loss = criterion_gan(discriminator(input_pointcloud), real_label)
    fake_loss = criterion_gan(discriminator(gen_imgs.detach()), fake_label)
    d_loss = (real_loss + fake_loss) / 2
    optimizer_D.zero_grad()
    d_loss.backward()
    grad_pack("discriminator", discriminator, out)
    optimizer_D.step()
    model.zero_grad()
    fake_points = points.detach()                       # what Form_dataset_cls hands to the classifier loader

    # ================= train_one_epoch body (train_autoaug.py:474-506) ============================================
    model.train()
    pts = fake_points
    target = label
    fps_idx = R.subsample.furthest_point_sample(pts[:, :, :3].contiguous(), 1200)
    fps_idx = fps_idx[:, keep]
    pts = torch.gather(pts, 1, fps_idx.unsqueeze(-1).long().expand(-1, -1, pts.shape[-1]))
    data = {"pos": pts[:, :, :3].contiguous(), "x": pts[:, :, :cfg.model.in_channels].transpose(1, 2).contiguous()}
    logits, loss = model.get_logits_loss(data, target)
    loss.backward()
    grad_pack("classifier", model, out)             # before clipping, as our probe sees them
    torch.nn.utils.clip_grad_norm_(model.parameters(), cfg.grad_norm_clip, norm_type=2)
    optimizer.step()
    model.zero_grad()

    path = os.path.join(HERE, "model_step.npz")
    np.savez_compressed(
        path, B=B, N=N, seed=SEED, label=label.numpy(),
        **{f"wolf_{k}": v.numpy() for k, v in wolf.items()}, gumbel=gumbel.numpy(),
        **{f"gen_{k}": v.numpy() for k, v in gen_noise.items()}, keep=keep.astype(np.int64),
        pointwolf=pointcloud_pointwolf.detach().numpy(), gen_imgs=gen_imgs.detach().numpy(),
        g_loss_raw=g_loss_raw.item(), feedback_loss=feedback_loss.item(), g_loss=g_loss.item(),
        d_loss=d_loss.item(), cls_loss=loss.item(), cls_logits=logits.detach().numpy(),
        post_g=dict(generator.named_parameters())[GNAMES["generator"][0]].detach().numpy(),
        post_c=dict(model.named_parameters())[GNAMES["classifier"][0]].detach().numpy(),
        **{k: np.asarray(v) for k, v in out.items()})
    print("wrote", path)
    for k in ("g_loss_raw", "feedback_loss", "g_loss", "d_loss", "cls_loss"):
        print(k, float(locals()[k] if k != "cls_loss" else loss))
    for k, v in out.items():
        if k.startswith("gnorm"):
            print(k, v)


if __name__ == "__main__":
    main()
