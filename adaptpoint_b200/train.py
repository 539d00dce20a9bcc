"""One AdaptPoint training step: imitator (generator) + discriminator update, then the classifier update.

Host-side mirror of the two hot loops of the reference
  train_gan        examples/classification/train_autoaug.py:119-237   (per-iteration body :131-207)
  train_one_epoch  examples/classification/train_autoaug.py:463-523   (per-iteration body :474-506)
  get_gan_model    examples/classification/train_autoaug.py:80-117    (Adam lr 1e-4 / 4e-4, betas .5/.999, BCE)
  get_feedback_loss_ver1  openpoints/function_adaptpoint/ganloss_cls.py:31-65
with the reference's config values from cfgs/scanobjectnn/pointnext-s_adaptpoint_1.yaml:38-74 and
cfgs/scanobjectnn/default.yaml:35-52.  Deliberate differences:
  * generated clouds stay on the device (the reference copies three (B,N,.) arrays to host numpy every
    iteration and rebuilds a DataLoader per epoch);
  * data parallelism is explicit: per-rank batch fixed, gradients of every updated network are averaged with
    ONE flat NCCL all-reduce per optimizer step (the reference wraps in DDP but calls `model.module....`, so
    its reducer never fires); the CLASSIFIER's BatchNorm becomes SyncBatchNorm when world_size > 1, as in the
    reference (main.py:27 / train_autoaug.py:276 convert the classifier only; get_gan_model wraps the imitator and
    the discriminator in plain DDP, so their BatchNorm statistics stay per rank -- here too);
  * logging / TensorBoard / h5 dumps are left out.
"""
import copy

import torch
import torch.distributed as dist
import torch.nn as nn

from . import fused, geometry, norm, ops, peer, tcconv
from .imitator import AdaptPoint_Augmentor, PointDiscriminator1, PointWOLF_classversion
from .models import SmoothCrossEntropy, build_model

ADAPTPOINT_CFG = dict(
    model="pointnext-s", num_points=1024, num_classes=15, in_channels=4, batch_size=32, loader_points=2048,
    lr=0.002, min_lr=1e-4, weight_decay=0.05, grad_norm_clip=10.0, label_smoothing=0.3, epochs=300,
    pointwolf=dict(w_num_anchor=4, w_sigma=0.5, w_R_range=10, w_S_range=3, w_T_range=0.25),
    adaptmodel_gan=dict(w_num_anchor=4, w_sigma=0.5, w_R_range=10, w_S_range=3, w_T_range=0.25),
    adaptmodel_dis=dict(num_classes=15, normal_channel=False),
    adaptpoint_params=dict(lr_generator=1e-4, lr_discriminator=4e-4, b1=0.5, b2=0.999, hardratio_s=3, hardratio=3),
    feedbackloss_ratio=1,
)


def _world():
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def allreduce_grads(params):
    """Average gradients over ranks with one flat all-reduce (NCCL over NVLink on the GPU box, gloo in tests)."""
    world = _world()
    if world == 1:
        return
    grads = [p.grad for p in params if p.grad is not None]
    if not grads:
        return
    from . import _cabi
    trace = _cabi.TRACE if grads[0].is_cuda else None
    if trace is not None:                  # bench.py's per-launch CUDA-event trace: flatten + collective + copy back
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    flat = torch._utils._flatten_dense_tensors(grads)
    if dist.get_backend() == "nccl":
        dist.all_reduce(flat, op=dist.ReduceOp.AVG)      # the division rides in the collective
    else:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        flat.div_(world)
    # one multi-tensor copy instead of one small kernel per parameter
    torch._foreach_copy_(grads, list(torch._utils._unflatten_dense_tensors(flat, grads)))
    if trace is not None:
        e1.record()
        trace.append(("nccl_allreduce_grads", (int(flat.numel()), world), e0, e1))


def hard_ratio(epoch, cfg):
    """rho of the perception-guidance loss, linear in the epoch (ganloss_cls.py:53-54)."""
    ap = cfg["adaptpoint_params"]
    return ap["hardratio_s"] + (ap["hardratio"] - ap["hardratio_s"]) * epoch / cfg["epochs"]


def cosine_lr(epoch, cfg):
    """Cosine decay from cfg['lr'] to cfg['min_lr'] over cfg['epochs'] (the reference's `sched: cosine`,
    warmup_epochs 0, stepped once per epoch)."""
    import math
    lo = cfg.get("min_lr", 1e-4)
    t = min(max(epoch, 0), cfg["epochs"]) / max(cfg["epochs"], 1)
    return lo + 0.5 * (cfg["lr"] - lo) * (1.0 + math.cos(math.pi * t))


def feedback_loss(classifier, criterion, data_real, data_fake, rho, pred_real=None):
    """|1 - exp(L_fake - rho * L_real)| with the classifier in eval mode but inside autograd
    (ganloss_cls.py:31-65).  `pred_real`: logits of the real batch when the caller already has them."""
    classifier.eval()
    pred_fake = classifier(data_fake)
    if pred_real is not None:
        pass
    elif data_real["pos"].requires_grad or data_real["x"].requires_grad or any(p.requires_grad for p in classifier.parameters()):
        pred_real = classifier(data_real)
    else:   # nothing upstream of the real batch takes a gradient: plain inference (the fused bf16 SA kernels when
        with torch.no_grad():   # the classifier was built with fused=True)
            pred_real = classifier(data_real)
    label = data_real["y"].long()
    loss_fake, loss_real = criterion(pred_fake, label), criterion(pred_real, label)
    return torch.abs(1 - torch.exp(loss_fake - rho * loss_real)).mean()


class AdaptPointTrainer:
    """Holds classifier / imitator / discriminator + their optimizers; `step(batch)` = one train_gan iteration
    followed by one train_one_epoch iteration on the generated batch."""

    def __init__(self, cfg=None, device="cuda", fused_eval=False, seed=0, conv_backend="tc", overlap=True,
                 host_gumbel=False):
        """conv_backend: 'tc' = every 1x1 convolution (fwd, dgrad, wgrad) on the tcgen05 kernels of libapb200
        (bf16/TF32 operands, fp32 accumulation); 'torch' = cuDNN fp32 (strict-parity runs).  host_gumbel: draw the
        mask branch's Gumbel noise with the host generator like every other draw (reproducible runs / tests); default
        is the reference's device-side draw (F.gumbel_softmax, generator_component4_15.py:714)."""
        self.cfg = copy.deepcopy(ADAPTPOINT_CFG if cfg is None else cfg)
        self.device = torch.device(device)
        self.conv_backend = conv_backend if self.device.type == "cuda" else "torch"
        torch.manual_seed(seed)
        self.classifier = build_model(self.cfg["model"], fused=fused_eval).to(self.device)
        self.generator = AdaptPoint_Augmentor(**self.cfg["adaptmodel_gan"]).to(self.device)
        self.discriminator = PointDiscriminator1(**self.cfg["adaptmodel_dis"]).to(self.device)
        if _world() > 1:
            if self.device.type != "cuda":
                pass                         # torch's SyncBatchNorm has no CPU path (gloo tests): per-rank statistics
            elif self.conv_backend == "tc":    # our BatchNorm kernels all-reduce their (C,2) statistics themselves
                with torch.cuda.device(self.device):
                    peer.enable()              # ... over NVLink peer memory when the ranks share a node (peer.py)
                norm.enable_sync(self.classifier)     # classifier only: the reference leaves G / D on per-rank BN
            else:
                self.classifier = nn.SyncBatchNorm.convert_sync_batchnorm(self.classifier)
            for net in (self.classifier, self.generator, self.discriminator):      # same init on every rank
                for t in list(net.parameters()) + list(net.buffers()):
                    dist.broadcast(t.data, 0)
        ap = self.cfg["adaptpoint_params"]
        cap = self.device.type == "cuda"      # device-side step counters: required inside a CUDA graph
        kw = {"capturable": True, "fused": True} if cap else {}   # one multi-tensor kernel per optimizer step
        # learning rates live in device scalars on CUDA: a captured graph reads them at replay time, so `set_lr` /
        # `set_epoch` (the reference's per-epoch cosine `scheduler.step(epoch)`, train_autoaug.py:431) keep working
        # after `capture_graph`
        mk = (lambda v: torch.tensor(float(v), dtype=torch.float32, device=self.device)) if cap else float
        self._lr = {"c": mk(self.cfg["lr"]), "g": mk(ap["lr_generator"]), "d": mk(ap["lr_discriminator"])}
        self.opt_c = torch.optim.AdamW(self.classifier.parameters(), lr=self._lr["c"],
                                       weight_decay=self.cfg["weight_decay"], **kw)
        self.opt_g = torch.optim.Adam(self.generator.parameters(), lr=self._lr["g"], betas=(ap["b1"], ap["b2"]),
                                      **kw)
        self.opt_d = torch.optim.Adam(self.discriminator.parameters(), lr=self._lr["d"],
                                      betas=(ap["b1"], ap["b2"]), **kw)
        self.criterion_gan = nn.BCELoss()
        self.criterion = SmoothCrossEntropy(label_smoothing=self.cfg["label_smoothing"])
        self.pointwolf = PointWOLF_classversion(**self.cfg["pointwolf"])
        self.host_gumbel = host_gumbel
        self.epoch = 0
        self._rho = torch.tensor(float(hard_ratio(0, self.cfg)), device=self.device)   # device scalar: graph-safe
        self._graph = None
        self.overlap = overlap and self.device.type == "cuda"
        self._side = None
        self._geo = None
        self._prepared = None
        self.grad_probe = None     # tests: callable(name, module) invoked after each backward, before the optimizer step

    def _probe(self, name, module):
        if self.grad_probe is not None:
            self.grad_probe(name, module)

    def _side_stream(self):
        if self._side is None:
            self._side = torch.cuda.Stream(device=self.device)
        return self._side

    def _geo_stream(self):
        if self._geo is None:
            self._geo = torch.cuda.Stream(device=self.device)
        return self._geo

    def _cls_chain(self, n):
        """Sample counts of the classifier's FPS chain on an n-point cloud (one per strided stage)."""
        chain = []
        for s in getattr(self.classifier.encoder, "strides", []):
            if s > 1:
                n = n // s
                chain.append(n)
        return chain

    def _gen_chain(self, n):
        chain = []
        for g in self.generator.predict_prob_layer.pointset_grouper_list:
            n = n // g.reduce
            chain.append(n)
        return chain

    def _cls_queries(self, n):
        """Ball queries of the classifier's strided stages on an n-point cloud, as SamplingPlan queries."""
        out, path = [], ()
        for stage in getattr(self.classifier.encoder, "encoder", []):
            sa = stage[0]
            s = getattr(sa, "stride", 1)
            if s > 1:
                n = n // s
                g = getattr(sa, "grouper", None)
                if getattr(g, "radius", None) is not None and getattr(g, "nsample", None) is not None:
                    out.append((path, path + (n,), g.radius, g.nsample))
                path = path + (n,)
        return out

    def _gen_queries(self, n):
        out, path = [], ()
        for g in self.generator.predict_prob_layer.pointset_grouper_list:
            n = n // g.reduce
            out.append((path, path + (n,), g.radi, g.kneighbors))
            path = path + (n,)
        return out

    def _resample_count(self, n_loader_points):
        npoints = self.cfg["num_points"]
        if n_loader_points <= npoints:
            return None
        return min({1024: 1200, 4096: 4800, 8192: 8192}[npoints], n_loader_points)

    # -- train_autoaug.py:131-207 ---------------------------------------------------------------------------
    def set_epoch(self, epoch, schedule_lr=True):
        """Per-epoch host hook: rho of the feedback loss (ganloss_cls.py:53-54) and, as the reference's
        `scheduler.step(epoch)` does for the classifier (cosine, cfgs/scanobjectnn/default.yaml:46-50), its lr."""
        self.epoch = epoch
        self._rho.fill_(float(hard_ratio(epoch, self.cfg)))
        if schedule_lr:
            self.set_lr(c=cosine_lr(epoch, self.cfg))

    def set_lr(self, c=None, g=None, d=None):
        """Change the learning rate of the classifier / imitator / discriminator optimizer; graph-safe (the captured
        optimizer kernels read the device scalars)."""
        for key, val, opt in (("c", c, self.opt_c), ("g", g, self.opt_g), ("d", d, self.opt_d)):
            if val is None:
                continue
            if isinstance(self._lr[key], torch.Tensor):
                self._lr[key].fill_(float(val))
            else:
                self._lr[key] = float(val)
                for grp in opt.param_groups:
                    grp["lr"] = float(val)

    def get_lr(self):
        return {k: float(v) for k, v in self._lr.items()}

    def draw_noise(self, B, n_loader_points):
        """All host-RNG draws of one step, in the order the reference makes them: PointWOLF (pointwolf.py:113-131),
        the imitator's dropout/axes (generator_component4_15.py:245-247,308), the resampling subset
        (train_autoaug.py:497, np.random.choice)."""
        npoints = self.cfg["num_points"]
        noise = {"wolf": self.pointwolf.draw(B), "gen": self.generator.draw(B)}
        if self.host_gumbel:
            noise["gen"]["gumbel"] = -torch.empty(B, n_loader_points, 2).exponential_().log()
        if n_loader_points > npoints:
            point_all = min({1024: 1200, 4096: 4800, 8192: 8192}[npoints], n_loader_points)
            noise["keep"] = torch.randperm(point_all)[:npoints]
        return noise

    def gan_iteration(self, points, label, noise=None):
        """points (B,N,C>=4) loader-shaped, label (B,).  Returns (generated batch (B,N,C), g_loss, d_loss)."""
        cfg = self.cfg
        G, D, C = self.generator, self.discriminator, self.classifier
        G.train(); D.train(); C.eval()
        points_clone = points.clone()
        inp = points[:, :, :3].contiguous()
        noise = noise or {}
        B = inp.size(0)
        nc = cfg["in_channels"]
        ratio = cfg.get("feedbackloss_ratio", 1)
        # nothing below writes into `points`, so the real batch's coordinates ARE `inp` (the reference clones because
        # it overwrites points[:, :, :3] with the generated cloud, :152)
        data_real = {"pos": inp, "y": label, "x": points_clone[:, :, :nc].transpose(1, 2).contiguous()}
        N = inp.shape[1]
        if inp.is_cuda and self.overlap:
            # one FPS chain for everything that samples the loader batch: imitator, classifier(real), both anchor sets
            geometry.SamplingPlan(inp, [self._gen_chain(N), self._cls_chain(N), [G.num_anchor],
                                        [self.pointwolf.num_anchor]], self._geo_stream(),
                                  self._gen_queries(N) + (self._cls_queries(N) if ratio > 0 else []))
        # Work that depends on the loader batch only -- PointWOLF (the reference runs it for its h5 dump, :143) and
        # the classifier's no-grad pass over the REAL batch for the feedback loss -- runs on a side stream beside the
        # imitator's forward pass (most of whose kernels leave SMs idle at B=32); captured as a parallel graph branch.
        pred_real, side = None, None
        if inp.is_cuda and self.overlap:
            main = torch.cuda.current_stream(inp.device)
            side = self._side_stream()
            side.wait_stream(main)
            with torch.cuda.stream(side):
                self.pointwolf(inp, noise.get("wolf"))
                if ratio > 0:
                    with torch.no_grad():
                        pred_real = C(data_real)
                    pred_real.record_stream(main)
        else:
            self.pointwolf(inp, noise.get("wolf"))
        real_label = torch.full((B, 1), 0.9, device=inp.device)
        fake_label = torch.full((B, 1), 0.1, device=inp.device)

        _, gen = G(inp, noise.get("gen"))
        # The discriminator only relays gradients to the imitator here: its own parameter gradients from this pass are
        # discarded by `optimizer_D.zero_grad()` (train_autoaug.py:183) before they are ever used, so they are not computed
        # (same reasoning as for the classifier below).
        dparams = [p for p in D.parameters() if p.requires_grad]
        for p in dparams:
            p.requires_grad_(False)
        try:
            g_loss_raw = self.criterion_gan(D(gen), real_label)
        finally:
            for p in dparams:
                p.requires_grad_(True)
        fake_pts = torch.cat([gen, points[:, :, 3:]], dim=-1)   # `points[:, :, :3] = gen_imgs` (:152)
        data_fake = {"pos": fake_pts[:, :, :3].contiguous(), "y": label,
                     "x": fake_pts[:, :, :nc].transpose(1, 2).contiguous()}
        self._prepared = None
        fake_det = fake_pts.detach()
        if inp.is_cuda and self.overlap:
            # geometry of the generated cloud (classifier(fake) chain + the resampling FPS) and the whole classifier
            # batch of the next iteration (resample, gather, its FPS chain) on the geometry stream, beside the
            # feedback loss and the two backward passes
            geo = self._geo_stream()
            point_all = self._resample_count(N)
            plan = geometry.SamplingPlan(data_fake["pos"], [self._cls_chain(N)] + ([[point_all]] if point_all else []),
                                         geo, self._cls_queries(N) if ratio > 0 else [])
            if point_all:
                with torch.cuda.stream(geo):
                    self._prepared = self._resample(fake_det, plan.nodes[(point_all,)][0], noise.get("keep"))
        g_loss = g_loss_raw
        if side is not None:
            torch.cuda.current_stream(inp.device).wait_stream(side)     # join the side branch
        if ratio > 0:
            # The classifier only relays gradients to the imitator here: its own parameter gradients are thrown away by
            # `model_pointcloud.zero_grad()` (train_autoaug.py:232) before they are ever used, so they are not computed.
            cparams = [p for p in C.parameters() if p.requires_grad]
            for p in cparams:
                p.requires_grad_(False)
            try:
                g_loss = g_loss_raw + feedback_loss(C, self.criterion, data_real, data_fake, self._rho,
                                                    pred_real) * ratio
            finally:
                for p in cparams:
                    p.requires_grad_(True)
        self.opt_g.zero_grad()
        g_loss.backward()
        allreduce_grads(list(G.parameters()))
        self._probe("generator", G)
        self.opt_g.step()

        real_loss = self.criterion_gan(D(inp), real_label)
        fake_loss = self.criterion_gan(D(gen.detach()), fake_label)
        d_loss = (real_loss + fake_loss) / 2
        self.opt_d.zero_grad()
        d_loss.backward()
        allreduce_grads(list(D.parameters()))
        self._probe("discriminator", D)
        self.opt_d.step()
        C.zero_grad(set_to_none=True)                           # model_pointcloud.zero_grad() (:232)
        return fake_det, g_loss.detach(), d_loss.detach()

    def _resample(self, points, fps_idx, keep=None):
        """Point resampling strategy (train_autoaug.py:485-499): a random npoints-subset of the FPS sample.  Runs on
        the current stream; on CUDA it also starts the classifier's FPS chain on the resampled cloud."""
        npoints = self.cfg["num_points"]
        if keep is None:                                        # np.random.choice(point_all, npoints, False)
            keep = torch.randperm(fps_idx.shape[1])[:npoints]
        fps_idx = fps_idx[:, keep.to(points.device)]
        out = torch.gather(points, 1, fps_idx.unsqueeze(-1).expand(-1, -1, points.shape[-1]))
        pos = out[:, :, :3].contiguous()
        if pos.is_cuda and self.overlap:
            geometry.SamplingPlan(pos, [self._cls_chain(npoints)], None, self._cls_queries(npoints))
        return {"source": points, "points": out, "pos": pos}

    # -- train_autoaug.py:474-506 ---------------------------------------------------------------------------
    def classifier_iteration(self, points, target, keep=None):
        cfg = self.cfg
        model = self.classifier
        model.train()
        npoints = cfg["num_points"]
        prepared, self._prepared = self._prepared, None
        if prepared is not None and prepared["source"] is points:      # resampled on the geometry stream already
            main = torch.cuda.current_stream(points.device)
            main.wait_stream(self._geo_stream())
            points, pos = prepared["points"], prepared["pos"]
            points.record_stream(main)
            pos.record_stream(main)
        else:
            if points.shape[1] > npoints:                       # point resampling strategy (:485-499)
                point_all = self._resample_count(points.size(1))
                fps_idx = ops.furthest_point_sample(points[:, :, :3].contiguous(), point_all).long()
                points = self._resample(points, fps_idx, keep)["points"]
            pos = points[:, :, :3].contiguous()
        data = {"pos": pos, "x": points[:, :, :cfg["in_channels"]].transpose(1, 2).contiguous()}
        logits, loss = model.get_logits_loss(data, target)
        loss.backward()
        allreduce_grads(list(model.parameters()))
        self._probe("classifier", model)
        if cfg.get("grad_norm_clip"):
            torch.nn.utils.clip_grad_norm_(model.parameters(), cfg["grad_norm_clip"], norm_type=2)
        self.opt_c.step()
        model.zero_grad()
        return logits.detach(), loss.detach()

    def _step_impl(self, points, label, noise):
        with tcconv.conv_backend(self.conv_backend):
            fake, g_loss, d_loss = self.gan_iteration(points, label, noise)
            _, loss = self.classifier_iteration(fake, label, noise.get("keep"))
            if self.overlap:     # every branch rejoins the main stream before the step ends (graph capture needs it)
                main = torch.cuda.current_stream(self.device)
                main.wait_stream(self._geo_stream())
                main.wait_stream(self._side_stream())
        return loss, g_loss, d_loss

    def step(self, points, label):
        """One full AdaptPoint step on a loader-shaped batch.  Returns (cls_loss, g_loss, d_loss) tensors.
        After `capture_graph`, the whole step (three forward/backward passes, gradient all-reduces, three optimizer
        updates: ~3000 kernel launches) is ONE CUDA-graph replay; only the host-RNG draws stay on the host."""
        noise = self.draw_noise(points.shape[0], points.shape[1])
        fused.invalidate_packed()    # parameters / BN statistics change below (a graph replay bumps no tensor version)
        if self._graph is None:
            return self._step_impl(points, label, noise)
        self._g_points.copy_(points, non_blocking=True)
        self._g_label.copy_(label, non_blocking=True)
        _copy_tree(self._g_noise, noise)
        self._graph.replay()
        return self._g_out

    def capture_graph(self, points, label, warmup=3):
        """Capture `step` for this batch shape into a CUDA graph (static input / noise buffers).  The warm-up
        iterations are real training steps."""
        if self.device.type != "cuda":
            raise RuntimeError("CUDA graphs need a CUDA device")
        self._graph = None
        self._g_points, self._g_label = points.clone(), label.clone()
        self._g_noise = _to_device_tree(self.draw_noise(points.shape[0], points.shape[1]), self.device)
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(warmup):
                _copy_tree(self._g_noise, self.draw_noise(points.shape[0], points.shape[1]))
                self._step_impl(self._g_points, self._g_label, self._g_noise)
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        ctx = peer.get()
        if ctx is not None:
            ctx.forget_order()       # nothing is in flight: the captured all-reduces need no edge to eager ones
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = self._step_impl(self._g_points, self._g_label, self._g_noise)
        self._g_out = out
        self._graph = graph
        return graph

    def release_graph(self):
        """Drop the captured graph (and the NCCL / peer-memory work it holds) so the process group can be torn down:
        reset the graph, drain the device.  The trainer keeps working eagerly afterwards."""
        if self._graph is not None:
            torch.cuda.synchronize(self.device)
            self._graph.reset()
            self._graph = None
            self._g_out = None
            torch.cuda.synchronize(self.device)


def _to_device_tree(tree, device):
    if isinstance(tree, dict):
        return {k: _to_device_tree(v, device) for k, v in tree.items()}
    return tree.to(device)


def _copy_tree(dst, src):
    """Copy host draws into the graph's static device buffers (stream-ordered before the replay)."""
    for k, v in src.items():
        if isinstance(v, dict):
            _copy_tree(dst[k], v)
        else:
            dst[k].copy_(v, non_blocking=True)
