"""In-run correctness checks of the cross-GPU paths against NCCL / torch.

These run inside a live job — ``bench.py`` calls them on every rank before its timed region and writes the
outcome into its JSON line, ``tests/_ddp_worker.py`` uses the same functions — because a multi-GPU box is
where multi-GPU truth has to be recorded.  Each check leaves the model, the optimizer state and the BN
buffers exactly as it found them.

Oracles: ``dist.all_reduce`` (NCCL) + the ``torch.optim.SGD`` formula for the gradient path
(/root/reference/distributed.py:262-263), ``torch.nn.SyncBatchNorm`` for the statistics exchange
(/root/reference/distributed_syncBN_amp.py:143-145), ``torch.nn.DataParallel`` for the single-process mode
(/root/reference/dataparallel.py:119).
"""
import torch
import torch.distributed as dist


def _rel(a, b):
    a, b = a.detach().float(), b.detach().float()
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()


def ranks_identical(ddp):
    """Every rank holds bit-identical parameters (and low-precision shadow) — what DDP promises after any step."""
    a = ddp.arena
    mine = a.flat_params.detach().clone()
    other = mine.clone()
    dist.broadcast(other, src=0, group=ddp.process_group)
    same = bool(torch.equal(mine, other))
    if a.flat_shadow is not None:
        ms = a.flat_shadow.detach().clone()
        os_ = ms.clone()
        dist.broadcast(os_, src=0, group=ddp.process_group)
        same = same and bool(torch.equal(ms, os_))
    flag = torch.tensor([1.0 if same else 0.0], device=a.device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=ddp.process_group)
    return bool(flag.item() > 0)


@torch.no_grad()
def gradient_path(ddp, opt, seed=0):
    """Our reduce-scatter (1/W, unscale) + fused SGD + weight all-gather on random per-rank gradients against
    NCCL all_reduce + the torch.optim.SGD formula on the same inputs.  Returns max relative errors."""
    a = ddp.arena
    dev, W, rank = a.device, ddp.world, ddp.rank
    keep = [t.detach().clone() for t in (a.flat_params, a.flat_grads, a.momentum)]
    keep_shadow = a.flat_shadow.detach().clone() if a.flat_shadow is not None else None
    st = opt._amp_state()
    keep_state = st.detach().clone()
    try:
        g = torch.Generator(device=dev).manual_seed(1000 + seed * 64 + rank)
        scale = float(st[0].item())
        local = torch.randn(a.total, device=dev, generator=g) * 1e-2
        a.flat_grads.copy_(local * scale)                  # what backward leaves there under a loss scale
        a.momentum.copy_(torch.randn(a.total, device=dev, generator=torch.Generator(device=dev).manual_seed(7)) * 1e-2)
        dist.broadcast(a.momentum, src=0, group=ddp.process_group)      # momentum shards agree wherever they live
        # ---- oracle: NCCL all-reduce, then torch.optim.SGD's arithmetic (dampening 0, no nesterov)
        avg = local.clone()
        dist.all_reduce(avg, group=ddp.process_group)
        avg /= W
        grp = opt.param_groups[0]
        lr, mu, wd = grp["lr"], grp["momentum"], grp["weight_decay"]
        d_p = avg + wd * keep[0]
        buf = mu * a.momentum + d_p
        expect = keep[0] - lr * buf
        torch.cuda.synchronize(dev)
        # ---- ours: the kernels the training step uses, launched the way the autograd hooks launch them
        main = torch.cuda.current_stream(dev)
        for seg in a.segments:
            ddp._launch_segment(seg)
        ddp._launched.clear()
        a.grads_zeroed = False
        opt.step()
        main.wait_stream(ddp._comm_stream)
        torch.cuda.synchronize(dev)
        out = {"param_rel": _rel(a.flat_params, expect)}
        if a.flat_shadow is not None:
            out["shadow_rel"] = _rel(a.flat_shadow.float(), expect.to(a.flat_shadow.dtype).float())
        # momentum lives only on the shard this rank owns
        worst = 0.0
        for seg in a.segments:
            sh = seg.length // W
            lo = seg.start + rank * sh
            worst = max(worst, _rel(a.momentum[lo:lo + sh], buf[lo:lo + sh]))
        out["momentum_rel"] = worst
        t = torch.tensor([out["param_rel"], out.get("shadow_rel", 0.0), out["momentum_rel"]], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=ddp.process_group)
        out["param_rel"], out["momentum_rel"] = float(t[0]), float(t[2])
        if "shadow_rel" in out:
            out["shadow_rel"] = float(t[1])
        return out
    finally:
        a.flat_params.copy_(keep[0]); a.flat_grads.copy_(keep[1]); a.momentum.copy_(keep[2])
        if keep_shadow is not None:
            a.flat_shadow.copy_(keep_shadow)
        st.copy_(keep_state)
        torch.cuda.synchronize(dev)
        dist.barrier(group=ddp.process_group)


def syncbn_layer(fabric, shapes=((64, 16), (256, 7), (512, 4)), dtypes=(torch.float32, torch.bfloat16, torch.float16), batch=6):
    """One SyncBatchNorm layer, ours (in-kernel NVLink exchange) against torch.nn.SyncBatchNorm: forward, dx,
    dgamma / dbeta and the running statistics, per rank.  Returns {"fp32_y": rel, ...} (max over ranks)."""
    from .syncbn import SyncBatchNorm, fused_bn
    device, rank = fabric.device, fabric.rank
    worst = {}
    for dtype in dtypes:
        for n, (C, HW) in enumerate(shapes):
            relu = n > 0
            torch.manual_seed(5 + rank)
            x = (torch.randn(batch, C, HW, HW, device=device) * (1 + rank) + rank).to(dtype).contiguous(memory_format=torch.channels_last)
            dy = torch.randn(batch, C, HW, HW, device=device).to(dtype).contiguous(memory_format=torch.channels_last)
            ours = SyncBatchNorm(C).to(device)
            ours._pdt_fabric = fabric
            ref = torch.nn.SyncBatchNorm(C).to(device)
            with torch.no_grad():
                ours.weight.uniform_(0.5, 1.5)
                ours.bias.uniform_(-0.5, 0.5)
                dist.broadcast(ours.weight, 0)
                dist.broadcast(ours.bias, 0)
            ref.load_state_dict(ours.state_dict())
            xo = x.clone().requires_grad_(True)
            xr = x.float().clone().requires_grad_(True)
            yo = fused_bn(ours, xo, relu=relu)
            yr = torch.relu(ref(xr)) if relu else ref(xr)
            yo.backward(dy)
            yr.backward(dy.float())
            torch.cuda.synchronize()
            key = {torch.float32: "fp32", torch.bfloat16: "bf16", torch.float16: "fp16"}[dtype]
            for name, p, q in (("y", yo, yr), ("dx", xo.grad, xr.grad), ("dgamma", ours.weight.grad, ref.weight.grad),
                               ("dbeta", ours.bias.grad, ref.bias.grad), ("rmean", ours.running_mean, ref.running_mean),
                               ("rvar", ours.running_var, ref.running_var)):
                k = key + "_" + name
                worst[k] = max(worst.get(k, 0.0), _rel(p, q))
    keys = sorted(worst)
    t = torch.tensor([worst[k] for k in keys], device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return {k: float(v) for k, v in zip(keys, t.tolist())}


def syncbn_tolerance(key):
    return 1e-5 if key.startswith("fp32") else (2e-2 if key.startswith("bf16") else 4e-3)


def dataparallel_step(device_ids, per_replica=64, image=64, classes=10, seed=11):
    """One SGD step of our single-process DataParallel against torch.nn.DataParallel over torchvision's
    ResNet-18 with the same weights, BatchNorm in eval mode (a trunk without batch statistics is well
    conditioned: what is compared is scatter / replica forward / gather / gradient reduce / SGD / weight push).
    Returns the loss difference and the relative error of the parameter UPDATE in the 2-norm."""
    import torch.nn.functional as F
    import torchvision
    from .. import optim
    from ..models import create_model
    from .dp import DataParallel
    n = len(device_ids)
    torch.manual_seed(seed)
    net = create_model("resnet18", num_classes=classes)
    tv = torchvision.models.resnet18(num_classes=classes)
    tv.load_state_dict(net.state_dict())
    dev0 = torch.device("cuda", device_ids[0])
    theirs = torch.nn.DataParallel(tv.to(dev0), device_ids=list(device_ids))
    ours = DataParallel(net, device_ids=list(device_ids)).to(dev0)
    o_ours = optim.SGD(ours.parameters(), 0.05, momentum=0.9, weight_decay=1e-4)
    o_ref = torch.optim.SGD(theirs.parameters(), 0.05, momentum=0.9, weight_decay=1e-4)
    g = torch.Generator().manual_seed(seed + 1)
    x = torch.randn(per_replica * n, 3, image, image, generator=g).to(dev0)
    y = torch.randint(0, classes, (per_replica * n,), generator=g).to(dev0)
    before = [p.detach().clone() for p in ours.module.parameters()]
    losses = []
    for model, opt in ((ours, o_ours), (theirs, o_ref)):
        model.train()
        for m in model.modules():          # BatchNorm on running statistics: no per-replica batch statistics
            if isinstance(m, torch.nn.modules.batchnorm._BatchNorm):
                m.eval()
        opt.zero_grad()
        loss = F.cross_entropy(model(x).float(), y)
        loss.backward()
        opt.step()
        losses.append(float(loss.item()))
    for d in device_ids:
        torch.cuda.synchronize(d)
    num = den = 0.0
    for p0, po, pr in zip(before, ours.module.parameters(), theirs.module.parameters()):
        do, dr = (po.detach() - p0).double(), (pr.detach().to(p0.device) - p0).double()
        num += float(((do - dr) ** 2).sum())
        den += float((dr ** 2).sum())
    replicas_equal = all(torch.equal(p.detach().cpu(), q.detach().cpu())
                         for rep in ours.replicas[1:] for p, q in zip(rep.parameters(), ours.module.parameters()))
    return {"loss_rel": abs(losses[0] - losses[1]) / abs(losses[1]), "update_rel_l2": (num / max(den, 1e-30)) ** 0.5,
            "replicas_equal": bool(replicas_equal), "devices": n, "per_replica": per_replica}
