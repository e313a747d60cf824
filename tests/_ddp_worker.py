"""Worker for test_ddp_gloo.py: 2 ranks on CPU/gloo (or N ranks on GPUs/NCCL with --cuda).
Compares our DistributedDataParallel + fused SGD against torch DDP + torch SGD, same seeds."""
import copy
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from pytorch_distributed_template_b200 import amp, env, optim  # noqa: E402
from pytorch_distributed_template_b200.models import create_model  # noqa: E402
from pytorch_distributed_template_b200.parallel import DistributedDataParallel, SyncBatchNorm  # noqa: E402


def syncbn_layer_check(device, rank):
    """One SyncBatchNorm layer, ours (in-kernel NVLink exchange) vs torch.nn.SyncBatchNorm, per rank
    (the same function bench.py runs on the live job)."""
    from pytorch_distributed_template_b200.parallel import Fabric, selfcheck
    return selfcheck.syncbn_layer(Fabric(device))


PER_RANK = 64


def main():
    cuda = "--cuda" in sys.argv
    syncbn = "--syncbn" in sys.argv
    if not cuda:
        os.environ["PDT_DEVICE"] = "cpu"
    # tight comparisons need identical arithmetic: TF32 convs pick different algorithms for NCHW
    # (torchvision oracle) and NHWC (ours) and BatchNorm amplifies the 1e-3 rounding differences
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    rank = int(os.environ["RANK"])
    device = env.init_distributed(int(os.environ["LOCAL_RANK"]))
    world = dist.get_world_size()
    if "--syncbn-layer" in sys.argv:
        out = {"rank": rank, "layer": syncbn_layer_check(device, rank)}
        with open(os.path.join(os.environ["PDT_RESULT_DIR"], "rank{}.json".format(rank)), "w") as f:
            json.dump(out, f)
        env.shutdown_distributed()
        return
    torch.manual_seed(100 + rank)                       # different init per rank: ctor must sync from rank 0
    net = create_model("resnet18", num_classes=10)
    if syncbn:
        net = SyncBatchNorm.convert_sync_batchnorm(net)
    net = net.to(device)
    ref_net = copy.deepcopy(net)
    if syncbn:     # oracle: torchvision's ResNet with torch's own SyncBatchNorm (CUDA only)
        import torchvision
        ref_net = torchvision.models.resnet18(num_classes=10)
        ref_net.load_state_dict(net.state_dict())
        ref_net = torch.nn.SyncBatchNorm.convert_sync_batchnorm(ref_net).to(device) if cuda else None

    ours = DistributedDataParallel(net, device_ids=[device.index] if cuda else None)
    out = {"rank": rank, "world": world, "segments": len(ours.arena.segments), "fabric": ours.fabric.describe()}
    # 1. constructor broadcast: every rank equals rank 0
    flat = ours.arena.flat_params.detach().clone()
    ref0 = flat.clone()
    dist.broadcast(ref0, src=0)
    out["init_synced"] = bool(torch.equal(flat, ref0))
    out["prefix_ok"] = all(k.startswith("module.") for k in ours.state_dict()) and \
        "conv1.weight" in ours.module.state_dict()

    res = {}
    if ref_net is not None:
        ref_net.load_state_dict(ours.module.state_dict())
        theirs = torch.nn.parallel.DistributedDataParallel(ref_net, device_ids=[device.index] if cuda else None)
        o_ours = optim.SGD(ours.parameters(), 0.01, momentum=0.9, weight_decay=1e-4)
        o_ref = torch.optim.SGD(theirs.parameters(), 0.01, momentum=0.9, weight_decay=1e-4)
        g = torch.Generator().manual_seed(7 + rank)
        worst_grad = worst_param = 0.0
        worst_name = None
        first_loss = []
        flat = lambda m: torch.cat([p.detach().float().reshape(-1) for p in m.parameters()])
        init = flat(theirs.module)
        for it in range(3):
            # 64 images per rank: with 8, the 20 BatchNorm layers amplify rounding differences between two correct
            # implementations into percent-level gradient differences and the bounds below would have to be useless
            x = torch.randn(PER_RANK, 3, 64, 64, generator=g).to(device)
            y = torch.randint(0, 10, (PER_RANK,), generator=g).to(device)
            for model, opt in ((ours, o_ours), (theirs, o_ref)):
                model.train()
                opt.zero_grad()
                loss = torch.nn.functional.cross_entropy(model(x), y)
                if it == 0:
                    first_loss.append(loss.item())
                loss.backward()
                if model is theirs or cuda is False:
                    pass
                opt.step()
            if it == 0 and not cuda:
                # torch backend leaves the averaged gradient in p.grad: compare directly
                for p, q in zip(ours.module.parameters(), theirs.module.parameters()):
                    d = (p.grad - q.grad).abs().max().item() / (q.grad.abs().max().item() + 1e-12)
                    worst_grad = max(worst_grad, d)
            # error of the accumulated UPDATE in the 2-norm, relative to the size of the update (a per-tensor
            # max-relative error is meaningless for tensors that start at zero, like BatchNorm biases)
            fo, ft = flat(ours.module), flat(theirs.module)
            d = ((fo - ft).norm() / ((ft - init).norm() + 1e-30)).item()
            if d > worst_param:
                worst_param, worst_name = d, (it, "update l2")
            if it == 0:
                first_param = worst_param
        res = {"worst_grad_rel": worst_grad, "worst_param_rel": worst_param, "worst_name": str(worst_name), "first_param_rel": first_param,
               "first_loss_rel": abs(first_loss[0] - first_loss[1]) / abs(first_loss[1])}
        # ranks agree with each other bit-for-bit
        mine = ours.arena.flat_params.detach().clone()
        other = mine.clone()
        dist.broadcast(other, src=0)
        res["ranks_identical"] = bool(torch.equal(mine, other))
        # running stats: rank 0's are broadcast before each forward (broadcast_buffers=True)
        rm = ours.module.bn1.running_mean.detach().clone()
        rm_ref = theirs.module.bn1.running_mean.detach().clone()
        res["bn_buffer_rel"] = ((rm - rm_ref).abs().max() / (rm_ref.abs().max() + 1e-12)).item()
    out.update(res)
    # scalar metric reduction
    t = torch.tensor([float(rank + 1), 2.0 * (rank + 1)], device=device)
    ours.fabric.allreduce_small_(t, mean=True)
    out["reduce_mean"] = t.tolist()
    res_dir = os.environ.get("PDT_RESULT_DIR")
    if res_dir:       # one file per rank: stdout lines of concurrent ranks can interleave
        with open(os.path.join(res_dir, "rank{}.json".format(rank)), "w") as f:
            json.dump(out, f)
    print("RESULT " + json.dumps(out), flush=True)
    env.shutdown_distributed()


if __name__ == "__main__":
    main()
