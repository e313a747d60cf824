"""Multi-GPU debug: where does our DDP/DP diverge from a manual all-reduce + SGD?"""
import copy, os, sys
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_template_b200 import env, optim
from pytorch_distributed_template_b200.models import create_model
from pytorch_distributed_template_b200.parallel import DistributedDataParallel

def rel(a, b):
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()

def main():
    rank = int(os.environ["RANK"]); device = env.init_distributed(int(os.environ["LOCAL_RANK"]))
    W = dist.get_world_size()
    torch.manual_seed(100 + rank)
    net = create_model("resnet18", num_classes=10).to(device)
    ours = DistributedDataParallel(net, device_ids=[device.index])
    ref = copy.deepcopy(net)            # after sync: same weights; plain module, manual all-reduce
    # deep copy shares nothing with the arena
    o = optim.SGD(ours.parameters(), 0.01, momentum=0.9, weight_decay=1e-4)
    g = torch.Generator().manual_seed(7 + rank)
    x = torch.randn(8, 3, 64, 64, generator=g).to(device); y = torch.randint(0, 10, (8,), generator=g).to(device)
    p0 = [p.detach().clone() for p in ref.parameters()]
    same0 = max(rel(p.detach(), q) for p, q in zip(ours.module.parameters(), p0))
    ours.train(); ref.train()
    o.zero_grad()
    torch.nn.functional.cross_entropy(ours(x), y).backward()
    torch.nn.functional.cross_entropy(ref(x), y).backward()
    torch.cuda.synchronize()
    ours.join_comm(); torch.cuda.synchronize()
    # local grads equal?  (our arena holds local grads except in the shard this rank reduced)
    loc = [q.grad.clone() for q in ref.parameters()]
    for q in ref.parameters():
        dist.all_reduce(q.grad); q.grad /= W
    a = ours.arena
    worst_own = worst_other = 0.0
    for seg in a.segments:
        sh = seg.length // W
        for p, q, l in zip(ours.module.parameters(), ref.parameters(), loc):
            if a.segment_of[id(p)] is not seg: continue
            off = a.offset[id(p)]; n = p.numel()
            flat_ours = a.flat_grads[off:off + n]
            avg = (q.grad.permute(0, 2, 3, 1) if q.dim() == 4 else q.grad).reshape(-1)
            lcl = (l.permute(0, 2, 3, 1) if l.dim() == 4 else l).reshape(-1)
            idx = torch.arange(off, off + n, device=device)
            own = (idx >= seg.start + rank * sh) & (idx < seg.start + (rank + 1) * sh)
            if own.any():
                worst_own = max(worst_own, ((flat_ours[own] - avg[own]).abs().max() / (avg.abs().max() + 1e-12)).item())
            if (~own).any():
                worst_other = max(worst_other, ((flat_ours[~own] - lcl[~own]).abs().max() / (lcl.abs().max() + 1e-12)).item())
    o.step(); torch.cuda.synchronize()
    # manual SGD step on ref
    worst_p = 0.0
    for p, q, q0 in zip(ours.module.parameters(), ref.parameters(), p0):
        gq = q.grad + 1e-4 * q0
        newq = q0 - 0.01 * gq
        worst_p = max(worst_p, rel(p.detach(), newq))
    sh = a.flat_shadow
    print("rank", rank, "init_same", same0, "own_shard_avg_err", worst_own, "other_shard_local_err", worst_other,
          "param_err_after_step", worst_p, flush=True)
    env.shutdown_distributed()

if __name__ == "__main__":
    main()
