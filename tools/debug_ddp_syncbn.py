"""2-GPU debug: DDP + SyncBN (ours, torch) against a single-GPU plain-BN run on the concatenated batch."""
import copy, os, sys
import torch, torch.distributed as dist, torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torchvision
from pytorch_distributed_template_b200 import env, optim
from pytorch_distributed_template_b200.models import create_model
from pytorch_distributed_template_b200.parallel import DistributedDataParallel, SyncBatchNorm

def rel(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-12)).item()

def main():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    rank = int(os.environ["RANK"]); device = env.init_distributed(int(os.environ["LOCAL_RANK"]))
    W = dist.get_world_size()
    torch.manual_seed(100)
    base = create_model("resnet18", num_classes=10)
    oracle = torchvision.models.resnet18(num_classes=10); oracle.load_state_dict(base.state_dict()); oracle = oracle.to(device)
    tvs = torchvision.models.resnet18(num_classes=10); tvs.load_state_dict(base.state_dict())
    tvs = torch.nn.SyncBatchNorm.convert_sync_batchnorm(tvs).to(device)
    theirs = torch.nn.parallel.DistributedDataParallel(tvs, device_ids=[device.index])
    ours_net = SyncBatchNorm.convert_sync_batchnorm(copy.deepcopy(base)).to(device)
    ours = DistributedDataParallel(ours_net, device_ids=[device.index])
    g = torch.Generator().manual_seed(7 + rank)
    x = torch.randn(8, 3, 64, 64, generator=g).to(device); y = torch.randint(0, 10, (8,), generator=g).to(device)
    xs = [torch.empty_like(x) for _ in range(W)]; ys = [torch.empty_like(y) for _ in range(W)]
    dist.all_gather(xs, x); dist.all_gather(ys, y)
    X, Y = torch.cat(xs), torch.cat(ys)
    for m in (oracle, theirs, ours):
        m.train()
    lo = F.cross_entropy(oracle(X), Y); lo.backward()
    lt = F.cross_entropy(theirs(x), y); lt.backward()
    lu = F.cross_entropy(ours(x), y); lu.backward()
    torch.cuda.synchronize()
    # our DDP (no fused optimizer bound) leaves the averaged gradient in p.grad
    wt = wu = 0.0; nt = nu = None
    for (k, po), pt, pu in zip(oracle.named_parameters(), theirs.module.parameters(), ours.module.parameters()):
        dt, du = rel(pt.grad, po.grad), rel(pu.grad, po.grad)
        if dt > wt: wt, nt = dt, k
        if du > wu: wu, nu = du, k
    names = [k for k, _ in oracle.named_parameters()]
    bad = [(k, round(rel(pu.grad, po.grad), 5)) for (k, po), pu in zip(oracle.named_parameters(), ours.module.parameters()) if rel(pu.grad, po.grad) > 1e-3]
    if rank == 0:
        print("BAD", len(bad), bad[:40], flush=True)
    print("rank", rank, "loss oracle/theirs/ours", lo.item(), lt.item(), lu.item(), "| torch-vs-oracle", wt, nt, "| ours-vs-oracle", wu, nu, flush=True)
    print("rank", rank, "bn1.running_mean ours-vs-oracle", rel(ours.module.bn1.running_mean, oracle.bn1.running_mean),
          "theirs-vs-oracle", rel(theirs.module.bn1.running_mean, oracle.bn1.running_mean), flush=True)
    env.shutdown_distributed()

if __name__ == "__main__":
    main()
