"""2-GPU debug: one SyncBatchNorm layer, ours (in-kernel exchange) vs torch.nn.SyncBatchNorm."""
import os, sys
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_template_b200 import env
from pytorch_distributed_template_b200.parallel import Fabric, SyncBatchNorm
from pytorch_distributed_template_b200.parallel.syncbn import fused_bn

def rel(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-12)).item()

def main():
    rank = int(os.environ["RANK"]); device = env.init_distributed(int(os.environ["LOCAL_RANK"]))
    fab = Fabric(device)
    for dtype in (torch.float32, torch.bfloat16):
        for C, HW in ((64, 16), (256, 7)):
            torch.manual_seed(5 + rank)
            x = (torch.randn(6, C, HW, HW, device=device) * (1 + rank) + rank).to(dtype).contiguous(memory_format=torch.channels_last)
            dy = torch.randn(6, C, HW, HW, device=device).to(dtype).contiguous(memory_format=torch.channels_last)
            ours = SyncBatchNorm(C).to(device); ours._pdt_fabric = fab
            ref = torch.nn.SyncBatchNorm(C).to(device)
            with torch.no_grad():
                ours.weight.uniform_(0.5, 1.5); ours.bias.uniform_(-0.5, 0.5)
                dist.broadcast(ours.weight, 0); dist.broadcast(ours.bias, 0)
            ref.load_state_dict(ours.state_dict())
            xo = x.clone().requires_grad_(True); xr = x.float().clone().requires_grad_(True)
            yo = fused_bn(ours, xo, relu=False); yr = ref(xr)
            yo.backward(dy); yr.backward(dy.float())
            torch.cuda.synchronize()
            print("rank", rank, str(dtype), C, "y", rel(yo, yr), "dx", rel(xo.grad, xr.grad), "dgamma", rel(ours.weight.grad, ref.weight.grad),
                  "dbeta", rel(ours.bias.grad, ref.bias.grad), "rmean", rel(ours.running_mean, ref.running_mean),
                  "rvar", rel(ours.running_var, ref.running_var), flush=True)
    env.shutdown_distributed()

if __name__ == "__main__":
    main()
