"""Cross-GPU kernels in isolation, one process per GPU (torchrun): the gradient path of one ResNet-18 step
(3 segments, 46.76 MB fp32) through
  nvls   k_reduce_scatter + k_sgd_allgather on the NVLS multicast mapping (multimem.ld_reduce / multimem.st)
  p2p    the same kernels on peer load/store
  nccl   dist.all_reduce of the same three buckets + torch.optim.SGD (foreach) — the reference's path
         (/root/reference/distributed.py:262-263)
and the SyncBatchNorm-sized exchange (1025 floats): our one-shot in-kernel all-reduce vs dist.all_gather.
CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks.  Roofline of the
gradient path: bytes that must cross NVLink per GPU / 770 GB/s (measured peer-copy bandwidth, B200_PROFILING.md).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/comm_bench.py
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_template_b200 import engine, env, optim  # noqa: E402
from pytorch_distributed_template_b200.models import create_model  # noqa: E402
from pytorch_distributed_template_b200.parallel import DistributedDataParallel  # noqa: E402


def timed(fn, device, iters=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / iters], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) * 1e3      # us


def main():
    device = env.init_distributed(int(os.environ.get("LOCAL_RANK", "0")))
    engine.work_stream(device)
    W, rank = dist.get_world_size(), dist.get_rank()
    ddp = DistributedDataParallel(create_model("resnet18").to(device), device_ids=[device.index], shadow_dtype=torch.float16)
    opt = optim.SGD(ddp.parameters(), 0.1, momentum=0.9, weight_decay=1e-4)
    a = ddp.arena
    a.flat_grads.normal_(0, 1e-3)
    nbytes = a.total * 4
    out = {"world": W, "arena_bytes": nbytes, "segments": [(s.length * 4) for s in a.segments], "fabric": ddp.fabric.describe()}

    def ours():
        for seg in a.segments:
            ddp._launch_segment(seg)
        ddp._launched.clear()
        opt.step()

    keep = a.flat_params.clone()
    planes = (("nvls", True), ("p2p", False)) if ddp.fabric.describe()["multicast"] else (("p2p", False),)
    for name, flag in planes:
        ddp.fabric.use_nvls = flag
        out[name + "_us"] = timed(ours, device)
        a.flat_params.copy_(keep)
    ddp.fabric.use_nvls = True

    # the reference's path: bucketed NCCL all-reduce + foreach SGD on the same sizes
    buckets = [torch.randn(s.length, device=device) * 1e-3 for s in a.segments]
    params = [torch.randn(s.length, device=device, requires_grad=True) for s in a.segments]
    for p, g in zip(params, buckets):
        p.grad = g
    topt = torch.optim.SGD(params, 0.1, momentum=0.9, weight_decay=1e-4)

    def nccl():
        for g in buckets:
            dist.all_reduce(g)
            g.div_(W)
        topt.step()

    out["nccl_allreduce_sgd_us"] = timed(nccl, device)
    # bytes that must cross NVLink per GPU for a reduce-scatter + all-gather of nbytes (+ the 16-bit shadow)
    link_bytes = (nbytes + nbytes + nbytes // 2) * (W - 1) / W
    out["roofline_us_at_770GBs"] = link_bytes / 770e9 * 1e6
    for k in ("nvls_us", "p2p_us", "nccl_allreduce_sgd_us"):
        if k in out:
            out[k.replace("_us", "_frac_of_roofline")] = out["roofline_us_at_770GBs"] / out[k]

    # SyncBatchNorm-sized exchange: 2 * 512 + 1 floats
    t = torch.randn(1025, device=device)
    gathered = torch.empty(1025 * W, device=device)
    out["small_allreduce_ours_us"] = timed(lambda: ddp.fabric.allreduce_small_(t), device, iters=200)
    out["small_allgather_nccl_us"] = timed(lambda: dist.all_gather_into_tensor(gathered, t), device, iters=200)
    if rank == 0:
        print("COMM_BENCH " + json.dumps(out), flush=True)
    env.shutdown_distributed()


if __name__ == "__main__":
    main()
