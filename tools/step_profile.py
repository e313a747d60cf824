"""One eager flagship training step between cudaProfilerStart/Stop, for
`ncu --profile-from-start off --metrics gpu__time_duration.sum ...` launch lists (shares, not absolutes).
The step is engine.StepRunner's (graph capture off so that ncu sees every kernel)."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--batch", type=int, default=1200)
    p.add_argument("--mode", default="ddp_amp")
    p.add_argument("--amp-dtype", default="fp16")
    p.add_argument("--warm", type=int, default=3)
    p.add_argument("--input", default="uint8")
    a = p.parse_args()
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT="29571", RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    from pytorch_distributed_template_b200 import engine, env
    device = env.init_distributed(0)
    engine.work_stream(device)
    args = argparse.Namespace(mode=a.mode, amp_dtype=a.amp_dtype, gpus=1, no_graph=True, warmup=a.warm)
    model, opt, scaler, make_runner, lp = bench.build_ours(args, a.mode, device, 1)
    runner = make_runner()
    if a.input == "uint8":
        x = torch.randint(0, 256, (a.batch, 3, 224, 224), dtype=torch.uint8, device=device)
    else:
        x = torch.randn(a.batch, 3, 224, 224, device=device)
    y = torch.randint(0, 1000, (a.batch,), device=device)
    for _ in range(a.warm):
        runner.step(x, y)
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    runner.step(x, y)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    env.shutdown_distributed()


if __name__ == "__main__":
    main()
