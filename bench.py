#!/usr/bin/env python
"""Headline benchmark: ResNet-18 training images/sec on N B200s (BASELINE.json metric).

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the unmodified reference, same metric/config

Config: ResNet-18, synthetic 224x224 ImageNet-shaped data, random init,
GLOBAL batch 1200 split over the GPUs (the reference divides ``-b`` by the world size,
distributed.py:143) -> strong scaling.  Default mode is DDP + amp (the tensor-core mode);
``--mode`` selects the other three cells of the reference's table.

Own arm, two measurements per run:
  value  K full training steps (forward, loss, backward with the in-kernel gradient
         reduce-scatter, fused SGD + weight all-gather, GradScaler update) replayed from a
         CUDA graph on device-resident synthetic batches; CUDA events, barrier +
         synchronize on both sides, max over ranks.
  e2e    the same K steps driven through the public API with, every step, a pinned-host ->
         device copy of that step's images/targets and a device -> host read of the loss.
Timing hygiene: W >= 3 warm-up steps; per-step working set (activations of >= 150 images,
several GB) far exceeds the 126 MB L2 and three distinct input batches rotate; nvidia-smi
clocks are sampled during the timed region.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=20)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--mode", default="ddp_amp", choices=["dp", "ddp", "ddp_amp", "ddp_amp_syncbn"])
    p.add_argument("--batch", type=int, default=1200, help="GLOBAL batch (reference default, distributed.py:55)")
    p.add_argument("--amp-dtype", default=os.environ.get("PDT_AMP_DTYPE", "bf16"))
    p.add_argument("--no-graph", action="store_true")
    p.add_argument("--image-size", type=int, default=224)
    p.add_argument("--input", default=os.environ.get("PDT_INPUT", "uint8"), choices=["uint8", "float"],
                   help="ours: what the loader ships to the GPU (uint8 pixels normalised on the device, or host-normalised fp32)")
    return p.parse_args()


class ClockSampler:
    """nvidia-smi clocks + throttle reasons while the timed region runs (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.gpu = gpu_index
        self.samples = []
        self.proc = None
        self.thread = None
        self.active = False

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "25"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._pump, daemon=True)
        self.thread.start()

    def _pump(self):
        for line in self.proc.stdout:
            if self.active:
                self.samples.append(line.strip())

    def mark(self, on):
        self.active = on

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            f = [x.strip() for x in s.split(",")]
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except Exception:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


def setup_env(args):
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("LOCAL_RANK", "0")
    os.environ.setdefault("WORLD_SIZE", "1")
    world = int(os.environ["WORLD_SIZE"])
    if args.impl == "reference":
        # the reference sizes the job with torch.cuda.device_count()
        n = world if args.mode != "dp" else args.gpus
        os.environ["CUDA_VISIBLE_DEVICES"] = ",".join(str(i) for i in range(n))
    return world


def max_over_ranks(x, device):
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=device)
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def emit(obj):
    print(json.dumps(obj), flush=True)


def baseline_number(mode):
    """train images/s derived from the reference's README table (3x TITAN Xp)."""
    return {"dp": 839.0, "ddp": 1389.0, "ddp_amp": 1369.0, "ddp_amp_syncbn": 784.0}[mode]


# --------------------------------------------------------------------------------------
def run_reference(args, world):
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import shim
    if not shim.available():
        emit({"impl": "reference", "unavailable": "baseline/_ref is empty (pip install of /root/reference not done)"})
        return
    rank = int(os.environ["RANK"])
    if args.mode == "dp" and rank != 0:
        return
    sampler = ClockSampler(int(os.environ["LOCAL_RANK"]))
    if rank == 0:
        sampler.start()
    res = shim.run_reference(args.mode, args.batch, args.steps, args.warmup, on_timed_region=sampler.mark)
    import torch
    dev = torch.device("cuda", torch.cuda.current_device())
    ms = max_over_ranks(res["device_ms"], dev) / args.steps
    wall = max_over_ranks(res["wall_ms"], dev) / args.steps
    sampler.stop()
    if rank == 0:
        per_rank = args.batch // (world if args.mode != "dp" else 1)
        img_bytes = per_rank * 3 * args.image_size * args.image_size * 4 + per_rank * 8
        value = args.batch / (ms / 1e3)
        emit({"impl": "reference", "metric": "resnet18_train_images_per_sec", "value": value, "unit": "images/s",
              "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
              "higher_is_better": True, "scaling": "strong", "vs_baseline": value / baseline_number(args.mode),
              "dtype": {"ddp_amp": "fp16", "ddp_amp_syncbn": "fp16"}.get(args.mode, "fp32(tf32 convs)"),
              "data": "synthetic", "clocks": sampler.summary(),
              "e2e": {"value": args.batch / (wall / 1e3), "unit": "images/s", "h2d_bytes_per_step": img_bytes,
                      "d2h_bytes_per_step": 4},
              "gpu_launches": 0,
              "config": {"model": "resnet18", "global_batch": args.batch, "image": args.image_size,
                         "mode": args.mode, "parallelism": "{}{}".format("dp" if args.mode == "dp" else "ddp", args.gpus),
                         "l2": "per-step working set >> 126 MB L2; 3 input batches rotate",
                         "path": "unmodified reference via baseline/shim.py (torch DDP + NCCL + cuDNN)"}})
    import torch.distributed as dist
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------
def build_ours(args, device, world):
    import torch
    from pytorch_distributed_template_b200 import amp, optim
    from pytorch_distributed_template_b200.models import create_model
    from pytorch_distributed_template_b200.ops import CrossEntropyLoss
    from pytorch_distributed_template_b200.parallel import DataParallel, DistributedDataParallel, SyncBatchNorm
    use_amp = args.mode in ("ddp_amp", "ddp_amp_syncbn")
    lp = {"bf16": torch.bfloat16, "fp16": torch.float16}[args.amp_dtype] if use_amp else None
    model = create_model("resnet18")
    if args.mode == "ddp_amp_syncbn":
        model = SyncBatchNorm.convert_sync_batchnorm(model)
    if args.mode == "dp":
        model = DataParallel(model, device_ids=list(range(args.gpus)), shadow_dtype=lp).to(device)
    else:
        model = DistributedDataParallel(model.to(device), device_ids=[device.index], shadow_dtype=lp)
    crit = CrossEntropyLoss().to(device)
    opt = optim.SGD(model.parameters(), 0.1, momentum=0.9, weight_decay=1e-4)
    scaler = amp.GradScaler(enabled=use_amp, device=device)
    if use_amp:
        scaler.bind(opt)
    model.train()

    def step(images, target):
        with amp.autocast(enabled=use_amp, dtype=lp):
            out = model(images)
            loss = crit(out, target)
        opt.zero_grad()
        scaler.scale(loss).backward()
        scaler.step(opt)
        scaler.update()
        return loss

    return model, step, lp


def run_ours(args, world):
    import torch
    import torch.distributed as dist
    from pytorch_distributed_template_b200 import env
    from pytorch_distributed_template_b200.ops import _native
    rank, local_rank = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    if args.mode == "dp":
        if rank != 0:
            return
        device = torch.device("cuda", 0)
        torch.cuda.set_device(0)
        world_for_batch = 1
    else:
        device = env.init_distributed(local_rank)
        world_for_batch = world
    _native.require()
    per_rank = args.batch // world_for_batch
    # Everything — construction, warm-up, capture, replay — runs on ONE non-default stream:
    # autograd's AccumulateGrad nodes remember the stream they were created on, and a
    # capture that has to synchronise with the legacy default stream is invalid.
    work_stream = torch.cuda.Stream(device)
    torch.cuda.set_stream(work_stream)
    model, step, lp = build_ours(args, device, world)

    g = torch.Generator().manual_seed(1234 + rank)
    S = args.image_size
    # what the framework's own loader hands over on a GPU (trainer.py / data.py): raw uint8 pixels, normalised on
    # the device inside the stem's layout kernel (--input float: fp32 tensors normalised on the host, like the
    # reference's pipeline)
    u8 = args.input == "uint8"
    if u8:
        host_imgs = [torch.randint(0, 256, (per_rank, 3, S, S), generator=g, dtype=torch.uint8).pin_memory() for _ in range(3)]
    else:
        host_imgs = [torch.randn(per_rank, 3, S, S, generator=g).pin_memory() for _ in range(3)]
    host_tgts = [torch.randint(0, 1000, (per_rank,), generator=g).pin_memory() for _ in range(3)]
    dev_imgs = [t.to(device) for t in host_imgs]
    dev_tgts = [t.to(device) for t in host_tgts]
    static_img, static_tgt = torch.empty_like(dev_imgs[0]), torch.empty_like(dev_tgts[0])

    def fence():
        torch.cuda.synchronize()
        if dist.is_initialized() and dist.get_world_size() > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- warm-up (eager) ----------------------------------------------------------------
    for i in range(max(3, args.warmup)):
        static_img.copy_(dev_imgs[i % 3]); static_tgt.copy_(dev_tgts[i % 3])
        loss = step(static_img, static_tgt)
    fence()

    # ---- capture the whole step ------------------------------------------------------------
    graph, static_loss, launches_per_step = None, None, None
    use_graph = (not args.no_graph) and args.mode != "dp" and os.environ.get("PDT_BENCH_GRAPH", "1") == "1"
    if use_graph:
        try:
            graph = torch.cuda.CUDAGraph()
            n0 = _native.launch_count
            with torch.cuda.graph(graph, stream=work_stream):
                static_loss = step(static_img, static_tgt)
            launches_per_step = _native.launch_count - n0
            fence()
            for _ in range(2):
                graph.replay()
            fence()
        except Exception as e:   # report, then measure eagerly
            if rank == 0:
                print("[bench] CUDA-graph capture failed, running eager: {!r}".format(e), file=sys.stderr)
            graph = None
            torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)

    # ---- measurement 1: device-timed steps -----------------------------------------------------
    fence()
    n0 = _native.launch_count
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.mark(True)
    start.record()
    for i in range(args.steps):
        static_img.copy_(dev_imgs[i % 3]); static_tgt.copy_(dev_tgts[i % 3])
        if graph is not None:
            graph.replay()
        else:
            loss = step(static_img, static_tgt)
    stop.record()
    torch.cuda.synchronize()
    sampler.mark(False)
    eager_launches = _native.launch_count - n0
    fence()
    ms = max_over_ranks(start.elapsed_time(stop), device) / args.steps
    final_loss = float((static_loss if graph is not None else loss).item())

    # ---- measurement 2: end to end through the public API ------------------------------------------
    # per step: H2D of that step's batch from pinned memory, the training step, D2H of the loss
    loss_host = torch.zeros(args.steps, dtype=torch.float32).pin_memory()
    copy_stream = torch.cuda.Stream(device)
    # staging buffers and events exist before the clock starts (a loader allocates them once per run)
    bufs = [(torch.empty_like(dev_imgs[0]), torch.empty_like(dev_tgts[0])) for _ in range(2)]
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]
    main = torch.cuda.current_stream(device)

    _skip = os.environ.get("PDT_E2E_SKIP", "")

    def stage(i):
        b = i % 2
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[b])
            src = dev_imgs if "h2d" in _skip else host_imgs       # PDT_E2E_SKIP: diagnostics only, result is flagged
            bufs[b][0].copy_(src[i % 3], non_blocking=True)
            bufs[b][1].copy_(host_tgts[i % 3], non_blocking=True)
            ready[b].record(copy_stream)

    def e2e_steps(nsteps):
        """nsteps training steps fed from pinned host memory; every step's H2D copy and loss read are inside."""
        for b in range(2):
            consumed[b].record(main)
        stage(0)
        for i in range(nsteps):
            b = i % 2
            if i + 1 < nsteps:
                stage(i + 1)
            main.wait_event(ready[b])
            if graph is not None:
                static_img.copy_(bufs[b][0]); static_tgt.copy_(bufs[b][1])
                consumed[b].record(main)
                graph.replay()
                l = static_loss
            else:
                l = step(bufs[b][0], bufs[b][1])
                consumed[b].record(main)
            if "d2h" not in _skip:
                loss_host[i:i + 1].copy_(l.detach().reshape(1), non_blocking=True)

    # untimed warm-up of this path too (copy stream, staging buffers, host pages), like the W steps above
    e2e_steps(min(args.warmup, args.steps))
    fence()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    e2e_steps(args.steps)
    e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    fence()
    e2e_wall = max_over_ranks(wall * 1e3, device) / args.steps
    e2e_dev = max_over_ranks(e0.elapsed_time(e1), device) / args.steps
    e2e_ms = max(e2e_wall, e2e_dev)
    sampler.stop()

    if rank == 0:
        value = args.batch / (ms / 1e3)
        h2d = per_rank * 3 * S * S * (1 if u8 else 4) + per_rank * 8
        launches = launches_per_step * args.steps if graph is not None else eager_launches
        fabric = getattr(model, "fabric", None)
        emit({"impl": "ours", "metric": "resnet18_train_images_per_sec", "value": value, "unit": "images/s",
              "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
              "higher_is_better": True, "scaling": "strong", "vs_baseline": value / baseline_number(args.mode),
              "dtype": (args.amp_dtype if lp is not None else "fp32(tf32 convs)"), "data": "synthetic",
              "clocks": sampler.summary(),
              "e2e": {"value": args.batch / (e2e_ms / 1e3), "unit": "images/s", "h2d_bytes_per_step": h2d,
                      "d2h_bytes_per_step": 4, "ms_per_step": e2e_ms, "wall_ms_per_step": e2e_wall,
                      **({"INVALID_diagnostic_skip": _skip} if _skip else {}),
                      "device_ms_per_step": e2e_dev},
              "gpu_launches": int(launches),
              "config": {"model": "resnet18", "global_batch": args.batch, "image": S, "mode": args.mode,
                         "parallelism": "{}{}".format("dp" if args.mode == "dp" else "ddp", args.gpus),
                         "cuda_graph": graph is not None, "final_loss": final_loss,
                         "input": "uint8 NCHW, (x/255-mean)/std fused into the stem kernel" if u8 else "fp32 NCHW, host-normalised",
                         "fabric": fabric.describe() if fabric is not None else None,
                         "l2": "per-step working set >> 126 MB L2; 3 input batches rotate"}})
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse()
    world = setup_env(args)
    if args.impl == "reference":
        run_reference(args, world)
    else:
        run_ours(args, world)


if __name__ == "__main__":
    main()
