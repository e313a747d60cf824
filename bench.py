#!/usr/bin/env python
"""Headline benchmark: ResNet-18 training images/sec on N B200s (BASELINE.json metric), all four modes.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the unmodified reference, same metric/config
    python bench.py --mode dp|ddp|ddp_amp|ddp_amp_syncbn ...   # one cell only

Config: ResNet-18, synthetic 224x224 ImageNet-shaped data, random init, GLOBAL batch 1200
split over the GPUs (the reference divides ``-b`` by the world size, distributed.py:143) -> strong scaling.

Without ``--mode`` a run produces
  * the headline cell: DDP + amp in fp16 (what the reference's autocast runs), reported as the top-level
    ``value`` / ``e2e``;
  * ``checks``: at N > 1, before anything is timed, every rank verifies the cross-GPU paths against NCCL / torch
    (parameters bit-identical across ranks after the warm-up steps; our reduce-scatter + fused SGD + all-gather
    against ``dist.all_reduce`` + the torch SGD formula; one SyncBatchNorm layer forward/backward against
    ``torch.nn.SyncBatchNorm``; in the dp cell one step against ``torch.nn.DataParallel``).  A failed check makes
    the process exit non-zero after printing the line;
  * ``config.grid``: the other cells of the reference's table (README.md:9-14) — DDP fp32, DDP + amp + SyncBN,
    DataParallel, plus DDP + amp in bf16 — each measured for a few steps in a child process launched with the same
    rank environment (img/s, ms/step, e2e).  ``--impl reference`` produces the same grid for the reference.

Every cell drives training through the package's public step API (``engine.StepRunner`` fed by
``engine.DeviceFeed``) — the objects ``trainer.train`` uses — so the timed step IS the trainer's step: forward,
loss + top-1, the cross-rank metric reduction and meter update, zero_grad, backward with the in-kernel gradient
reduce-scatter, fused SGD + weight all-gather, GradScaler update, replayed from a CUDA graph.
  value  K steps on device-resident synthetic batches (three rotate); CUDA events, barrier + synchronize on both
         sides, max over ranks.
  e2e    K steps with, every step, a pinned-host -> device copy of that step's images/targets (uint8 pixels,
         normalised on the device; ``e2e_float_input`` repeats it with host-normalised fp32 like the reference's
         loader) and a device -> host read of the step's loss/accuracy.
Timing hygiene: W >= 3 warm-up steps; per-step working set (activations of >= 150 images, several GB) far
exceeds the 126 MB L2 and three distinct input batches rotate; nvidia-smi clocks are sampled during the timed region.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GRID = (("ddp", []), ("ddp_amp_syncbn", []), ("ddp_amp_bf16", ["--amp-dtype", "bf16"]), ("dp", []))


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=20)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--mode", default=None, choices=["dp", "ddp", "ddp_amp", "ddp_amp_syncbn"],
                   help="one cell only; default: headline ddp_amp + the grid of the other cells")
    p.add_argument("--batch", type=int, default=1200, help="GLOBAL batch (reference default, distributed.py:55)")
    p.add_argument("--amp-dtype", default=os.environ.get("PDT_AMP_DTYPE", "fp16"), choices=["fp16", "bf16"])
    p.add_argument("--no-graph", action="store_true")
    p.add_argument("--no-grid", action="store_true")
    p.add_argument("--checks", default="full", choices=["full", "basic", "none"])
    p.add_argument("--grid-steps", type=int, default=10)
    p.add_argument("--cell-out", default=None, help="(internal) write the result JSON here instead of stdout")
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


def emit(obj, args):
    line = json.dumps(obj)
    if args.cell_out:
        with open(args.cell_out, "w") as f:
            f.write(line)
    else:
        print(line, flush=True)


def baseline_number(mode):
    """train images/s derived from the reference's README table (3x TITAN Xp)."""
    return {"dp": 839.0, "ddp": 1389.0, "ddp_amp": 1369.0, "ddp_amp_syncbn": 784.0}[mode]


# --------------------------------------------------------------------------------------
def run_reference(args, world, mode):
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import shim
    if not shim.available():
        return {"impl": "reference", "unavailable": "baseline/_ref is empty (pip install of /root/reference not done)"}
    rank = int(os.environ["RANK"])
    if mode == "dp" and rank != 0:
        return None
    sampler = ClockSampler(int(os.environ["LOCAL_RANK"]))
    if rank == 0:
        sampler.start()
    res = shim.run_reference(mode, args.batch, args.steps, args.warmup, on_timed_region=sampler.mark)
    import torch
    dev = torch.device("cuda", torch.cuda.current_device())
    ms = max_over_ranks(res["device_ms"], dev) / args.steps
    wall = max_over_ranks(res["wall_ms"], dev) / args.steps
    sampler.stop()
    out = None
    if rank == 0:
        per_rank = args.batch // (world if mode != "dp" else 1)
        img_bytes = per_rank * 3 * args.image_size * args.image_size * 4 + per_rank * 8
        value = args.batch / (ms / 1e3)
        out = {"impl": "reference", "metric": "resnet18_train_images_per_sec", "value": value, "unit": "images/s",
               "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
               "higher_is_better": True, "scaling": "strong", "vs_baseline": value / baseline_number(mode),
               "dtype": {"ddp_amp": "fp16", "ddp_amp_syncbn": "fp16"}.get(mode, "fp32(tf32 convs)"),
               "data": "synthetic", "clocks": sampler.summary(),
               "e2e": {"value": args.batch / (wall / 1e3), "unit": "images/s", "h2d_bytes_per_step": img_bytes,
                       "d2h_bytes_per_step": 4},
               "gpu_launches": 0,
               "config": {"model": "resnet18", "global_batch": args.batch, "image": args.image_size,
                          "mode": mode, "parallelism": "{}{}".format("dp" if mode == "dp" else "ddp", args.gpus),
                          "l2": "per-step working set >> 126 MB L2; 3 input batches rotate",
                          "path": "unmodified reference via baseline/shim.py (torch DDP + NCCL + cuDNN)"}}
    import torch.distributed as dist
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    return out


# --------------------------------------------------------------------------------------
class HostBatches:
    """What a DataLoader(pin_memory=True) hands over: pinned host (images, target) pairs, three distinct ones rotating."""

    def __init__(self, images, targets, n):
        self.images, self.targets, self.n = images, targets, n

    def __len__(self):
        return self.n

    def __iter__(self):
        for i in range(self.n):
            yield self.images[i % len(self.images)], self.targets[i % len(self.targets)]


def build_ours(args, mode, device, world):
    import torch
    from pytorch_distributed_template_b200 import amp, engine, optim
    from pytorch_distributed_template_b200.models import create_model
    from pytorch_distributed_template_b200.ops import CrossEntropyLoss
    from pytorch_distributed_template_b200.parallel import DataParallel, DistributedDataParallel, SyncBatchNorm
    use_amp = mode in ("ddp_amp", "ddp_amp_syncbn")
    lp = {"bf16": torch.bfloat16, "fp16": torch.float16}[args.amp_dtype] if use_amp else None
    model = create_model("resnet18")
    if mode == "ddp_amp_syncbn":
        model = SyncBatchNorm.convert_sync_batchnorm(model)
    if mode == "dp":
        model = DataParallel(model, device_ids=list(range(args.gpus)), shadow_dtype=lp).to(device)
    else:
        model = DistributedDataParallel(model.to(device), device_ids=[device.index], shadow_dtype=lp)
    crit = CrossEntropyLoss().to(device)
    opt = optim.SGD(model.parameters(), 0.1, momentum=0.9, weight_decay=1e-4)
    scaler = amp.GradScaler(enabled=use_amp, device=device)
    if use_amp:
        scaler.bind(opt)
    model.train()

    def make_runner():
        return engine.StepRunner(model, crit, opt, scaler, device, use_amp=use_amp, amp_dtype=lp,
                                 graph=False if args.no_graph else None, eager_steps=max(3, args.warmup) - 1)
    return model, opt, scaler, make_runner, lp


def run_checks(args, mode, model, opt, world):
    """Cross-GPU self-checks on the live job (module docstring).  Returns (dict, all_pass)."""
    if world <= 1 or args.checks == "none" or mode == "dp":
        return None, True
    from pytorch_distributed_template_b200.parallel import selfcheck
    out, ok = {}, True
    out["ranks_identical_after_warmup"] = selfcheck.ranks_identical(model)
    ok &= out["ranks_identical_after_warmup"]
    if args.checks == "full":
        gp = selfcheck.gradient_path(model, opt)
        shadow_tol = 1e-2 if args.amp_dtype == "bf16" else 2e-3
        gp["pass"] = bool(gp["param_rel"] <= 1e-5 and gp["momentum_rel"] <= 1e-4 and gp.get("shadow_rel", 0.0) <= shadow_tol)
        gp["oracle"] = "dist.all_reduce (NCCL) / world + torch.optim.SGD formula; bounds 1e-5 param, 1e-4 momentum"
        out["grad_reduce_sgd_vs_nccl"] = gp
        ok &= gp["pass"]
        sb = selfcheck.syncbn_layer(model.fabric)
        worst_key = max(sb, key=lambda k: sb[k] / selfcheck.syncbn_tolerance(k))
        sb_ok = all(v <= selfcheck.syncbn_tolerance(k) for k, v in sb.items())
        out["syncbn_layer_vs_torch"] = {"pass": bool(sb_ok), "worst": worst_key, "worst_rel": sb[worst_key],
                                        "fp32_max_rel": max(v for k, v in sb.items() if k.startswith("fp32")),
                                        "bf16_max_rel": max(v for k, v in sb.items() if k.startswith("bf16")),
                                        "fp16_max_rel": max(v for k, v in sb.items() if k.startswith("fp16")),
                                        "oracle": "torch.nn.SyncBatchNorm at the same world size; bounds 1e-5 fp32, 2e-2 bf16, 4e-3 fp16"}
        ok &= sb_ok
        out["ranks_identical_after_checks"] = selfcheck.ranks_identical(model)
        ok &= out["ranks_identical_after_checks"]
    return out, bool(ok)


def run_ours(args, world, mode):
    import torch
    import torch.distributed as dist
    from pytorch_distributed_template_b200 import engine, env
    from pytorch_distributed_template_b200.ops import _native
    rank, local_rank = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    checks, checks_ok = None, True
    if mode == "dp":
        if rank != 0:
            return None, True
        device = torch.device("cuda", 0)
        torch.cuda.set_device(0)
        world_for_batch = 1
        _native.require()
        if args.gpus > 1 and args.checks != "none":
            from pytorch_distributed_template_b200.parallel import selfcheck
            dpc = selfcheck.dataparallel_step(list(range(args.gpus)))
            dpc["pass"] = bool(dpc["loss_rel"] <= 1e-2 and dpc["update_rel_l2"] <= 3e-2 and dpc["replicas_equal"])
            dpc["oracle"] = "torch.nn.DataParallel over torchvision resnet18, eval-mode BN, 64 images per replica; bounds 1e-2 loss, 3e-2 update in the 2-norm (TF32 convs on both sides)"
            checks, checks_ok = {"dataparallel_step_vs_torch": dpc}, dpc["pass"]
    else:
        device = env.init_distributed(local_rank)
        world_for_batch = world
    _native.require()
    per_rank = args.batch // world_for_batch
    # Everything — construction, warm-up, capture, replay — runs on ONE non-default stream (engine.work_stream)
    engine.work_stream(device)
    model, opt, scaler, make_runner, lp = build_ours(args, mode, device, world)
    runner = make_runner()

    g = torch.Generator().manual_seed(1234 + rank)
    S = args.image_size
    u8 = args.input == "uint8"

    def host_batches(as_u8):
        if as_u8:
            imgs = [torch.randint(0, 256, (per_rank, 3, S, S), generator=g, dtype=torch.uint8).pin_memory() for _ in range(3)]
        else:
            imgs = [torch.randn(per_rank, 3, S, S, generator=g).pin_memory() for _ in range(3)]
        tgts = [torch.randint(0, 1000, (per_rank,), generator=g).pin_memory() for _ in range(3)]
        return imgs, tgts

    host_imgs, host_tgts = host_batches(u8)
    dev_imgs = [t.to(device) for t in host_imgs]
    dev_tgts = [t.to(device) for t in host_tgts]

    def fence():
        for d in range(args.gpus if mode == "dp" else 0):
            torch.cuda.synchronize(d)
        torch.cuda.synchronize()
        if dist.is_initialized() and dist.get_world_size() > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- warm-up: the runner's eager iterations, then the capture and two replays --------------------
    for i in range(max(3, args.warmup)):
        runner.step(dev_imgs[i % 3], dev_tgts[i % 3])
    fence()
    if mode != "dp":
        c, checks_ok = run_checks(args, mode, model, opt, world)
        checks = c
    for i in range(2):
        runner.step(dev_imgs[i % 3], dev_tgts[i % 3])
    fence()

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
        runner.step(dev_imgs[i % 3], dev_tgts[i % 3])
    stop.record()
    fence()
    sampler.mark(False)
    eager_launches = _native.launch_count - n0
    ms = max_over_ranks(start.elapsed_time(stop), device) / args.steps
    final_loss = float(runner.result[0].item())
    graphed = runner.graphs > 0

    # ---- measurement 2: end to end through the public API ------------------------------------------
    # per step: H2D of that step's batch from pinned memory (engine.DeviceFeed), the training step, D2H of the result
    def e2e(rn, imgs, tgts, nsteps, sink):
        # the device -> host read of every step's (loss, acc1) is part of the step itself: StepRunner copies its 8-byte
        # result into pinned host memory (runner.result_host) as the last node of the step / of the captured graph
        feed = engine.DeviceFeed(HostBatches(imgs, tgts, nsteps), rn)
        for i, (di, dt) in enumerate(feed):
            rn.step(di, dt)
        torch.cuda.current_stream(device).synchronize()
        sink[min(nsteps, sink.shape[0]) - 1].copy_(rn.result_host)

    def timed_e2e(rn, imgs, tgts):
        sink = torch.zeros(args.steps, 2, dtype=torch.float32).pin_memory()
        e2e(rn, imgs, tgts, max(4, min(args.warmup, args.steps)), sink)     # both staging buffers get their graph
        fence()
        t0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        e2e(rn, imgs, tgts, args.steps, sink)
        e1.record()
        fence_local()
        wall = time.perf_counter() - t0
        fence()
        w = max_over_ranks(wall * 1e3, device) / args.steps
        d = max_over_ranks(e0.elapsed_time(e1), device) / args.steps
        return max(w, d), w, d

    def fence_local():
        for d in range(args.gpus if mode == "dp" else 0):
            torch.cuda.synchronize(d)
        torch.cuda.synchronize()

    sampler.mark(True)
    e2e_ms, e2e_wall, e2e_dev = timed_e2e(runner, host_imgs, host_tgts)
    sampler.mark(False)
    e2e_float = None
    if u8 and args.mode is None and not args.cell_out:
        # the same with the reference's input format: fp32 NCHW batches normalised on the host (4x the H2D bytes)
        f_imgs, f_tgts = host_batches(False)
        runner_f = make_runner()
        runner_f.eager_left = 1
        fm, fw, fd = timed_e2e(runner_f, f_imgs, f_tgts)
        e2e_float = {"value": args.batch / (fm / 1e3), "unit": "images/s", "ms_per_step": fm,
                     "h2d_bytes_per_step": per_rank * 3 * S * S * 4 + per_rank * 8, "d2h_bytes_per_step": 8,
                     "input": "fp32 NCHW, host-normalised (the reference's loader format)"}
    sampler.stop()

    out = None
    if rank == 0:
        value = args.batch / (ms / 1e3)
        h2d = per_rank * 3 * S * S * (1 if u8 else 4) + per_rank * 8
        launches = runner.launches_per_step * args.steps if graphed else eager_launches
        fabric = getattr(model, "fabric", None)
        out = {"impl": "ours", "metric": "resnet18_train_images_per_sec", "value": value, "unit": "images/s",
               "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
               "higher_is_better": True, "scaling": "strong", "vs_baseline": value / baseline_number(mode),
               "dtype": (args.amp_dtype if lp is not None else "fp32(tf32 convs)"), "data": "synthetic",
               "clocks": sampler.summary(),
               "e2e": {"value": args.batch / (e2e_ms / 1e3), "unit": "images/s", "h2d_bytes_per_step": h2d,
                       "d2h_bytes_per_step": 8, "ms_per_step": e2e_ms, "wall_ms_per_step": e2e_wall,
                       "device_ms_per_step": e2e_dev},
               "gpu_launches": int(launches),
               "config": {"model": "resnet18", "global_batch": args.batch, "image": S, "mode": mode,
                          "parallelism": "{}{}".format("dp" if mode == "dp" else "ddp", args.gpus),
                          "cuda_graph": graphed, "graph_capture_error": runner.capture_error, "final_loss": final_loss,
                          "step_api": "engine.StepRunner.step fed by engine.DeviceFeed (what trainer.train runs)",
                          "d2h": "every step copies its (loss, acc1) to pinned host memory (StepRunner.result_host), as a node of the step graph",
                          "loss_scale": scaler.get_scale() if lp is not None else None,
                          "skipped_steps": scaler.skipped_steps() if lp is not None else None,
                          "input": "uint8 NCHW, (x/255-mean)/std fused into the stem kernel" if u8 else "fp32 NCHW, host-normalised",
                          "fabric": fabric.describe() if fabric is not None else None,
                          "l2": "per-step working set >> 126 MB L2; 3 input batches rotate"}}
        if e2e_float is not None:
            out["e2e_float_input"] = e2e_float
        if checks is not None:
            checks["all_pass"] = bool(checks_ok)
            out["checks"] = checks
    if dist.is_initialized():
        ok_t = torch.tensor([1.0 if checks_ok else 0.0], device=device)
        dist.all_reduce(ok_t, op=dist.ReduceOp.MIN)
        checks_ok = bool(ok_t.item() > 0)
        if out is not None and "checks" in out:
            out["checks"]["all_pass"] = checks_ok
        dist.barrier()
        dist.destroy_process_group()
    return out, checks_ok


# --------------------------------------------------------------------------------------
def run_grid(args, world):
    """The other cells, each in a child process with this rank's environment (fresh CUDA context, fresh process
    group on its own port).  Every rank runs its own child for the DDP cells; the single-process dp cell runs
    on rank 0 alone, last, while the other ranks are done."""
    rank = int(os.environ["RANK"])
    base_port = int(os.environ.get("MASTER_PORT", "29533"))
    grid = {}
    steps = max(3, min(args.steps, args.grid_steps))
    warm = max(3, min(args.warmup, 5))
    for k, (name, extra) in enumerate(GRID):
        mode = name if name in ("dp", "ddp", "ddp_amp", "ddp_amp_syncbn") else "ddp_amp"
        if (mode == "dp" and rank != 0) or (args.impl == "reference" and extra):
            continue          # dp is one process; the reference has no bf16 variant
        env = dict(os.environ)
        env["MASTER_PORT"] = str(base_port + 11 + k)
        if mode == "dp":
            env.update(RANK="0", LOCAL_RANK="0", WORLD_SIZE="1")
        for v in ("TORCHELASTIC_RUN_ID", "TORCHELASTIC_RESTART_COUNT", "TORCHELASTIC_MAX_RESTARTS", "TORCHELASTIC_USE_AGENT_STORE"):
            env.pop(v, None)
        with tempfile.TemporaryDirectory() as d:
            cell = os.path.join(d, "cell.json")
            cmd = [sys.executable, os.path.abspath(__file__), "--impl", args.impl, "--mode", mode, "--gpus", str(args.gpus),
                   "--steps", str(steps), "--warmup", str(warm), "--batch", str(args.batch), "--image-size",
                   str(args.image_size), "--no-grid", "--checks", "basic" if args.checks != "none" else "none",
                   "--cell-out", cell] + extra
            if args.impl == "ours":
                cmd += ["--input", args.input] + (["--no-graph"] if args.no_graph else [])
            t0 = time.time()
            try:
                r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
                rc, tail = r.returncode, (r.stdout[-1500:] + r.stderr[-1500:])
            except subprocess.TimeoutExpired:
                rc, tail = -9, "timeout"
            if rank == 0:
                if os.path.exists(cell):
                    c = json.load(open(cell))
                    keep = {k2: c.get(k2) for k2 in ("value", "ms_per_step", "dtype", "steps", "warmup", "gpu_launches", "unavailable")
                            if k2 in c}
                    if "e2e" in c:
                        keep["e2e"] = {k2: c["e2e"].get(k2) for k2 in ("value", "ms_per_step", "h2d_bytes_per_step") if k2 in c["e2e"]}
                    cfg = c.get("config", {})
                    for k2 in ("cuda_graph", "final_loss", "skipped_steps", "parallelism"):
                        if k2 in cfg:
                            keep[k2] = cfg[k2]
                    if "checks" in c:
                        keep["checks"] = c["checks"]
                    keep["exit_code"] = rc
                    keep["seconds"] = round(time.time() - t0, 1)
                    grid[name] = keep
                else:
                    grid[name] = {"failed": True, "exit_code": rc, "tail": tail[-600:]}
    return grid


def main():
    args = parse()
    world = setup_env(args)
    mode = args.mode or "ddp_amp"
    if args.impl == "reference":
        out, ok = run_reference(args, world, mode), True
    else:
        out, ok = run_ours(args, world, mode)
    grid = None
    if args.mode is None and not args.no_grid and not (out or {}).get("unavailable"):
        grid = run_grid(args, world)
    if out is not None:
        if grid is not None:
            out.setdefault("config", {})["grid"] = grid
            # a cell whose own checks failed fails the run; a cell that could not run is recorded, not fatal
            ok = ok and all((c.get("checks") or {}).get("all_pass", True) for c in grid.values())
        emit(out, args)
    if not ok:
        sys.exit(3)


if __name__ == "__main__":
    main()
