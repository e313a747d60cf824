"""One epoch loop for the three entrypoints (SURVEY layer L2, §3.1-3.5).

The reference triplicates ``main_worker`` / ``train`` / ``validate``
(/root/reference/dataparallel.py:99-281, distributed.py:108-334,
distributed_syncBN_amp.py:111-349); here ``run(args, mode)`` serves all three
with the same log lines, TensorBoard scalars and checkpoint dict.

Hot-loop differences: the reference drains the stream twice per
iteration (``dist.barrier()`` + ``loss.item()``).  Here loss and accuracy are
reduced across ranks by one 8-byte in-kernel all-reduce (which also IS the
barrier) into device-resident meters that are read only when a line is printed,
and the whole iteration is one ``engine.StepRunner.step`` — after a few eager
iterations a single CUDA-graph replay fed by ``engine.DeviceFeed`` (the same
objects ``bench.py`` times).

Environment knobs (the CLI stays the reference's): PDT_SYNTHETIC, PDT_MAX_STEPS,
PDT_MAX_VAL_STEPS, PDT_AMP_DTYPE, PDT_DEVICE, PDT_MODEL_IMPL, PDT_OUTPUT_EXISTS,
PDT_SYNTH_TRAIN_SIZE, PDT_SYNTH_VAL_SIZE, PDT_IMAGE_SIZE, PDT_NUM_CLASSES, PDT_GRAPH,
PDT_GRAPH_WARMUP, PDT_NVTX, PDT_INPUT, PDT_NVLS, PDT_SPIN_TIMEOUT_S (INTEGRATION.md lists them all).
"""
import os
import random
import time
import warnings

import numpy as np
import torch
import torch.distributed as dist
import torch.nn as nn

from . import amp, data, engine, env, optim
from .config import LOGGER_NAME, normalise_steps
from .models import create_model
from .ops import CrossEntropyLoss
from .parallel import DataParallel, DistributedDataParallel, SyncBatchNorm
from .utils import (AverageMeter, ddp_print, get_learning_rate, get_logger,
                    make_summary_writer, output_process, save_checkpoint, write_settings)


def seed_everything(args):
    if args.seed is not None:
        random.seed(args.seed)
        np.random.seed(args.seed)       # the reference calls np.random(args.seed) and crashes
        torch.manual_seed(args.seed)
        if torch.cuda.is_available():
            torch.cuda.manual_seed_all(args.seed)
        torch.backends.cudnn.deterministic = True
        torch.backends.cudnn.benchmark = False
        # the native weight-gradient kernels honour the same promise: ordered split-K instead of atomics
        os.environ.setdefault("PDT_DETERMINISTIC", "1")
        warnings.warn("You have chosen to seed training.This will turn on the cudnn deterministic setting,"
                      "which can slow down your training considerably!"
                      "You may see unexpected behavior when restarting from checkpoint.")
    else:
        torch.backends.cudnn.benchmark = True


class Context:
    """Everything ``train`` / ``validate`` need, built once by ``setup``."""
    pass


def setup(args, mode):
    ctx = Context()
    ctx.mode, ctx.args = mode, args
    distributed = mode in ("ddp", "ddp_amp")
    ctx.local_rank = env.resolve_local_rank(args) if distributed else 0
    args.local_rank = ctx.local_rank if distributed else getattr(args, "local_rank", 0)
    args.outpath = args.outpath + "_" + args.arch
    args.step = normalise_steps(args.step)
    ctx.logger = ctx.writer = None
    ctx.is_main = ctx.local_rank == 0
    if ctx.is_main:
        output_process(args.outpath)
        ctx.logger = get_logger(args.outpath, LOGGER_NAME[mode])
        ctx.writer = make_summary_writer(args.outpath)

    if distributed:
        ctx.device = env.init_distributed(ctx.local_rank)
        args.nprocs = dist.get_world_size()       # the reference uses device_count()
    else:
        ctx.device = torch.device("cuda", 0) if env.use_cuda() else torch.device("cpu")
        args.nprocs = max(1, torch.cuda.device_count()) if ctx.device.type == "cuda" else 1
    engine.work_stream(ctx.device)      # everything from the first forward on runs on this stream (graph capture)
    if ctx.is_main:
        write_settings(args)
        ctx.logger.info(args)

    use_amp = bool(getattr(args, "use_amp", False)) and ctx.device.type == "cuda"
    ctx.use_amp = use_amp
    ctx.amp_dtype = amp.default_dtype() if use_amp else None
    num_classes = env.env_int("PDT_NUM_CLASSES", 0) or None
    if args.pretrained:
        ddp_print("=> using pre-trained model: {}".format(args.arch), ctx.logger, ctx.local_rank)
    else:
        ddp_print("=> creating model: {}".format(args.arch), ctx.logger, ctx.local_rank)
    model = create_model(args.arch, pretrained=args.pretrained, num_classes=num_classes)

    if mode == "ddp_amp":
        if dist.is_available() and dist.is_initialized() and args.sync_batchnorm:
            ddp_print("=> using sync BN", ctx.logger, ctx.local_rank)
            model = SyncBatchNorm.convert_sync_batchnorm(model)
        else:
            ddp_print("=> not use sync BN", ctx.logger, ctx.local_rank)

    if distributed:
        model = model.to(ctx.device)
        args.batch_size = int(args.batch_size / args.nprocs)
        model = DistributedDataParallel(model, device_ids=[ctx.local_rank] if ctx.device.type == "cuda" else None,
                                        shadow_dtype=ctx.amp_dtype)
    else:
        model = DataParallel(model, shadow_dtype=ctx.amp_dtype)
        model = model.to(ctx.device)
    ctx.model = model

    ctx.criterion = CrossEntropyLoss().to(ctx.device)
    ctx.optimizer = optim.SGD(model.parameters(), args.lr, momentum=args.momentum, weight_decay=args.weight_decay)
    if args.lr_scheduler == "steplr":
        ctx.lr_scheduler = optim.MultiStepLR(ctx.optimizer, milestones=args.step, gamma=args.gamma)
        ddp_print("lr_scheduler: SGD MultiStepLR !!!", ctx.logger, ctx.local_rank)
    else:
        ddp_print("invalid lr_scheduler={}".format(args.lr_scheduler), ctx.logger, ctx.local_rank)
        raise AssertionError("invalid lr_scheduler={}".format(args.lr_scheduler))
    ctx.scaler = amp.GradScaler(enabled=use_amp, device=ctx.device)
    if use_amp:
        ctx.scaler.bind(ctx.optimizer)

    # On a GPU the loader ships raw uint8 pixels and (x / 255 - mean) / std runs on the device (fused into the
    # stem's layout kernel): a quarter of the H2D bytes of the reference's host-side ToTensor + Normalize.
    # PDT_INPUT=float keeps the host-side pipeline; models that cannot take uint8 get it automatically.
    inner = getattr(model, "module", model)
    ctx.uint8_input = (ctx.device.type == "cuda" and os.environ.get("PDT_INPUT", "uint8") == "uint8"
                       and getattr(inner, "accepts_uint8", False))
    train_set, val_set = data.build_datasets(args, uint8=ctx.uint8_input)
    pin = ctx.device.type == "cuda"
    if distributed:
        ctx.train_sampler = data.ShardSampler(train_set)
        ctx.val_sampler = data.ShardSampler(val_set)      # val is sharded AND shuffled, like distributed.py:177
        ctx.train_loader = data.build_loader(train_set, args.batch_size, args.workers, sampler=ctx.train_sampler, pin=pin)
        ctx.val_loader = data.build_loader(val_set, args.batch_size, args.workers, sampler=ctx.val_sampler, pin=pin)
    else:
        ctx.train_sampler = ctx.val_sampler = None
        ctx.train_loader = data.build_loader(train_set, args.batch_size, args.workers, shuffle=True, pin=pin)
        ctx.val_loader = data.build_loader(val_set, args.batch_size, args.workers, shuffle=False, pin=pin)
    ctx.max_steps = env.env_int("PDT_MAX_STEPS", 0)
    ctx.max_val_steps = env.env_int("PDT_MAX_VAL_STEPS", ctx.max_steps)
    ctx.fabric = getattr(model, "fabric", None)
    ctx.runner = engine.StepRunner(model, ctx.criterion, ctx.optimizer, ctx.scaler, ctx.device, use_amp=use_amp,
                                   amp_dtype=ctx.amp_dtype)
    return ctx


def train(ctx, epoch):
    args, log, runner = ctx.args, ctx.logger, ctx.runner
    batch_times, data_times = AverageMeter("Time", ":6.3f"), AverageMeter("Data", ":6.3f")
    ctx.model.train()
    runner.meter.reset()
    loader = engine.DeviceFeed(ctx.train_loader, runner)
    n_iter = len(loader)
    end = time.time()
    for i, (images, target) in enumerate(loader):
        data_times.update(time.time() - end)
        runner.step(images, target)       # forward, metrics, zero_grad, backward (+ gradient exchange), SGD, scaler
        batch_times.update(time.time() - end)
        end = time.time()
        if i % args.print_freq == 0:
            m = runner.meter.read()
            ddp_print("Train epoch: [{:d}/{:d}][{:d}/{:d}]\tlr={:.6f}\tce_loss={:.4f}\ttop1_acc={:.4f}\tdata_time={:6.3f}s"
                      "\tbatch_time={:6.3f}s".format(epoch, args.epochs, i, n_iter, get_learning_rate(ctx.optimizer),
                                                    m["loss"], m["acc1"], data_times.avg, batch_times.avg), log, ctx.local_rank)
        if ctx.max_steps and i + 1 >= ctx.max_steps:
            break
    m = runner.meter.read()
    l, a = m["loss"], m["acc1"]
    ddp_print("||==> Train epoch: [{:d}/{:d}]\tlr={:.6f}\tce_loss={:.4f}\ttop1_acc={:.4f}\tbatch_time={:6.3f}s"
              .format(epoch, args.epochs, get_learning_rate(ctx.optimizer), l, a, batch_times.avg), log, ctx.local_rank)
    if ctx.is_main:
        ctx.writer.add_scalar("lr", get_learning_rate(ctx.optimizer), epoch)
        ctx.writer.add_scalar("Train_ce_loss", l, epoch)
        ctx.writer.add_scalar("Train_top1_accuracy", a, epoch)
    return l, a


def validate(ctx, epoch):
    args, log, runner = ctx.args, ctx.logger, ctx.runner
    batch_times = AverageMeter("Time", ":6.3f")
    meter = engine.Meter(ctx.device)
    ctx.model.eval()
    with torch.no_grad():
        loader = data.PinnedPrefetcher(ctx.val_loader, ctx.device)
        n_iter = len(loader)
        end = time.time()
        for i, (images, target) in enumerate(loader):
            runner.eval_step(images, target, meter)
            batch_times.update(time.time() - end)
            end = time.time()
            if i % args.print_freq == 0:
                m = meter.read()
                ddp_print("Val epoch: [{:d}/{:d}][{:d}/{:d}]\tce_loss={:.4f}\ttop1_acc={:.4f}\tbatch_time={:6.3f}s"
                          .format(epoch, args.epochs, i, n_iter, m["loss"], m["acc1"], batch_times.avg), log, ctx.local_rank)
            if ctx.max_val_steps and i + 1 >= ctx.max_val_steps:
                break
        m = meter.read()
        l, a = m["loss"], m["acc1"]
        ddp_print("||==> Val epoch: [{:d}/{:d}]\tce_loss={:.4f}\ttop1_acc={:.4f}\tbatch_time={:6.3f}s"
                  .format(epoch, args.epochs, l, a, batch_times.avg), log, ctx.local_rank)
        if ctx.is_main:
            ctx.writer.add_scalar("Val_ce_loss", l, epoch)
            ctx.writer.add_scalar("Val_top1_accuracy", a, epoch)
    return a


def run(args, mode):
    """``main_worker`` of all three scripts."""
    seed_everything(args)
    ctx = setup(args, mode)
    args = ctx.args
    try:
        if args.evaluate:
            validate(ctx, epoch=-1)
            return 0
        best_acc1, best_acc1_index = 0, 0
        total_start = time.time()
        for epoch in range(args.start_epoch, args.epochs):
            if ctx.train_sampler is not None:
                ctx.train_sampler.set_epoch(epoch)
                ctx.val_sampler.set_epoch(epoch)
            epoch_start = time.time()
            ctx.lr_scheduler.step(epoch)
            train(ctx, epoch)
            acc1 = validate(ctx, epoch)
            is_best = acc1 > best_acc1
            if is_best:
                best_acc1_index, best_acc1 = epoch, acc1
            ddp_print("||==> Epoch=[{:d}/{:d}]\tbest_acc1={:.4f}\tbest_acc1_index={}\ttime_cost={:.4f}s"
                      .format(epoch, args.epochs, best_acc1, best_acc1_index, time.time() - epoch_start),
                      ctx.logger, ctx.local_rank)
            if ctx.is_main:
                save_checkpoint({"epoch": epoch + 1, "arch": args.arch,
                                 "state_dict": ctx.model.module.state_dict(),
                                 "best_acc1": best_acc1}, is_best, args.outpath)
            if mode != "dp" and dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
                # Rank 0 may spend a long time writing the checkpoint; the peers wait HERE on the host instead
                # of inside the next epoch's first device-side exchange, whose spin budget is finite.
                dist.barrier()
        ddp_print("||==> total_time_cost={:.4f}s".format(time.time() - total_start), ctx.logger, ctx.local_rank)
        if env.env_flag("PDT_REPORT_GRAPHS") and ctx.is_main:      # not part of the reference's log format: opt-in
            print("[pdt] training step: {} CUDA graphs, {} kernels per replay{}".format(
                ctx.runner.graphs, ctx.runner.launches_per_step,
                "" if ctx.runner.capture_error is None else " (capture failed: {})".format(ctx.runner.capture_error)), flush=True)
        if ctx.is_main:
            ctx.writer.close()
        return ctx
    finally:
        if mode != "dp":
            env.shutdown_distributed()
