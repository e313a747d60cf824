"""T5: N training steps of a mode against the UNMODIFIED reference run through baseline/shim.py —
same initial weights, same synthetic batches, same hyper-parameters — comparing the logged loss curve
(/root/reference/distributed.py:237-273 is the loop being matched; its log line is at :270-272).
World size 1 on the driver's single-GPU box; bench.py's in-run checks cover W > 1."""
import os
import re
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_WORKER = r'''
import json, os, sys, torch
sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, "baseline"))
mode, steps, batch, lr, init, out = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], sys.argv[5], sys.argv[6]
import shim, torchvision
def factory():
    m = torchvision.models.resnet18()
    m.load_state_dict(torch.load(init))
    return m
res = shim.run_reference(mode, batch, steps, 0, extra_argv=["--lr", lr], model_factory=factory, print_freq=1)
json.dump(res, open(out, "w"))
'''


def _reference_curve(tmp_path, mode, steps, batch, lr, init):
    out = tmp_path / "ref_{}.json".format(mode)
    script = tmp_path / "ref_worker.py"
    script.write_text(_WORKER.format(root=ROOT))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29741", RANK="0", WORLD_SIZE="1", LOCAL_RANK="0",
               CUDA_VISIBLE_DEVICES="0")
    r = subprocess.run([sys.executable, str(script), mode, str(steps), str(batch), str(lr), str(init), str(out)], env=env,
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    import json
    log = open(os.path.join(json.load(open(out))["outpath"], "experiment.log")).read()
    curve = [float(v) for v in re.findall(r"Train epoch: \[\d+/\d+\]\[\d+/\d+\]\s+lr=\S+\s+ce_loss=(\S+)", log)]
    assert len(curve) == steps, (len(curve), log[-1500:])
    return curve


def _our_curve(mode, steps, batch, lr, init):
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import shim
    from pytorch_distributed_template_b200 import amp, engine, env, optim
    from pytorch_distributed_template_b200.models import create_model
    from pytorch_distributed_template_b200.ops import CrossEntropyLoss
    from pytorch_distributed_template_b200.parallel import DistributedDataParallel
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT="29742", RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    dev = env.init_distributed(0)
    try:
        engine.work_stream(dev)
        use_amp = mode == "ddp_amp"
        lp = torch.float16 if use_amp else None
        net = create_model("resnet18")
        net.load_state_dict(torch.load(init))
        model = DistributedDataParallel(net.to(dev), device_ids=[0], shadow_dtype=lp)
        crit = CrossEntropyLoss()
        opt = optim.SGD(model.parameters(), lr, momentum=0.9, weight_decay=1e-4)
        scaler = amp.GradScaler(enabled=use_amp, device=dev, optimizer=opt if use_amp else None)
        runner = engine.StepRunner(model, crit, opt, scaler, dev, use_amp=use_amp, amp_dtype=lp, eager_steps=2)
        data = shim.TimedSyntheticLoader(batch, 0, steps)          # the very tensors the reference arm is fed
        model.train()
        curve = []
        for i in range(steps):
            runner.step(data.images[i % 3].to(dev), data.targets[i % 3].to(dev))
            curve.append(runner.meter.read()["loss"])              # running average, what the reference logs
        return curve
    finally:
        env.shutdown_distributed()


@pytest.mark.parametrize("mode,rtol", [("ddp", 1e-2), ("ddp_amp", 3e-2)])
def test_loss_curve_matches_the_reference_for_20_steps(native_lib, tmp_path, mode, rtol):
    import torchvision
    if not os.path.exists(os.path.join(ROOT, "baseline", "_ref", "distributed_syncBN_amp.py")):
        pytest.skip("baseline/_ref not installed")
    steps, batch, lr = 20, 32, 0.02
    torch.manual_seed(1)
    init = tmp_path / "init.pt"
    torch.save(torchvision.models.resnet18().state_dict(), init)
    ref = _reference_curve(tmp_path, mode, steps, batch, lr, init)
    ours = _our_curve(mode, steps, batch, lr, init)
    worst = max(abs(a - b) / abs(b) for a, b in zip(ours, ref))
    assert worst <= rtol, (worst, list(zip(ours, ref)))
    assert ours[0] == pytest.approx(ref[0], rel=2e-3)              # first step: same weights, same batch
