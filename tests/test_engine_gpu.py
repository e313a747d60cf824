"""engine.StepRunner / DeviceFeed: the CUDA-graph step is the product's step.

* graph replay and eager execution produce IDENTICAL parameters (deterministic mode: no floating-point atomics);
* two seeded runs are bit-identical (the reference's --seed promise, /root/reference/distributed.py:91-99);
* the training scripts reach the graph (a user of start.sh gets the replayed step, not only bench.py).
"""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(graph, steps=7, amp_dtype=torch.float16, seed=0):
    from pytorch_distributed_template_b200 import amp, engine, optim
    from pytorch_distributed_template_b200.models import create_model
    from pytorch_distributed_template_b200.ops import CrossEntropyLoss
    from pytorch_distributed_template_b200.parallel import DistributedDataParallel
    dev = torch.device("cuda", 0)
    engine.work_stream(dev)
    torch.manual_seed(seed)
    model = DistributedDataParallel(create_model("resnet18", num_classes=100).to(dev), device_ids=[0], shadow_dtype=amp_dtype)
    crit = CrossEntropyLoss()
    opt = optim.SGD(model.parameters(), 0.05, momentum=0.9, weight_decay=1e-4)
    scaler = amp.GradScaler(enabled=amp_dtype is not None, device=dev, optimizer=opt, init_scale=1024.0)
    runner = engine.StepRunner(model, crit, opt, scaler, dev, use_amp=amp_dtype is not None, amp_dtype=amp_dtype, graph=graph,
                               eager_steps=2)
    g = torch.Generator().manual_seed(99)
    batches = [(torch.randint(0, 256, (32, 3, 64, 64), generator=g, dtype=torch.uint8).pin_memory(),
                torch.randint(0, 100, (32,), generator=g).pin_memory()) for _ in range(steps)]
    model.train()
    losses = []
    for img, tgt in engine.DeviceFeed(batches, runner):
        res = runner.step(img, tgt)
        losses.append(res.clone())
    torch.cuda.synchronize()
    m = runner.meter.read()
    return (model.arena.flat_params.detach().clone(), torch.stack(losses).cpu(), runner, m)


def test_graph_replay_equals_eager_and_seeded_runs_are_bit_identical(native_lib, monkeypatch):
    monkeypatch.setenv("PDT_DETERMINISTIC", "1")
    p_eager, l_eager, r_eager, m_eager = _run(graph=False)
    p_graph, l_graph, r_graph, m_graph = _run(graph=True)
    assert r_eager.graphs == 0 and r_graph.graphs == 2, (r_graph.graphs, r_graph.capture_error)   # one per staging buffer
    assert r_graph.launches_per_step and r_graph.launches_per_step > 50
    assert torch.equal(l_eager, l_graph)
    assert torch.equal(p_eager, p_graph)
    assert m_eager == m_graph and m_graph["count"] == 7 * 32
    p_again, l_again, _, _ = _run(graph=True)
    assert torch.equal(p_graph, p_again) and torch.equal(l_graph, l_again)
    assert torch.isfinite(l_graph).all()


def test_training_script_replays_the_step_from_a_graph(native_lib, tmp_path):
    env = dict(os.environ, PDT_SYNTHETIC="1", PDT_SYNTH_TRAIN_SIZE="512", PDT_SYNTH_VAL_SIZE="64", PDT_IMAGE_SIZE="64",
               PDT_NUM_CLASSES="10", PDT_OUTPUT_EXISTS="delete", MASTER_ADDR="127.0.0.1", MASTER_PORT="29743", RANK="0",
               WORLD_SIZE="1", LOCAL_RANK="0", PDT_REPORT_GRAPHS="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "distributed_syncBN_amp.py"), "--epochs", "1", "-b", "64", "-j", "2",
                        "-p", "2", "--outpath", str(tmp_path / "out")], cwd=ROOT, env=env, capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "||==> Train epoch: [0/1]" in r.stdout
    # (-j 2: the DataLoader's pin-memory thread allocates pinned memory while the step is being captured)
    assert "[pdt] training step: 2 CUDA graphs" in r.stdout + r.stderr, r.stdout[-1500:] + r.stderr[-1500:]
