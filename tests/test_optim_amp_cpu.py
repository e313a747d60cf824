"""T0: fused-SGD arithmetic (torch backend), GradScaler state machine, arena layout."""
import copy

import pytest
import torch

from pytorch_distributed_template_b200 import amp, optim
from pytorch_distributed_template_b200.models import create_model
from pytorch_distributed_template_b200.parallel import Fabric, ParamArena


def test_sgd_matches_torch_sgd():
    torch.manual_seed(0)
    a = create_model("resnet18", num_classes=10)
    b = copy.deepcopy(a)
    oa = optim.SGD(a.parameters(), 0.1, momentum=0.9, weight_decay=1e-4)
    ob = torch.optim.SGD(b.parameters(), 0.1, momentum=0.9, weight_decay=1e-4)
    # identical gradient streams into both optimizers: isolates the update arithmetic from
    # the (chaotic at lr=0.1) training dynamics
    g = torch.Generator().manual_seed(1)
    for it in range(3):
        for p, q in zip(a.parameters(), b.parameters()):
            grad = torch.randn(q.shape, generator=g) * 0.1
            p.grad, q.grad = grad.clone(), grad.clone()
        oa.step()
        ob.step()
        for (k, p), q in zip(a.named_parameters(), b.parameters()):
            assert torch.allclose(p, q, rtol=1e-5, atol=1e-6), (it, k)
    x, y = torch.randn(4, 3, 32, 32), torch.tensor([1, 2, 3, 4])     # and one real backward through the arena views
    for m, o in ((a, oa), (b, ob)):
        o.zero_grad()
        torch.nn.functional.cross_entropy(m(x), y).backward()
        o.step()
    for (k, p), q in zip(a.named_parameters(), b.parameters()):
        assert torch.allclose(p, q, rtol=1e-3, atol=1e-4), k
    sd = oa.state_dict()
    assert len(sd["state"]) == 62 and "momentum_buffer" in sd["state"][0]


def test_arena_layout_and_segments():
    net = create_model("resnet18")
    arena = ParamArena(net.parameters(), Fabric(torch.device("cpu"), group="solo"), shadow_dtype=torch.bfloat16)
    assert len(arena.params) == 62
    # reverse registration order: fc first (what backward produces first)
    first = arena.segments[0].params[0]
    assert first is net.fc.bias
    assert len(arena.segments) == 3                       # 1 MiB first bucket, then 25 MiB (reducer.hpp:30-31)
    covered = 0
    for s in arena.segments:
        assert s.start == covered and s.length % 4 == 0
        covered = s.end
    assert covered == arena.total and arena.total >= 11689512
    w = net.layer1[0].conv1.weight
    assert w.shape == (64, 64, 3, 3) and w.is_contiguous(memory_format=torch.channels_last)
    assert w.data_ptr() == arena.flat_params[arena.offset[id(w)]:].data_ptr()
    assert torch.equal(arena.shadow_view(w).float(), w.detach().bfloat16().float())


def test_grad_scaler_state_machine():
    p = torch.nn.Parameter(torch.ones(8))
    opt = optim.SGD([p], 0.5, momentum=0.0, weight_decay=0.0)
    sc = amp.GradScaler(init_scale=4.0, growth_interval=2, enabled=True, device=torch.device("cpu"), optimizer=opt)
    for i in range(2):                                    # two clean steps -> scale doubles
        opt.zero_grad()
        sc.scale((p * 2).sum()).backward()
        sc.step(opt)
        sc.update()
    assert sc.get_scale() == 8.0
    assert torch.allclose(p.detach(), torch.full((8,), 1 - 2 * 0.5 * 2.0))
    before = p.detach().clone()
    opt.zero_grad()
    sc.scale((p * float("inf")).sum()).backward()        # non-finite gradient: skip, halve
    sc.step(opt)
    sc.update()
    assert torch.equal(p.detach(), before) and sc.get_scale() == 4.0 and sc.skipped_steps() == 1
    off = amp.GradScaler(enabled=False)
    t = torch.tensor(3.0)
    assert off.scale(t) is t and off.get_scale() == 1.0
    sd = sc.state_dict()
    assert set(sd) == {"scale", "growth_factor", "backoff_factor", "growth_interval", "_growth_tracker"}


def test_batched_counters_are_per_thread():
    """num_batches_tracked increments deferred by ops.bn.batched_counters: applied once on exit, nested contexts
    hand their list to the outer one, and worker threads (DataParallel replicas) never see each other's list."""
    import threading
    import torch
    from pytorch_distributed_template_b200.ops import bn as bnops

    a, b, c = (torch.zeros((), dtype=torch.long) for _ in range(3))
    with bnops.batched_counters():
        bnops._bump(a)
        with bnops.batched_counters():
            bnops._bump(b)
        assert a.item() == 0 and b.item() == 0          # deferred to the outermost exit
    assert a.item() == 1 and b.item() == 1
    bnops._bump(c)                                       # outside any context: immediate
    assert c.item() == 1

    counters = [torch.zeros((), dtype=torch.long) for _ in range(4)]
    barrier = threading.Barrier(4)

    def worker(t):
        with bnops.batched_counters():
            barrier.wait()                               # all four contexts are open at the same time
            for _ in range(5):
                bnops._bump(t)
            barrier.wait()
        # every thread applied exactly its own increments

    threads = [threading.Thread(target=worker, args=(t,)) for t in counters]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert [int(t) for t in counters] == [5, 5, 5, 5]
    assert getattr(bnops._tls, "pending", None) is None
