"""2-GPU single-process debug of our DataParallel against a manual computation."""
import copy, os, sys
import torch, torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_template_b200 import optim
from pytorch_distributed_template_b200.models import create_model
from pytorch_distributed_template_b200.parallel import DataParallel

def rel(a, b):
    return ((a.float().cpu() - b.float().cpu()).abs().max() / (b.float().cpu().abs().max() + 1e-12)).item()

def main():
    n = 2
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(21)
    net = create_model("resnet18", num_classes=10)
    manual = [copy.deepcopy(net).to("cuda:%d" % d) for d in range(n)]   # plain replicas, manual sum of grads
    ours = DataParallel(net, device_ids=list(range(n))).cuda()
    opt = optim.SGD(ours.parameters(), 0.01, momentum=0.9, weight_decay=1e-4)
    g = torch.Generator().manual_seed(3)
    x = torch.randn(16, 3, 64, 64, generator=g).cuda(); y = torch.randint(0, 10, (16,), generator=g).cuda()
    ours.train()
    out = ours(x)
    # manual forward
    outs = []
    for d, m in enumerate(manual):
        m.train()
        outs.append(m(x[d * 8:(d + 1) * 8].to("cuda:%d" % d)))
    mout = torch.cat([o.to("cuda:0") for o in outs])
    print("forward rel", rel(out, mout), "first half", rel(out[:8], mout[:8]), "second half", rel(out[8:], mout[8:]))
    for d in range(n):
        wr = max(rel(p.detach(), q.detach()) for p, q in zip(ours.replicas[d].parameters(), manual[d].parameters()))
        br = max(rel(p.detach().float(), q.detach().float()) for p, q in zip(ours.replicas[d].buffers(), manual[d].buffers()))
        print("replica", d, "weights vs manual (before step)", wr, "buffers", br)
    with torch.no_grad():
        xs1 = x[8:16].to("cuda:1")
        ours.replicas[1].eval(); manual[1].eval()
        print("replica 1 eval fwd on main thread vs manual", rel(ours.replicas[1](xs1), manual[1](xs1)))
        ours.replicas[1].train(); manual[1].train()
        from pytorch_distributed_template_b200.parallel.dp import _Scatter
        sc = _Scatter.apply(ours, x)
        torch.cuda.synchronize(0); torch.cuda.synchronize(1)
        print("scatter slice 1 rel", rel(sc[1], x[8:16]))
    loss = F.cross_entropy(out, y); mloss = F.cross_entropy(mout, y)
    opt.zero_grad(); loss.backward(); mloss.backward()
    for d in range(n): torch.cuda.synchronize(d)
    # local grads of each replica
    for d in range(n):
        w = max(rel(p.grad, q.grad) for p, q in zip(ours.replicas[d].parameters(), manual[d].parameters()))
        print("replica", d, "local grad rel", w)
    p0 = [p.detach().clone() for p in manual[0].parameters()]
    gsum = [sum(m_p.grad.to("cuda:0") for m_p in ps) for ps in zip(*[m.parameters() for m in manual])]
    opt.step()
    for d in range(n): torch.cuda.synchronize(d)
    a0 = ours.arenas[0]
    wg = 0.0
    for p, gs in zip(ours.module.parameters(), gsum):
        fo = a0.grad_view(p)
        wg = max(wg, rel(fo, gs))
    print("arena0 grads (mixed: own shard = sum, other = local) rel vs sum", wg)
    wp = 0.0
    for p, q0, gs in zip(ours.module.parameters(), p0, gsum):
        newq = q0 - 0.01 * (gs + 1e-4 * q0)
        wp = max(wp, rel(p.detach(), newq))
    print("param after step rel vs manual", wp)
    for d in range(1, n):
        wr = max(rel(p.detach(), q.detach()) for p, q in zip(ours.replicas[d].parameters(), ours.module.parameters()))
        print("replica", d, "weights vs module", wr)
    # per-shard check of the summed gradient
    W = n; tot = a0.total; sh = tot // W
    full = torch.zeros(tot, device="cuda:0")
    for p, gs in zip(ours.module.parameters(), gsum):
        off = a0.offset[id(p)]
        full[off:off + p.numel()] = (gs.permute(0, 2, 3, 1) if gs.dim() == 4 else gs).reshape(-1)
    for r in range(W):
        mine = ours.arenas[r].flat_grads[r * sh:(r + 1) * sh].to("cuda:0")
        print("shard", r, "reduced rel", rel(mine, full[r * sh:(r + 1) * sh]))

if __name__ == "__main__":
    main()
