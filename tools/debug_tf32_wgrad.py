import os, sys, torch, torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["PDT_WGRAD_TF32"] = "1"
from pytorch_distributed_template_b200.ops import conv_native
torch.backends.cudnn.allow_tf32 = False
dev = torch.device("cuda", 0)
for (cin, cout, hw, k, stride, pad) in [(64, 64, 8, 1, 1, 0), (32, 64, 8, 1, 1, 0), (64, 64, 56, 3, 1, 1), (128, 128, 28, 3, 1, 1)]:
    torch.manual_seed(0)
    x = torch.randn(4, cin, hw, hw, device=dev).contiguous(memory_format=torch.channels_last)
    P = (hw + 2 * pad - k) // stride + 1
    dy = torch.randn(4, cout, P, P, device=dev).contiguous(memory_format=torch.channels_last)
    dw = conv_native.wgrad(x, dy, (cout, cin, k, k), stride, pad)
    wf = torch.zeros(cout, cin, k, k, device=dev, requires_grad=True)
    F.conv2d(x, wf, None, stride, pad).backward(dy)
    ref = wf.grad
    nz = (dw != 0).float().mean().item()
    a = dw.permute(0, 2, 3, 1).reshape(cout, -1); b = ref.permute(0, 2, 3, 1).reshape(cout, -1)
    cos = F.cosine_similarity(a.reshape(-1), b.reshape(-1), dim=0).item()
    # which (k-row, m-col) blocks of 32 are right?
    blk = []
    for mb in range(b.shape[1] // 32):
        e = (a[:, mb*32:(mb+1)*32] - b[:, mb*32:(mb+1)*32]).abs().max().item() / (b.abs().max().item())
        blk.append(round(e, 3))
    print((cin, cout, hw, k), "nonzero frac", nz, "cos", cos, "norm ratio", (a.norm() / b.norm()).item(), "per-32col-block err", blk[:20], flush=True)
