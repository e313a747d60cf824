import copy, os, sys, threading
import torch, torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pytorch_distributed_template_b200 import ops
from pytorch_distributed_template_b200.models import create_model

def rel(a, b):
    return ((a.float().cpu() - b.float().cpu()).abs().max() / (b.float().cpu().abs().max() + 1e-12)).item()

def bn_check(dev, tag):
    torch.manual_seed(1)
    x = torch.randn(8, 64, 32, 32, device=dev).contiguous(memory_format=torch.channels_last)
    bn = torch.nn.BatchNorm2d(64).to(dev); ref = torch.nn.BatchNorm2d(64).to(dev)
    y = ops.bn_act(x.clone().requires_grad_(True), bn, None, True)
    yr = F.relu(ref(x))
    torch.cuda.synchronize(dev)
    print(tag, "bn y rel", rel(y, yr), "rmean rel", rel(bn.running_mean, ref.running_mean), "cur dev", torch.cuda.current_device(), flush=True)

def main():
    torch.backends.cudnn.allow_tf32 = False
    d1 = torch.device("cuda", 1)
    bn_check(torch.device("cuda", 0), "main dev0")
    with torch.cuda.device(1):
        bn_check(d1, "main dev1 (ctx mgr)")
    bn_check(d1, "main dev1 (current device 0!)")
    def work():
        torch.cuda.set_device(1)
        bn_check(d1, "thread dev1")
    t = threading.Thread(target=work); t.start(); t.join()
    # whole model on device 1 from a thread
    torch.manual_seed(3)
    net = create_model("resnet18", num_classes=10)
    a = copy.deepcopy(net).to(d1); b = copy.deepcopy(net).to(d1)
    x = torch.randn(8, 3, 64, 64, device=d1)
    os.environ["PDT_MODEL_DEBUG"] = "1"
    with torch.cuda.device(1):
        yb = b(x)
    res = {}
    def work2():
        torch.cuda.set_device(1)
        res["y"] = a(x)
    t = threading.Thread(target=work2); t.start(); t.join()
    torch.cuda.synchronize(1)
    print("model thread vs main on dev1", rel(res["y"], yb), flush=True)

main()
