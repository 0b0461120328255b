import torch, torch.nn.functional as F
from commefficient_b200 import models
from commefficient_b200.ops import nn as nnops
from commefficient_b200.utils import FlatParams
DEV="cuda"
def rel(a,b): return float((a-b).norm()/b.norm())
torch.manual_seed(0)
m = models.ResNet9(num_classes=10)
fp = FlatParams(m, DEV, layout="nhwc"); m._flat_params = fp
for n in (24, 500):
    x = torch.randn(n, 3, 32, 32, device=DEV).contiguous(memory_format=torch.channels_last)
    tgt = torch.randint(0, 10, (n,), device=DEV)
    def run(native_on, prec):
        nnops.set_native(native_on); nnops.set_precision(prec)
        fp.zero_grad()
        if prec == "bf16" and not native_on:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss, _ = m.ce_per_sample(x, tgt)
        else:
            loss, _ = m.ce_per_sample(x, tgt)
        loss.sum().backward()
        return loss.detach().clone(), fp.grad.clone()
    l0,g0 = run(False,"fp32")
    for name,(nat,prec) in {"lib_tf32":(False,"tf32"),"native_tf32":(True,"tf32"),"native_bf16":(True,"bf16"),"lib_bf16":(False,"bf16")}.items():
        l,g = run(nat,prec)
        per = max(rel(g[o:o+s], g0[o:o+s]) for o,s in zip(fp.offsets, fp.sizes))
        print(n, name, "loss err %.2e grad err %.2e worst layer %.2e" % (rel(l,l0), rel(g,g0), per))
