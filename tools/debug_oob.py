import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from masterthesis_b200 import ops
from masterthesis_b200.ops import ConvGeom
DEV='cuda'
def run(N,C,H,W,Cout,k=1, poison_a=True, poison_w=True, pre=0):
    g = ConvGeom(C, Cout, k, k, 1, 0)
    n = N*H*W*C
    big = torch.full((pre + n + 4096,), float('nan'), dtype=torch.bfloat16, device=DEV)
    if not poison_a: big.zero_()
    x = big[pre:pre+n].view(N,H,W,C); x.copy_(torch.randn(N,H,W,C, device=DEV))
    w = torch.randn(Cout, C, k, k, device=DEV)*0.05
    wf, wb = ops.pack_weight(w, g)
    if poison_w:
        bigw = torch.full((wf.numel()+4096,), float('nan'), dtype=torch.bfloat16, device=DEV)
        bigw[:wf.numel()].copy_(wf); wf = bigw[:wf.numel()]
    y = ops.conv_fwd(x, wf, None, g)
    torch.cuda.synchronize()
    bad = (~torch.isfinite(y.float())).nonzero()
    print(f"N{N} C{C} {H}x{W} Cout{Cout} k{k} pa={poison_a} pw={poison_w} pre={pre}: bad={bad.shape[0]}",
          (bad[0].tolist(), bad[-1].tolist()) if bad.shape[0] else '')
run(2,32,8,8,32)
run(2,32,8,8,32, poison_w=False)
run(2,32,8,8,32, poison_a=False)
run(2,64,8,8,32)
run(2,32,8,8,32, pre=64)
run(1,32,16,8,32)
run(2,16,8,8,32)
run(2,32,16,16,64)
run(2,96,8,8,32)
run(2,32,10,10,32,k=3)
