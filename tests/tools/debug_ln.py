import sys, torch
sys.path.insert(0, '.')
import torch.nn.functional as F
from masterthesis_b200 import ops
from tests.kernel_cases import bf, to_nhwc, from_nhwc
DEV='cuda'
for (N,C,H,W) in ((32,64,256,256),(8,64,256,256),(32,64,128,128),(4,64,256,256)):
    torch.manual_seed(0)
    x = bf(torch.randn(N, C, H, W, device=DEV) * 2 + 0.5)
    x_ref = x.clone().requires_grad_(True)
    lw = (torch.rand(C, device=DEV) + 0.5); lb = torch.randn(C, device=DEV)
    y_ref = F.relu(F.layer_norm(x_ref, x_ref.shape[1:], lw.view(C,1,1).expand(C,H,W), lb.view(C,1,1).expand(C,H,W)))
    g = bf(torch.randn_like(y_ref)); y_ref.backward(g)
    xn = to_nhwc(x)
    ss, mr = ops.norm_finalize(ops.plane_stats(xn), 'layer', H*W, None, lw, lb)
    dx = ops.norm_bwd(to_nhwc(g), xn, ss, mr, 'layer', act='relu', ln_w=lw)[0]
    d = (from_nhwc(dx, C) - x_ref.grad).abs()
    per_n = d.flatten(1).max(1).values
    print((N,C,H,W), 'max err', d.max().item(), 'ref max', x_ref.grad.abs().max().item(), 'per-n', [round(v,4) for v in per_n.tolist()][:8])
    # fp64 reference of the same formula
    xd = x.double(); mean = xd.mean((1,2,3), keepdim=True); var = xd.var((1,2,3), unbiased=False, keepdim=True)
    rstd = (var + 1e-5).rsqrt(); xh = (xd - mean) * rstd
    A = lw.double().view(1,C,1,1); y = A * xh + lb.double().view(1,C,1,1)
    gp = g.double() * (y > 0)
    mg = (A*gp).mean((1,2,3), keepdim=True); mgx = (A*gp*xh).mean((1,2,3), keepdim=True)
    dx64 = rstd * (A*gp - mg - xh*mgx)
    print('   torch fp32 vs fp64 formula', (x_ref.grad.double()-dx64).abs().max().item(), ' ours vs fp64', (from_nhwc(dx,C).double()-dx64).abs().max().item())
