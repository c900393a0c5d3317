"""Micro-benchmark of the per-plane HBM-bound kernels on the hot shape (N=16, 64x64, 256 ch)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from masterthesis_b200 import ops
N, H, W, C = int(os.environ.get("MT_N", "16")), 64, 64, 256
x = torch.randn(N, H, W, C, device='cuda').to(torch.bfloat16)
gp = torch.randn(N, H + 2, W + 2, C, device='cuda').to(torch.bfloat16)
xp = torch.randn(N, H + 2, W + 2, C, device='cuda').to(torch.bfloat16)
flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device='cuda')
MB = x.numel() * 2 / 1e6
def timeit(name, fn, mbytes, reps=20, cold=True):
    for _ in range(3): fn()
    ts = []
    for _ in range(reps):
        if cold: flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort(); us = ts[len(ts) // 2]
    print(f"{name:28s} {'cold' if cold else 'warm'} {us:7.1f} us  {mbytes / us:6.2f} TB/s ({mbytes:.0f} MB)", flush=True)
stats = ops.plane_stats(x)
ss, mr = ops.norm_finalize(stats, 'instance', H * W)
L = ops._lib.lib()
bparts = L.mt_norm_bwd_parts(N, H, W, C, 1, 1)
sums = torch.empty(N, bparts, C, 2, device='cuda'); coef = torch.randn(N, C, 3, device='cuda'); dx = torch.empty_like(x)
for cold in (True, False):
    timeit("plane_stats", lambda: ops.plane_stats(x), MB, cold=cold)
    timeit("affine_act_fwd(+res,pad)", lambda: ops.affine_act_fwd(x, ss, act='relu', res_pad=xp, res_p=1, out_p=1), 3 * MB, cold=cold)
    timeit("affine_act_fwd", lambda: ops.affine_act_fwd(x, ss, act='relu', out_p=1), 2 * MB, cold=cold)
    timeit("norm_bwd_reduce(fold)", lambda: ops.check(L.mt_norm_bwd_reduce(gp.data_ptr(), 1, x.data_ptr(), ss.data_ptr(), N, H, W, C, 1, 0.01, sums.data_ptr(), bparts, ops._stream()), "r"), 2 * MB, cold=cold)
    timeit("norm_bwd_apply(fold)", lambda: ops.check(L.mt_norm_bwd_apply(gp.data_ptr(), 1, x.data_ptr(), ss.data_ptr(), coef.data_ptr(), N, H, W, C, 1, 0.01, dx.data_ptr(), ops._stream()), "a"), 3 * MB, cold=cold)
    timeit("reflect_fold", lambda: ops.reflect_fold(gp, 1), 2 * MB, cold=cold)
    timeit("torch copy (ref)", lambda: dx.copy_(x), 2 * MB, cold=cold)
