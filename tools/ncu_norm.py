"""The HBM-bound norm kernels on the shapes they see inside one training iteration, for an ncu
capture (`ncu --set full -k regex:affine_act_fwd|norm_bwd_apply|plane_reduce ...`) and, without
ncu, CUDA-event timings with an L2 flush between launches.

Shapes (batch 8 => N = 16 images per call):
  ln256 : LayerNorm + ReLU behind the 128->64 transposed conv, [16,256,256,64]  (134 MB per tensor)
  ln128 : LayerNorm + ReLU behind the 256->128 transposed conv, [16,128,128,128] (67 MB)
  ad64  : AdaIN (+ReLU / +residual) of the residual blocks, [16,64,64,256]       (34 MB)
Algorithmic bytes: forward 2+2 B per element (bf16 read + bf16 write; +2 with a residual),
backward reduce 2+2 (dy, x), backward apply 2+2+2 (dy, x, dx).
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from masterthesis_b200 import _lib  # noqa: E402
if os.environ.get("MT_LIB"):            # a kernel variant built to another file (tuning experiments)
    from pathlib import Path
    _lib.LIB_PATH = Path(os.environ["MT_LIB"]).resolve()
from masterthesis_b200 import ops  # noqa: E402

REPS = int(os.environ.get("MT_REPS", "5"))
flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device='cuda')


def timed(name, fn, mbytes):
    ts = []
    for _ in range(REPS):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    us = ts[len(ts) // 2]
    print(f"{name:34s} {us:8.1f} us  {mbytes / us * 1e3:7.0f} GB/s  ({mbytes:.0f} MB algorithmic)", flush=True)


def run(tag, N, H, W, C, mode):
    x = torch.randn(N, H, W, C, device='cuda').to(torch.bfloat16)
    g = torch.randn(N, H, W, C, device='cuda').to(torch.bfloat16)
    MB = x.numel() * 2 / 1e6
    stats = ops.plane_stats(x)
    gb = torch.randn(N, 2 * C, device='cuda') * 0.3 if mode == 'adain' else None
    lw = torch.rand(C, device='cuda') + 0.5 if mode == 'layer' else None
    lb = torch.randn(C, device='cuda') if mode == 'layer' else None
    ss, mr = ops.norm_finalize(stats, mode, H * W, gb, lw, lb)
    L = ops._lib.lib()
    bparts = L.mt_norm_bwd_parts(N, H, W, C, 1, 0)
    sums = torch.empty(N, bparts, C, 2, device='cuda')
    coef = torch.randn(N, C, 3, device='cuda') * 0.1
    dx = torch.empty_like(x)
    out = torch.empty_like(x)
    timed(f"{tag} affine+relu fwd", lambda: ops.affine_act_fwd(x, ss, act='relu', out=out), 2 * MB)
    timed(f"{tag} bwd reduce", lambda: ops.check(L.mt_norm_bwd_reduce(
        g.data_ptr(), 0, x.data_ptr(), ss.data_ptr(), N, H, W, C, 1, 0.01, sums.data_ptr(), bparts, ops._stream()), "r"), 2 * MB)
    timed(f"{tag} bwd apply", lambda: ops.check(L.mt_norm_bwd_apply(
        g.data_ptr(), 0, x.data_ptr(), ss.data_ptr(), coef.data_ptr(), N, H, W, C, 1, 0.01, dx.data_ptr(),
        ops._stream()), "a"), 3 * MB)
    if mode == 'adain':
        xp = torch.randn(N, H + 2, W + 2, C, device='cuda').to(torch.bfloat16)
        timed(f"{tag} affine+residual fwd (pad 1)", lambda: ops.affine_act_fwd(x, ss, res_pad=xp, res_p=1, out_p=1), 3 * MB)
    timed(f"{tag} torch copy_ (yardstick)", lambda: out.copy_(x), 2 * MB)


if __name__ == "__main__":
    run("ln256 [16,256,256,64]", 16, 256, 256, 64, 'layer')
    run("ln128 [16,128,128,128]", 16, 128, 128, 128, 'layer')
    run("ad64  [16,64,64,256]", 16, 64, 64, 256, 'adain')
