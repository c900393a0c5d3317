"""Micro-benchmark of single conv launches (CUDA events, L2 flushed between launches)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from masterthesis_b200 import ops
from masterthesis_b200.ops import ConvGeom
DEV = 'cuda'
def bench(N, Cin, Cout, H, W, k, stride, rpad, reps=20, kind='fwd'):
    g = ConvGeom(Cin, Cout, k, k, stride, 0)
    x = torch.randn(N, H + 2 * rpad, W + 2 * rpad, g.Cin_s, device=DEV).to(torch.bfloat16)
    w = torch.randn(Cout, Cin, k, k, device=DEV) * 0.05
    wf, wb = ops.pack_weight(w, g)
    Ho, Wo = g.out_hw(H + 2 * rpad, W + 2 * rpad)
    dy = torch.randn(N, Ho, Wo, g.Cout_s, device=DEV).to(torch.bfloat16)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=DEV)
    def run():
        if kind == 'fwd': ops.conv_fwd(x, wf, None, g)
        elif kind == 'dgrad': ops.conv_dgrad(dy, wb, g, x.shape[1], x.shape[2])
        else: ops.conv_wgrad(x, dy, g)
    for _ in range(3): run()
    torch.cuda.synchronize()
    # back-to-back launches: the queue stays full, so host launch latency is hidden (L2-warm numbers)
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    fl = 2.0 * N * Ho * Wo * Cin * Cout * k * k
    print(f"{kind:5s} N{N} {Cin}->{Cout} {H}x{W} k{k}s{stride}: {ms*1e3:8.1f} us  {fl/ms/1e9:8.1f} TFLOP/s  dbg={os.environ.get('MT_IGEMM_DBG','0')}", flush=True)
if __name__ == '__main__':
    for kind in sys.argv[1:] or ['fwd']:
        bench(16, 256, 256, 64, 64, 3, 1, 1, kind=kind)
        bench(8, 256, 256, 64, 64, 3, 1, 1, kind=kind)
        bench(32, 256, 256, 64, 64, 3, 1, 1, kind=kind)
