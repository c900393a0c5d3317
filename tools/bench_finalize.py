"""norm_finalize timing vs number of partial-sum slots (100 calls captured in a CUDA graph)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from masterthesis_b200 import ops
for N, P, C in ((16, 1, 256), (16, 2, 256), (16, 32, 256), (16, 128, 128), (16, 512, 64), (8, 512, 64), (8, 128, 128)):
    st = torch.randn(N, P, C, 2, device='cuda') if P > 1 else torch.randn(N, C, 2, device='cuda')
    for _ in range(3): ops.norm_finalize(st, 'instance', 4096)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(100): keep = ops.norm_finalize(st, 'instance', 4096)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    print(f"N={N} parts={P} C={C}: {e0.elapsed_time(e1) * 10:.2f} us per launch")
