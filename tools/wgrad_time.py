"""Kernel-only time of one weight-gradient class: 20 back-to-back launches of mt_conv2d_wgrad (no unpack)
between two events.  MT_SHAPE = Cin,Cout,k,stride,Hin(padded)   MT_N images.  MT_IGEMM_DBG bits: 1 no TMA
traffic, 2 no MMAs, 4 no epilogue stores."""
import os, sys, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from masterthesis_b200 import ops, _lib
from masterthesis_b200.ops import ConvGeom
N = int(os.environ.get("MT_N", "16"))
Cin, Cout, k, stride, H = [int(v) for v in os.environ.get("MT_SHAPE", "256,256,3,1,66").split(",")]
g = ConvGeom(Cin, Cout, k, k, stride, 0)
L = _lib.lib()
x = torch.randn(N, H, H, g.Cin_s, device='cuda').to(torch.bfloat16)
Ho, Wo = g.out_hw(H, H)
dy = torch.randn(N, Ho, Wo, g.Cout_s, device='cuda').to(torch.bfloat16)
d = g.desc(N, H, H)
n = L.mt_packed_wgrad_elems(g.Cout, g.Cin, g.kh, g.kw, 0, 0)
splits = L.mt_conv2d_wgrad_splits(ctypes.byref(d))
packed = torch.empty(splits * n, dtype=torch.float32, device='cuda')
st = torch.cuda.current_stream().cuda_stream
def run():
    rc = L.mt_conv2d_wgrad(ctypes.byref(d), x.data_ptr(), dy.data_ptr(), packed.data_ptr(), splits, st)
    assert rc == 0, rc
for _ in range(3): run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): run()
e1.record(); torch.cuda.synchronize()
fl = 2.0 * N * Ho * Wo * Cout * Cin * k * k
us = e0.elapsed_time(e1) * 1000 / 20
print(f"wgrad {os.environ.get('MT_SHAPE', '256,256,3,1,66')} N={N} dbg={os.environ.get('MT_IGEMM_DBG', '0')} splits={splits}: "
      f"{us:.1f} us per launch, {fl / us / 1e6:.0f} TFLOP/s")
