"""clock64 timeline of CTA 0 of the hot conv (needs MT_IGEMM_DBG with bit 32 set)."""
import os, sys, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from masterthesis_b200 import ops, _lib
from masterthesis_b200.ops import ConvGeom
N = int(os.environ.get("MT_N", "16"))
# MT_SHAPE = Cin,Cout,k,stride,Hin,kind   kind: conv | stem | convT | dgrad
Cin, Cout, k, stride, H, kind = (os.environ.get("MT_SHAPE", "256,256,3,1,66,conv").split(",") + ["conv"])[:6]
Cin, Cout, k, stride, H = int(Cin), int(Cout), int(k), int(stride), int(H)
if kind == "convT":
    g = ConvGeom(Cin, Cout, 3, 3, 2, 1, transposed=True)
    w = torch.randn(Cin, Cout, 3, 3, device='cuda') * 0.05
else:
    g = ConvGeom(Cin, Cout, k, k, stride, 0, False, kind == "stem")
    w = torch.randn(Cout, Cin, k, k, device='cuda') * 0.05
if kind == "stem":
    x = ops.image_to_nhwc(torch.randn(N, Cin, H - 2 * (k // 2), H - 2 * (k // 2), device='cuda'), None, k // 2)
else:
    x = torch.randn(N, H, H, g.Cin_s, device='cuda').to(torch.bfloat16)
wf, wb = ops.pack_weight(w, g)
if kind == "dgrad":
    Ho, Wo = g.out_hw(H, H)
    dy = torch.randn(N, Ho, Wo, g.Cout_s, device='cuda').to(torch.bfloat16)
fwd = (lambda: ops.conv_fwd_stats(x, wf, None, g)) if os.environ.get("MT_STATS") else (lambda: ops.conv_fwd(x, wf, None, g))
if kind == "dgrad":
    fwd = lambda: ops.conv_dgrad(dy, wb, g, H, H)
for _ in range(3): fwd()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): fwd()
e1.record(); torch.cuda.synchronize()
print(f"avg of 10 back-to-back: {e0.elapsed_time(e1)*100:.1f} us")
e0.record(); fwd(); e1.record(); torch.cuda.synchronize()
buf = (ctypes.c_longlong * 64)()
_lib.lib().mt_debug_timestamps.argtypes = [ctypes.c_void_p]
_lib.check(_lib.lib().mt_debug_timestamps(ctypes.cast(buf, ctypes.c_void_p)), "ts")
t = list(buf)
print(f"shape={os.environ.get('MT_SHAPE')} dbg={os.environ.get('MT_IGEMM_DBG')} N={N} kernel {e0.elapsed_time(e1)*1e3:.1f} us")
print("setup", t[7], "| producer done", t[6], "| kernel end", t[0])
for i in range(8):
    print(f"tile {i}: mma {t[8+2*i]:7d} .. {t[9+2*i]:7d} ({t[9+2*i]-t[8+2*i]:6d})   epilogue {t[24+2*i]:7d} .. {t[25+2*i]:7d} ({t[25+2*i]-t[24+2*i]:6d})")
print("epilogue of tile 2, warp 2 (last slab): enter", t[40], "| box free", t[41], "| slab in smem", t[42], "| barrier", t[43], "| store issued", t[44], "| stats", t[45])
print("ring offset flag (negative = mismatch):", t[63])
