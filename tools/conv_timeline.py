"""clock64 timeline of CTA 0 of the hot conv (needs MT_IGEMM_DBG with bit 32 set)."""
import os, sys, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from masterthesis_b200 import ops, _lib
from masterthesis_b200.ops import ConvGeom
N = int(os.environ.get("MT_N", "16"))
g = ConvGeom(256, 256, 3, 3, 1, 0)
x = torch.randn(N, 66, 66, 256, device='cuda').to(torch.bfloat16)
w = torch.randn(256, 256, 3, 3, device='cuda') * 0.05
wf, wb = ops.pack_weight(w, g)
fwd = (lambda: ops.conv_fwd_stats(x, wf, None, g)) if os.environ.get("MT_STATS") else (lambda: ops.conv_fwd(x, wf, None, g))
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
print(f"dbg={os.environ.get('MT_IGEMM_DBG')} N={N} kernel {e0.elapsed_time(e1)*1e3:.1f} us")
print("setup", t[7], "| producer done", t[6], "| kernel end", t[0])
for i in range(8):
    print(f"tile {i}: mma {t[8+2*i]:7d} .. {t[9+2*i]:7d} ({t[9+2*i]-t[8+2*i]:6d})   epilogue {t[24+2*i]:7d} .. {t[25+2*i]:7d} ({t[25+2*i]-t[24+2*i]:6d})")
print("ring offset flag (negative = mismatch):", t[63])
