"""In-situ per-kernel times of the graph-replayed training iteration (torch.profiler / CUPTI):
unlike an ncu launch list the kernels run back to back, warm, at the clocks of the real step."""
import os, sys, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from torch.profiler import profile, ProfilerActivity
import bench
from masterthesis_b200 import models
B = int(os.environ.get("MT_B", "8"))
torch.manual_seed(1234)
m = models.AdaINModel(bench.make_args(B)); m.initialize()
b = {k: v.cuda() for k, v in bench.synth_batch(B, 100, False).items()}
m.set_inputs(b)
for _ in range(5):
    m.optimize_parameters(0)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        m.optimize_parameters(0)
    torch.cuda.synchronize()
agg = collections.defaultdict(lambda: [0, 0.0])
t0, t1 = None, None
for e in prof.events():
    if e.device_type.name != 'CUDA' or e.device_time <= 0:
        continue
    n = e.name
    for cut in ('(', ):
        n = n.split(cut)[0]
    agg[n][0] += 1; agg[n][1] += e.device_time
tot = sum(v[1] for v in agg.values())
print(f"sum of kernel time {tot / 3e3:.3f} ms per iteration")
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(os.environ.get('MT_TOP', '45'))]:
    print(f"{n[:78]:78s} {c // 3:5d} {t / 3e3:8.3f} ms {100 * t / tot:5.1f}%  avg {t / c:7.1f} us")
