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

# ---- exposure: how much of one replayed iteration each kernel spends running ALONE (no other
# kernel of any stream overlapping it), and how long the GPU runs nothing at all.  Kernels with a
# large exclusive time and a small grid are the serial points of the graph.
if os.environ.get('MT_EXCL', '1') == '1':
    evs = sorted(((e.time_range.start, e.time_range.end, e.name.split('(')[0]) for e in prof.events()
                  if e.device_type.name == 'CUDA' and e.device_time > 0), key=lambda t: t[0])
    if evs:
        pts = []
        for i, (s, t, n) in enumerate(evs):
            pts.append((s, 1, i)); pts.append((t, -1, i))
        pts.sort()
        active, last = set(), pts[0][0]
        excl = collections.defaultdict(float)
        idle = busy = multi = 0.0
        for x, d, i in pts:
            dt = x - last
            if dt > 0:
                if not active:
                    idle += dt
                elif len(active) == 1:
                    excl[evs[next(iter(active))][2]] += dt; busy += dt
                else:
                    multi += dt; busy += dt
            last = x
            if d > 0: active.add(i)
            else: active.discard(i)
        span = evs[-1][1] - evs[0][0]
        print(f"\nspan {span / 3e3:.3f} ms per iteration: >=2 kernels {multi / 3e3:.3f} ms, exactly one kernel "
              f"{(busy - multi) / 3e3:.3f} ms, no kernel {idle / 3e3:.3f} ms")
        print("exclusive (alone on the GPU) time by kernel:")
        for n, t in sorted(excl.items(), key=lambda kv: -kv[1])[:30]:
            print(f"{n[:78]:78s} {t / 3e3:8.3f} ms")
