"""Is the training iteration launch-bound?  Measures, per iteration at batch MT_B (default 8):
host time to ENQUEUE the iteration (no sync inside) vs device time between CUDA events, and dumps
the per-class conv table (MT_BENCH_DUMP)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import bench
from masterthesis_b200 import models, ops

B = int(os.environ.get("MT_B", "8"))
torch.manual_seed(1234)
m = models.AdaINModel(bench.make_args(B)); m.initialize()
b = {k: v.cuda() for k, v in bench.synth_batch(B, 100, False).items()}
m.set_inputs(b)
for _ in range(3):
    m.optimize_parameters(0)
torch.cuda.synchronize()

# host enqueue time: run the two update_* calls without the final loss read-back
enq, dev = [], []
for i in range(5):
    m._scalars = {}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter(); e0.record()
    m.update_discriminator(m.img, m.c_org)
    t1 = time.perf_counter()
    m.update_generator(m.img, m.c_org)
    e1.record(); t2 = time.perf_counter()
    torch.cuda.synchronize()
    enq.append(((t1 - t0) * 1e3, (t2 - t1) * 1e3)); dev.append(e0.elapsed_time(e1))
print("B", B, "host enqueue ms (D phase, G phases):", [f"{a:.1f}+{b_:.1f}" for a, b_ in enq])
print("device ms:", [f"{d:.1f}" for d in dev])
l0 = ops.launches(); m.optimize_parameters(0); print("mt launches / iteration:", ops.launches() - l0)
os.environ.setdefault("MT_BENCH_DUMP", os.path.join(ROOT, "gpurun_out", f"conv_classes_b{B}.txt"))
roof = bench.conv_roofline(m, b, bench.peaks())
print("conv ms/iter", roof["all_convs"]["conv_ms_per_iter"], "TFLOP/s", roof["all_convs"]["achieved"])
