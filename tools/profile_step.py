"""One training iteration (batch 8, 256x256) bracketed by cudaProfilerStart/Stop, for
  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv ...
  ncu --profile-from-start off --set full -k regex:igemm ... """
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import bench
from masterthesis_b200 import models
B = int(os.environ.get("MT_B", "8"))
torch.manual_seed(1234)
m = models.AdaINModel(bench.make_args(B)); m.initialize()
b = {k: v.cuda() for k, v in bench.synth_batch(B, 100, False).items()}
m.set_inputs(b)
for _ in range(int(os.environ.get("MT_WARM", "2"))):
    m.optimize_parameters(0)
torch.cuda.synchronize()
torch.cuda.profiler.start()
m.optimize_parameters(0)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("done", m.loss.total_g)
