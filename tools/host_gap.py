"""Host-side time per replayed iteration: how long the GPU idles between the loss read-back of one iteration and
the launch of the next (GraphHyper.advance / upload, python bookkeeping), and the cost of set_inputs from pinned
host memory."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import bench
from masterthesis_b200 import models
B = 8
torch.manual_seed(1234)
m = models.AdaINModel(bench.make_args(B)); m.initialize()
host = bench.synth_batch(B, 100, True)
res = {k: v.cuda() for k, v in host.items()}
m.set_inputs(res)
for i in range(5):
    m.optimize_parameters(i)
torch.cuda.synchronize()
g = m._graphs['main']
T = {k: 0.0 for k in ('advance', 'upload', 'replay_launch', 'wait+tolist', 'named', 'set_inputs_host', 'set_inputs_resident')}
K = 30
for i in range(K):
    t0 = time.perf_counter(); g['ctx'].advance(); t1 = time.perf_counter()
    g['ctx'].upload(); t2 = time.perf_counter()
    g['graph'].replay(); t3 = time.perf_counter()
    vals = g['vec'].tolist(); t4 = time.perf_counter()
    list(m._named_losses(g['keys'], vals)); t5 = time.perf_counter()
    T['advance'] += t1 - t0; T['upload'] += t2 - t1; T['replay_launch'] += t3 - t2; T['wait+tolist'] += t4 - t3; T['named'] += t5 - t4
    torch.cuda.synchronize(); t6 = time.perf_counter(); m.set_inputs(host); torch.cuda.synchronize(); t7 = time.perf_counter()
    m.set_inputs(res); torch.cuda.synchronize(); t8 = time.perf_counter()
    T['set_inputs_host'] += t7 - t6; T['set_inputs_resident'] += t8 - t7
for k, v in T.items():
    print(f"{k:22s} {v / K * 1e3:8.3f} ms")
