"""Other configurations of SURVEY 8d on one GPU: MT_B pairs per iteration, MT_SIZE crop, MT_MODEL
(AdaINModel | BaseModel), MT_FLAGS (comma list: concat,reparam,use_dis_content,ms_dis,dis_sn,use_dropout,up_type=nearest,...).
Prints ms / iteration and images/s (6 timed iterations after the warm-up / graph captures) and peak memory."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import bench
from masterthesis_b200 import models
B, size = int(os.environ.get("MT_B", "8")), int(os.environ.get("MT_SIZE", "256"))
flags = [f for f in os.environ.get("MT_FLAGS", "").split(",") if f]
bench.SIZE = size
args = bench.make_args(B)
args.crop_size = size
for f in flags:          # `flag` sets a boolean, `key=value` a string option (up_type=nearest, dis_norm=instance, ...)
    if '=' in f:
        k, v = f.split('=', 1)
        setattr(args, k, v)
    else:
        setattr(args, f, True)
torch.manual_seed(1234)
m = getattr(models, os.environ.get("MT_MODEL", "AdaINModel"))(args); m.initialize()
b = {k: v.cuda() for k, v in bench.synth_batch(B, 100, False).items()}
m.set_inputs(b)
# warm-up until every iteration kind is captured (2 eager iterations + the capture each; with
# --use_dis_content two kinds alternate on the d_iter schedule, the main one every third iteration)
WARM = 12 if getattr(args, 'use_dis_content', False) else 4
for i in range(WARM):
    m.optimize_parameters(i)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
K = 6
for i in range(K):
    m.optimize_parameters(WARM + i)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / K
loss = {k: round(float(v), 4) for k, v in m.loss.items()}
print(f"{os.environ.get('MT_MODEL', 'AdaINModel')} flags={flags} B={B} size={size}: {ms:.2f} ms/iter, "
      f"{2 * B / ms * 1e3:.1f} images/s, peak mem {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB, "
      f"finite={all(map(lambda v: v == v and abs(v) < 1e9, loss.values()))} total_g={loss.get('total_g')}")
