"""Run-to-run determinism of the eager sampling path on small shapes: MT_NET selects the part."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT)
import torch
from oracle import ref_import, synth
from masterthesis_b200 import models
dim, size, N, nd, seed = int(os.environ.get("MT_DIM", "16")), int(os.environ.get("MT_SIZE", "64")), 2, 4, 9
args = ref_import.default_args(dim=dim, crop_size=size, batch_size=N, num_domains=nd, init_type='normal', mode='test')
torch.manual_seed(seed)
m = models.AdaINModel(args); m.initialize(); m.use_graph = False
for n in m.model: m.model[n].eval()
def rel(x, y): return ((x.float() - y.float()).abs().max() / y.float().abs().max().clamp_min(1e-12)).item()
bad = {"Ec": 0, "G": 0}
K = int(os.environ.get("MT_CALLS", "40"))
with torch.no_grad():
    for call in range(K):
        a = synth.synth_images(N, size, seed + call, 'src').cuda(); z = synth.synth_tensor('zr', (N, 8), seed + call).cuda(); c = synth.synth_onehot(N, nd, seed + call).cuda()
        zc1 = m.Ec(a); zc2 = m.Ec(a)
        if rel(zc1, zc2) > 1e-3: bad["Ec"] += 1
        y1 = m.G(zc1, z, c); y2 = m.G(zc1, z, c)
        if rel(y1, y2) > 1e-3: bad["G"] += 1
print({k: os.environ.get(k) for k in ("MT_PHASE_FORK", "MT_CONV_NO_STATS")}, "mismatching pairs of", K, ":", bad)
