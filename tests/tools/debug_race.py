"""Find the first op whose output differs between two identical eager Ec forwards."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT)
import torch
from oracle import ref_import, synth
from masterthesis_b200 import models, ops
dim, size, N, nd, seed = 16, 64, 2, 4, 9
args = ref_import.default_args(dim=dim, crop_size=size, batch_size=N, num_domains=nd, init_type='normal', mode='test')
torch.manual_seed(seed)
m = models.AdaINModel(args); m.initialize(); m.use_graph = False
for n in m.model: m.model[n].eval()
names = ["conv_fwd", "conv_fwd_stats", "norm_finalize", "reflect_border_", "plane_stats", "image_to_nhwc"]
log = []
def wrap(name):
    f = getattr(ops, name)
    def g(*a, **k):
        out = f(*a, **k)
        outs = out if isinstance(out, tuple) else (out,)
        if name == "reflect_border_":
            outs = (a[0],)
        torch.cuda.synchronize()
        log.append((name, [o.detach().float().clone() if torch.is_tensor(o) else None for o in outs],
                    [tuple(x.shape) for x in a if torch.is_tensor(x)]))
        return out
    setattr(ops, name, g)
for n_ in names: wrap(n_)
sync_mode = os.environ.get("MT_SYNC", "1")
def run(a):
    log.clear()
    with torch.no_grad():
        m.Ec(a)
    return list(log)
found = 0
for call in range(60):
    a = synth.synth_images(N, size, seed + call, 'src').cuda()
    l1, l2 = run(a), run(a)
    for i, ((n1, o1, s1), (n2, o2, s2)) in enumerate(zip(l1, l2)):
        d = max(((x - y).abs().max().item() if x is not None else 0.0) for x, y in zip(o1, o2))
        if d > 1e-4:
            print(f"call {call}: first differing op #{i} {n1} input shapes {s1} max abs diff {d:.4g}")
            for j, (x, y) in enumerate(zip(o1, o2)):
                if x is not None:
                    nz = ((x - y).abs() > 1e-4).nonzero()
                    print("   output", j, tuple(x.shape), "differing elements", nz.shape[0], "first", nz[0].tolist() if nz.shape[0] else None,
                          "last", nz[-1].tolist() if nz.shape[0] else None)
            found += 1
            break
    if found >= 4: break
print("done, found", found)
