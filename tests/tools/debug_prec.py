"""Is the CUDA-vs-fp32 error intrinsic to bf16 storage?  Compare (a) CUDA path, (b) the oracle with
bf16 rounding emulated at the same storage points, against the fp32 oracle."""
import sys, os, types
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F
from oracle import synth, nets as onets
from masterthesis_b200.models.core import networks
from masterthesis_b200.models.core.blocks import from_padded
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
DEV = 'cuda'
Q = lambda t: t.to(torch.bfloat16).float()
def l2(a, b): return ((a - b).norm() / b.norm()).item()
def mx(a, b): return ((a - b).abs().max() / b.abs().max()).item()

class FQ:  # F with bf16 storage emulation
    def __getattr__(self, n): return getattr(F, n)
    def conv2d(self, x, w, b=None, **k): return Q(F.conv2d(Q(x), Q(w), b, **k))
    def conv_transpose2d(self, x, w, b=None, **k): return Q(F.conv_transpose2d(Q(x), Q(w), b, **k))

def emu(fn, *a, **k):
    oldF, oldact = onets.F, onets._act
    onets.F = FQ(); onets._act = lambda x, act: Q(oldact(x, act))
    try: return fn(*a, **k)
    finally: onets.F, onets._act = oldF, oldact

dim, size, N, nd, seed = 64, 256, 1, 4, 5
Ec = networks.ContentEncoder(3, dim)
shapes = {k: list(v.shape) for k, v in Ec.state_dict().items()}
P = synth.synth_params(shapes, seed, 'content_encoder.')
Ec.load_state_dict(P); Ec.to(DEV).eval()
P = {k: v.to(DEV) for k, v in P.items()}
img = synth.synth_images(N, size, seed).to(DEV)
with torch.no_grad():
    ref = onets.content_encoder(P, img)
    e = emu(onets.content_encoder, P, img)
    mine = from_padded(Ec.run(img), dim * 4, 1)
print('Ec  emu-vs-fp32  l2 %.4f max %.4f | cuda-vs-fp32 l2 %.4f max %.4f | cuda-vs-emu l2 %.4f max %.4f | ref std %.3f absmax %.3f' % (
    l2(e, ref), mx(e, ref), l2(mine, ref), mx(mine, ref), l2(mine, e), mx(mine, e), ref.std().item(), ref.abs().max().item()))
# gradient of sum(z_c * proj) wrt first conv weight and image
proj = synth.synth_tensor('p', ref.shape, seed).to(DEV)
def grads(fn):
    Pg = {k: v.clone().requires_grad_(True) for k, v in P.items()}
    x = img.clone().requires_grad_(True)
    (fn(Pg, x) * proj).sum().backward()
    return x.grad, Pg['model.0.block.1.weight'].grad, Pg['model.6.model.1.block.1.weight'].grad
gr = grads(onets.content_encoder)
ge = grads(lambda p, x: emu(onets.content_encoder, p, x))
Ec.train(False)
x = img.clone().requires_grad_(True)
z = Ec.run(x)
(from_padded(z, dim * 4, 1) * proj).sum().backward()
gm = (x.grad, Ec.model[0].conv.weight.grad, Ec.model[6].model[1].conv.weight.grad)
cos = lambda a, b: (torch.dot(a.flatten().double(), b.flatten().double()) / (a.double().norm() * b.double().norm())).item()
for name, a, b, c in zip(('dimg', 'dW_stem', 'dW_last'), gr, ge, gm):
    print(name, 'cos emu-fp32 %.4f  cuda-fp32 %.4f  cuda-emu %.4f' % (cos(b, a), cos(c, a), cos(c, b)))
