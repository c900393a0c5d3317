"""TEST INFRASTRUCTURE — where does the gradient distance between a bf16-storage implementation and the fp32
reference come from?  Runs the fp32 oracle's first training iteration (dim 64, 256x256, batch 8: the
benchmarked geometry) on the GPU under several storage emulations and prints, per optimizer step, the cosine
of the gradients against the fp32 run:

  full      : weights, conv outputs and activations rounded to bf16 in forward AND their gradients rounded to
              bf16 in backward (tests/emu.py: what any bf16-storage / fp32-accumulate implementation does)
  fwd_only  : the same forward rounding, gradients kept in fp32 (straight-through rounding)
  bwd_only  : fp32 forward, gradients rounded to bf16 at the same points
  weights   : only the GEMM weights rounded to bf16

Usage: python tests/tools/precision_experiment.py [dim size B]      (prints one RESULT json line)
"""
import contextlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from oracle import nets as onets, step as ostep, synth  # noqa: E402

DEV = 'cuda'


class _RoundFwd(torch.autograd.Function):        # bf16 rounding in forward, identity in backward
    @staticmethod
    def forward(ctx, t):
        return t.to(torch.bfloat16).float()

    @staticmethod
    def backward(ctx, g):
        return g


class _RoundBwd(torch.autograd.Function):        # identity in forward, bf16 rounding of the gradient
    @staticmethod
    def forward(ctx, t):
        return t.view_as(t)

    @staticmethod
    def backward(ctx, g):
        return g.to(torch.bfloat16).float()


def _both(t):
    return t.to(torch.bfloat16).float()            # autograd rounds the gradient on the way back as well


@contextlib.contextmanager
def storage(q_act, q_w):
    """q_act: rounding applied to conv outputs / post-activation tensors, q_w: to the GEMM weights"""
    old_f, old_act = onets.F, onets._act

    class FQ:
        def __getattr__(self, n):
            return getattr(F, n)

        def conv2d(self, x, w, b=None, **k):
            return q_act(F.conv2d(q_act(x), q_w(w), b, **k))

        def conv_transpose2d(self, x, w, b=None, **k):
            return q_act(F.conv_transpose2d(q_act(x), q_w(w), b, **k))

    onets.F = FQ()
    onets._act = lambda x, act: q_act(old_act(x, act))
    try:
        yield
    finally:
        onets.F, onets._act = old_f, old_act


IDENT = lambda t: t  # noqa: E731
VARIANTS = {
    'full': (_both, _both),
    'fwd_only': (_RoundFwd.apply, _RoundFwd.apply),
    'bwd_only': (_RoundBwd.apply, IDENT),
    'weights': (IDENT, _RoundFwd.apply),
}


def run(dim=64, size=256, B=8, nd=4, seed=2):
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    from masterthesis_b200 import models
    from oracle import ref_import
    args = ref_import.default_args(dim=dim, crop_size=size, batch_size=B, num_domains=nd, init_type=None)
    m = models.AdaINModel(args)          # only for the state_dict shapes
    shapes = {n: {k: list(v.shape) for k, v in m.model[n].state_dict().items()} for n in m.model}
    params = {n: synth.synth_params(shapes[n], seed, n + '.') for n in shapes}
    img = torch.cat((synth.synth_images(B, size, seed, 'x1'), synth.synth_images(B, size, seed, 'x2'))).to(DEV)
    c = torch.cat((synth.synth_onehot(B, nd, seed, 'y'), synth.synth_onehot(B, nd, seed, 'y', offset=1))).to(DEV)

    def one(ctx):
        st = ostep.AdaINStep({n: {k: v.to(DEV) for k, v in p.items()} for n, p in params.items()}, B)
        st.Ec = lambda x, draw: onets.content_encoder(
            st.P['content_encoder'], x, draw((x.size(0), dim * 4, x.size(2) // 4, x.size(3) // 4)).to(DEV))
        src = synth.Draws(seed)
        grads = []
        for n in st.opt:
            o = st.opt[n]

            def rec(orig=o.step, n=n, o=o):
                grads.append((n, torch.cat([p.grad.detach().flatten() for p in o.params.values() if p.grad is not None])))
                return orig()
            o.step = rec
        with ctx:
            st.optimize_parameters(img, c, lambda shape: src(shape).to(DEV))
        return grads

    ref = one(contextlib.nullcontext())
    out = {}
    for name, (qa, qw) in VARIANTS.items():
        g = one(storage(qa, qw))
        for i, ((n, a), (_, b)) in enumerate(zip(g, ref)):
            a, b = a.double(), b.double()
            out[f'{name}.step{i}_{n}'] = float(torch.dot(a, b) / (a.norm() * b.norm()))
    return out


if __name__ == '__main__':
    a = [int(v) for v in sys.argv[1:4]]
    print('RESULT ' + json.dumps(run(*a)))
