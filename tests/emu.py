"""TEST INFRASTRUCTURE — the fp32 oracle with bf16 STORAGE emulated (weights, conv outputs and
post-activation tensors rounded to bf16; everything else fp32).  It answers "how far from the
fp32 reference is ANY bf16-storage / fp32-accumulate implementation on this input?", which is the
yardstick for the chained (30+ layer) gradient checks where rounding noise is amplified by
ReLU-mask flips: the CUDA path must be as close to fp32 as this emulation is."""
import contextlib

import torch
import torch.nn.functional as F

from oracle import nets as onets


def Q(t):
    return t.to(torch.bfloat16).float()


class _FQ:
    def __getattr__(self, n):
        return getattr(F, n)

    def conv2d(self, x, w, b=None, **k):
        return Q(F.conv2d(Q(x), Q(w), b, **k))

    def conv_transpose2d(self, x, w, b=None, **k):
        return Q(F.conv_transpose2d(Q(x), Q(w), b, **k))


@contextlib.contextmanager
def bf16_storage():
    old_f, old_act = onets.F, onets._act
    onets.F = _FQ()
    onets._act = lambda x, act: Q(old_act(x, act))
    try:
        yield
    finally:
        onets.F, onets._act = old_f, old_act
