"""The decoder norms (reference src/models/core/norm.py:5-33).  Inside the networks their arithmetic runs
through the enclosing block's fused plan (csrc/ew_norm.cu via functional.ConvNormAct / ResBlock); called
on their own, as the reference's modules can be, `forward` runs the same kernels un-fused on NCHW fp32
tensors (layout conversion at the edge)."""
import torch
import torch.nn as nn

from ... import functional as Fn


def _nhwc(x):
    from .blocks import to_padded
    return to_padded(x, 0)


def _nchw(y, C):
    from .blocks import from_padded
    return from_padded(y, C)


class LayerNorm(nn.Module):
    """Layer norm over (C,H,W) with per-channel affine; eps is accepted and ignored exactly as in
    the reference (norm.py:6-21 never stores it, so F.layer_norm's 1e-5 applies)."""

    def __init__(self, n_out, eps=1e-5, affine=True):
        super().__init__()
        self.n_out = n_out
        self.affine = affine
        if affine:
            self.weight = nn.Parameter(torch.ones(n_out, 1, 1))
            self.bias = nn.Parameter(torch.zeros(n_out, 1, 1))

    def forward(self, x):
        """F.layer_norm(x, x.shape[1:], weight.expand, bias.expand) (norm.py:16-21); x NCHW fp32"""
        C = x.shape[1]
        w, b = (self.weight, self.bias) if self.affine else (None, None)
        return _nchw(Fn.NormAct.apply(_nhwc(x), 'layer', w, b, None, C), C)


class AdaptiveInstanceNorm(nn.Module):
    """(1 + w) * InstanceNorm(x) + b with (w, b) = chunk(fc(s), 2)   (norm.py:23-33)."""

    def __init__(self, num_features, style_dim):
        super().__init__()
        self.norm = nn.InstanceNorm2d(num_features, affine=False)
        self.fc = nn.Linear(style_dim, num_features * 2)

    def modulation(self, s):
        """fc(s): [N, 2C] fp32 = (weight | bias); consumed by the AdaIN kernels."""
        return self.fc(s)

    def forward(self, x, s):
        """x NCHW fp32, s [N, style_dim] -> (1 + w) * IN(x) + b   (norm.py:29-33)"""
        C = x.shape[1]
        gb = self.modulation(s)
        if C % 8:      # the kernels address (weight | bias) at the stored channel count
            Cs = (C + 7) // 8 * 8
            z = gb.new_zeros(gb.shape[0], Cs - C)
            gb = torch.cat((gb[:, :C], z, gb[:, C:], z), 1)
        return _nchw(Fn.NormAct.apply(_nhwc(x), 'adain', gb, None, None, C), C)
