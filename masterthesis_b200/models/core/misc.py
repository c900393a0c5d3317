"""GaussianNoiseLayer (reference src/models/core/misc.py:18-26): x + N(0,1) while `.training`.  Inside
ContentEncoder the add is fused into the producing block's apply kernel (`draw` supplies the tensor);
called on its own, `forward` runs the same kernel on an NCHW fp32 tensor."""
import torch
import torch.nn as nn

from ... import functional as Fn


class GaussianNoiseLayer(nn.Module):
    def __init__(self):
        super().__init__()

    def draw(self, N, H, W, C, device, rng=None):
        """fp32 NHWC noise or None (eval mode).  `rng(shape_nchw)` replays reference draws (tests)."""
        if not self.training:
            return None
        if rng is not None:
            return rng((N, C, H, W)).to(device).permute(0, 2, 3, 1).contiguous()
        return torch.randn((N, H, W, C), device=device, dtype=torch.float32)

    def forward(self, x, rng=None):
        """x NCHW fp32 -> x + noise in training mode, x itself otherwise (misc.py:22-26)"""
        if not self.training:
            return x
        from .blocks import from_padded, to_padded
        N, C, H, W = x.shape
        xn = to_padded(x, 0)
        noise = self.draw(N, H, W, xn.shape[3], x.device, None)
        if rng is not None:
            noise.zero_()
            noise[..., :C] = rng((N, C, H, W)).to(x.device).permute(0, 2, 3, 1)
        return from_padded(Fn.AddNoise.apply(xn, noise), C)
