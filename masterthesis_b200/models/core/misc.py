"""GaussianNoiseLayer descriptor (reference src/models/core/misc.py:18-26): x + N(0,1) while
`.training`.  The add is fused into the producing block's apply kernel; this module only carries
the training flag and draws the noise tensor."""
import torch
import torch.nn as nn


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
