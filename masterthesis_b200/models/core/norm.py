"""Parameter holders for the decoder norms (reference src/models/core/norm.py:5-33).  The
arithmetic runs in csrc/ew_norm.cu through the enclosing block's fused plan."""
import torch
import torch.nn as nn


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


class AdaptiveInstanceNorm(nn.Module):
    """(1 + w) * InstanceNorm(x) + b with (w, b) = chunk(fc(s), 2)   (norm.py:23-33)."""

    def __init__(self, num_features, style_dim):
        super().__init__()
        self.norm = nn.InstanceNorm2d(num_features, affine=False)
        self.fc = nn.Linear(style_dim, num_features * 2)

    def modulation(self, s):
        """fc(s): [N, 2C] fp32 = (weight | bias); consumed by the AdaIN kernels."""
        return self.fc(s)
