"""The translation networks with the reference's constructor signatures, module trees and
state_dict keys (src/models/core/networks.py:8-466), executed as fused CUDA plans.

`forward(...)` keeps the reference contract (NCHW fp32 tensors in and out); `run(...)` is the native
entry the models use: images stay NCHW fp32 (3 channels), the content code stays a padded NHWC
bf16 buffer so that it feeds the decoder's first convolution without a layout round trip.
"""
import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from ... import functional as Fn
from .blocks import (AdaINResnetBlock, Conv2d, ConvBlock, DecResnetBlock, DownResnetBlock, ResnetBlock,
                     UpsampleBlock, from_padded, to_padded)
from .functions import get_activation_layer, get_norm_layer
from .misc import GaussianNoiseLayer

CONTENT_PAD = 1  # border carried by the content code z_c (its consumers are 3x3 / 7x7-pad-1 convs)


class ContentEncoder(nn.Module):
    """7x7 stem -> num_downs x (3x3 s2) -> n_blocks x ResnetBlock -> + N(0,1)
    (reference networks.py:8-43)."""

    def __init__(self, input_dim: int, dim: int = 64, num_downs: int = 2, n_blocks: int = 4,
                 norm_layer: str = 'instance', padding_type: str = 'reflect', bias: bool = True):
        super().__init__()
        layers = [ConvBlock(input_dim, dim, 7, 1, 3, padding_type=padding_type, norm_layer=norm_layer,
                            activation='lrelu', bias=bias)]
        for _ in range(num_downs):
            layers.append(ConvBlock(dim, dim * 2, 3, 2, 1, padding_type=padding_type,
                                    norm_layer=norm_layer, activation='relu', bias=bias))
            dim *= 2
        for _ in range(n_blocks):
            layers.append(ResnetBlock(dim, dim, norm_layer=norm_layer, activation='relu'))
        layers.append(GaussianNoiseLayer())
        layers[0].stem = True
        self.model = nn.Sequential(*layers)
        self.output_dim = dim
        self.num_downs, self.n_blocks = num_downs, n_blocks

    def run(self, img, rng=None):
        """img NCHW fp32 -> z_c padded NHWC bf16 [N, H/4+2, W/4+2, C]."""
        if self.n_blocks < 1:
            raise NotImplementedError("ContentEncoder needs >= 1 ResnetBlock on the CUDA path")
        x = to_padded(img, 3)
        for i in range(1 + self.num_downs):
            x = self.model[i].run(x, out_pad=1)
        first = 1 + self.num_downs
        for i in range(self.n_blocks):
            noise = None
            if i == self.n_blocks - 1:
                N, Hp, Wp, C = x.shape
                noise = self.model[-1].draw(N, Hp - 2, Wp - 2, C, x.device, rng)
            x = self.model[first + i].run(x, out_pad=CONTENT_PAD, noise=noise)
        return x

    def forward(self, x):
        return from_padded(self.run(x), self.output_dim, CONTENT_PAD)


class StyleEncoder(nn.Module):
    """cat(x, one-hot maps) -> 7x7 conv + act -> num_downs x (4x4 s2 conv + act) -> GAP -> 1x1 conv
    (reference networks.py:45-85; BaseModel without --reparam).  All ConvBlocks are bias-free."""

    def __init__(self, input_dim: int, output_dim: int = 8, dim: int = 64, num_downs: int = 4,
                 num_domains: int = 2, padding_type: str = 'reflect', activation: str = 'relu'):
        super().__init__()
        layers = [ConvBlock(input_dim + num_domains, dim, 7, 1, padding=3, padding_type=padding_type,
                            activation=activation)]
        max_filter_size = 256
        out_dim = dim
        for _ in range(num_downs):
            in_dim = min(max_filter_size, dim)
            out_dim = min(max_filter_size, dim * 2)
            layers.append(ConvBlock(in_dim, out_dim, 4, 2, padding=1, padding_type=padding_type,
                                    activation=activation))
            dim *= 2
        layers.append(nn.AdaptiveAvgPool2d(1))
        layers.append(Conv2d(out_dim, output_dim, 1, 1, 0))
        layers[0].stem = True
        self.model = nn.Sequential(*layers)
        self.num_downs, self.out_nch = num_downs, out_dim

    def run(self, img, c, rng=None):
        if self.model[0].rpad != 3:
            raise NotImplementedError("StyleEncoder is accelerated for reflect padding")
        x = to_padded(img, 3, onehot=c)
        for i in range(1 + self.num_downs):
            x = self.model[i].run(x, out_pad=0 if i == self.num_downs else 1)
        feat = Fn.GlobalAvgPool.apply(x)[:, :self.out_nch]
        head = self.model[-1]
        return Fn.Linear.apply(feat, head.weight, head.bias)

    def forward(self, x, c):
        return self.run(x, c)


class ReparameterizedStyleEncoder(nn.Module):
    """cat(x, one-hot maps) -> 4x4 s2 stem -> DownResnetBlocks -> act -> GAP -> fc, fcVar ->
    z = mu + eps * exp(logvar / 2)   (reference networks.py:87-146)."""

    def __init__(self, input_dim: int, output_dim: int = 8, dim: int = 64, n_blocks: int = 4,
                 num_domains: int = 2, norm_layer: str = None, activation: str = None, bias: bool = True):
        super().__init__()
        if isinstance(norm_layer, str):
            norm_layer = get_norm_layer(norm_layer)
        if isinstance(activation, str):
            activation = get_activation_layer(activation)
        max_filter_size = 256
        layers = [ConvBlock(input_dim + num_domains, dim, 4, 2, 1, padding_type='reflect', bias=bias)]
        out_dim = dim
        for _ in range(1, n_blocks):
            in_dim = min(max_filter_size, dim)
            out_dim = min(max_filter_size, dim * 2)
            layers.append(DownResnetBlock(in_dim, out_dim, norm_layer, activation, bias=bias))
            dim *= 2
        layers.append(activation())
        layers.append(nn.AdaptiveAvgPool2d(1))
        layers[0].stem = True
        self.model = nn.Sequential(*layers)
        self.out_nch = out_dim
        self.fc = nn.Linear(out_dim, output_dim)
        self.fcVar = nn.Linear(out_dim, output_dim)
        self.n_blocks = n_blocks

    def reparameterize(self, mu, logvar, rng=None):
        std = logvar.mul(0.5).exp()
        if rng is not None:
            eps = rng(tuple(std.shape)).to(mu.device)
        else:
            eps = torch.randn(std.shape, device=mu.device, dtype=mu.dtype)
        return eps * std + mu

    def run(self, img, c, rng=None):
        if not isinstance(self.model[-2], nn.LeakyReLU):
            raise NotImplementedError("style encoder is accelerated for activation='lrelu' (the models' setting)")
        x = to_padded(img, 1, onehot=c)
        last = self.n_blocks == 1
        a = self.model[0].run(x, out_pad=0 if last else 1, act_override='lrelu')
        for i in range(1, self.n_blocks):
            a = self.model[i].run(a, out_pad=0 if i == self.n_blocks - 1 else 1)
        feat = Fn.GlobalAvgPool.apply(a)[:, :self.out_nch]
        mu = Fn.Linear.apply(feat, self.fc.weight, self.fc.bias)
        logvar = Fn.Linear.apply(feat, self.fcVar.weight, self.fcVar.bias)
        return self.reparameterize(mu, logvar, rng), mu, logvar

    def forward(self, x, c):
        return self.run(x, c)


class AdaINDecoder(nn.Module):
    """style MLP -> n_blocks x AdaINResnetBlock -> num_ups x (ConvT 3x3 s2 + LN + ReLU) -> ConvT 1x1
    + tanh   (reference networks.py:207-270)."""

    def __init__(self, output_dim: int, dim: int = 256, n_blocks: int = 4, num_domains: int = 2,
                 num_ups: int = 2, latent_dim: int = 8, up_type: str = 'transpose',
                 res_norm: str = 'adain', dropout: bool = False, norm_layer: str = 'layer',
                 activation: str = 'relu', bias: bool = True):
        super().__init__()
        self.dim_add = dim
        self.adain = 'adain' in res_norm
        self.dec1 = nn.ModuleList()
        for _ in range(n_blocks):
            if self.adain:
                self.dec1.append(AdaINResnetBlock(dim, self.dim_add, style_dim=self.dim_add, dropout=dropout))
            else:
                self.dec1.append(ResnetBlock(dim, self.dim_add, dropout=dropout))
        ups = []
        for _ in range(num_ups):
            ups.append(UpsampleBlock(dim, dim // 2, 3, 2, 1, 1, norm_layer=norm_layer,
                                     activation=activation, up_type=up_type, bias=bias))
            dim = dim // 2
        if 'transpose' in up_type:
            ups.append(UpsampleBlock(dim, output_dim, 1, 1, 0, activation='tanh', up_type='transpose'))
        else:
            ups.append(ConvBlock(dim, output_dim, 7, 1, 3, activation='tanh'))
        self.dec2 = nn.Sequential(*ups)
        if self.adain:
            self.linear = nn.Sequential(nn.Linear(latent_dim + num_domains, 256), nn.ReLU(inplace=True),
                                        nn.Linear(256, 256), nn.ReLU(inplace=True),
                                        nn.Linear(256, self.dim_add))

    def run(self, z_c, z, c):
        """z_c padded NHWC bf16 (CONTENT_PAD), z [N,latent], c [N,nd] -> NCHW fp32 image."""
        out = z_c
        n = len(self.dec1)
        if self.adain:
            # style MLP + the fc of every block's AdaIN as one plan of our own kernels
            params = [p for i in (0, 2, 4) for p in (self.linear[i].weight, self.linear[i].bias)]
            params += [p for dec in self.dec1 for p in (dec.norm.fc.weight, dec.norm.fc.bias)]
            outs = Fn.StyleMLP.apply(torch.cat([c, z], 1), n, *params)
            for i, dec in enumerate(self.dec1):
                out = dec.run(out, None, out_pad=0 if i == n - 1 else 1, gb=outs[1 + i])
        else:
            for i, dec in enumerate(self.dec1):
                out = dec.run(out, out_pad=0 if i == n - 1 else 1)
        if n == 0:
            out = out[:, CONTENT_PAD:-CONTENT_PAD, CONTENT_PAD:-CONTENT_PAD, :].contiguous()
        return _run_ups(self.dec2, out)

    def forward(self, x, z, c):
        return self.run(to_padded(x, CONTENT_PAD), z, c)


def _run_head(head, out):
    """image head: ConvTranspose2d 1x1 + tanh (up_type 'transpose') or the 7x7 ConvBlock + tanh"""
    return head.run_image_head(out) if isinstance(head, ConvBlock) else head.run(out)


def _run_ups(dec2, out):
    n = len(dec2)
    for i, up in enumerate(dec2):
        out = _run_head(up, out) if i == n - 1 else up.run(out)
    return out


def _make_ups(dim, output_dim, num_ups, up_type, norm_layer, activation, bias):
    ups = []
    for _ in range(num_ups):
        ups.append(UpsampleBlock(dim, dim // 2, 3, 2, 1, 1, norm_layer=norm_layer, activation=activation,
                                 up_type=up_type, bias=bias))
        dim = dim // 2
    if 'transpose' in up_type:
        ups.append(UpsampleBlock(dim, output_dim, 1, 1, 0, activation='tanh', up_type='transpose'))
    else:
        ups.append(ConvBlock(dim, output_dim, 7, 1, 3, activation='tanh'))
    return nn.Sequential(*ups)


class Decoder(nn.Module):
    """style MLP -> n_blocks x DecResnetBlock (one 256-slice of the MLP output each) -> the same
    up-sampling tail as AdaINDecoder   (reference networks.py:148-205; BaseModel's default decoder)."""

    def __init__(self, output_dim: int, dim: int = 256, n_blocks: int = 4, num_domains: int = 2,
                 num_ups: int = 2, latent_dim: int = 8, up_type: str = 'transpose', dropout: bool = False,
                 norm_layer: str = 'layer', activation: str = 'relu', bias: bool = True):
        super().__init__()
        self.dim_add = dim
        self.dec1 = nn.ModuleList([DecResnetBlock(dim, self.dim_add, dropout=dropout) for _ in range(n_blocks)])
        self.dec2 = _make_ups(dim, output_dim, num_ups, up_type, norm_layer, activation, bias)
        self.linear = nn.Sequential(nn.Linear(latent_dim + num_domains, 256), nn.ReLU(inplace=True),
                                    nn.Linear(256, 256), nn.ReLU(inplace=True),
                                    nn.Linear(256, self.dim_add * n_blocks))

    def run(self, z_c, z, c):
        params = [p for i in (0, 2, 4) for p in (self.linear[i].weight, self.linear[i].bias)]
        s = Fn.StyleMLP.apply(torch.cat([c, z], 1), 0, *params)[0]
        out = z_c
        n = len(self.dec1)
        for i, (dec, zi) in enumerate(zip(self.dec1, torch.split(s, self.dim_add, dim=1))):
            out = dec.run(out, zi.contiguous(), out_pad=0 if i == n - 1 else 1)
        if n == 0:
            out = out[:, CONTENT_PAD:-CONTENT_PAD, CONTENT_PAD:-CONTENT_PAD, :].contiguous()
        return _run_ups(self.dec2, out)

    def forward(self, x, z, c):
        return self.run(to_padded(x, CONTENT_PAD), z, c)


class DecoderConcat(nn.Module):
    """ResnetBlock -> cat(c, z maps) -> n_blocks x ResnetBlock -> cat z -> ConvT+LN+ReLU -> cat z ->
    ConvT+LN+ReLU -> cat z -> ConvT 1x1 + tanh   (reference networks.py:272-333; BaseModel --concat).
    The odd channel counts (268/276/138/146/73/81 at latent 8, 4 domains) are stored rounded up to 8."""

    def __init__(self, output_dim: int, dim: int = 256, n_blocks: int = 3, num_domains: int = 2,
                 latent_dim: int = 8, up_type: str = 'transpose', dropout: bool = False,
                 norm_layer: str = 'layer', activation: str = 'relu', bias: bool = True):
        super().__init__()
        self.dec_share = ResnetBlock(dim, dim)
        nch = dim + latent_dim + num_domains
        self.dec1 = nn.Sequential(*[ResnetBlock(nch, nch, dropout=dropout) for _ in range(n_blocks)])
        self._c0, self._c1 = dim, nch
        nch = nch + latent_dim
        self.dec2 = UpsampleBlock(nch, nch // 2, 3, 2, 1, 1, norm_layer=norm_layer, activation=activation,
                                  up_type=up_type, bias=bias)
        nch = nch // 2
        self._c2 = nch
        nch = nch + latent_dim
        self.dec3 = UpsampleBlock(nch, nch // 2, 3, 2, 1, 1, norm_layer=norm_layer, activation=activation,
                                  up_type=up_type, bias=bias)
        nch = nch // 2
        self._c3 = nch
        nch = nch + latent_dim
        if 'transpose' in up_type:
            self.dec4 = UpsampleBlock(nch, output_dim, 1, 1, 0, activation='tanh', up_type='transpose')
        else:
            self.dec4 = ConvBlock(nch, output_dim, 7, 1, 3, activation='tanh')

    def run(self, z_c, z, c):
        n = len(self.dec1)
        out = self.dec_share.run(z_c, out_pad=0)
        out = Fn.ConcatChannels.apply(out, torch.cat([c, z], 1), self._c0, 1 if n else 0)
        for i, blk in enumerate(self.dec1):
            out = blk.run(out, out_pad=0 if i == n - 1 else 1)
        out = self.dec2.run(Fn.ConcatChannels.apply(out, z, self._c1, 0))
        out = self.dec3.run(Fn.ConcatChannels.apply(out, z, self._c2, 0))
        return _run_head(self.dec4, Fn.ConcatChannels.apply(out, z, self._c3, 0))

    def forward(self, x, z, c):
        return self.run(to_padded(x, CONTENT_PAD), z, c)


class ContentDiscriminator(nn.Module):
    """3 x (refl1 -> 7x7 s2 conv + IN + LReLU) -> 4x4 conv + LReLU -> 1x1 conv -> GAP: domain logits of a
    content code   (reference networks.py:386-419, --use_dis_content)."""

    def __init__(self, dim: int = 256, num_domains: int = 3, norm_layer: str = 'instance',
                 activation: str = 'lrelu', padding_type: str = 'reflect', bias: bool = True):
        super().__init__()
        layers = [ConvBlock(dim, dim, kernel_size=7, stride=2, padding=1, padding_type=padding_type,
                            norm_layer=norm_layer, activation=activation, bias=bias) for _ in range(3)]
        layers.append(ConvBlock(dim, dim, kernel_size=4, stride=1, padding=0, padding_type=padding_type,
                                activation=activation, bias=bias))
        layers.append(Conv2d(dim, num_domains, 1, 1, 0))
        self.model = nn.Sequential(*layers)
        self.pool = nn.AdaptiveAvgPool2d(1)
        self.dim = dim

    def run(self, z_c):
        """z_c: padded NHWC bf16 content code (CONTENT_PAD = the reflection pad of the first conv)."""
        x = z_c
        for i in range(4):
            x = self.model[i].run(x, out_pad=1 if i < 2 else 0)
        feat = Fn.GlobalAvgPool.apply(x)[:, :self.dim]     # GAP commutes with the 1x1 conv + bias
        head = self.model[4]
        return Fn.Linear.apply(feat, head.weight, head.bias)

    def forward(self, x):
        return self.run(to_padded(x, CONTENT_PAD))


class MultiScaleDiscriminator(nn.Module):
    """n_layers x (4x4 s2 zero-padded conv + LReLU) applied at num_scales image scales (AvgPool2d(3,2,1,
    count_include_pad=False) between them), heads `dis` (1x1 -> 1) and `cls` (1x1 -> nd, GAP)
    (reference networks.py:421-466, --ms_dis)."""

    def __init__(self, input_dim: int, dim: int = 64, n_layers: int = 6, num_domains: int = 2,
                 norm_layer: str = None, activation: str = 'lrelu', padding_type: str = None,
                 num_scales: int = 3, sn: bool = False):
        super().__init__()
        self.num_scales = num_scales
        self.downsample = nn.AvgPool2d(3, stride=2, padding=[1, 1], count_include_pad=False)
        layers = [ConvBlock(input_dim, dim, 4, 2, 1, norm_layer=None, activation=activation,
                            padding_type=padding_type, sn=sn)]
        for _ in range(n_layers - 1):
            layers.append(ConvBlock(dim, dim * 2, 4, 2, 1, norm_layer=norm_layer, activation=activation,
                                    padding_type=padding_type, sn=sn))
            dim *= 2
        self.model = nn.Sequential(*layers)
        self.dis = Conv2d(dim, 1, 1, 1, 0)
        self.cls = Conv2d(dim, num_domains, 1, 1, 0)
        self.pool = nn.AdaptiveAvgPool2d(1)
        self.output_dim = dim

    def run(self, img):
        outputs = []
        x = img
        for _ in range(self.num_scales):
            h = to_padded(x, self.model[0].rpad)
            n = len(self.model)
            for i, blk in enumerate(self.model):
                h = blk.run(h, out_pad=0 if i == n - 1 else self.model[i + 1].rpad)
            hf = h.float()[..., :self.output_dim]
            dis = torch.einsum('nhwc,c->nhw', hf, self.dis.weight[0, :, 0, 0]).unsqueeze(1) + self.dis.bias.view(1, 1, 1, 1)
            cls = F.linear(hf.mean(dim=(1, 2)), self.cls.weight.flatten(1), self.cls.bias)
            outputs.append((dis, cls))
            x = self.downsample(x)
        return outputs

    def forward(self, x):
        return self.run(x)


class Discriminator(nn.Module):
    """6 x (3x3 s2 + LeakyReLU) -> {1x1 conv with padding=1 (real/fake map), k x k conv + GAP
    (domain logits)}   (reference networks.py:335-384)."""

    def __init__(self, input_dim: int, dim: int = 64, n_layers: int = 6, num_domains: int = 2,
                 norm_layer: str = None, activation: str = 'lrelu', padding_type: str = 'reflect',
                 bias: bool = True, sn: bool = False, image_size: int = 256):
        super().__init__()
        layers = [ConvBlock(input_dim, dim, kernel_size=3, stride=2, padding=1, padding_type=padding_type,
                            norm_layer=norm_layer, sn=sn, activation=activation, bias=bias)]
        nch = dim
        for _ in range(n_layers - 2):
            layers.append(ConvBlock(nch, nch * 2, kernel_size=3, stride=2, padding=1,
                                    padding_type=padding_type, norm_layer=norm_layer, sn=sn,
                                    activation=activation, bias=bias))
            nch *= 2
        layers.append(ConvBlock(nch, nch, kernel_size=3, stride=2, padding=1, padding_type=padding_type,
                                sn=sn, activation=activation, bias=bias))
        layers[0].stem = True
        self.model = nn.Sequential(*layers)
        self.conv1 = Conv2d(nch, 1, kernel_size=1, stride=1, padding=1, bias=False)
        kernal_size = int(image_size / np.power(2, n_layers))
        self.conv2 = Conv2d(nch, num_domains, kernel_size=kernal_size, bias=False)
        self.pool = nn.AdaptiveAvgPool2d(1)
        self.output_dim = nch

    def run(self, img):
        x = to_padded(img, 1)
        n = len(self.model)
        for i, blk in enumerate(self.model):
            x = blk.run(x, out_pad=0 if i == n - 1 else 1)
        w2 = self.conv2.weight
        if x.shape[1] == w2.shape[2] and x.shape[2] == w2.shape[3] and w2.shape[0] <= 8 and x.shape[1] <= 12:
            # both heads in one kernel each way (mt_dis_head_fwd / _bwd): the k x k conv covers the whole map
            return Fn.DisHead.apply(x, self.conv1.weight, w2)
        # other image sizes than crop_size: the general form on the tiny [N,S,S,C] feature map
        h = x.float()[..., :self.output_dim]
        out = torch.einsum('nhwc,c->nhw', h, self.conv1.weight[0, :, 0, 0])
        out = F.pad(out, (1, 1, 1, 1)).unsqueeze(1)           # conv1 has padding=1: exact-zero border
        if h.shape[1] == w2.shape[2] and h.shape[2] == w2.shape[3]:
            cls = torch.einsum('nhwc,dchw->nd', h, w2)
        else:
            cls = F.conv2d(h.permute(0, 3, 1, 2), w2).mean(dim=(2, 3))
        return out, cls

    def forward(self, x):
        return self.run(x)
