"""Blocks of the translation networks, with the reference's constructor signatures, module tree
and state_dict keys (src/models/core/blocks.py:10-208) — executed as fused CUDA plans.

Every block has two entry points:
  run(...)      native: padded NHWC bf16 buffers in/out (what the networks use);
  forward(x)    reference-compatible: NCHW fp32 in/out (layout conversion at the edge).
"""
import math
import os

import torch
import torch.nn as nn
import torch.nn.functional as F

from ... import functional as Fn
from ... import ops
from ...ops import ConvGeom
from .functions import (act_name, get_activation_layer, get_norm_layer, get_padding_layer,
                        spectral_norm)
from .norm import AdaptiveInstanceNorm, LayerNorm


# ------------------------------------------------------------------------- parameter holders
class HostContext:
    """Host-side state of ONE model instance: the pack epoch (bumped before a CUDA-graph capture so that every
    weight is re-packed inside it), the stream on which weights are packed ahead of use, and the replayed
    dropout draws of the parity harness.  `Model` owns one and binds it to the modules of its networks
    (`bind_context`); modules used on their own share `_DEFAULT_CTX`.  (Round 1 kept these as module globals,
    so two models in one process shared them.)"""

    def __init__(self):
        self.pack_epoch = 0
        self.pack_streams = {}
        self.dropout_rng = None


_DEFAULT_CTX = HostContext()


def context_of(module) -> HostContext:
    return module.__dict__.get('_mt_ctx') or _DEFAULT_CTX


def bind_context(ctx: HostContext, *nets):
    """make every module of `nets` use `ctx` (Model.initialize)"""
    for net in nets:
        for m in net.modules():
            m.__dict__['_mt_ctx'] = ctx


# DecResnetBlock: fold the tiled-z half of its 1x1 convs into a per-image bias (MT_FOLD_Z=0: materialise the concat)
FOLD_TILED_Z = os.environ.get('MT_FOLD_Z', '1') != '0'


def invalidate_packs(ctx: HostContext = None):
    """Drop every cached bf16 weight pack of the context's modules (Model does this before capturing an
    iteration in a CUDA graph: a replayed iteration must re-pack every weight at its first use, because the
    weights were stepped by the previous replay)."""
    (ctx or _DEFAULT_CTX).pack_epoch += 1


# Weight packs ahead of their first use.  A pack is a ~8 us, 64-block kernel: at its point of use it sits
# alone on the GPU in front of the convolution that needs it (78 of them per iteration).  Right after
# an optimizer step (and at the start of an iteration) the networks' weights are re-packed on ONE
# separate stream instead, layer by layer in module order, each pack followed by an event that the
# consuming convolution waits for: the packs of the later layers run under the earlier layers' work.
PREPACK = os.environ.get('MT_PREPACK', '1') == '1'


def _pack_stream(ctx, device):
    s = ctx.pack_streams.get(device)
    if s is None:
        s = ctx.pack_streams[device] = torch.cuda.Stream(device=device)
    return s


def prepack(*nets):
    """Re-pack, on the pack stream, every conv weight of `nets` whose cached pack is stale (same
    geometry as its last use).  Call on the stream that just wrote the weights."""
    if not PREPACK or ops.POISON or ops.PROFILE is not None:
        return
    todo = []
    ctx = None
    for net in nets:
        for m in net.modules():
            if isinstance(m, _ConvParams) and not m.spectral and m._pack is not None:
                w = m.weight
                ctx = context_of(m)
                (_, _, zpad, stem, single, _), _ = m._pack
                key = (w._version, w.data_ptr(), zpad, stem, single, ctx.pack_epoch)
                if m._pack[0] != key:
                    todo.append((m, key, zpad, stem, single))
    if not todo:
        return
    cur = torch.cuda.current_stream()
    side = _pack_stream(ctx, todo[0][0].weight.device)
    side.wait_stream(cur)
    with torch.cuda.stream(side):
        for m, key, zpad, stem, single in todo:
            m._pack = (key, ops.pack_weight(m.weight, m.geom(zpad, stem, single)))
            ev = torch.cuda.Event()
            ev.record(side)
            m._pack_event = ev


def join_prepack(ctx: HostContext = None):
    """current stream waits for every pack issued so far (before the packed weights are overwritten by
    the next optimizer step, and at the end of an iteration so that a captured graph has no open branch)"""
    for dev, s in (ctx or _DEFAULT_CTX).pack_streams.items():
        torch.cuda.current_stream(dev).wait_stream(s)


class _ConvParams(nn.Module):
    transposed = False

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, bias=True):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.stride, self.padding = kernel_size, stride, padding
        shape = ((in_channels, out_channels) if self.transposed else (out_channels, in_channels))
        self.weight = nn.Parameter(torch.empty(*shape, kernel_size, kernel_size))
        self.bias = nn.Parameter(torch.empty(out_channels)) if bias else None
        self.reset_parameters()
        self._pack = None
        self._pack_event = None

    def reset_parameters(self):
        # torch's nn.Conv2d / nn.ConvTranspose2d default init
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            fan_in = self.weight.shape[1] * self.kernel_size * self.kernel_size
            bound = 1 / math.sqrt(fan_in) if fan_in > 0 else 0
            nn.init.uniform_(self.bias, -bound, bound)

    def geom(self, zpad=0, stem=False, single=False) -> ConvGeom:
        """`single`: the call has odd spatial extents, for which the merged-parity / paired-pixel launch
        modes do not exist (plain operand layouts, MtConv2d.single)"""
        return ConvGeom(self.in_channels, self.out_channels, self.kernel_size, self.kernel_size,
                        self.stride, zpad, self.transposed, stem, single)

    def packed(self, zpad=0, stem=False, single=False):
        """bf16 GEMM operands of the fp32 master weight, refreshed when the weight changes."""
        w = self.weight
        key = (w._version, w.data_ptr(), zpad, stem, single, context_of(self).pack_epoch)
        if self._pack is None or self._pack[0] != key:
            self._pack = (key, ops.pack_weight(w, self.geom(zpad, stem, single)))
            self._pack_event = None
        elif self._pack_event is not None:
            # packed ahead of time on the pack stream (prepack below): every consuming stream orders
            # itself behind that launch; waiting on the same event again is free
            torch.cuda.current_stream().wait_event(self._pack_event)
        return self._pack[1]

    # spectral normalisation ------------------------------------------------
    spectral = False

    def enable_spectral_norm(self, eps=1e-12):
        """torch.nn.utils.spectral_norm(module, 'weight', n_power_iterations=1, eps, dim) as the
        reference wraps its convs (functions.py:113-121, blocks.py:34-35,71-72): the parameter becomes
        `weight_orig`, the power-iteration vectors are the buffers `weight_u` / `weight_v` (same
        state_dict keys and order: bias, weight_orig, weight_u, weight_v)."""
        w = self._parameters.pop('weight')
        self.register_parameter('weight_orig', w)
        h, wd = self._sn_matrix(w).shape
        self.register_buffer('weight_u', F.normalize(w.new_empty(h).normal_(0, 1), dim=0, eps=eps))
        self.register_buffer('weight_v', F.normalize(w.new_empty(wd).normal_(0, 1), dim=0, eps=eps))
        self.sn_eps = eps
        self.spectral = True
        return self

    def _sn_matrix(self, w):
        # dim = 1 for transposed convs (functions.py:115-119), 0 otherwise
        return w.transpose(0, 1).flatten(1) if self.transposed else w.flatten(1)

    def sn_weight(self):
        """W / sigma(W).  Training mode runs ONE power iteration per call (in place on the buffers,
        outside the graph), exactly once per reference forward; sigma = u . (W v) is differentiated
        through W only.  The matrix-vector products are two tiny GEMVs per conv on the fp32 master
        weight (host framework ops); the convolution itself consumes the re-packed bf16 operand."""
        w, u, v = self.weight_orig, self.weight_u, self.weight_v
        wm = self._sn_matrix(w)
        if self.training:
            with torch.no_grad():
                v.copy_(F.normalize(torch.mv(wm.t(), u), dim=0, eps=self.sn_eps))
                u.copy_(F.normalize(torch.mv(wm, v), dim=0, eps=self.sn_eps))
        sigma = torch.dot(u.clone(), torch.mv(wm, v.clone()))
        return w / sigma

    def operands(self, zpad=0, stem=False, single=False):
        """(weight as autograd sees it, bf16 forward operand, bf16 backward operand)"""
        if not self.spectral:
            return (self.weight,) + tuple(self.packed(zpad, stem, single))
        w = self.sn_weight()
        return (w,) + tuple(ops.pack_weight(w, self.geom(zpad, stem, single)))

    def extra_repr(self):
        return (f"{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}, "
                f"stride={self.stride}, padding={self.padding}, bias={self.bias is not None}")


class Conv2d(_ConvParams):
    """nn.Conv2d parameters (weight [Cout,Cin,k,k], bias [Cout])."""
    transposed = False


class ConvTranspose2d(_ConvParams):
    """nn.ConvTranspose2d parameters (weight [Cin,Cout,k,k], bias [Cout])."""
    transposed = True

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, output_padding=0,
                 bias=True):
        self.output_padding = output_padding
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, bias)


# ------------------------------------------------------------------------- dropout masks
def set_dropout_rng(rng, ctx: HostContext = None):
    """rng(shape_nchw) -> N(0,1) CPU tensor: replayed draws of the parity harness (keep = draw > 0,
    the rule oracle/make_golden.py installs in the reference); None: device Philox (torch.rand)."""
    (ctx or _DEFAULT_CTX).dropout_rng = rng


def dropout_mask(x, C, p, rng=None):
    """bf16 mask of x's NHWC shape holding 0 or 1/(1-p)   (nn.Dropout(p), blocks.py:133-134,152,190)"""
    N, H, W, Cs = x.shape
    if rng is not None:
        keep = torch.zeros((N, H, W, Cs), dtype=torch.bool, device=x.device)
        keep[..., :C] = (rng((N, C, H, W)) > 0).to(x.device).permute(0, 2, 3, 1)
    else:
        keep = torch.rand((N, H, W, Cs), device=x.device) >= p
    return keep.to(torch.bfloat16) * (1.0 / (1.0 - p))


def apply_dropout(h, C, module):
    """h: dense NHWC bf16; module: the block's nn.Dropout (identity in eval mode)"""
    if not module.training or module.p == 0:
        return h
    return Fn.MaskMul.apply(h, dropout_mask(h, C, module.p, context_of(module).dropout_rng))


# ------------------------------------------------------------------------- layout edges
def to_padded(x, pad, onehot=None):
    """NCHW fp32 -> padded NHWC bf16 (reflection border).  Small channel counts use the image
    packer (autograd-aware); wide tensors go through torch ops (compatibility path only)."""
    if x.shape[1] + (0 if onehot is None else onehot.shape[1]) <= 8:
        return Fn.ImageToNHWC.apply(x, onehot, pad)
    assert onehot is None
    if pad:
        x = F.pad(x, (pad,) * 4, mode='reflect')
    N, C, H, W = x.shape
    out = x.permute(0, 2, 3, 1)
    if ops.cs(C) != C:
        out = F.pad(out, (0, ops.cs(C) - C))
    return out.to(torch.bfloat16).contiguous()


def from_padded(x, C, pad=0):
    """padded NHWC bf16 -> NCHW fp32 (compatibility path only)."""
    if pad:
        x = x[:, pad:-pad, pad:-pad, :]
    return x[..., :C].permute(0, 3, 1, 2).float().contiguous()


def _find(seq, cls):
    for m in seq:
        if isinstance(m, cls):
            return m
    return None


class ConvBlock(nn.Module):
    """[ReflectionPad2d] -> Conv2d -> [norm] -> [activation]   (reference blocks.py:10-46)."""

    def __init__(self, input_dim: int, output_dim: int, kernel_size: int, stride: int = 1,
                 padding: int = 0, bias: bool = False, norm_layer: str = None, activation: str = None,
                 padding_type: str = None, sn: bool = False):
        super().__init__()
        layers = []
        pad_cls = get_padding_layer(padding_type)
        act_cls = get_activation_layer(activation)
        norm_cls = get_norm_layer(norm_layer)
        self.rpad = 0
        if pad_cls is not None:
            layers.append(pad_cls(padding))
            self.rpad, padding = padding, 0
        conv = Conv2d(input_dim, output_dim, kernel_size, stride, padding, bias=bias)
        layers.append(spectral_norm(conv) if sn else conv)
        if norm_cls is not None:
            layers.append(norm_cls(output_dim))
        if act_cls is not None:
            layers.append(act_cls())
        self.block = nn.Sequential(*layers)
        self.zpad = padding
        self.stem = False   # set by the networks on layers fed by the image packer (slack buffer)

    # structural lookups -------------------------------------------------
    @property
    def conv(self):
        return _find(self.block, Conv2d)

    @property
    def norm(self):
        for m in self.block:
            if isinstance(m, (nn.InstanceNorm2d, LayerNorm, nn.BatchNorm2d)):
                return m
        return None

    @property
    def act(self):
        last = self.block[-1]
        return None if isinstance(last, (Conv2d, nn.InstanceNorm2d, LayerNorm, nn.BatchNorm2d)) else act_name(last)

    def _check(self):
        if isinstance(self.block[0], nn.ReplicationPad2d):
            raise NotImplementedError("replicate padding has no CUDA path")
        if isinstance(self.norm, nn.BatchNorm2d):
            raise NotImplementedError("batch norm has no CUDA path (not a default of either model)")

    def run(self, x, out_pad=0, noise=None, act_override='keep'):
        """x: NHWC bf16 with `self.rpad` reflection border.  Returns a buffer padded by out_pad."""
        self._check()
        conv = self.conv
        # odd extents (e.g. the reference's 540x960 sampling size: 135- and 67-wide feature maps): the
        # pixel-pair launch modes need an even output width / even padded input, so such calls use the plain
        # single-pixel launches with their own operand packs
        _, Hp, Wp, _ = x.shape
        Wo = (Wp + 2 * self.zpad - conv.kernel_size) // conv.stride + 1
        single = bool((Hp | Wp | Wo) & 1)
        g = conv.geom(self.zpad, self.stem, single)
        w, wf, wb = conv.operands(self.zpad, self.stem, single)
        act = self.act if act_override == 'keep' else act_override
        norm = self.norm
        if norm is None:
            assert noise is None
            return Fn.ConvAct.apply(x, w, conv.bias, wf, wb, g, act, out_pad)
        if isinstance(norm, LayerNorm):
            mode, p1, p2 = 'layer', (norm.weight if norm.affine else None), (norm.bias if norm.affine else None)
        else:
            mode, p1, p2 = 'instance', None, None
        return Fn.ConvNormAct.apply(x, w, conv.bias, wf, wb, g, mode, p1, p2, act, out_pad, noise)

    def run_image_head(self, x):
        """The decoders' image head when up_type != 'transpose' (networks.py:187-188,252-253,314-315:
        ConvBlock(dim, 3, 7, 1, 3, activation='tanh'), zero padding, no bias): dense NHWC bf16 in, NCHW
        fp32 image out.  The conv stores its pre-activation; tanh is applied on the 3-channel fp32
        image at the layout edge (host framework op, like the NHWC->NCHW conversion it is fused with)."""
        assert self.act == 'tanh' and self.norm is None and self.rpad == 0
        y = self.run(x, act_override=None)
        return torch.tanh(from_padded(y, self.conv.out_channels))

    def forward(self, x):
        y = self.run(to_padded(x, self.rpad))
        return from_padded(y, self.conv.out_channels)


class UpsampleBlock(nn.Module):
    """ConvTranspose2d -> [norm] -> [activation]   (reference blocks.py:48-91, up_type 'transpose')."""

    def __init__(self, input_dim: int, output_dim: int, kernel_size: int, stride: int = 1,
                 padding: int = 0, output_padding: int = 0, bias: bool = False, norm_layer: str = None,
                 activation: str = None, padding_type: str = None, sn: bool = False,
                 up_type: str = 'transpose'):
        super().__init__()
        get_padding_layer(padding_type)
        act_cls = get_activation_layer(activation)
        norm_cls = get_norm_layer(norm_layer)
        layers = []
        self.up_type = up_type
        if 'transpose' in up_type:
            convt = ConvTranspose2d(input_dim, output_dim, kernel_size, stride, padding,
                                    output_padding, bias=bias)
            layers.append(spectral_norm(convt) if sn else convt)
        elif 'nearest' in up_type:
            # blocks.py:74-76: note that the inner ConvBlock gets the padding CLASS (or None) as its
            # padding_type string, so it never pads by reflection: zero padding `padding`
            layers += [nn.Upsample(scale_factor=2, mode='nearest'),
                       ConvBlock(input_dim, output_dim, kernel_size, 1, padding, padding_type=None,
                                 bias=bias, sn=sn)]
        elif 'pixelshuffle' in up_type:
            # blocks.py:77-79 builds ConvBlock(input_dim, output_dim) + PixelShuffle(2) and then a norm
            # layer for output_dim channels on a tensor that has output_dim / 4: the reference itself
            # raises in its first forward (F.layer_norm shape mismatch; DecoderConcat: 138 % 4 != 0), so
            # there is no behaviour to reproduce
            raise NotImplementedError("up_type 'pixelshuffle' cannot run in the reference either "
                                      "(norm layer built for 4x the channels PixelShuffle leaves)")
        else:
            raise NotImplementedError(f"Mode {up_type} is not supported at the moment")
        if norm_cls is not None:
            layers.append(norm_cls(output_dim))
        if act_cls is not None:
            layers.append(act_cls())
        self.block = nn.Sequential(*layers)

    @property
    def conv(self):
        return self.block[1].conv if 'nearest' in self.up_type else self.block[0]

    @property
    def norm(self):
        for m in self.block:
            if isinstance(m, (nn.InstanceNorm2d, LayerNorm)):
                return m
        return None

    @property
    def act(self):
        last = self.block[-1]
        return None if isinstance(last, (ConvTranspose2d, ConvBlock, nn.InstanceNorm2d, LayerNorm)) else act_name(last)

    @property
    def is_rgb_head(self):
        c = self.conv
        return ('transpose' in self.up_type and not c.spectral and c.kernel_size == 1 and c.stride == 1 and c.padding == 0 and c.bias is None
                and self.norm is None and self.act == 'tanh' and c.out_channels <= 4)

    def run(self, x, out_pad=0):
        """x: dense NHWC bf16.  The 1x1 -> tanh head returns the NCHW fp32 image."""
        c = self.conv
        if self.is_rgb_head:
            return Fn.RgbHead.apply(x, c.weight)
        if 'nearest' in self.up_type:
            # Upsample(x2) -> zero-padded stride-1 conv (TMA out-of-bounds fill supplies the zeros)
            x = Fn.UpsampleNearest2.apply(x)
            g = c.geom(self.block[1].zpad)
            w, wf, wb = c.operands(self.block[1].zpad)
        else:
            if not (c.kernel_size == 3 and c.stride == 2 and c.padding == 1 and c.output_padding == 1):
                raise NotImplementedError("only ConvTranspose2d(k=3,s=2,p=1,op=1) and the 1x1 tanh head are accelerated")
            g = c.geom(1)
            w, wf, wb = c.operands(1)
        norm = self.norm
        if norm is None:
            return Fn.ConvAct.apply(x, w, c.bias, wf, wb, g, self.act, out_pad)
        if isinstance(norm, LayerNorm):
            mode, p1, p2 = 'layer', (norm.weight if norm.affine else None), (norm.bias if norm.affine else None)
        else:
            mode, p1, p2 = 'instance', None, None
        return Fn.ConvNormAct.apply(x, w, c.bias, wf, wb, g, mode, p1, p2, self.act, out_pad, None)

    def forward(self, x):
        y = self.run(to_padded(x, 0))
        return y if self.is_rgb_head else from_padded(y, self.conv.out_channels)


class DownResnetBlock(nn.Module):
    """[norm] -> act -> ConvBlock(3x3,+act) -> ConvBlock(3x3) -> AvgPool2  +  AvgPool2 -> Conv1x1
    (reference blocks.py:93-119).  The first activation is in-place in the reference, so BOTH
    branches consume act(x); the native path therefore takes a = act(x) as its input."""

    def __init__(self, input_dim: int, output_dim: int, norm_layer: str = 'instance',
                 activation: str = 'lrelu', padding_type: str = 'reflect', bias: bool = True):
        super().__init__()
        if isinstance(norm_layer, str):
            norm_layer = get_norm_layer(norm_layer)
        if isinstance(activation, str):
            activation = get_activation_layer(activation)
        conv = []
        if norm_layer is not None:
            conv.append(norm_layer(input_dim))
        conv.append(activation())
        conv.append(ConvBlock(input_dim, input_dim, 3, 1, padding=1, padding_type=padding_type,
                              norm_layer=norm_layer, activation=activation, bias=bias))
        conv.append(ConvBlock(input_dim, output_dim, 3, 1, padding=1, padding_type=padding_type, bias=bias))
        conv.append(nn.AvgPool2d(kernel_size=2, stride=2))
        self.conv = nn.Sequential(*conv)
        self.shortcut = nn.Sequential(nn.AvgPool2d(kernel_size=2, stride=2),
                                      Conv2d(input_dim, output_dim, 1, 1, 0, bias=bias))
        self._has_norm = norm_layer is not None

    def run(self, a, out_pad=1):
        """a = LeakyReLU(x) padded by 1.  Returns LeakyReLU(block(x)) padded by out_pad."""
        if self._has_norm or not isinstance(self.conv[0], nn.LeakyReLU):
            raise NotImplementedError("DownResnetBlock is accelerated for norm_layer=None, activation='lrelu' "
                                      "(the only configuration the models build, networks.py:117-121)")
        cA, cB, cS = self.conv[1].conv, self.conv[2].conv, self.shortcut[1]
        packs = (cA.packed(), cB.packed(), cS.packed())
        geoms = (cA.geom(), cB.geom(), cS.geom())
        return Fn.DownBlock.apply(a, cA.weight, cA.bias, cB.weight, cB.bias, cS.weight, cS.bias,
                                  packs, geoms, out_pad)

    def forward(self, x):
        # reference semantics on raw x: returns block(x) WITHOUT the consumer's activation, so the
        # compatibility path undoes the fused LeakyReLU (it is invertible).
        a = to_padded(F.leaky_relu(x, 0.01), 1)
        y = from_padded(self.run(a, out_pad=0), self.shortcut[1].out_channels)
        return torch.where(y > 0, y, y / 0.01)


class ResnetBlock(nn.Module):
    """x + ConvBlock(norm, act) -> ConvBlock(norm) [-> Dropout]   (reference blocks.py:121-138)."""

    def __init__(self, input_dim: int, output_dim: int, dropout: bool = False,
                 norm_layer: str = 'instance', padding_type: str = 'reflect', activation: str = 'relu'):
        super().__init__()
        model = [ConvBlock(input_dim, output_dim, 3, 1, 1, padding_type=padding_type,
                           norm_layer=norm_layer, activation=activation),
                 ConvBlock(output_dim, output_dim, 3, 1, 1, padding_type=padding_type, norm_layer=norm_layer)]
        if dropout:
            model.append(nn.Dropout(0.5))
        self.model = nn.Sequential(*model)
        self.dropout = dropout

    def run(self, x, out_pad=1, noise=None):
        b1, b2 = self.model[0], self.model[1]
        if b1.rpad != 1:
            raise NotImplementedError("ResnetBlock is accelerated for reflect padding")
        drop = self.dropout and self.model[2].training
        if drop or not isinstance(b1.norm, nn.InstanceNorm2d):
            # x + Dropout(block(x)) (the mask sits between the second norm and the residual add), or a
            # norm other than InstanceNorm (--enc_norm layer: each ConvBlock has its own affine
            # LayerNorm): the block runs un-fused: ConvBlock, ConvBlock, [mask], residual add (+ noise,
            # + reflection border)
            h = b2.run(b1.run(x, out_pad=1), out_pad=0)
            if drop:
                h = apply_dropout(h, b2.conv.out_channels, self.model[2])
            return Fn.ResidualPad.apply(x, h, 1, out_pad, noise)
        c1, c2 = b1.conv, b2.conv
        (wf1, wb1), (wf2, wb2) = c1.packed(), c2.packed()
        return Fn.ResBlock.apply(x, c1.weight, c2.weight, wf1, wb1, wf2, wb2, c1.geom(), 'instance',
                                 None, b1.act, out_pad, noise)

    def forward(self, x):
        return from_padded(self.run(to_padded(x, 1), out_pad=0), self.model[1].conv.out_channels)


class AdaINResnetBlock(nn.Module):
    """x + AdaIN(conv2(relu(AdaIN(conv1(x), z))), z) with ONE shared AdaptiveInstanceNorm
    (reference blocks.py:140-167)."""

    def __init__(self, input_dim: int, output_dim: int, dropout: bool = False, style_dim: int = 256,
                 padding_type: str = 'reflect', activation: str = 'relu'):
        super().__init__()
        self.activation = get_activation_layer(activation)()
        self.conv1 = ConvBlock(input_dim, output_dim, 3, 1, 1, padding_type=padding_type)
        self.conv2 = ConvBlock(output_dim, output_dim, 3, 1, 1, padding_type=padding_type)
        self.norm = AdaptiveInstanceNorm(output_dim, style_dim)
        self.dropout = nn.Dropout(0.5) if dropout else nn.Identity()
        self._dropout = dropout

    def run(self, x, z, out_pad=1, gb=None):
        """gb: the block's AdaIN modulation fc(z) [N, 2C] when the decoder computed it with the style MLP
        (functional.StyleMLP); else it is formed here from the style code z"""
        if gb is None:
            gb = self.norm.modulation(z)
        c1, c2 = self.conv1.conv, self.conv2.conv
        if self._dropout and self.dropout.training:
            # the mask sits between the second AdaIN and the residual add: un-fused plan
            (wf1, wb1), (wf2, wb2) = c1.packed(), c2.packed()
            h = Fn.ConvNormAct.apply(x, c1.weight, None, wf1, wb1, c1.geom(), 'adain', gb, None,
                                     act_name(self.activation), 1, None)
            h = Fn.ConvNormAct.apply(h, c2.weight, None, wf2, wb2, c2.geom(), 'adain', gb, None, None, 0, None)
            h = apply_dropout(h, c2.out_channels, self.dropout)
            return Fn.ResidualPad.apply(x, h, 1, out_pad)
        (wf1, wb1), (wf2, wb2) = c1.packed(), c2.packed()
        return Fn.ResBlock.apply(x, c1.weight, c2.weight, wf1, wb1, wf2, wb2, c1.geom(), 'adain',
                                 gb, act_name(self.activation), out_pad, None)

    def forward(self, x, z):
        return from_padded(self.run(to_padded(x, 1), z, out_pad=0), self.conv2.conv.out_channels)


class DecResnetBlock(nn.Module):
    """conv3 -> IN -> cat(z) -> 1x1+ReLU -> 1x1+ReLU -> conv3 -> IN -> cat(z) -> 1x1+ReLU -> 1x1+ReLU
    -> + x   (reference blocks.py:169-208, the residual block of BaseModel's default Decoder)."""

    def __init__(self, n_channel: int, add_channel: int, norm_layer: str = 'instance',
                 padding_type: str = 'reflect', stride: int = 1, dropout: bool = False):
        super().__init__()
        self.conv1 = ConvBlock(n_channel, n_channel, 3, stride=stride, padding=1, padding_type=padding_type)
        self.conv2 = ConvBlock(n_channel, n_channel, 3, stride=stride, padding=1, padding_type=padding_type)
        self.norm = get_norm_layer(norm_layer)(n_channel)
        c = n_channel + add_channel
        self.block1 = nn.Sequential(Conv2d(c, c, 1, stride, 0), nn.ReLU(), Conv2d(c, n_channel, 1, 1, 0), nn.ReLU())
        self.block2 = nn.Sequential(Conv2d(c, c, 1, 1, 0), nn.ReLU(), Conv2d(c, n_channel, 1, 1, 0), nn.ReLU())
        self.dropout = nn.Dropout(0.5) if dropout else nn.Identity()
        self._dropout, self._stride = dropout, stride
        self.n_channel = n_channel

    def _packed_h(self, m):
        """bf16 operands of W[:, :n_channel] of a 1x1 conv over cat(h, z): the part that multiplies h"""
        w = m.weight
        key = (w._version, w.data_ptr(), context_of(self).pack_epoch)
        cache = self.__dict__.setdefault('_h_packs', {})
        rec = cache.get(id(m))
        if rec is None or rec[0] != key:
            g = ConvGeom(self.n_channel, m.out_channels, 1, 1, 1, 0)
            rec = cache[id(m)] = (key, g, ops.pack_weight(w.detach()[:, :self.n_channel].contiguous(), g))
        return rec[1], rec[2]

    def _half(self, x, z, conv, block, out_pad):
        c = conv.conv
        wf, wb = c.packed()
        h = Fn.ConvNormAct.apply(x, c.weight, c.bias, wf, wb, c.geom(), 'instance', None, None, None, 0, None)
        m = block[0]
        N, H, W, _ = h.shape
        if FOLD_TILED_Z and not m.spectral and m.kernel_size == 1 and self._stride == 1:
            # the tiled-z half of K folded into a per-image bias (Fn.ConvZAct): no concatenated tensor
            gh, (wf, wb) = self._packed_h(m)
            if ops.image_bias_ok(gh, N, H, W):
                h = Fn.ConvZAct.apply(h, z, m.weight, m.bias, wf, wb, gh, 'relu', 0)
                m = block[2]
                wf, wb = m.packed()
                return Fn.ConvAct.apply(h, m.weight, m.bias, wf, wb, m.geom(), 'relu', out_pad)
        h = Fn.ConcatChannels.apply(h, z, self.n_channel, 0)
        for i, pad in ((0, 0), (2, out_pad)):
            m = block[i]
            wf, wb = m.packed()
            h = Fn.ConvAct.apply(h, m.weight, m.bias, wf, wb, m.geom(), 'relu', pad)
        return h

    def run(self, x, z, out_pad=1):
        """x padded by 1; z [N, add_channel] fp32.  Returns x + block(x, z) padded by out_pad."""
        if self._stride != 1 or not isinstance(self.norm, nn.InstanceNorm2d) or self.conv1.rpad != 1:
            raise NotImplementedError("DecResnetBlock is accelerated for stride 1, instance norm, reflect padding")
        h = self._half(x, z, self.conv1, self.block1, 1)
        h = self._half(h, z, self.conv2, self.block2, 0)
        if self._dropout:
            h = apply_dropout(h, self.n_channel, self.dropout)
        return Fn.ResidualPad.apply(x, h, 1, out_pad)

    def forward(self, x, z):
        return from_padded(self.run(to_padded(x, 1), z, out_pad=0), self.n_channel)
