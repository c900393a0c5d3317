"""torch.autograd glue over the CUDA entry points.

Each Function is one *block* of the reference graph (blocks.py) executed as a static plan of our
kernels in forward and backward; autograd only stitches blocks, concatenations and losses
together.  Activations crossing Function boundaries are padded NHWC bf16 buffers whose
reflection border is materialised for the consumer; their gradients have the same padded shape
and are folded by the producer's backward.
"""
from __future__ import annotations

import torch
from torch.autograd import Function

from . import ops


import os as _os
# residual blocks: the skip-path gradient is added by the data-gradient conv's epilogue (MT_FUSE_RES_GRAD=0:
# separate mt_add_interior pass, the round-1 form)
FUSE_RESIDUAL_GRAD = _os.environ.get('MT_FUSE_RES_GRAD', '1') == '1'


def _c(t):
    return t if t is None or t.is_contiguous() else t.contiguous()


def _grad_buf(p):
    """the parameter's gradient buffer if our kernels may add into it directly (a leaf with an fp32
    contiguous .grad — FlatAdam's bucket view — and no higher-order graph being recorded), else None"""
    if p is None or not p.is_leaf:
        return None
    gr = p.grad
    if gr is not None and gr.is_contiguous() and gr.dtype == torch.float32 and not torch.is_grad_enabled():
        return gr
    return None


def _bias_grad(dy, bias):
    """dL/d(bias) = column sums of dy: added into bias.grad in place when possible (returns None, so
    autograd launches no accumulation kernel), else returned"""
    return ops.bias_grad(dy, bias.shape[0], into=_grad_buf(bias))


class ImageToNHWC(Function):
    """NCHW fp32 image (+ tiled one-hot code) -> padded NHWC bf16 (networks.py:136-140 + ReflectionPad2d)."""

    @staticmethod
    def forward(ctx, img, onehot, pad):
        ctx.meta = (img.shape[1], img.shape[2], img.shape[3], pad)
        return ops.image_to_nhwc(img, onehot, pad)

    @staticmethod
    def backward(ctx, dout):
        Ci, H, W, pad = ctx.meta
        dimg = ops.image_grad_from_nhwc(_c(dout), Ci, H, W, pad) if ctx.needs_input_grad[0] else None
        return dimg, None, None


def _wgrad(weight, x, dy, geom, needed):
    """dL/dW of one conv site: added into weight.grad on the side stream when the parameter already
    has a gradient buffer (FlatAdam's bucket view), else returned for autograd to accumulate."""
    if not needed:
        return None
    gr = _grad_buf(weight)                           # spectral-normalised weights are graph nodes
    if gr is not None:
        ops.conv_wgrad_into(x, dy, geom, gr)
        return None
    return ops.conv_wgrad(x, dy, geom)


class ConvAct(Function):
    """[pre-padded input] -> Conv2d/ConvTranspose2d (+bias) (+activation) -> padded output
    (ConvBlock without norm: blocks.py:10-46; used by Discriminator, style encoder, shortcuts)."""

    @staticmethod
    def forward(ctx, x, weight, bias, wf, wb, geom, act, out_pad):
        out = ops.conv_fwd(x, wf, bias, geom, act=act, out_pad=out_pad)
        ops.reflect_border_(out, out_pad)
        ctx.geom, ctx.act, ctx.out_pad = geom, act, out_pad
        ctx.bias = bias
        ctx.weight = weight
        ctx.save_for_backward(x, out if (act is not None) else None, wb)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, out, wb = ctx.saved_tensors
        g, act, op = ctx.geom, ctx.act, ctx.out_pad
        dout = _c(dout)
        if act is None:
            dy = ops.reflect_fold(dout, op)
        else:
            dy = ops.act_bwd_fold(dout, out, op, act)
        dx = ops.conv_dgrad(dy, wb, g, x.shape[1], x.shape[2]) if ctx.needs_input_grad[0] else None
        dw = _wgrad(ctx.weight, x, dy, g, ctx.needs_input_grad[1])
        db = _bias_grad(dy, ctx.bias) if (ctx.bias is not None and ctx.needs_input_grad[2]) else None
        return dx, dw, db, None, None, None, None, None


class ConvZAct(Function):
    """activation(conv1x1(cat(h, z tiled over the plane)) + b) without the concatenation: the first 1x1 conv of
    DecResnetBlock's block1 / block2 (reference blocks.py:180-190, 199-205).  The tiled half of K is a per-image
    constant, so the layer is a convolution of h with W[:, :Cx] plus the per-image bias W[:, Cx:] z_n + b
    (mt_linear_fwd -> mt_conv2d_fwd_image_bias).  Backward: the usual data / weight gradient on (h, W[:, :Cx]);
    the per-image column sums of dy (mt_colsum_images) are the gradient of the per-image bias, from which
    dz, dW[:, Cx:] and db follow as a Linear backward."""

    @staticmethod
    def forward(ctx, h, z, weight, bias, wf_h, wb_h, geom_h, act, out_pad):
        Cx = geom_h.Cin
        wz = weight.detach()[:, Cx:, 0, 0].contiguous()                     # [Cout, Cv] fp32
        bias_n = ops.linear_fwd(z.detach().float().contiguous(), wz, bias.detach() if bias is not None else None)
        out = ops.conv_fwd_image_bias(h, wf_h, bias_n, geom_h, act=act, out_pad=out_pad)
        ops.reflect_border_(out, out_pad)
        ctx.geom, ctx.act, ctx.out_pad = geom_h, act, out_pad
        ctx.has_bias = bias is not None
        ctx.save_for_backward(h, out if (act is not None) else None, wb_h, z, wz)
        ctx.wshape = tuple(weight.shape)
        return out

    @staticmethod
    def backward(ctx, dout):
        h, out, wb_h, z, wz = ctx.saved_tensors
        g, act, op = ctx.geom, ctx.act, ctx.out_pad
        dout = _c(dout)
        dy = ops.reflect_fold(dout, op) if act is None else ops.act_bwd_fold(dout, out, op, act)
        dh = ops.conv_dgrad(dy, wb_h, g, h.shape[1], h.shape[2]) if ctx.needs_input_grad[0] else None
        dw = db = dz = None
        need_w = ctx.needs_input_grad[2]
        if need_w or ctx.needs_input_grad[1] or ctx.needs_input_grad[3]:
            cs = ops.colsum_images(dy, g.Cout)                                 # d(per-image bias) [N, Cout]
            dz, dwz, db = ops.linear_bwd(cs, None, z.float().contiguous(), wz, act=None,
                                         need_dx=ctx.needs_input_grad[1], has_bias=ctx.has_bias,
                                         need_params=need_w or ctx.needs_input_grad[3])
            if need_w:
                dwh = ops.conv_wgrad(h, dy, g)                                 # [Cout, Cx, 1, 1]
                dw = torch.cat((dwh.view(g.Cout, g.Cin), dwz), dim=1).view(ctx.wshape)
            if not ctx.needs_input_grad[3]:
                db = None
            if dz is not None:
                dz = dz.to(z.dtype)
        return dh, dz, dw, db, None, None, None, None, None


class ConvNormAct(Function):
    """conv -> {InstanceNorm | AdaIN | LayerNorm} -> activation (-> + noise) -> padded output
    (ConvBlock with norm: blocks.py:38-42; UpsampleBlock: blocks.py:82-90)."""

    @staticmethod
    def forward(ctx, x, weight, bias, wf, wb, geom, mode, p1, p2, act, out_pad, noise):
        raw, stats = ops.conv_fwd_stats(x, wf, bias, geom)
        N, H, W, C = raw.shape
        gb = p1 if mode == "adain" else None
        lw, lb = (p1, p2) if mode == "layer" else (None, None)
        # a biased conv in front of a LayerNorm: keep the per-channel plane sums for the bias gradient
        want_cs = mode == "layer" and bias is not None and bias.requires_grad
        fin = ops.norm_finalize(stats, mode, H * W, _c(gb), lw, lb, C_valid=geom.Cout, csum=want_cs)
        ss, mr, csum = fin if want_cs else (fin[0], fin[1], None)
        out = ops.affine_act_fwd(raw, ss, act=act, noise=noise, out_p=out_pad)   # incl. reflection border
        ctx.geom, ctx.mode, ctx.act, ctx.out_pad = geom, mode, act, out_pad
        ctx.bias, ctx.p2 = bias, p2
        ctx.weight = weight
        ctx.save_for_backward(x, raw, ss, mr, p1, wb, csum)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, raw, ss, mr, p1, wb, csum = ctx.saved_tensors
        g, mode, act, op = ctx.geom, ctx.mode, ctx.act, ctx.out_pad
        bias, p2 = ctx.bias, ctx.p2
        need_db = bias is not None and ctx.needs_input_grad[2]
        ln = mode == "layer" and p1 is not None
        dln_into = None
        if ln and _grad_buf(p1) is not None and _grad_buf(p2) is not None:
            dln_into = (_grad_buf(p1), _grad_buf(p2))
        dy, dgb, dlw, dlb, db = ops.norm_bwd(
            _c(dout), raw, ss, mr, mode, act=act, gb=_c(p1) if mode == "adain" else None,
            ln_w=p1 if ln else None, C_valid=g.Cout, g_pad=op, csum=csum,
            dln_into=dln_into, dbias_into=_grad_buf(bias) if need_db else None,
            want_dbias=need_db and csum is not None)
        if need_db and db is None and csum is None and _grad_buf(bias) is None:
            # bias in front of InstanceNorm / AdaIN: sum_hw dx is exactly zero for every plane
            db = torch.zeros_like(bias)
        dx = ops.conv_dgrad(dy, wb, g, x.shape[1], x.shape[2]) if ctx.needs_input_grad[0] else None
        dw = _wgrad(ctx.weight, x, dy, g, ctx.needs_input_grad[1])
        dp1 = dgb if mode == "adain" else (dlw.view_as(p1) if ln and dlw is not None else None)
        dp2 = dlb.view_as(p2) if (ln and dlb is not None) else None
        return dx, dw, db, None, None, None, None, dp1, dp2, None, None, None


class NormAct(Function):
    """{InstanceNorm2d | AdaptiveInstanceNorm | LayerNorm} (+ activation) on a dense NHWC bf16 tensor: the
    stand-alone form behind the reference-compatible `forward` of the norm modules (norm.py:16-21,29-33,
    functions.py:17); the networks use the fused ConvNormAct / ResBlock plans instead."""

    @staticmethod
    def forward(ctx, x, mode, p1, p2, act, C_valid):
        N, H, W, C = x.shape
        gb = _c(p1) if mode == "adain" else None
        lw, lb = (p1, p2) if mode == "layer" else (None, None)
        ss, mr = ops.norm_finalize(ops.plane_stats(x), mode, H * W, gb, lw, lb, C_valid=C_valid)
        out = ops.affine_act_fwd(x, ss, act=act)
        ctx.mode, ctx.act, ctx.C_valid = mode, act, C_valid
        ctx.save_for_backward(x, ss, mr, p1)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, ss, mr, p1 = ctx.saved_tensors
        mode = ctx.mode
        ln = mode == "layer" and p1 is not None
        dx, dgb, dlw, dlb, _ = ops.norm_bwd(_c(dout), x, ss, mr, mode, act=ctx.act,
                                            gb=_c(p1) if mode == "adain" else None,
                                            ln_w=p1 if ln else None, C_valid=ctx.C_valid)
        dp1 = dgb if mode == "adain" else (dlw.view_as(p1) if ln else None)
        dp2 = dlb.view_as(p1) if ln else None
        return dx, None, dp1, dp2, None, None


class AddNoise(Function):
    """x + noise (GaussianNoiseLayer, misc.py:22-26) on a dense NHWC bf16 tensor; noise fp32 NHWC."""

    @staticmethod
    def forward(ctx, x, noise):
        return ops.affine_act_fwd(x, None, noise=noise)

    @staticmethod
    def backward(ctx, dout):
        return dout, None


class ResBlock(Function):
    """x + norm(conv3x3(relu(norm(conv3x3(x)))))  with norm = InstanceNorm2d (ResnetBlock,
    blocks.py:121-138) or one shared AdaptiveInstanceNorm (AdaINResnetBlock, blocks.py:140-167).
    Optional Gaussian noise on the block output (ContentEncoder tail, networks.py:38, misc.py:25)."""

    @staticmethod
    def forward(ctx, x, w1, w2, wf1, wb1, wf2, wb2, geom, mode, gb, act, out_pad, noise):
        gbc = _c(gb) if mode == "adain" else None
        raw1, st1 = ops.conv_fwd_stats(x, wf1, None, geom)
        N, H, W, C = raw1.shape
        ss1, mr1 = ops.norm_finalize(st1, mode, H * W, gbc)
        h = ops.affine_act_fwd(raw1, ss1, act=act, out_p=1)                       # incl. reflection border
        raw2, st2 = ops.conv_fwd_stats(h, wf2, None, geom)
        ss2, mr2 = ops.norm_finalize(st2, mode, H * W, gbc)
        out = ops.affine_act_fwd(raw2, ss2, act=None, res_pad=x, res_p=1, noise=noise, out_p=out_pad)
        ctx.geom, ctx.mode, ctx.act, ctx.out_pad = geom, mode, act, out_pad
        ctx.w1, ctx.w2 = w1, w2
        ctx.save_for_backward(x, raw1, h, raw2, ss1, mr1, ss2, mr2, gbc, wb1, wb2)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, raw1, h, raw2, ss1, mr1, ss2, mr2, gb, wb1, wb2 = ctx.saved_tensors
        g, mode, act, op = ctx.geom, ctx.mode, ctx.act, ctx.out_pad
        dout = _c(dout)
        # the adjoint of the reflection borders is applied inside the norm-backward kernels (g_pad)
        dy2, dgb2 = ops.norm_bwd(dout, raw2, ss2, mr2, mode, act=None, gb=gb, g_pad=op)[:2]
        dw2 = _wgrad(ctx.w2, h, dy2, g, ctx.needs_input_grad[2])
        dh = ops.conv_dgrad(dy2, wb2, g, h.shape[1], h.shape[2])
        dy1, dgb1 = ops.norm_bwd(dh, raw1, ss1, mr1, mode, act=act, gb=gb, g_pad=1)[:2]
        dw1 = _wgrad(ctx.w1, x, dy1, g, ctx.needs_input_grad[1])
        dx = None
        if ctx.needs_input_grad[0]:
            if op == 1 and FUSE_RESIDUAL_GRAD:
                # residual path: out_pad = pad(x_int + h) and x is padded like out, so adding the PADDED
                # dout is the same as adding fold(dout) to the interior once x's producer folds dx.  The
                # data-gradient conv adds its result straight onto dout (TMA reduce-add epilogue): dout is
                # this node's own incoming gradient buffer (produced by the next block's backward or by
                # autograd's accumulation), nothing else reads it afterwards
                dx = ops.conv_dgrad(dy1, wb1, g, x.shape[1], x.shape[2], add_into=dout)
            else:
                dx = ops.conv_dgrad(dy1, wb1, g, x.shape[1], x.shape[2])
                if op == 1:
                    ops.add_interior_(dx, dout, 0)
                else:
                    ops.add_interior_(dx, ops.reflect_fold(dout, op), 1)
        dgb = (dgb1 + dgb2) if mode == "adain" else None
        return dx, dw1, dw2, None, None, None, None, None, None, dgb, None, None, None


class DownBlock(Function):
    """DownResnetBlock (blocks.py:93-119) on a = LeakyReLU(x) (the in-place activation that both
    branches see), fused with the LeakyReLU that consumes the block's sum:
        out = lrelu( avgpool2(convB(lrelu(convA(a)))) + conv1x1(avgpool2(a)) )"""

    @staticmethod
    def forward(ctx, a, wA, bA, wB, bB, wS, bS, packs, geoms, out_pad):
        (wfA, wbA), (wfB, wbB), (wfS, wbS) = packs
        gA, gB, gS = geoms
        b = ops.conv_fwd(a, wfA, bA, gA, act="lrelu", out_pad=1)
        ops.reflect_border_(b, 1)
        c = ops.conv_fwd(b, wfB, bB, gB)
        pa = ops.avgpool2_fwd(a, 1)
        d = ops.conv_fwd(pa, wfS, bS, gS)
        out = ops.pool_add_act_fwd(c, d, "lrelu", out_pad)
        ops.reflect_border_(out, out_pad)
        ctx.geoms, ctx.out_pad = geoms, out_pad
        ctx.ws = (wA, wB, wS)
        ctx.bs = (bA, bB, bS)
        ctx.save_for_backward(a, b, pa, out, wbA, wbB, wbS)
        return out

    @staticmethod
    def backward(ctx, dout):
        a, b, pa, out, wbA, wbB, wbS = ctx.saved_tensors
        gA, gB, gS = ctx.geoms
        need = ctx.needs_input_grad
        dd, dc = ops.pool_add_act_bwd(_c(dout), out, ctx.out_pad, "lrelu")
        bA, bB, bS = ctx.bs
        dwB = _wgrad(ctx.ws[1], b, dc, gB, need[3])
        # dc is dd spread over the 2x2 windows with weight 1/4 (exact in bf16): its column sums equal dd's,
        # so ONE pass over the quarter-size dd yields both bias gradients
        dbB = _bias_grad(dd, bB) if (bB is not None and need[4]) else None
        db_pad = ops.conv_dgrad(dc, wbB, gB, b.shape[1], b.shape[2])
        dyA = ops.act_bwd_fold(db_pad, b, 1, "lrelu")
        dwA = _wgrad(ctx.ws[0], a, dyA, gA, need[1])
        dbA = _bias_grad(dyA, bA) if (bA is not None and need[2]) else None
        dwS = _wgrad(ctx.ws[2], pa, dd, gS, need[5])
        dbS = _bias_grad(dd, bS) if (bS is not None and need[6]) else None
        da = None
        if need[0]:
            da = ops.conv_dgrad(dyA, wbA, gA, a.shape[1], a.shape[2])
            dpa = ops.conv_dgrad(dd, wbS, gS, pa.shape[1], pa.shape[2])
            ops.avgpool2_bwd_add_(dpa, da, 1)
        return da, dwA, dbA, dwB, dbB, dwS, dbS, None, None, None


class GlobalAvgPool(Function):
    """nn.AdaptiveAvgPool2d(1) on a dense NHWC bf16 tensor -> fp32 [N, C]  (networks.py:124)."""

    @staticmethod
    def forward(ctx, x):
        ctx.hw = (x.shape[1], x.shape[2])
        return ops.gap_fwd(x)

    @staticmethod
    def backward(ctx, dout):
        return ops.gap_bwd(dout.float(), *ctx.hw)


class RgbHead(Function):
    """ConvTranspose2d(C -> 3, k=1, bias=False) + Tanh, NHWC bf16 -> NCHW fp32 (networks.py:248-251)."""

    @staticmethod
    def forward(ctx, x, weight):
        Cin, Co = weight.shape[0], weight.shape[1]
        w2 = weight.detach().reshape(Cin, Co).float()
        if x.shape[3] != Cin:                       # stored channels are padded to a multiple of 8
            w2 = torch.cat((w2, w2.new_zeros(x.shape[3] - Cin, Co)))
        w2 = w2.contiguous()
        y = ops.rgb_head_fwd(x, w2, Co)
        ctx.save_for_backward(x, y, w2)
        ctx.wshape = weight.shape
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, w2 = ctx.saved_tensors
        dx, dw = ops.rgb_head_bwd(dy.float(), y, x, w2, w2.shape[1])
        return dx, dw[:ctx.wshape[0]].reshape(ctx.wshape)


class SumSqInterior(Function):
    """sum(z^2) over the interior of a padded buffer (torch.pow(mu,2) of adain_model.py:396-399)."""

    @staticmethod
    def forward(ctx, x, pad):
        ctx.pad = pad
        ctx.save_for_backward(x)
        return ops.sumsq_interior(x, pad)

    @staticmethod
    def backward(ctx, dout):
        (x,) = ctx.saved_tensors
        scale = (dout.float() * 2.0).reshape(())
        return ops.scale_interior(x, scale.contiguous(), ctx.pad), None


class ConcatChannels(Function):
    """torch.cat([x, v.expand(H, W)], dim=1) of the reference's decoders (blocks.py:198-205 DecResnetBlock,
    networks.py:320-332 DecoderConcat) on NHWC bf16: x [N,H,W,Cs] holds Cx logical channels, v [N,Cv]
    fp32 is broadcast over the plane into channels Cx..Cx+Cv-1; the result is padded by `out_pad`
    (reflection border) for a following 3x3 convolution.  Layout glue: torch copies, no arithmetic."""

    @staticmethod
    def forward(ctx, x, v, Cx, out_pad):
        N, H, W, Cs = x.shape
        Cv = v.shape[1]
        Ct = ops.cs(Cx + Cv)
        out = ops.empty_act(N, H + 2 * out_pad, W + 2 * out_pad, Ct, x.device)
        inner = out[:, out_pad:out_pad + H, out_pad:out_pad + W, :]
        inner[..., :Cx] = x[..., :Cx]
        inner[..., Cx:Cx + Cv] = v.to(torch.bfloat16)[:, None, None, :]
        if Ct > Cx + Cv:
            inner[..., Cx + Cv:] = 0
        ops.reflect_border_(out, out_pad)
        ctx.meta = (Cx, Cv, Cs, out_pad)
        return out

    @staticmethod
    def backward(ctx, dout):
        Cx, Cv, Cs, op = ctx.meta
        g = ops.reflect_fold(_c(dout), op)
        dx = None
        if ctx.needs_input_grad[0]:
            if Cs == Cx:
                dx = g[..., :Cx].contiguous()
            else:
                dx = torch.zeros(g.shape[:3] + (Cs,), dtype=g.dtype, device=g.device)
                dx[..., :Cx] = g[..., :Cx]
        dv = g[..., Cx:Cx + Cv].float().sum(dim=(1, 2)) if ctx.needs_input_grad[1] else None
        return dx, dv, None, None


class ResidualPad(Function):
    """out = x + y written into a buffer padded by `out_pad`; x is padded by x_pad, y dense
    (the `out += residual` of DecResnetBlock, blocks.py:206)."""

    @staticmethod
    def forward(ctx, x, y, x_pad, out_pad, noise=None):
        # `noise`: optional fp32 NHWC tensor added on top (GaussianNoiseLayer behind the block, misc.py:25)
        out = ops.affine_act_fwd(y, None, act=None, res_pad=x, res_p=x_pad, noise=noise, out_p=out_pad)
        ctx.meta = (x_pad, out_pad, x.shape)
        return out

    @staticmethod
    def backward(ctx, dout):
        x_pad, op, xshape = ctx.meta
        G = ops.reflect_fold(_c(dout), op)
        if G is dout:
            G = G.clone()
        dx = None
        if ctx.needs_input_grad[0]:
            if x_pad:
                dx = torch.zeros(xshape, dtype=G.dtype, device=G.device)
                ops.add_interior_(dx, G, x_pad)
            else:
                dx = G.clone()      # never hand ONE buffer to two producers: a fused ResBlock adds into its gradient
        return dx, G, None, None, None


class UpsampleNearest2(Function):
    """nn.Upsample(scale_factor=2, mode='nearest') of UpsampleBlock(up_type='nearest') (blocks.py:74-76)
    on a dense NHWC bf16 tensor."""

    @staticmethod
    def forward(ctx, x):
        return ops.upsample2_fwd(_c(x))

    @staticmethod
    def backward(ctx, dout):
        return ops.upsample2_bwd(_c(dout))


class MaskMul(Function):
    """nn.Dropout(p) in training mode (blocks.py:133-134,152,190): x * m with m in {0, 1/(1-p)} drawn by
    the caller (bf16, same shape as x)."""

    @staticmethod
    def forward(ctx, x, m):
        ctx.save_for_backward(m)
        return ops.mul(_c(x), m)

    @staticmethod
    def backward(ctx, dout):
        (m,) = ctx.saved_tensors
        return ops.mul(_c(dout), m), None


class Linear(Function):
    """nn.Linear (+ReLU) on fp32 [N, In] rows by our own kernels (fc / fcVar of the style encoder
    networks.py:127-128, the 1x1 conv heads behind a global average pool networks.py:77,402-403: a weight
    [Out, In, 1, 1] is the same memory as [Out, In])."""

    @staticmethod
    def forward(ctx, x, weight, bias, act=None):
        x = _c(x.float())
        w2 = _c(weight.detach()).view(weight.shape[0], -1)
        y = ops.linear_fwd(x, w2, bias, act)
        ctx.act, ctx.weight, ctx.bias = act, weight, bias
        ctx.save_for_backward(x, y if act is not None else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y = ctx.saved_tensors
        w, b = ctx.weight, ctx.bias
        need_w = ctx.needs_input_grad[1]
        dx, dw, db = ops.linear_bwd(_c(dy.float()), y, x, _c(w.detach()).view(w.shape[0], -1), ctx.act,
                                    ctx.needs_input_grad[0], _grad_buf(w) if need_w else None,
                                    _grad_buf(b) if (need_w and b is not None) else None,
                                    has_bias=b is not None, need_params=need_w)
        return dx, (None if dw is None else dw.view(w.shape)), db, None


class DisHead(Function):
    """The two heads of Discriminator (networks.py:373-376,380-383) on the last feature map h [N,S,S,Cs] bf16:
    out = Conv2d(C, 1, 1, padding=1, bias=False)(h) -> [N,1,S+2,S+2] (exact-zero border) and
    cls = AdaptiveAvgPool2d(1)(Conv2d(C, nd, k=S, bias=False)(h)) -> [N, nd]; one kernel each way."""

    @staticmethod
    def forward(ctx, h, w1, w2):
        ctx.set_materialize_grads(False)
        C, nd = w1.shape[1], w2.shape[0]
        out, cls = ops.dis_head_fwd(_c(h), _c(w1.detach()).view(-1), _c(w2.detach()), C, nd)
        ctx.w1, ctx.w2 = w1, w2
        ctx.save_for_backward(h)
        return out, cls

    @staticmethod
    def backward(ctx, dout, dcls):
        (h,) = ctx.saved_tensors
        w1, w2 = ctx.w1, ctx.w2
        C, nd = w1.shape[1], w2.shape[0]
        N, S = h.shape[0], h.shape[1]
        if dout is None:
            dout = torch.zeros((N, 1, S + 2, S + 2), dtype=torch.float32, device=h.device)
        if dcls is None:
            dcls = torch.zeros((N, nd), dtype=torch.float32, device=h.device)
        need_w = ctx.needs_input_grad[1] or ctx.needs_input_grad[2]
        into = None
        if need_w and _grad_buf(w1) is not None and _grad_buf(w2) is not None:
            into = (_grad_buf(w1), _grad_buf(w2))
        dh, dw1, dw2 = ops.dis_head_bwd(_c(dout.float()), _c(dcls.float()), h, _c(w1.detach()).view(-1), _c(w2.detach()),
                                        C, nd, 0, ctx.needs_input_grad[0], into, need_w)
        return dh, (None if dw1 is None else dw1.view(w1.shape)), dw2


class StyleMLP(Function):
    """The decoder's style network as one plan (networks.py:256-261, norm.py:27 of every AdaINResnetBlock):
    s = L2(relu(L1(relu(L0(cz))))) and gb_k = fc_k(s) for each of the K AdaIN blocks.
    apply(cz, K, w0, b0, w1, b1, w2, b2, hw_0, hb_0, ..., hw_{K-1}, hb_{K-1}) -> (s, gb_0, ..., gb_{K-1})"""

    @staticmethod
    def forward(ctx, cz, K, *params):
        ctx.set_materialize_grads(False)
        cz = _c(cz.float())
        mlp = [(_c(params[2 * i]), params[2 * i + 1]) for i in range(3)]
        heads = [(_c(params[6 + 2 * k]), params[7 + 2 * k]) for k in range(K)]
        h1, h2, s, gb = ops.style_mlp_fwd(cz, mlp, heads)
        ctx.K, ctx.params = K, params
        ctx.save_for_backward(cz, h1, h2, s)
        return (s,) + tuple(gb)

    @staticmethod
    def backward(ctx, ds, *dgb):
        cz, h1, h2, s = ctx.saved_tensors
        K, params = ctx.K, ctx.params
        mlp_w = [params[2 * i].detach() for i in range(3)]
        head_w = [params[6 + 2 * k].detach() for k in range(K)]
        bufs = [_grad_buf(p) for p in params]
        need_params = any(ctx.needs_input_grad[2:])
        into = None
        if need_params and all(b is not None for b in bufs):
            into = ([bufs[0], bufs[2], bufs[4]], [bufs[1], bufs[3], bufs[5]],
                    [bufs[6 + 2 * k] for k in range(K)], [bufs[7 + 2 * k] for k in range(K)])
        dcz, d_mw, d_mb, d_hw, d_hb = ops.style_mlp_bwd(cz, h1, h2, s, mlp_w, head_w, list(dgb), ds, into,
                                                        ctx.needs_input_grad[0])
        if into is not None or not need_params:
            return (dcz, None) + (None,) * len(params)
        out = []
        for i in range(3):
            out += [d_mw[i], d_mb[i]]
        for k in range(K):
            out += [d_hw[k], d_hb[k]]
        return (dcz, None) + tuple(out)
