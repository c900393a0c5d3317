"""Op-level parity cases: each CUDA entry point vs the same op in plain PyTorch fp32 on the same
bf16-rounded inputs (reference call sites: src/models/core/blocks.py:30-42,69-73, norm.py:5-33).
Used by tests/test_kernels_gpu.py and tools/gpu_probe.py.  Returns relative errors
max|a-b| / max|b|; bf16 storage + fp32 accumulate bounds them by ~2^-8.
"""
import torch
import torch.nn.functional as F

from masterthesis_b200 import ops
from masterthesis_b200.ops import ConvGeom

DEV = "cuda"


def rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-12)).item()


def bf(x):
    return x.to(torch.bfloat16).float()


def to_nhwc(x, Cs=None):
    """NCHW fp32 -> NHWC bf16 with channels zero-padded to the stored count."""
    N, C, H, W = x.shape
    Cs = Cs or ops.cs(C)
    out = torch.zeros(N, H, W, Cs, dtype=torch.bfloat16, device=x.device)
    out[..., :C] = x.permute(0, 2, 3, 1).to(torch.bfloat16)
    return out


def from_nhwc(x, C):
    return x[..., :C].permute(0, 3, 1, 2).float()


def conv_case(N, Cin, Cout, H, W, k, stride, rpad, zpad=0, bias=True, act=None, out_pad=0, seed=0,
              stem=False):
    """Conv2d forward + dgrad + wgrad.  rpad: reflection padding materialised in the input."""
    torch.manual_seed(seed)
    g = ConvGeom(Cin, Cout, k, k, stride, zpad, False, stem)
    x = bf(torch.randn(N, Cin, H, W, device=DEV))
    xp = F.pad(x, (rpad,) * 4, mode="reflect") if rpad else x
    w = bf(torch.randn(Cout, Cin, k, k, device=DEV) * 0.05)
    b = torch.randn(Cout, device=DEV) if bias else None
    xp_ref = xp.clone().requires_grad_(True)
    w_ref = w.clone().requires_grad_(True)
    y_ref = F.conv2d(xp_ref, w_ref, b, stride=stride, padding=zpad)
    if act == "lrelu":
        y_ref = F.leaky_relu(y_ref, 0.01)
    elif act == "relu":
        y_ref = F.relu(y_ref)
    dy = bf(torch.randn_like(y_ref))
    y_ref.backward(dy)

    if g.window:
        xn = ops.image_to_nhwc(x, None, rpad)          # slack-padded buffer for the row-window TMA
    else:
        xn = to_nhwc(xp)
    wf, wb = ops.pack_weight(w, g)
    y = ops.conv_fwd(xn, wf, b, g, act=act, out_pad=out_pad)
    if out_pad:
        y = y[:, out_pad:-out_pad, out_pad:-out_pad, :].contiguous()
    res = {"fwd": rel(from_nhwc(y, Cout), y_ref.detach())}
    if act is None:
        dyn = to_nhwc(dy)
        dx = ops.conv_dgrad(dyn, wb, g, xn.shape[1], xn.shape[2])
        res["dgrad"] = rel(from_nhwc(dx, Cin), xp_ref.grad)
        dw = ops.conv_wgrad(xn, dyn, g)
        res["wgrad"] = rel(dw, w_ref.grad)
    return res


def convT_case(N, Cin, Cout, H, W, bias=True, act=None, seed=0):
    torch.manual_seed(seed)
    g = ConvGeom(Cin, Cout, 3, 3, 2, 1, transposed=True)
    x = bf(torch.randn(N, Cin, H, W, device=DEV))
    w = bf(torch.randn(Cin, Cout, 3, 3, device=DEV) * 0.05)
    b = torch.randn(Cout, device=DEV) if bias else None
    x_ref = x.clone().requires_grad_(True)
    w_ref = w.clone().requires_grad_(True)
    y_ref = F.conv_transpose2d(x_ref, w_ref, b, stride=2, padding=1, output_padding=1)
    dy = bf(torch.randn_like(y_ref))
    y_ref.backward(dy)
    xn = to_nhwc(x)
    wf, wb = ops.pack_weight(w, g)
    y = ops.conv_fwd(xn, wf, b, g)
    dyn = to_nhwc(dy)
    dx = ops.conv_dgrad(dyn, wb, g, H, W)
    dw = ops.conv_wgrad(xn, dyn, g)
    return {"fwd": rel(from_nhwc(y, Cout), y_ref.detach()),
            "dgrad": rel(from_nhwc(dx, Cin), x_ref.grad),
            "wgrad": rel(dw, w_ref.grad)}


def norm_case(mode, N, C, H, W, act, residual=False, noise=False, seed=0):
    """The reference runs in float64: torch's fp32 CUDA F.layer_norm backward is itself off by 40 % of the
    gradient range at [32, 64, 256, 256] (32 rows of 4.2 M normalised elements; checked against the
    closed-form float64 gradient in tests/tools/debug_ln.py), so it cannot be the yardstick there."""
    torch.manual_seed(seed)
    x = bf(torch.randn(N, C, H, W, device=DEV) * 2 + 0.5)
    x_ref = x.double().requires_grad_(True)
    gb = lw = lb = None
    params = []
    if mode == "adain":
        gb = (torch.randn(N, 2 * C, device=DEV) * 0.5).requires_grad_(True)
        params = [gb]
        gbd = gb.double()
        y_ref = (1 + gbd[:, :C, None, None]) * F.instance_norm(x_ref) + gbd[:, C:, None, None]
    elif mode == "instance":
        y_ref = F.instance_norm(x_ref)
    else:
        lw = (torch.rand(C, device=DEV) + 0.5).requires_grad_(True)
        lb = torch.randn(C, device=DEV).requires_grad_(True)
        params = [lw, lb]
        # F.layer_norm over (C,H,W) with per-channel affine (norm.py:16-21), written out in float64
        mean = x_ref.mean((1, 2, 3), keepdim=True)
        var = x_ref.var((1, 2, 3), unbiased=False, keepdim=True)
        y_ref = (x_ref - mean) * (var + 1e-5).rsqrt() * lw.double().view(1, C, 1, 1) + lb.double().view(1, C, 1, 1)
    if act == "relu":
        y_ref = F.relu(y_ref)
    elif act == "lrelu":
        y_ref = F.leaky_relu(y_ref, 0.01)
    res_t = nz = None
    if residual:
        res_t = bf(torch.randn(N, C, H, W, device=DEV))
        y_ref = y_ref + res_t
    if noise:
        nz = torch.randn(N, C, H, W, device=DEV)
        y_ref = y_ref + nz
    g = bf(torch.randn(y_ref.shape, device=DEV))
    y_ref.backward(g.double())

    xn = to_nhwc(x)
    stats = ops.plane_stats(xn)
    ss, mr, csum = ops.norm_finalize(stats, mode, H * W, None if gb is None else gb.detach(),
                                     None if lw is None else lw.detach(), None if lb is None else lb.detach(),
                                     csum=True)
    res_pad = None
    if residual:
        res_pad = to_nhwc(F.pad(res_t, (1, 1, 1, 1), mode="reflect"))
    out = ops.affine_act_fwd(xn, ss, act=act, res_pad=res_pad, res_p=1,
                             noise=None if nz is None else nz.permute(0, 2, 3, 1).contiguous(),
                             out_p=1)          # writes the reflection border too
    out_ref = F.pad(y_ref.detach(), (1, 1, 1, 1), mode="reflect")
    r = {"fwd": rel(from_nhwc(out, C), out_ref)}
    dx, dgb, dlw, dlb, dbias = ops.norm_bwd(to_nhwc(g), xn, ss, mr, mode, act=act,
                                            gb=None if gb is None else gb.detach(),
                                            ln_w=None if lw is None else lw.detach(),
                                            csum=csum if mode == "layer" else None, want_dbias=mode == "layer")
    r["dx"] = rel(from_nhwc(dx, C), x_ref.grad)
    if mode == "layer":
        # gradient of a bias added in front of the norm = sum_{n,h,w} dx, from the saved plane sums
        r["dbias"] = rel(dbias, x_ref.grad.sum((0, 2, 3)))
        # accumulate-into form: += into existing buffers
        acc_w, acc_b, acc_d = torch.ones_like(dlw), torch.ones_like(dlb), torch.ones_like(dbias)
        ops.norm_bwd(to_nhwc(g), xn, ss, mr, mode, act=act, ln_w=lw.detach(), csum=csum,
                     dln_into=(acc_w, acc_b), dbias_into=acc_d)
        r["acc_into"] = max(rel(acc_w - 1, dlw), rel(acc_b - 1, dlb), rel(acc_d - 1, dbias))
    if mode == "adain":
        r["dgb"] = rel(dgb, gb.grad)
    if mode == "layer":
        r["dlw"] = rel(dlw, lw.grad)
        r["dlb"] = rel(dlb, lb.grad)
    return r


def pad_case(N, C, H, W, p, seed=0):
    """reflect border / fold / add_interior / act_bwd_fold / image pack + grad."""
    torch.manual_seed(seed)
    x = bf(torch.randn(N, C, H, W, device=DEV))
    x_ref = x.clone().requires_grad_(True)
    xp_ref = F.pad(x_ref, (p,) * 4, mode="reflect")
    gpad = bf(torch.randn_like(xp_ref))
    xp_ref.backward(gpad)
    buf = torch.zeros(N, H + 2 * p, W + 2 * p, C, dtype=torch.bfloat16, device=DEV)
    buf[:, p:H + p, p:W + p, :] = to_nhwc(x)
    ops.reflect_border_(buf, p)
    r = {"border": rel(from_nhwc(buf, C), xp_ref.detach())}
    dense = ops.reflect_fold(to_nhwc(gpad), p)
    r["fold"] = rel(from_nhwc(dense, C), x_ref.grad)
    acc = to_nhwc(gpad).clone()
    ops.add_interior_(acc, to_nhwc(x), p)
    ref = gpad.clone()
    ref[:, :, p:H + p, p:W + p] += x
    r["add_interior"] = rel(from_nhwc(acc, C), ref)
    dy = ops.act_bwd_fold(to_nhwc(gpad), buf, p, "lrelu")
    r["act_bwd_fold"] = rel(from_nhwc(dy, C), x_ref.grad * torch.where(x > 0, 1.0, 0.01))
    # image edge
    img = torch.rand(N, 3, H, W, device=DEV) * 2 - 1
    onehot = torch.eye(4, device=DEV)[torch.arange(N) % 4]
    packed = ops.image_to_nhwc(img, onehot, p)
    full = torch.cat([img, onehot[:, :, None, None].expand(N, 4, H, W)], 1)
    r["image_pack"] = rel(from_nhwc(packed, 7), F.pad(bf(full), (p,) * 4, mode="reflect"))
    img_ref = img.clone().requires_grad_(True)
    F.pad(img_ref, (p,) * 4, mode="reflect").backward(gpad[:, :3])
    gp8 = to_nhwc(gpad[:, :3], 8)
    r["image_grad"] = rel(ops.image_grad_from_nhwc(gp8, 3, H, W, p), img_ref.grad)
    return r


def fused_stats_case(N, Cin, Cout, H, W, k=3, stride=1, transposed=False, bias=True, seed=0):
    """Plane statistics fused into the conv epilogue (mt_conv2d_fwd_stats) vs the separate pass
    over the stored output (mt_plane_stats) and vs torch on that output; the conv output itself must
    be bit-identical to mt_conv2d_fwd.  Shapes with tile overhang exercise the zero-staged rows."""
    torch.manual_seed(seed)
    rp = 0 if transposed else k // 2
    g = ConvGeom(Cin, Cout, k, k, 2 if transposed else stride, 1 if transposed else 0, transposed)
    x = bf(torch.randn(N, Cin, H, W, device=DEV))
    xn = to_nhwc(F.pad(x, (rp,) * 4, mode="reflect") if rp else x)
    w = bf(torch.randn(*((Cin, Cout) if transposed else (Cout, Cin)), k, k, device=DEV) * 0.05)
    b = torch.randn(Cout, device=DEV) if bias else None
    wf, _ = ops.pack_weight(w, g)
    y0 = ops.conv_fwd(xn, wf, b, g)
    y, st = ops.conv_fwd_stats(xn, wf, b, g)
    fused = st.dim() == 4
    tot = st.sum(1) if fused else st
    sep = ops.plane_stats(y).sum(1)
    yf = y.float()
    ref = torch.stack((yf.sum((1, 2)), (yf * yf).sum((1, 2))), -1)
    return {"same_output": 0.0 if torch.equal(y0, y) else 1.0,
            "fused": 0.0 if fused else 1.0,
            "vs_separate": rel(tot, sep) * 1e2,       # both fp32 sums of the same bf16 values: 1e-4
            "vs_torch": rel(tot, ref) * 1e2}


def upsample_case(N, C, H, W, seed=0):
    """nn.Upsample(x2, nearest) forward and its adjoint (UpsampleBlock 'nearest', blocks.py:74-76)."""
    torch.manual_seed(seed)
    x = bf(torch.randn(N, C, H, W, device=DEV)).requires_grad_(True)
    y_ref = F.interpolate(x, scale_factor=2, mode="nearest")
    dy = bf(torch.randn_like(y_ref))
    y_ref.backward(dy)
    y = ops.upsample2_fwd(to_nhwc(x.detach()))
    dx = ops.upsample2_bwd(to_nhwc(dy))
    return {"fwd_exact": 0.0 if torch.equal(from_nhwc(y, C), y_ref.detach()) else 1.0,
            "bwd": rel(from_nhwc(dx, C), x.grad)}


def mask_mul_case(N, C, H, W, seed=0):
    """Dropout(0.5) as a mask multiply (blocks.py:133-134,152,190): exact in bf16 (x * 0 or x * 2)."""
    torch.manual_seed(seed)
    x = to_nhwc(bf(torch.randn(N, C, H, W, device=DEV)))
    m = (torch.rand(x.shape, device=DEV) >= 0.5).to(torch.bfloat16) * 2.0
    y = ops.mul(x, m)
    return {"exact": 0.0 if torch.equal(y, x * m) else 1.0}


def misc_case(N=3, C=72, H=12, W=20, p=1, seed=0):
    """AvgPool2d(2), the DownResnetBlock sum + LeakyReLU, AdaptiveAvgPool2d(1), the RGB head, the
    content regulariser and the fused Adam step, each against plain PyTorch fp32 (blocks.py:113,118-119,
    networks.py:124,248-251, adain_model.py:396-399,58-61)."""
    torch.manual_seed(seed)
    r = {}
    x = bf(torch.randn(N, C, H, W, device=DEV))
    xp = to_nhwc(F.pad(x, (p,) * 4, mode="reflect"))
    # avgpool2 forward / backward-add
    r["avgpool2"] = rel(from_nhwc(ops.avgpool2_fwd(xp, p), C), F.avg_pool2d(x, 2))
    dpo = bf(torch.randn(N, C, H // 2, W // 2, device=DEV))
    base = bf(torch.randn(N, C, H + 2 * p, W + 2 * p, device=DEV))
    acc = to_nhwc(base).clone()
    ops.avgpool2_bwd_add_(to_nhwc(dpo), acc, p)
    ref = base.clone()
    ref[:, :, p:H + p, p:W + p] += 0.25 * F.interpolate(dpo, scale_factor=2, mode="nearest")
    r["avgpool2_bwd_add"] = rel(from_nhwc(acc, C), bf(ref))
    # pool_add_act forward / backward
    c_ = bf(torch.randn(N, C, H, W, device=DEV)).requires_grad_(True)
    d_ = bf(torch.randn(N, C, H // 2, W // 2, device=DEV)).requires_grad_(True)
    y_ref = F.leaky_relu(F.avg_pool2d(c_, 2) + d_, 0.01)
    yp_ref = F.pad(y_ref, (p,) * 4, mode="reflect")
    g = bf(torch.randn_like(yp_ref))
    yp_ref.backward(g)
    out = ops.pool_add_act_fwd(to_nhwc(c_.detach()), to_nhwc(d_.detach()), "lrelu", p)
    ops.reflect_border_(out, p)
    r["pool_add_act_fwd"] = rel(from_nhwc(out, C), yp_ref.detach())
    dd, dc = ops.pool_add_act_bwd(to_nhwc(g), out, p, "lrelu")
    r["pool_add_act_dd"] = rel(from_nhwc(dd, C), d_.grad)
    r["pool_add_act_dc"] = rel(from_nhwc(dc, C), c_.grad)
    # global average pooling
    r["gap_fwd"] = rel(ops.gap_fwd(to_nhwc(x))[:, :C], x.mean((2, 3)))
    dg = torch.randn(N, ops.cs(C), device=DEV)
    r["gap_bwd"] = rel(from_nhwc(ops.gap_bwd(dg, H, W), C), bf(dg[:, :C, None, None].expand(N, C, H, W) / (H * W)))
    # RGB head
    xr = bf(torch.randn(N, C, H, W, device=DEV)).requires_grad_(True)
    w = (torch.randn(C, 3, device=DEV) * 0.2).requires_grad_(True)
    y_ref = torch.tanh(torch.einsum("nchw,co->nohw", xr, w))
    dy = torch.randn_like(y_ref)
    y_ref.backward(dy)
    w2 = torch.cat((w.detach(), w.new_zeros(ops.cs(C) - C, 3))).contiguous()
    y = ops.rgb_head_fwd(to_nhwc(xr.detach()), w2, 3)
    r["rgb_head_fwd"] = rel(y, y_ref.detach())
    dx, dw = ops.rgb_head_bwd(dy, y, to_nhwc(xr.detach()), w2, 3)
    r["rgb_head_dx"] = rel(from_nhwc(dx, C), xr.grad)
    r["rgb_head_dw"] = rel(dw[:C], w.grad)
    # bias gradient = column sums of a dense gradient tensor
    db = ops.bias_grad(to_nhwc(x), C)
    r["bias_grad"] = rel(db, x.sum((0, 2, 3)))
    acc = torch.full((C,), 2.0, device=DEV)
    ops.bias_grad(to_nhwc(x), C, into=acc)
    r["bias_grad_into"] = rel(acc - 2.0, x.sum((0, 2, 3)))
    # content regulariser
    r["sumsq_interior"] = rel(ops.sumsq_interior(xp, p), (x * x).sum())
    sc = torch.tensor(0.37, device=DEV)
    dxp = ops.scale_interior(xp, sc, p)
    ref = torch.zeros_like(F.pad(x, (p,) * 4))
    ref[:, :, p:H + p, p:W + p] = 0.37 * x
    r["scale_interior"] = rel(from_nhwc(dxp, C), ref)
    return r


def dgrad_acc_case(N=3, C=256, H=20, W=24, seed=0):
    """mt_conv2d_dgrad_acc: dx += dgrad (the epilogue's TMA reduce-add store) vs dgrad followed by a bf16 add."""
    torch.manual_seed(seed)
    g = ConvGeom(C, C, 3, 3, 1, 0)
    w = bf(torch.randn(C, C, 3, 3, device=DEV) * 0.05)
    _, wb = ops.pack_weight(w, g)
    dy = to_nhwc(bf(torch.randn(N, C, H, W, device=DEV)))
    base = to_nhwc(bf(torch.randn(N, C, H + 2, W + 2, device=DEV)))
    plain = ops.conv_dgrad(dy, wb, g, H + 2, W + 2)
    ref = (plain.float() + base.float()).to(torch.bfloat16)
    acc = base.clone()
    out = ops.conv_dgrad(dy, wb, g, H + 2, W + 2, add_into=acc)
    return {"same_buffer": 0.0 if out.data_ptr() == acc.data_ptr() else 1.0,
            "exact": 0.0 if torch.equal(out, ref) else rel(out, ref)}


def adam_case(n=10007, steps=3, seed=0):
    """FlatAdam (mt_adam_flat) vs torch.optim.Adam with coupled weight decay (adain_model.py:58-61)."""
    from masterthesis_b200.optim import FlatAdam
    torch.manual_seed(seed)
    p0 = torch.randn(n, device=DEV)
    a = [torch.nn.Parameter(p0[:5000].clone().view(50, 100)), torch.nn.Parameter(p0[5000:].clone())]
    b = [torch.nn.Parameter(p0[:5000].clone().view(50, 100)), torch.nn.Parameter(p0[5000:].clone())]
    oa = FlatAdam(a, lr=1e-3, betas=(0.5, 0.999), weight_decay=1e-4)
    ob = torch.optim.Adam(b, lr=1e-3, betas=(0.5, 0.999), weight_decay=1e-4)
    for s in range(steps):
        oa.zero_grad(); ob.zero_grad()
        for pa, pb in zip(a, b):
            g = torch.randn_like(pb) * (0.1 + s)
            pa.grad.copy_(g); pb.grad = g.clone()
        oa.step(); ob.step()
    torch.cuda.synchronize()
    # max |delta| relative to the total update (the weights themselves barely move in 3 steps)
    num = max((pa.detach() - pb.detach()).abs().max().item() for pa, pb in zip(a, b))
    den = max((pb.detach().flatten() - q).abs().max().item() for pb, q in zip(b, (p0[:5000], p0[5000:])))
    return {"adam_update": num / den}


def big_mean_norm_case(seed=0):
    """InstanceNorm on planes whose mean is ~1e2 standard deviations away from zero (a conv bias in front
    of InstanceNorm, networks.py:29-33): bounds the E[x^2] - E[x]^2 formulation against torch's two-pass
    variance.  Inputs are bf16, so |mean| / std is limited to what bf16 can resolve (~2^7)."""
    torch.manual_seed(seed)
    N, C, H, W = 2, 64, 32, 32
    x = bf(torch.randn(N, C, H, W, device=DEV) + 100.0 * torch.randn(N, C, 1, 1, device=DEV).sign())
    xn = to_nhwc(x)
    ss, mr = ops.norm_finalize(ops.plane_stats(xn), "instance", H * W)
    out = ops.affine_act_fwd(xn, ss)
    return {"fwd": rel(from_nhwc(out, C), F.instance_norm(x))}


def module_forward_case(seed=0):
    """The reference-compatible `forward` of the norm / noise modules on NCHW fp32 tensors (norm.py:16-21,
    29-33, misc.py:22-26) vs torch, with gradients."""
    from masterthesis_b200.models.core.misc import GaussianNoiseLayer
    from masterthesis_b200.models.core.norm import AdaptiveInstanceNorm, LayerNorm
    torch.manual_seed(seed)
    N, C, H, W = 2, 40, 12, 20
    r = {}
    x = bf(torch.randn(N, C, H, W, device=DEV) * 2 + 0.3)
    ln = LayerNorm(C).to(DEV)
    with torch.no_grad():
        ln.weight.copy_(torch.rand(C, 1, 1) + 0.5); ln.bias.copy_(torch.randn(C, 1, 1))
    xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    ya = ln(xa)
    yb = F.layer_norm(xb, xb.shape[1:], ln.weight.detach().expand(C, H, W), ln.bias.detach().expand(C, H, W))
    g = bf(torch.randn_like(yb))
    ya.backward(g); yb.backward(g)
    r["ln_fwd"], r["ln_dx"] = rel(ya.detach(), yb.detach()), rel(xa.grad, xb.grad)
    ad = AdaptiveInstanceNorm(C, 16).to(DEV)
    s_ = torch.randn(N, 16, device=DEV)
    xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    ya = ad(xa, s_)
    h = ad.fc(s_).detach()
    yb = (1 + h[:, :C, None, None]) * F.instance_norm(xb) + h[:, C:, None, None]
    ya.backward(g); yb.backward(g)
    r["adain_fwd"], r["adain_dx"] = rel(ya.detach(), yb.detach()), rel(xa.grad, xb.grad)
    nz = GaussianNoiseLayer().to(DEV)
    draw = torch.randn(N, C, H, W)
    r["noise_fwd"] = rel(nz(x, rng=lambda shape: draw), bf(x + draw.to(DEV)))
    nz.eval()
    r["noise_eval_identity"] = 0.0 if nz(x) is x else 1.0
    return r


def linear_case(N=32, In=12, Out=256, act="relu", seed=0):
    """mt_linear_fwd / _bwd vs F.linear (+ReLU) with autograd, incl. the accumulate-into-.grad form."""
    torch.manual_seed(seed)
    x = torch.randn(N, In, device=DEV, requires_grad=True)
    w = (torch.randn(Out, In, device=DEV) * 0.2).requires_grad_(True)
    b = torch.randn(Out, device=DEV).requires_grad_(True)
    y_ref = F.linear(x, w, b)
    if act == "relu":
        y_ref = F.relu(y_ref)
    dy = torch.randn_like(y_ref)
    y_ref.backward(dy)
    y = ops.linear_fwd(x.detach(), w.detach(), b.detach(), act)
    dx, dw, db = ops.linear_bwd(dy, y if act else None, x.detach(), w.detach(), act)
    r = {"fwd": rel(y, y_ref.detach()), "dx": rel(dx, x.grad), "dw": rel(dw, w.grad), "db": rel(db, b.grad)}
    gw, gb_ = torch.ones_like(w), torch.ones_like(b)
    ops.linear_bwd(dy, y if act else None, x.detach(), w.detach(), act, True, gw, gb_)
    r["into"] = max(rel(gw - 1, w.grad), rel(gb_ - 1, b.grad))
    return r


def conv_z_case(N=4, Cx=256, Cv=256, Cout=512, H=16, W=16, seed=0):
    """functional.ConvZAct (mt_linear_fwd + mt_conv2d_fwd_image_bias; backward through mt_colsum_images and
    mt_linear_bwd) vs the reference's formulation act(conv1x1(cat(h, z tiled), W) + b) (blocks.py:180-190,
    199-205) in fp32 with autograd: output, dh, dz, dW (both halves), db.  Gradients are compared for the
    activation-free form and for ReLU with the mask our forward produced (a pre-activation within bf16 rounding
    of zero may land on either side; a single flipped unit is a 5-10 % max-norm error in dh, which says nothing
    about the kernels)."""
    from masterthesis_b200 import functional as Fn
    torch.manual_seed(seed)
    h = bf(torch.randn(N, Cx, H, W, device=DEV))
    z = torch.randn(N, Cv, device=DEV)
    w = bf(torch.randn(Cout, Cx + Cv, 1, 1, device=DEV) * 0.05)
    b = torch.randn(Cout, device=DEV) * 0.1
    g = ConvGeom(Cx, Cout, 1, 1, 1, 0)
    assert ops.image_bias_ok(g, N, H, W)
    wf, wb = ops.pack_weight(w[:, :Cx].contiguous(), g)
    res = {}
    for act in (None, 'relu'):
        h_ref, z_ref = h.clone().requires_grad_(True), z.clone().requires_grad_(True)
        w_ref, b_ref = w.clone().requires_grad_(True), b.clone().requires_grad_(True)
        cat = torch.cat((h_ref, z_ref[:, :, None, None].expand(N, Cv, H, W)), 1)
        pre = F.conv2d(cat, w_ref, b_ref)
        hn = to_nhwc(h).requires_grad_(True)
        zz = z.clone().requires_grad_(True)
        ww, bb = w.clone().requires_grad_(True), b.clone().requires_grad_(True)
        y = Fn.ConvZAct.apply(hn, zz, ww, bb, wf, wb, g, act, 0)
        yo = from_nhwc(y.detach(), Cout)
        if act is None:
            y_ref = pre
        else:
            res["fwd_relu_vs_fp32"] = rel(yo, F.relu(pre).detach())
            y_ref = pre * (yo > 0).float()
        dy = bf(torch.randn_like(y_ref))
        y_ref.backward(dy)
        y.backward(to_nhwc(dy))
        ops.join_side()
        torch.cuda.synchronize()
        t = "" if act is None else "_relu"
        res.update({"fwd" + t: rel(yo, y_ref.detach()), "dh" + t: rel(from_nhwc(hn.grad, Cx), h_ref.grad),
                    "dz" + t: rel(zz.grad, z_ref.grad), "dw_h" + t: rel(ww.grad[:, :Cx], w_ref.grad[:, :Cx]),
                    "dw_z" + t: rel(ww.grad[:, Cx:], w_ref.grad[:, Cx:]), "db" + t: rel(bb.grad, b_ref.grad)})
    return res


def style_mlp_case(N=32, K=4, C=256, seed=0):
    """functional.StyleMLP (mt_style_mlp_fwd / _bwd) vs the nn.Sequential of the reference decoder
    (networks.py:256-261) followed by the fc of K AdaIN blocks (norm.py:27), with autograd."""
    from masterthesis_b200 import functional as Fn
    torch.manual_seed(seed)
    mlp = torch.nn.Sequential(torch.nn.Linear(12, 256), torch.nn.ReLU(), torch.nn.Linear(256, 256), torch.nn.ReLU(),
                              torch.nn.Linear(256, C)).to(DEV)
    fcs = [torch.nn.Linear(C, 2 * C).to(DEV) for _ in range(K)]
    cz = torch.randn(N, 12, device=DEV, requires_grad=True)
    s_ref = mlp(cz)
    gb_ref = [fc(s_ref) for fc in fcs]
    proj = [torch.randn_like(g) for g in gb_ref]
    ps = torch.randn_like(s_ref) * (0.0 if K else 1.0)       # the AdaIN decoder uses s only through the heads
    ((s_ref * ps).sum() + sum((g * p).sum() for g, p in zip(gb_ref, proj))).backward()
    ref_grads = [p.grad.clone() for p in list(mlp.parameters()) + [q for fc in fcs for q in fc.parameters()]]
    dcz_ref = cz.grad.clone()
    for p in list(mlp.parameters()) + [q for fc in fcs for q in fc.parameters()]:
        p.grad = None
    cz.grad = None
    params = [p for i in (0, 2, 4) for p in (mlp[i].weight, mlp[i].bias)] + [p for fc in fcs for p in (fc.weight, fc.bias)]
    outs = Fn.StyleMLP.apply(cz, K, *params)
    r = {"s": rel(outs[0].detach(), s_ref.detach())}
    if K:
        r["gb"] = max(rel(a.detach(), b_.detach()) for a, b_ in zip(outs[1:], gb_ref))
    if K:
        sum((g * p).sum() for g, p in zip(outs[1:], proj)).backward()     # s itself receives NO gradient
    else:
        (outs[0] * ps).sum().backward()
    mine = [p.grad for p in list(mlp.parameters()) + [q for fc in fcs for q in fc.parameters()]]
    r["dparams"] = max(rel(a, b_) for a, b_ in zip(mine, ref_grads))
    r["dcz"] = rel(cz.grad, dcz_ref)
    return r


def dis_head_case(N=6, S=4, C=1024, nd=4, seed=0):
    """functional.DisHead (mt_dis_head_fwd / _bwd) vs Conv2d(C,1,1,padding=1) and Conv2d(C,nd,S)+GAP."""
    from masterthesis_b200 import functional as Fn
    torch.manual_seed(seed)
    h = bf(torch.randn(N, C, S, S, device=DEV))
    w1 = (torch.randn(1, C, 1, 1, device=DEV) * 0.05).requires_grad_(True)
    w2 = (torch.randn(nd, C, S, S, device=DEV) * 0.02).requires_grad_(True)
    h_ref = h.clone().requires_grad_(True)
    out_ref = F.conv2d(h_ref, w1, padding=1)
    cls_ref = F.conv2d(h_ref, w2).mean((2, 3))
    do, dc = torch.randn_like(out_ref), torch.randn_like(cls_ref)
    ((out_ref * do).sum() + (cls_ref * dc).sum()).backward()
    ref = (w1.grad.clone(), w2.grad.clone())
    w1.grad = w2.grad = None
    hn = to_nhwc(h).requires_grad_(True)
    out, cls = Fn.DisHead.apply(hn, w1, w2)
    r = {"out": rel(out.detach(), out_ref.detach()), "cls": rel(cls.detach(), cls_ref.detach())}
    ((out * do).sum() + (cls * dc).sum()).backward()
    r["dh"] = rel(from_nhwc(hn.grad, C), h_ref.grad)
    r["dw1"], r["dw2"] = rel(w1.grad, ref[0]), rel(w2.grad, ref[1])
    return r


def loss_case(seed=0):
    """mt_loss_bce_logits / mt_loss_l1 / mt_loss_kl_normal: value and gradient vs torch."""
    torch.manual_seed(seed)
    r = {}
    buf = torch.zeros(4, device=DEV)
    x = (torch.randn(16, 1, 6, 6, device=DEV) * 3).requires_grad_(True)
    for name, tgt in (("bce_const1", 1.0), ("bce_const0", 0.0), ("bce_tensor", (torch.rand(16, 1, 6, 6, device=DEV) > 0.5).float())):
        x.grad = None
        t = torch.full_like(x, tgt) if not torch.is_tensor(tgt) else tgt
        ref = F.binary_cross_entropy_with_logits(x, t)
        (ref * 5.0).backward()
        dx = ops.loss_bce_logits(x.detach(), tgt, 5.0, buf[0:1])
        r[name] = max(rel(buf[0], ref.detach()), rel(dx, x.grad))
    a = (torch.rand(4, 3, 64, 64, device=DEV) * 2 - 1).requires_grad_(True)
    b = torch.rand(4, 3, 64, 64, device=DEV) * 2 - 1
    ref = F.l1_loss(a, b) * 10.0
    ref.backward()
    da = ops.loss_l1(a.detach(), b, 10.0, buf[1:2])
    r["l1"] = max(rel(buf[1], ref.detach()), rel(da, a.grad))
    a2 = torch.randn(7, 9, device=DEV).requires_grad_(True)      # odd element count: the scalar tail
    b2 = torch.randn(7, 9, device=DEV)
    ref = F.l1_loss(a2, b2) * 2.0
    ref.backward()
    da2 = ops.loss_l1(a2.detach(), b2, 2.0, buf[1:2])
    r["l1_tail"] = max(rel(buf[1], ref.detach()), rel(da2, a2.grad))
    mu = torch.randn(16, 8, device=DEV).requires_grad_(True)
    lv = (torch.randn(16, 8, device=DEV) * 0.5).requires_grad_(True)
    ref = torch.sum(1 + lv - mu.pow(2) - lv.exp()) * (-0.5)
    (ref * 0.01).backward()
    dmu, dlv = ops.loss_kl_normal(mu.detach(), lv.detach(), 0.01, buf[2:3])
    r["kl"] = max(rel(buf[2], ref.detach()), rel(dmu, mu.grad), rel(dlv, lv.grad))
    return r


CASES = {
    "linear_relu": linear_case,
    "linear_plain_n3": lambda: linear_case(3, 256, 8, None),
    "conv_z_fold": lambda: conv_z_case(),
    "conv_z_fold_n16": lambda: conv_z_case(16, 256, 256, 512, 64, 64),
    "linear_wide": lambda: linear_case(16, 256, 1024, None),
    "style_mlp": style_mlp_case,
    "style_mlp_base": lambda: style_mlp_case(16, 0, 1024),
    "dis_head": dis_head_case,
    "dis_head_s8": lambda: dis_head_case(4, 8, 128, 3),
    "losses": loss_case,
    "module_forwards": module_forward_case,
    "misc_ops": misc_case,
    "misc_ops_c256": lambda: misc_case(2, 256, 16, 16),
    # the warp-tiled RGB head (C / 8 in {4, 8, 16}); 12 x 20 planes: a warp's 32 pixels straddle two images
    "misc_ops_c64": lambda: misc_case(3, 64, 12, 20),
    "misc_ops_c32": lambda: misc_case(2, 32, 10, 14),
    "misc_ops_c128": lambda: misc_case(2, 128, 16, 16),
    # the residual blocks' skip-path gradient added by the data-gradient conv (TMA reduce-add epilogue)
    "dgrad_accumulate": lambda: dgrad_acc_case(),
    "adam_flat": adam_case,
    "in_big_mean": big_mean_norm_case,
    # ---- the launches of the benchmarked configuration (BASELINE config 3: dim 64, 256x256, batch 8):
    # 16 / 32 images per call, i.e. 3.5 - 7 work items per CTA pair of the persistent conv kernel (TMEM
    # double buffer and ring phases wrap across items) and split-K weight gradients over 16 / 32 images
    "hot_n16": lambda: conv_case(16, 256, 256, 64, 64, 3, 1, 1, bias=False),
    "hot_n32": lambda: conv_case(32, 256, 256, 64, 64, 3, 1, 1, bias=False),
    "hot_stats_n16": lambda: fused_stats_case(16, 256, 256, 64, 64, bias=False),
    "hot_stats_n32": lambda: fused_stats_case(32, 256, 256, 64, 64, bias=False),
    "convT_256_128_n16": lambda: convT_case(16, 256, 128, 64, 64),
    "convT_128_64_n16": lambda: convT_case(16, 128, 64, 128, 128),
    "convT_128_64_n32": lambda: convT_case(32, 128, 64, 128, 128),
    "stem7x7_n16": lambda: conv_case(16, 3, 64, 256, 256, 7, 1, 3, stem=True),
    "stem4x4_s2_n16": lambda: conv_case(16, 7, 64, 256, 256, 4, 2, 1, stem=True),
    "stem3x3_s2_n32": lambda: conv_case(32, 3, 64, 256, 256, 3, 2, 1, stem=True),
    "down1_n16": lambda: conv_case(16, 64, 128, 256, 256, 3, 2, 1),
    "down2_n16": lambda: conv_case(16, 128, 256, 128, 128, 3, 2, 1),
    "dis_l2_n32": lambda: conv_case(32, 64, 128, 128, 128, 3, 2, 1, act="lrelu", out_pad=1),
    "dis_l5_n32": lambda: conv_case(32, 512, 1024, 16, 16, 3, 2, 1),
    "dis_l6_n32": lambda: conv_case(32, 1024, 1024, 8, 8, 3, 2, 1),
    "es_conv_n16": lambda: conv_case(16, 64, 128, 128, 128, 3, 1, 1),
    # the other 3x3 stride-1 convs of the style encoder's DownResnetBlocks (blocks.py:108-111): halo kernel with
    # 64-column CTA pairs / two K chunks; and extents that are no multiple of its 8 x 16 pixel tile
    "es_conv64_n16": lambda: conv_case(16, 64, 64, 128, 128, 3, 1, 1),
    "es_conv128_n16": lambda: conv_case(16, 128, 128, 64, 64, 3, 1, 1),
    "halo_ragged_64": lambda: conv_case(3, 64, 64, 36, 20, 3, 1, 1),
    "halo_ragged_128": lambda: conv_case(3, 128, 128, 20, 44, 3, 1, 1, act="relu"),
    "halo_ragged_64_128": lambda: conv_case(2, 64, 128, 52, 28, 3, 1, 1),
    "adain_res_n32": lambda: norm_case("adain", 32, 256, 64, 64, None, residual=True),
    "in_relu_n16": lambda: norm_case("instance", 16, 256, 64, 64, "relu"),
    "ln_relu_n32": lambda: norm_case("layer", 32, 64, 256, 256, "relu"),
    # optional variants: nearest up-sampling tail, 7x7 zero-padded image head, dropout
    "upsample2": lambda: upsample_case(2, 40, 12, 20),
    "mask_mul": lambda: mask_mul_case(2, 24, 16, 16),
    "conv3x3_zpad1": lambda: conv_case(2, 128, 64, 32, 32, 3, 1, 0, zpad=1),
    "conv7x7_zpad3_rgb": lambda: conv_case(2, 64, 3, 32, 32, 7, 1, 0, zpad=3, bias=False),
    "fused_stats_hot": lambda: fused_stats_case(2, 256, 256, 64, 64, bias=False),
    "fused_stats_overhang": lambda: fused_stats_case(3, 64, 136, 24, 40),
    "fused_stats_s2": lambda: fused_stats_case(2, 64, 128, 64, 64, k=3, stride=2),
    "fused_stats_convT": lambda: fused_stats_case(2, 128, 64, 32, 32, transposed=True),
    "fused_stats_convT_overhang": lambda: fused_stats_case(1, 256, 128, 12, 20, transposed=True),
    # the hot shape: ResnetBlock / AdaINResnetBlock conv (blocks.py:131-134,150-151)
    "conv3x3_256": lambda: conv_case(2, 256, 256, 64, 64, 3, 1, 1, bias=False),
    "conv3x3_s2": lambda: conv_case(2, 64, 128, 64, 64, 3, 2, 1),
    "conv3x3_s2_act_pad": lambda: conv_case(2, 64, 128, 32, 32, 3, 2, 1, act="lrelu", out_pad=1),
    "stem7x7": lambda: conv_case(2, 3, 64, 64, 64, 7, 1, 3, stem=True),
    # extents that are no multiple of any pixel tile of the rows-in-K data gradient (narrow tiles: the gather
    # adds more than two tile slots per column)
    "stem7x7_ragged": lambda: conv_case(3, 3, 64, 60, 44, 7, 1, 3, stem=True),
    "stem4x4_s2": lambda: conv_case(2, 7, 64, 64, 64, 4, 2, 1, stem=True),
    "stem3x3_s2": lambda: conv_case(2, 3, 64, 64, 64, 3, 2, 1, stem=True),
    "c3_generic": lambda: conv_case(2, 3, 64, 32, 32, 3, 2, 1),
    "conv1x1": lambda: conv_case(2, 128, 256, 16, 16, 1, 1, 0),
    "dis_tail": lambda: conv_case(3, 1024, 1024, 8, 8, 3, 2, 1),
    "odd_channels": lambda: conv_case(1, 32, 48, 20, 12, 3, 1, 1),
    "conv4x4_s2_zpad": lambda: conv_case(2, 64, 128, 32, 32, 4, 2, 0, zpad=1),
    "conv7x7_s2_256": lambda: conv_case(1, 256, 256, 32, 32, 7, 2, 1),
    "convT_256_128": lambda: convT_case(2, 256, 128, 16, 16),
    "convT_128_64": lambda: convT_case(1, 128, 64, 24, 40, bias=False),
    # odd channel counts of DecoderConcat (stored 56 -> 32): merged-parity columns straddle 64-wide slabs
    # merged-parity stride-2 data gradients: 4x4 (no zero blocks), 3x3 into few channels, odd channels
    # paired-pixel data gradient into <= 8 channels (stride 1): 3x3 with 8 channels, 1x1 with 5
    "conv3x3_c8_pair": lambda: conv_case(2, 8, 16, 16, 12, 3, 1, 1),
    "conv1x1_c5_pair": lambda: conv_case(2, 5, 24, 8, 10, 1, 1, 0),
    "conv4x4_s2_refl": lambda: conv_case(2, 64, 128, 32, 32, 4, 2, 1),
    "conv3x3_s2_odd": lambda: conv_case(2, 44, 52, 20, 12, 3, 2, 1),
    "convT_odd": lambda: convT_case(2, 52, 26, 12, 20),
    "convT_odd_wide": lambda: convT_case(1, 276, 138, 16, 16),
    "in_relu": lambda: norm_case("instance", 2, 256, 32, 32, "relu"),
    "in_res_noise": lambda: norm_case("instance", 2, 128, 16, 24, None, residual=True, noise=True),
    "adain_relu": lambda: norm_case("adain", 3, 256, 16, 16, "relu"),
    "adain_res": lambda: norm_case("adain", 2, 64, 16, 16, None, residual=True),
    "ln_relu": lambda: norm_case("layer", 2, 128, 32, 32, "relu"),
    "in_lrelu_c1024": lambda: norm_case("instance", 2, 1024, 8, 8, "lrelu"),
    "small_1x1_c32": lambda: conv_case(2, 32, 32, 8, 8, 1, 1, 0),
    "small_1x1_c64": lambda: conv_case(2, 64, 64, 8, 8, 1, 1, 0),
    "small_3x3_c32": lambda: conv_case(2, 32, 32, 8, 8, 3, 1, 1),
    "small_1x1_c32_n1": lambda: conv_case(1, 32, 32, 8, 16, 1, 1, 0),
    "pad1": lambda: pad_case(2, 64, 16, 24, 1),
    "pad3": lambda: pad_case(2, 8, 32, 32, 3),
}

TOL = 1e-2
