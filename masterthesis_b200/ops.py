"""Tensor-level wrappers over the C ABI (include/mt_b200.h).

Everything here takes/returns torch CUDA tensors purely as device-memory handles: allocation and
stream come from torch, arithmetic happens in libmt_b200.so.  Activations are NHWC bf16 with the
stored channel count rounded up to 8; a "padded" buffer is [N, H+2p, W+2p, C] with the reflection
border materialised.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass

import torch

from . import _lib
from ._lib import MtConv2d, check

ACT = {None: 0, "none": 0, "relu": 1, "lrelu": 2, "tanh": 3}
NORM = {"instance": 0, "adain": 1, "layer": 2}
LRELU_SLOPE = 0.01  # nn.LeakyReLU default (reference functions.py:34)
EPS = 1e-5

_launches = 0  # number of kernels of OURS launched (bench.py reports it)

# MT_DEBUG_POISON=1: every buffer this module allocates is NaN-filled and every wrapper checks its
# results, so that a read of unwritten memory shows up at the op that consumed it (debug only).
import os as _os
POISON = _os.environ.get("MT_DEBUG_POISON", "0") == "1"


class _T:
    """torch.empty / torch.empty_like that poison their result in debug mode"""

    @staticmethod
    def empty(*a, **k):
        t = torch.empty(*a, **k)
        if POISON and t.is_floating_point():
            t.fill_(float("nan"))
        return t

    @staticmethod
    def empty_like(x):
        t = torch.empty_like(x)
        if POISON and t.is_floating_point():
            t.fill_(float("nan"))
        return t


def _checked(fn):
    if not POISON:
        return fn
    import functools

    @functools.wraps(fn)
    def wrap(*a, **k):
        out = fn(*a, **k)
        outs = out if isinstance(out, tuple) else (out,)
        pad = k.get("out_pad", k.get("out_p", 0))
        if fn.__name__ == "pool_add_act_fwd":
            pad = a[3]
        if pad:
            outs = tuple(o[:, pad:-pad, pad:-pad, :] for o in outs)
        for i, o in enumerate(outs):
            if torch.is_tensor(o) and o.is_floating_point() and not fn.__name__.startswith("empty"):
                if not torch.isfinite(o.float()).all():
                    bad = (~torch.isfinite(o.float())).nonzero()
                    raise RuntimeError(f"[poison] {fn.__name__} output {i} shape {tuple(o.shape)} has "
                                       f"{bad.shape[0]} non-finite values, first at {bad[0].tolist()}, "
                                       f"last at {bad[-1].tolist()}")
        return out
    return wrap


def launches() -> int:
    return _launches


def _count(n=1):
    global _launches
    _launches += n


def _stream():
    return torch.cuda.current_stream().cuda_stream


# Optional per-launch timing of the implicit-GEMM kernels (bench.py's roofline leg): when a list is
# installed here, every conv launch is bracketed by CUDA events on the launching stream and
# (kind, label, algorithmic_flops, start_event, end_event) is appended.
PROFILE = None


class _timed:
    def __init__(self, kind, g, N, Hin, Win):
        self.on = PROFILE is not None
        if self.on:
            Ho, Wo = g.out_hw(Hin, Win)
            px = N * (Hin * Win if g.transposed else Ho * Wo)
            self.flops = 2.0 * px * g.Cin * g.Cout * g.kh * g.kw     # logical channels
            self.kind = kind
            self.label = (f"{'convT' if g.transposed else 'conv'}{g.kh}x{g.kw}s{g.stride}"
                          f"_{g.Cin}->{g.Cout}@{Ho}x{Wo}")

    def __enter__(self):
        if self.on:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *a):
        if self.on:
            self.e1.record()
            PROFILE.append((self.kind, self.label, self.flops, self.e0, self.e1))
        return False


class _timed_bytes:
    """the same for an HBM-bound kernel: (kind, label, algorithmic_bytes, start_event, end_event)"""

    def __init__(self, kind, shape, nbytes):
        self.on = PROFILE is not None
        if self.on:
            self.kind, self.nbytes = kind, float(nbytes)
            self.label = "x".join(str(int(v)) for v in shape)

    def __enter__(self):
        if self.on:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *a):
        if self.on:
            self.e1.record()
            PROFILE.append((self.kind, self.label, self.nbytes, self.e0, self.e1))
        return False


def _p(t):
    return None if t is None else t.data_ptr()


def cs(c: int) -> int:
    """stored channel count"""
    return (c + 7) // 8 * 8


def _require_cuda(t: torch.Tensor):
    if not t.is_cuda:
        raise RuntimeError("masterthesis_b200 has no CPU path: tensors must live on a CUDA device")


def empty_act(N, H, W, C, device, slack=False):
    """bf16 NHWC buffer; with `slack` 256 zeroed elements follow (row-window TMA over-read)."""
    n = N * H * W * C
    if not slack:
        return _T.empty((N, H, W, C), dtype=torch.bfloat16, device=device)
    flat = _T.empty(n + 256, dtype=torch.bfloat16, device=device)
    flat[n:].zero_()
    return flat[:n].view(N, H, W, C)


@dataclass(frozen=True)
class ConvGeom:
    """Static geometry of one Conv2d / ConvTranspose2d call site."""
    Cin: int
    Cout: int
    kh: int
    kw: int
    stride: int = 1
    zpad: int = 0          # logical zero padding (reflection padding lives in the input buffer)
    transposed: bool = False
    stem: bool = False     # input is an image buffer with slack -> row-window mode if eligible
    single: bool = False   # odd spatial extents: no merged-parity / paired-pixel column modes (MtConv2d.single)

    @property
    def Cin_s(self):
        return cs(self.Cin)

    @property
    def Cout_s(self):
        return cs(self.Cout)

    @property
    def merged_convT(self) -> bool:
        """ConvTranspose2d(k3, s2, p1) forward in the merged-parity mode (include/mt_b200.h)"""
        return bool(self.transposed and _lib.lib().mt_convT_merged_mode(self.kh, self.kw, self.zpad))

    @property
    def window_pair(self) -> bool:
        """row-window stem computing two adjacent output pixels per GEMM row (include/mt_b200.h)"""
        return bool(self.window and not self.single and
                    _lib.lib().mt_conv_window_pair_mode(self.Cin_s, self.kw, self.stride))

    @property
    def rowk_dgrad(self) -> bool:
        """many-tap stride-1 conv into <= 4 channels (7x7 image stems): data gradient with the filter rows in K
        and the filter columns in N (include/mt_b200.h).  Takes precedence over pair_s1_dgrad."""
        return bool(not self.transposed and _lib.lib().mt_conv_rowk_dgrad_mode(
            self.kh, self.kw, self.stride, self.zpad, self.Cin, self.Cout_s))

    @property
    def pair_s1_dgrad(self) -> bool:
        """stride-1 conv into <= 8 channels whose data gradient computes pixel pairs (include/mt_b200.h)"""
        return bool(not self.transposed and not self.single and not self.rowk_dgrad and
                    _lib.lib().mt_conv_s1_dgrad_pair_mode(self.kh, self.kw, self.stride, self.zpad, self.Cin_s))

    @property
    def merged_s2_dgrad(self) -> bool:
        """stride-2 conv whose data gradient runs in the merged-parity mode (include/mt_b200.h)"""
        return bool(not self.transposed and not self.single and
                    _lib.lib().mt_conv_s2_dgrad_merged_mode(self.kh, self.kw, self.stride, self.zpad, self.Cin))

    @property
    def window(self) -> bool:
        return (self.stem and not self.transposed and self.zpad == 0 and self.Cin_s <= 16
                and self.Cin_s * self.kw <= 64)

    def out_hw(self, Hin, Win):
        if self.transposed:
            return Hin * 2, Win * 2
        return ((Hin + 2 * self.zpad - self.kh) // self.stride + 1,
                (Win + 2 * self.zpad - self.kw) // self.stride + 1)

    def desc(self, N, Hin, Win, out_pad=0, act=None) -> MtConv2d:
        Ho, Wo = self.out_hw(Hin, Win)
        return MtConv2d(N, Hin, Win, self.Cin, self.Cin_s, Ho, Wo, self.Cout, self.Cout_s,
                        self.kh, self.kw, self.stride, self.zpad, Ho + 2 * out_pad, Wo + 2 * out_pad,
                        out_pad, ACT[act], LRELU_SLOPE, int(self.window), int(self.single))


# ----------------------------------------------------------------------------- weights
@_checked
def pack_weight(w: torch.Tensor, g: ConvGeom):
    """fp32 master weight (OIHW, or IOHW for transposed) -> (w_fwd, w_bwd) bf16 GEMM operands."""
    _require_cuda(w)
    L = _lib.lib()
    w = w.detach().contiguous().float()
    if g.rowk_dgrad:
        # the data-gradient operand is the rows-in-K pack; the forward operand follows the usual rules
        if g.window_pair:
            wf = _T.empty(L.mt_packed_conv_window_pair_elems(g.Cout, g.kh), dtype=torch.bfloat16, device=w.device)
            check(L.mt_pack_conv_window_pair_weight(_p(w), g.Cout, g.Cin, g.kh, g.kw, g.stride, _p(wf), _stream()),
                  "mt_pack_conv_window_pair_weight")
        else:
            wf = _T.empty(L.mt_packed_weight_elems(g.Cout, g.Cin, g.kh, g.kw, int(g.window), 0),
                          dtype=torch.bfloat16, device=w.device)
            check(L.mt_pack_conv_weight(_p(w), g.Cout, g.Cin, g.kh, g.kw, 0, int(g.window), _p(wf), None,
                                        _stream()), "mt_pack_conv_weight")
        wb = _T.empty(L.mt_packed_conv_rowk_elems(g.Cout, g.kh), dtype=torch.bfloat16, device=w.device)
        check(L.mt_pack_conv_rowk_weight(_p(w), g.Cout, g.Cin, g.kh, g.kw, _p(wb), _stream()),
              "mt_pack_conv_rowk_weight")
        _count(2)
        return wf, wb
    if g.window_pair:
        # paired row-window stem: its own forward operand; the data-gradient operand by the usual rules
        wf = _T.empty(L.mt_packed_conv_window_pair_elems(g.Cout, g.kh), dtype=torch.bfloat16, device=w.device)
        check(L.mt_pack_conv_window_pair_weight(_p(w), g.Cout, g.Cin, g.kh, g.kw, g.stride, _p(wf), _stream()),
              "mt_pack_conv_window_pair_weight")
        if g.merged_s2_dgrad:
            wb = _T.empty(L.mt_packed_conv_s2_dgrad_elems(g.Cout, g.Cin, g.kh), dtype=torch.bfloat16, device=w.device)
            check(L.mt_pack_conv_s2_dgrad_weight(_p(w), g.Cout, g.Cin, g.kh, _p(wb), _stream()),
                  "mt_pack_conv_s2_dgrad_weight")
        elif g.pair_s1_dgrad:
            wb = _T.empty(L.mt_packed_conv_s1_dgrad_pair_elems(g.Cout, g.kh), dtype=torch.bfloat16, device=w.device)
            check(L.mt_pack_conv_s1_dgrad_pair_weight(_p(w), g.Cout, g.Cin, g.kh, _p(wb), _stream()),
                  "mt_pack_conv_s1_dgrad_pair_weight")
        else:
            wb = _T.empty(L.mt_packed_weight_elems(g.Cout, g.Cin, g.kh, g.kw, 1, 1), dtype=torch.bfloat16,
                          device=w.device)
            check(L.mt_pack_conv_weight(_p(w), g.Cout, g.Cin, g.kh, g.kw, 0, 1, None, _p(wb), _stream()),
                  "mt_pack_conv_weight")
        _count(2)
        return wf, wb
    if g.pair_s1_dgrad:
        # stride-1 conv into <= 8 channels: forward operand as usual, paired-pixel data-gradient operand
        nf = L.mt_packed_weight_elems(g.Cout, g.Cin, g.kh, g.kw, int(g.window), 0)
        wf = _T.empty(nf, dtype=torch.bfloat16, device=w.device)
        wb = _T.empty(L.mt_packed_conv_s1_dgrad_pair_elems(g.Cout, g.kh), dtype=torch.bfloat16, device=w.device)
        check(L.mt_pack_conv_weight(_p(w), g.Cout, g.Cin, g.kh, g.kw, 0, int(g.window), _p(wf), None,
                                    _stream()), "mt_pack_conv_weight")
        check(L.mt_pack_conv_s1_dgrad_pair_weight(_p(w), g.Cout, g.Cin, g.kh, _p(wb), _stream()),
              "mt_pack_conv_s1_dgrad_pair_weight")
        _count(2)
        return wf, wb
    if g.merged_s2_dgrad:
        # stride-2 conv without zero padding: forward operand as usual, data-gradient operand of the
        # merged-parity mode
        nf = L.mt_packed_weight_elems(g.Cout, g.Cin, g.kh, g.kw, int(g.window), 0)
        wf = _T.empty(nf, dtype=torch.bfloat16, device=w.device)
        wb = _T.empty(L.mt_packed_conv_s2_dgrad_elems(g.Cout, g.Cin, g.kh), dtype=torch.bfloat16, device=w.device)
        check(L.mt_pack_conv_weight(_p(w), g.Cout, g.Cin, g.kh, g.kw, 0, int(g.window), _p(wf), None,
                                    _stream()), "mt_pack_conv_weight")
        check(L.mt_pack_conv_s2_dgrad_weight(_p(w), g.Cout, g.Cin, g.kh, _p(wb), _stream()),
              "mt_pack_conv_s2_dgrad_weight")
        _count(2)
        return wf, wb
    nb = L.mt_packed_weight_elems(g.Cout, g.Cin, g.kh, g.kw, int(g.window), 1)
    wb = _T.empty(nb, dtype=torch.bfloat16, device=w.device)
    if g.merged_convT:
        # ConvTranspose2d(k3, s2, p1): forward operand of the merged-parity mode, [6][2*Cout_s][Cin_s]
        wf = _T.empty(L.mt_packed_convT3x3s2_elems(g.Cout, g.Cin), dtype=torch.bfloat16, device=w.device)
        check(L.mt_pack_convT3x3s2_weight(_p(w), g.Cout, g.Cin, _p(wf), _stream()), "mt_pack_convT3x3s2_weight")
        check(L.mt_pack_conv_weight(_p(w), g.Cout, g.Cin, g.kh, g.kw, 1, 0, None, _p(wb), _stream()),
              "mt_pack_conv_weight")
        _count(2)
        return wf, wb
    nf = L.mt_packed_weight_elems(g.Cout, g.Cin, g.kh, g.kw, int(g.window), 0)
    wf = _T.empty(nf, dtype=torch.bfloat16, device=w.device)
    check(L.mt_pack_conv_weight(_p(w), g.Cout, g.Cin, g.kh, g.kw, int(g.transposed), int(g.window),
                                _p(wf), _p(wb), _stream()), "mt_pack_conv_weight")
    _count()
    return wf, wb


# ----------------------------------------------------------------------------- convolutions
@_checked
def conv_fwd(x, wf, bias, g: ConvGeom, act=None, out_pad=0, out=None):
    """x: [N,Hin,Win,Cin_s] -> [N,Ho+2op,Wo+2op,Cout_s]; border of `out` is NOT filled here."""
    _require_cuda(x)
    N, Hin, Win, C = x.shape
    assert C == g.Cin_s and x.dtype == torch.bfloat16 and x.is_contiguous()
    d = g.desc(N, Hin, Win, out_pad, act)
    if out is None:
        out = empty_act(N, d.out_H, d.out_W, g.Cout_s, x.device)
    fn = _lib.lib().mt_conv_transpose2d_fwd if g.transposed else _lib.lib().mt_conv2d_fwd
    with _timed("fwd", g, N, Hin, Win):
        check(fn(ctypes.byref(d), _p(x), _p(wf), _p(bias), _p(out), _stream()), "conv_fwd")
    _count((2 if g.merged_convT else 4) if g.transposed else 1)
    return out


@_checked
def conv_fwd_stats(x, wf, bias, g: ConvGeom):
    """conv_fwd (no activation, dense output) plus the plane statistics of its stored output,
    [N, Cout_s, 2] fp32 (sum, sum of squares): fused into the conv epilogue where the tile shape
    allows it, else by the separate mt_plane_stats pass."""
    _require_cuda(x)
    N, Hin, Win, C = x.shape
    assert C == g.Cin_s and x.dtype == torch.bfloat16 and x.is_contiguous()
    d = g.desc(N, Hin, Win, 0, None)
    out = empty_act(N, d.out_H, d.out_W, g.Cout_s, x.device)
    L = _lib.lib()
    parts = L.mt_conv_stats_parts(ctypes.byref(d), int(g.transposed))
    stats = _T.empty((N, parts, g.Cout_s, 2), dtype=torch.float32, device=x.device) if parts else None
    fn = L.mt_conv_transpose2d_fwd_stats if g.transposed else L.mt_conv2d_fwd_stats
    with _timed("fwd", g, N, Hin, Win):
        check(fn(ctypes.byref(d), _p(x), _p(wf), _p(bias), _p(out), _p(stats), parts, _stream()),
              "conv_fwd_stats")
    _count((2 if g.merged_convT else 4) if g.transposed else 1)
    if stats is None:
        stats = plane_stats(out)
    return out, stats


@_checked
def conv_dgrad(dy, wb, g: ConvGeom, Hin, Win, add_into=None):
    """dy: dense [N,Ho,Wo,Cout_s] -> dx [N,Hin,Win,Cin_s] (gradient of the padded input buffer).
    add_into: a buffer of dx's shape that the result is ADDED to in place (and returned): the skip-path
    gradient of a residual block, so that no separate add pass runs (stride-1 Conv2d only)."""
    _require_cuda(dy)
    N = dy.shape[0]
    assert dy.is_contiguous() and dy.shape[3] == g.Cout_s
    d = g.desc(N, Hin, Win)
    assert (d.Hout, d.Wout) == tuple(dy.shape[1:3]), (d.Hout, d.Wout, dy.shape)
    if add_into is not None:
        assert not g.transposed and g.stride == 1 and not g.pair_s1_dgrad
        assert tuple(add_into.shape) == (N, Hin, Win, g.Cin_s) and add_into.is_contiguous()
        with _timed("dgrad", g, N, Hin, Win):
            check(_lib.lib().mt_conv2d_dgrad_acc(ctypes.byref(d), _p(dy), _p(wb), _p(add_into), 1, _stream()),
                  "conv_dgrad_acc")
        _count()
        return add_into
    dx = empty_act(N, Hin, Win, g.Cin_s, dy.device)
    if g.rowk_dgrad:
        L = _lib.lib()
        scratch = _T.empty(L.mt_conv_rowk_scratch_elems(ctypes.byref(d)), dtype=torch.float32, device=dy.device)
        with _timed("dgrad", g, N, Hin, Win):
            check(L.mt_conv2d_dgrad_rowk(ctypes.byref(d), _p(dy), _p(wb), _p(scratch), _p(dx), _stream()),
                  "conv_dgrad_rowk")
        _count(2)
        return dx
    fn = _lib.lib().mt_conv_transpose2d_dgrad if g.transposed else _lib.lib().mt_conv2d_dgrad
    with _timed("dgrad", g, N, Hin, Win):
        check(fn(ctypes.byref(d), _p(dy), _p(wb), _p(dx), _stream()), "conv_dgrad")
    _count((2 if g.merged_s2_dgrad else 4) if (g.stride == 2 and not g.transposed) else 1)
    return dx


def _wgrad_launch(x, dy, g: ConvGeom, dw, accumulate, stream):
    N, Hin, Win, _ = x.shape
    L = _lib.lib()
    d = g.desc(N, Hin, Win)
    if g.window_pair:
        n = L.mt_packed_conv_window_pair_elems(g.Cout, g.kh)
    else:
        n = L.mt_packed_wgrad_elems(g.Cout, g.Cin, g.kh, g.kw, int(g.transposed), int(g.window))
    # split-K partial results, one slab per split (plain stores: no zero fill, no atomics), added in split
    # order by the unpack kernel
    splits = (L.mt_conv_transpose2d_wgrad_splits if g.transposed else L.mt_conv2d_wgrad_splits)(ctypes.byref(d))
    packed = _T.empty(splits * n, dtype=torch.float32, device=x.device)
    fn = L.mt_conv_transpose2d_wgrad if g.transposed else L.mt_conv2d_wgrad
    with _timed("wgrad", g, N, Hin, Win):
        check(fn(ctypes.byref(d), _p(x), _p(dy), _p(packed), splits, stream), "conv_wgrad")
    if g.window_pair:
        check(L.mt_unpack_conv_window_pair_wgrad(_p(packed), splits, g.Cout, g.Cin, g.kh, g.kw, g.stride,
                                                 int(accumulate), _p(dw), stream), "unpack_wgrad")
    else:
        check(L.mt_unpack_conv_wgrad(_p(packed), splits, g.Cout, g.Cin, g.kh, g.kw, int(g.transposed),
                                     int(g.window), int(accumulate), _p(dw), stream), "unpack_wgrad")
    _count(2)
    return packed


@_checked
def conv_wgrad(x, dy, g: ConvGeom):
    """-> fp32 weight gradient in the master layout (OIHW, or IOHW for transposed)."""
    _require_cuda(x)
    shape = (g.Cin, g.Cout, g.kh, g.kw) if g.transposed else (g.Cout, g.Cin, g.kh, g.kw)
    dw = _T.empty(shape, dtype=torch.float32, device=x.device)
    _wgrad_launch(x, dy, g, dw, False, _stream())
    return dw


# Weight gradients off the critical path.  Nothing in a backward pass consumes a weight gradient
# before the optimizer does, while the data gradient is what the next backward node waits for: the
# wgrad kernel (tensor-bound) therefore runs on a side stream, next to the HBM-bound norm-backward
# kernels of the main stream, and adds its result straight into the parameter's .grad (a view of the
# optimizer's flat bucket).  The main stream joins before anything reads the bucket (join_side(),
# called after every .backward()).  Operands are kept alive on the host side until the join: no
# record_stream(), so the scheme is the same under CUDA-graph capture (fork / join become graph edges).
SIDE_WGRAD = _os.environ.get("MT_WGRAD_SIDE", "1") == "1"
_side_streams = {}     # device -> THE side stream: every wgrad of the process runs on it, in order,
                       # so that two backward streams adding into the same .grad cannot interleave
_side_keep = {}        # (device, origin stream handle) -> [(x, dy, packed)] alive until that origin joined


def _side_key(device):
    cur = torch.cuda.current_stream(device)
    return (device, cur.cuda_stream), cur


def conv_wgrad_into(x, dy, g: ConvGeom, grad):
    """grad (fp32, master layout, contiguous) += dL/dW, asynchronously on the side stream."""
    _require_cuda(x)
    assert grad.is_contiguous() and grad.dtype == torch.float32
    if PROFILE is not None or POISON or not SIDE_WGRAD:
        _wgrad_launch(x, dy, g, grad, True, _stream())
        return
    key, cur = _side_key(x.device)
    side = _side_streams.get(x.device)
    if side is None:
        side = _side_streams[x.device] = torch.cuda.Stream(device=x.device)
    side.wait_stream(cur)
    with torch.cuda.stream(side):
        packed = _wgrad_launch(x, dy, g, grad, True, side.cuda_stream)
    _side_keep.setdefault(key, []).append((x, dy, packed))


def join_side(all_streams=False):
    """Make the current stream wait for the side-stream weight gradients forked from it; with
    `all_streams` also for those forked by backward nodes that ran on other streams (a network whose
    forward was issued on an auxiliary stream has its backward there)."""
    if not _side_keep:
        return
    key, cur = _side_key(torch.device("cuda", torch.cuda.current_device()))
    waited = False
    for k, pending in _side_keep.items():
        if pending and (all_streams or k == key):
            if not waited:
                cur.wait_stream(_side_streams[k[0]])
                waited = True
            pending.clear()


# ----------------------------------------------------------------------------- layout
@_checked
def image_to_nhwc(img, onehot, pad, Cs=None):
    _require_cuda(img)
    N, Ci, H, W = img.shape
    nd = 0 if onehot is None else onehot.shape[1]
    Cs = Cs or cs(Ci + nd)
    out = empty_act(N, H + 2 * pad, W + 2 * pad, Cs, img.device, slack=True)
    img = img.contiguous().float()
    oh = None if onehot is None else onehot.contiguous().float()
    check(_lib.lib().mt_image_to_nhwc(_p(img), _p(oh), N, Ci, nd, H, W, pad, Cs, _p(out), _stream()),
          "image_to_nhwc")
    _count()
    return out


@_checked
def image_grad_from_nhwc(dpad, Ci, H, W, pad):
    N, _, _, Cs = dpad.shape
    dimg = _T.empty((N, Ci, H, W), dtype=torch.float32, device=dpad.device)
    check(_lib.lib().mt_image_grad_from_nhwc(_p(dpad), N, Ci, H, W, pad, Cs, _p(dimg), _stream()),
          "image_grad_from_nhwc")
    _count()
    return dimg


@_checked
def reflect_border_(buf, pad):
    if pad == 0:
        return buf
    N, Hp, Wp, C = buf.shape
    check(_lib.lib().mt_reflect_border(_p(buf), N, Hp - 2 * pad, Wp - 2 * pad, C, pad, _stream()),
          "reflect_border")
    _count()
    return buf


@_checked
def reflect_fold(dpad, pad):
    N, Hp, Wp, C = dpad.shape
    if pad == 0:
        return dpad
    H, W = Hp - 2 * pad, Wp - 2 * pad
    dense = empty_act(N, H, W, C, dpad.device)
    check(_lib.lib().mt_reflect_fold(_p(dpad), N, H, W, C, pad, _p(dense), _stream()), "reflect_fold")
    _count()
    return dense


def add_interior_(dpad, dense, pad):
    N, H, W, C = dense.shape
    check(_lib.lib().mt_add_interior(_p(dpad), _p(dense), N, H, W, C, pad, _stream()), "add_interior")
    _count()
    return dpad


# ----------------------------------------------------------------------------- norms
@_checked
def plane_stats(x):
    """[N, parts, C, 2] per-chunk (sum, sum of squares) of x [N,H,W,C]; norm_finalize adds the parts."""
    N, H, W, C = x.shape
    L = _lib.lib()
    parts = L.mt_plane_stats_parts(N, H * W, C)
    stats = _T.empty((N, parts, C, 2), dtype=torch.float32, device=x.device)
    check(L.mt_plane_stats(_p(x), N, H * W, C, _p(stats), parts, _stream()), "plane_stats")
    _count()
    return stats


@_checked
def norm_finalize(stats, mode, HW, gb=None, ln_w=None, ln_b=None, C_valid=None, csum=False):
    """-> (ss, mr) [, csum [N, C] fp32: the per-channel plane sums (kept for the bias gradient of a biased
    convolution in front of a LayerNorm)]"""
    parts = stats.shape[1] if stats.dim() == 4 else 1      # per-tile partial sums of a fused conv
    N, C = stats.shape[0], stats.shape[-2]
    ss = _T.empty((N, C, 2), dtype=torch.float32, device=stats.device)
    mr = _T.empty_like(ss)
    cs_ = _T.empty((N, C), dtype=torch.float32, device=stats.device) if csum else None
    check(_lib.lib().mt_norm_finalize(_p(stats), parts, NORM[mode], N, C, C_valid or C, HW, EPS, _p(gb), _p(ln_w), _p(ln_b),
                                      _p(ss), _p(mr), _p(cs_), _stream()), "norm_finalize")
    _count()
    return (ss, mr, cs_) if csum else (ss, mr)


@_checked
def affine_act_fwd(x, ss, act=None, res_pad=None, res_p=0, noise=None, out_p=0, out=None):
    N, H, W, C = x.shape
    if out is None:
        out = empty_act(N, H + 2 * out_p, W + 2 * out_p, C, x.device)
    # algorithmic bytes: one bf16 read + one bf16 write per element (+ a bf16 residual read,
    # + an fp32 noise read)
    nb = x.numel() * (4 + (2 if res_pad is not None else 0) + (4 if noise is not None else 0))
    with _timed_bytes("hbm.norm_apply_fwd", x.shape, nb):
        check(_lib.lib().mt_affine_act_fwd(_p(x), _p(ss), N, H, W, C, ACT[act], LRELU_SLOPE, _p(res_pad),
                                           res_p, _p(noise), _p(out), out_p, _stream()), "affine_act_fwd")
    _count()
    return out


@_checked
def norm_bwd(g, x, ss, mr, mode, act=None, gb=None, ln_w=None, C_valid=None, g_pad=0, csum=None,
             dln_into=None, dbias_into=None, want_dbias=False):
    """x dense [N,H,W,C]; g dense, or (g_pad > 0) the consumer's padded gradient buffer
    [N,H+2g_pad,W+2g_pad,C] that is folded (adjoint of the reflection border) on the fly.
    LayerNorm affine gradients are added into `dln_into` = (weight.grad, bias.grad) when given, else
    returned; the gradient of the bias of the convolution in front of a LayerNorm (`csum` from the forward
    finalize) is added into `dbias_into` when given, returned when `want_dbias`.
    Returns (dx, dgb | None, dln_w | None, dln_b | None, dbias | None)."""
    N, H, W, C = x.shape
    assert tuple(g.shape) == (N, H + 2 * g_pad, W + 2 * g_pad, C) and g.is_contiguous()
    L = _lib.lib()
    dev = x.device
    Cv = C_valid or C
    parts = L.mt_norm_bwd_parts(N, H, W, C, ACT[act], g_pad)
    sums = _T.empty((N, parts, C, 2), dtype=torch.float32, device=dev)
    with _timed_bytes("hbm.norm_bwd_reduce", x.shape, x.numel() * 4):        # dy + x
        check(L.mt_norm_bwd_reduce(_p(g), g_pad, _p(x), _p(ss), N, H, W, C, ACT[act], LRELU_SLOPE, _p(sums),
                                   parts, _stream()), "norm_bwd_reduce")
    coef = _T.empty((N, C, 3), dtype=torch.float32, device=dev)
    dgb = _T.empty((N, 2 * C), dtype=torch.float32, device=dev) if mode == "adain" else None
    need_ln = mode == "layer" and ln_w is not None
    need_db = mode == "layer" and csum is not None and (dbias_into is not None or want_dbias)
    red = _T.empty((N, C, 2), dtype=torch.float32, device=dev) if (need_ln or need_db) else None
    check(L.mt_norm_bwd_finalize(_p(sums), parts, _p(mr), NORM[mode], N, C, Cv, H * W, _p(gb), _p(ln_w), _p(coef),
                                 _p(dgb), _p(red), _stream()), "norm_bwd_finalize")
    dlw = dlb = dbias = None
    if red is not None:
        # the accumulating form needs every requested output to be a .grad buffer; otherwise fresh tensors
        acc = (not need_ln or dln_into is not None) and (not need_db or dbias_into is not None)
        if need_ln:
            dlw, dlb = dln_into if acc else (_T.empty(Cv, dtype=torch.float32, device=dev),
                                             _T.empty(Cv, dtype=torch.float32, device=dev))
        if need_db:
            dbias = dbias_into if acc else _T.empty(Cv, dtype=torch.float32, device=dev)
        check(L.mt_norm_bwd_params(_p(red), _p(coef), _p(csum), N, C, Cv, H * W, _p(dlw), _p(dlb), _p(dbias),
                                   int(acc), _stream()), "norm_bwd_params")
        _count()
        if acc:
            dlw = dlb = dbias = None
    dx = _T.empty_like(x)
    with _timed_bytes("hbm.norm_bwd_apply", x.shape, x.numel() * 6):         # dy + x + dx
        check(L.mt_norm_bwd_apply(_p(g), g_pad, _p(x), _p(ss), _p(coef), N, H, W, C, ACT[act], LRELU_SLOPE,
                                  _p(dx), _stream()), "norm_bwd_apply")
    _count(3)
    return dx, dgb, dlw, dlb, dbias


@_checked
def act_bwd_fold(dpad, out_pad, pad, act):
    N, Hp, Wp, C = out_pad.shape
    H, W = Hp - 2 * pad, Wp - 2 * pad
    dy = empty_act(N, H, W, C, dpad.device)
    check(_lib.lib().mt_act_bwd_fold(_p(dpad), _p(out_pad), N, H, W, C, pad, ACT[act], LRELU_SLOPE,
                                     _p(dy), _stream()), "act_bwd_fold")
    _count()
    return dy


# ----------------------------------------------------------------------------- optional variants
@_checked
def upsample2_fwd(x):
    """nn.Upsample(scale_factor=2, mode='nearest') on dense NHWC bf16 (blocks.py:75)."""
    _require_cuda(x)
    N, H, W, C = x.shape
    assert x.is_contiguous() and x.dtype == torch.bfloat16
    out = empty_act(N, 2 * H, 2 * W, C, x.device)
    check(_lib.lib().mt_upsample2_fwd(_p(x), N, H, W, C, _p(out), _stream()), "upsample2_fwd")
    _count()
    return out


@_checked
def upsample2_bwd(dy):
    _require_cuda(dy)
    N, Ho, Wo, C = dy.shape
    assert dy.is_contiguous() and Ho % 2 == 0 and Wo % 2 == 0
    dx = empty_act(N, Ho // 2, Wo // 2, C, dy.device)
    check(_lib.lib().mt_upsample2_bwd(_p(dy), N, Ho // 2, Wo // 2, C, _p(dx), _stream()), "upsample2_bwd")
    _count()
    return dx


@_checked
def mul(x, m):
    """x * m on bf16 tensors of the same shape (dropout mask; its own adjoint)."""
    _require_cuda(x)
    assert x.shape == m.shape and x.is_contiguous() and m.is_contiguous()
    assert x.dtype == torch.bfloat16 and m.dtype == torch.bfloat16
    out = _T.empty_like(x)
    check(_lib.lib().mt_mul(_p(x), _p(m), x.numel(), _p(out), _stream()), "mul")
    _count()
    return out


# ----------------------------------------------------------------------------- pooling / heads
@_checked
def avgpool2_fwd(x_pad, pad):
    N, Hp, Wp, C = x_pad.shape
    H, W = Hp - 2 * pad, Wp - 2 * pad
    out = empty_act(N, H // 2, W // 2, C, x_pad.device)
    check(_lib.lib().mt_avgpool2_fwd(_p(x_pad), pad, N, H, W, C, _p(out), _stream()), "avgpool2_fwd")
    _count()
    return out


def avgpool2_bwd_add_(dpooled, dpad, pad):
    N, Hp, Wp, C = dpad.shape
    check(_lib.lib().mt_avgpool2_bwd_add(_p(dpooled), _p(dpad), pad, N, Hp - 2 * pad, Wp - 2 * pad, C,
                                         _stream()), "avgpool2_bwd_add")
    _count()
    return dpad


@_checked
def pool_add_act_fwd(c, d, act, out_p):
    N, H, W, C = c.shape
    out = empty_act(N, H // 2 + 2 * out_p, W // 2 + 2 * out_p, C, c.device)
    check(_lib.lib().mt_pool_add_act_fwd(_p(c), _p(d), N, H, W, C, ACT[act], LRELU_SLOPE, _p(out), out_p,
                                         _stream()), "pool_add_act_fwd")
    _count()
    return out


@_checked
def pool_add_act_bwd(dout_pad, out_pad, out_p, act):
    N, Hp, Wp, C = out_pad.shape
    Ho, Wo = Hp - 2 * out_p, Wp - 2 * out_p
    dd = empty_act(N, Ho, Wo, C, out_pad.device)
    dc = empty_act(N, Ho * 2, Wo * 2, C, out_pad.device)
    check(_lib.lib().mt_pool_add_act_bwd(_p(dout_pad), _p(out_pad), N, Ho * 2, Wo * 2, C, out_p, ACT[act],
                                         LRELU_SLOPE, _p(dd), _p(dc), _stream()), "pool_add_act_bwd")
    _count()
    return dd, dc


@_checked
def gap_fwd(x):
    N, H, W, C = x.shape
    out = _T.empty((N, C), dtype=torch.float32, device=x.device)
    check(_lib.lib().mt_gap_fwd(_p(x), N, H * W, C, _p(out), _stream()), "gap_fwd")
    _count()
    return out


@_checked
def gap_bwd(dout, H, W):
    N, C = dout.shape
    dx = empty_act(N, H, W, C, dout.device)
    check(_lib.lib().mt_gap_bwd(_p(dout.contiguous()), N, H * W, C, _p(dx), _stream()), "gap_bwd")
    _count()
    return dx


@_checked
def rgb_head_fwd(x, w, Co):
    """x [N,H,W,C] bf16 dense, w fp32 [C, Co] -> NCHW fp32 tanh image."""
    N, H, W, C = x.shape
    y = _T.empty((N, Co, H, W), dtype=torch.float32, device=x.device)
    check(_lib.lib().mt_rgb_head_fwd(_p(x), _p(w), N, H * W, C, Co, _p(y), _stream()), "rgb_head_fwd")
    _count()
    return y


@_checked
def rgb_head_bwd(dy, y, x, w, Co):
    N, H, W, C = x.shape
    dx = _T.empty_like(x)
    dw = _T.empty((C, Co), dtype=torch.float32, device=x.device)
    L = _lib.lib()
    parts = L.mt_rgb_head_bwd_parts(N, H * W, C)
    scratch = _T.empty((parts, C, 4), dtype=torch.float32, device=x.device)
    check(L.mt_rgb_head_bwd(_p(dy.contiguous()), _p(y), _p(x), _p(w), N, H * W, C, Co, _p(dx),
                            _p(dw), _p(scratch), parts, _stream()), "rgb_head_bwd")
    _count(2)
    return dx, dw


@_checked
def sumsq_interior(x_pad, pad, out=None):
    N, Hp, Wp, C = x_pad.shape
    if out is None:
        out = _T.empty((), dtype=torch.float32, device=x_pad.device)
    L = _lib.lib()
    parts = L.mt_sumsq_interior_parts(N, Hp - 2 * pad, Wp - 2 * pad, C)
    scratch = _T.empty(parts, dtype=torch.float32, device=x_pad.device)
    check(L.mt_sumsq_interior(_p(x_pad), N, Hp - 2 * pad, Wp - 2 * pad, C, pad, _p(out), _p(scratch), parts,
                              _stream()), "sumsq_interior")
    _count(2)
    return out


@_checked
def scale_interior(x_pad, scale, pad):
    """scale: fp32 0-dim device tensor."""
    N, Hp, Wp, C = x_pad.shape
    dx = _T.empty_like(x_pad)
    check(_lib.lib().mt_scale_interior(_p(x_pad), _p(scale), N, Hp - 2 * pad, Wp - 2 * pad, C, pad, _p(dx),
                                       _stream()), "scale_interior")
    _count()
    return dx


@_checked
def conv_fwd_image_bias(x, wf, bias_n, g: ConvGeom, act=None, out_pad=0):
    """conv_fwd with a per-image bias bias_n [N, Cout] fp32 (mt_conv2d_fwd_image_bias); raises _lib.MtError
    -1014 when a pixel tile would span images (tiny planes)"""
    _require_cuda(x)
    N, Hin, Win, C = x.shape
    assert C == g.Cin_s and x.dtype == torch.bfloat16 and x.is_contiguous()
    assert tuple(bias_n.shape) == (N, g.Cout)
    d = g.desc(N, Hin, Win, out_pad, act)
    out = empty_act(N, d.out_H, d.out_W, g.Cout_s, x.device)
    check(_lib.lib().mt_conv2d_fwd_image_bias(ctypes.byref(d), _p(x), _p(wf), _p(_f32c(bias_n)), _p(out), _stream()),
          "conv_fwd_image_bias")
    _count()
    return out


def image_bias_ok(g: ConvGeom, N, Hin, Win):
    """whether mt_conv2d_fwd_image_bias can run this launch (one image per pixel tile)"""
    d = g.desc(N, Hin, Win)
    return bool(_lib.lib().mt_conv2d_image_bias_ok(ctypes.byref(d)))


def colsum_images(dy, C):
    """per-image column sums of a dense bf16 tensor [N,H,W,C_s] -> fp32 [N, C] (mt_plane_stats + mt_colsum_images)"""
    st = plane_stats(dy)
    N, parts, Cs, _ = st.shape
    out = _T.empty((N, C), dtype=torch.float32, device=dy.device)
    check(_lib.lib().mt_colsum_images(_p(st), N, parts, Cs, C, _p(out), _stream()), "colsum_images")
    _count()
    return out


def bias_grad(dy, C, into=None):
    """sum over (n, h, w) of a dense bf16 tensor [N,H,W,C_s] -> its first C channels (the bias gradient of
    the convolution that produced the tensor dy is the gradient of, blocks.py:36): per-chunk column sums
    (mt_plane_stats) added in a fixed order.  `into` (the bias's .grad, fp32 [C]): += in place, returns
    None; else a fresh fp32 [C] tensor."""
    st = plane_stats(dy)
    N, parts, Cs, _ = st.shape
    out = into if into is not None else _T.empty(C, dtype=torch.float32, device=dy.device)
    check(_lib.lib().mt_colsum_into(_p(st), N * parts, Cs, C, _p(out), int(into is not None), _stream()),
          "colsum_into")
    _count()
    return None if into is not None else out


# ----------------------------------------------------------------------------- small fp32 layers, heads, losses
def _ptr_array(tensors):
    """host array of device pointers (the `const float* const*` arguments of the C ABI)"""
    arr = (ctypes.c_void_p * len(tensors))(*[None if t is None else t.data_ptr() for t in tensors])
    return arr


def _f32c(t):
    assert t.dtype == torch.float32 and t.is_contiguous(), (t.dtype, t.is_contiguous())
    return t


@_checked
def linear_fwd(x, w, b=None, act=None):
    """nn.Linear (+ReLU): x [N, In] fp32, w [Out, In], b [Out] -> [N, Out]"""
    _require_cuda(x)
    N, In = x.shape
    Out = w.shape[0]
    y = _T.empty((N, Out), dtype=torch.float32, device=x.device)
    check(_lib.lib().mt_linear_fwd(_p(_f32c(x)), _p(_f32c(w)), _p(b), N, In, Out, ACT[act], _p(y), _stream()),
          "linear_fwd")
    _count()
    return y


def linear_bwd(dy, y, x, w, act=None, need_dx=True, dw_into=None, db_into=None, has_bias=True, need_params=True):
    """-> (dx | None, dw | None, db | None); dw / db are None when added into the given .grad buffers
    (both buffers must then be given, or the layer has no bias) or when not needed"""
    N, In = x.shape
    Out = w.shape[0]
    dev = x.device
    acc = need_params and dw_into is not None and (db_into is not None or not has_bias)
    dx = _T.empty((N, In), dtype=torch.float32, device=dev) if need_dx else None
    dw = db = None
    if need_params:
        dw = dw_into if acc else _T.empty((Out, In), dtype=torch.float32, device=dev)
        db = (db_into if acc else _T.empty((Out,), dtype=torch.float32, device=dev)) if has_bias else None
    check(_lib.lib().mt_linear_bwd(_p(_f32c(dy)), _p(y), _p(_f32c(x)), _p(_f32c(w)), N, In, Out, ACT[act], _p(dx),
                                   _p(dw), _p(db), int(acc), 0, _stream()), "linear_bwd")
    _count()
    return dx, (None if acc else dw), (None if acc or not has_bias else db)


def style_mlp_fwd(cz, mlp, heads):
    """mlp: [(w, b)] x 3, heads: [(w, b)] x K -> (h1, h2, s, [gb_k])   (mt_style_mlp_fwd)"""
    _require_cuda(cz)
    N, in_dim = cz.shape
    hid, style_dim = mlp[0][0].shape[0], mlp[2][0].shape[0]
    K = len(heads)
    head_out = heads[0][0].shape[0] if K else 0
    dev = cz.device
    h1 = _T.empty((N, hid), dtype=torch.float32, device=dev)
    h2 = _T.empty((N, hid), dtype=torch.float32, device=dev)
    s = _T.empty((N, style_dim), dtype=torch.float32, device=dev)
    gb = [_T.empty((N, head_out), dtype=torch.float32, device=dev) for _ in range(K)]
    check(_lib.lib().mt_style_mlp_fwd(
        _p(_f32c(cz)), N, in_dim, hid, style_dim, _ptr_array([_f32c(w) for w, _ in mlp]),
        _ptr_array([b for _, b in mlp]), K, head_out, _ptr_array([_f32c(w) for w, _ in heads]),
        _ptr_array([b for _, b in heads]), _p(h1), _p(h2), _p(s), _ptr_array(gb), _stream()), "style_mlp_fwd")
    _count(3 + K)
    return h1, h2, s, gb


def style_mlp_bwd(cz, h1, h2, s, mlp_w, head_w, dgb, ds_in, grads_into, need_dcz):
    """dgb: list of K tensors or None; ds_in: [N, style_dim] or None.  grads_into: None, or
    (d_mlp_w[3], d_mlp_b[3], d_head_w[K], d_head_b[K]) .grad buffers to add into.
    -> (dcz | None, d_mlp_w, d_mlp_b, d_head_w, d_head_b) with the gradient lists None when accumulated."""
    N, in_dim = cz.shape
    hid, style_dim = mlp_w[0].shape[0], mlp_w[2].shape[0]
    K = len(head_w)
    head_out = head_w[0].shape[0] if K else 0
    dev = cz.device
    acc = grads_into is not None
    if acc:
        d_mw, d_mb, d_hw, d_hb = grads_into
    else:
        d_mw = [_T.empty(tuple(w.shape), dtype=torch.float32, device=dev) for w in mlp_w]
        d_mb = [_T.empty((w.shape[0],), dtype=torch.float32, device=dev) for w in mlp_w]
        # a head without incoming gradient keeps a zero gradient
        d_hw = [(_T.empty if dgb[k] is not None else torch.zeros)(tuple(w.shape), dtype=torch.float32, device=dev)
                for k, w in enumerate(head_w)]
        d_hb = [(_T.empty if dgb[k] is not None else torch.zeros)((w.shape[0],), dtype=torch.float32, device=dev)
                for k, w in enumerate(head_w)]
    ds = _T.empty((N, style_dim), dtype=torch.float32, device=dev)
    dh2 = _T.empty((N, hid), dtype=torch.float32, device=dev)
    dh1 = _T.empty((N, hid), dtype=torch.float32, device=dev)
    dcz = _T.empty((N, in_dim), dtype=torch.float32, device=dev) if need_dcz else None
    dgb = [None if g is None else _f32c(g.contiguous()) for g in (dgb or [])]
    check(_lib.lib().mt_style_mlp_bwd(
        _p(cz), _p(h1), _p(h2), _p(s), N, in_dim, hid, style_dim, _ptr_array(list(mlp_w)), K, head_out,
        _ptr_array(list(head_w)), _ptr_array(dgb), _p(None if ds_in is None else _f32c(ds_in.contiguous())),
        _ptr_array(d_mw), _ptr_array(d_mb), _ptr_array(d_hw), _ptr_array(d_hb), _p(ds), _p(dh2), _p(dh1), _p(dcz),
        int(acc), _stream()), "style_mlp_bwd")
    _count(3 + K)
    if acc:
        return dcz, None, None, None, None
    return dcz, d_mw, d_mb, d_hw, d_hb


@_checked
def dis_head_fwd(h, w1, w2, C, nd):
    """h [N,S,S,Cs] bf16 dense -> (out [N,1,S+2,S+2] fp32, cls [N,nd] fp32)   (networks.py:373-383)"""
    N, S, S2, Cs = h.shape
    assert S == S2 and h.is_contiguous() and tuple(w2.shape) == (nd, C, S, S), (h.shape, w2.shape)
    out = _T.empty((N, 1, S + 2, S + 2), dtype=torch.float32, device=h.device)
    cls = _T.empty((N, nd), dtype=torch.float32, device=h.device)
    check(_lib.lib().mt_dis_head_fwd(_p(h), _p(_f32c(w1)), _p(_f32c(w2)), N, S, Cs, C, nd, _p(out), _p(cls),
                                     _stream()), "dis_head_fwd")
    _count()
    return out, cls


def dis_head_bwd(dout, dcls, h, w1, w2, C, nd, cls_first=0, need_dh=True, dw_into=None, need_dw=True):
    """-> (dh | None, dw1 | None, dw2 | None); dw_into = (w1.grad, w2.grad) adds in place"""
    N, S, _, Cs = h.shape
    dev = h.device
    dh = _T.empty_like(h) if need_dh else None
    acc = dw_into is not None
    dw1 = dw2 = None
    if need_dw:
        dw1, dw2 = dw_into if acc else (_T.empty((C,), dtype=torch.float32, device=dev),
                                        _T.empty(tuple(w2.shape), dtype=torch.float32, device=dev))
    check(_lib.lib().mt_dis_head_bwd(_p(_f32c(dout)), _p(_f32c(dcls)), _p(h), _p(w1), _p(w2), N, S, Cs, C, nd,
                                     int(cls_first), _p(dh), _p(dw1), _p(dw2), int(acc), _stream()), "dis_head_bwd")
    _count()
    return dh, (None if acc else dw1), (None if acc else dw2)


def loss_bce_logits(x, target, gscale, loss_out, want_grad=True, dx_out=None):
    """mean BCE-with-logits of x (fp32, any shape, contiguous) against `target` (same-shape tensor or a python
    constant); loss_out: a 1-element fp32 device view that receives the mean.  -> dx = gscale * dloss/dx
    (written into `dx_out`, a contiguous view of x's shape, when given)"""
    n = x.numel()
    t_ptr, t_const = (_p(_f32c(target)), 0.0) if torch.is_tensor(target) else (None, float(target))
    dx = (dx_out if dx_out is not None else _T.empty_like(x)) if want_grad else None
    assert dx is None or (dx.is_contiguous() and dx.numel() == n)
    check(_lib.lib().mt_loss_bce_logits(_p(_f32c(x)), t_ptr, t_const, n, float(gscale), _p(loss_out), _p(dx),
                                        _stream()), "loss_bce_logits")
    _count()
    return dx


def loss_l1(a, b, weight, loss_out, want_grad=True):
    """loss_out[0] = weight * mean|a - b|; -> da = weight / n * sign(a - b)"""
    n = a.numel()
    assert b.numel() == n
    L = _lib.lib()
    parts = L.mt_loss_l1_parts(n)
    scratch = _T.empty(parts, dtype=torch.float32, device=a.device)
    da = _T.empty_like(a) if want_grad else None
    check(L.mt_loss_l1(_p(_f32c(a)), _p(_f32c(b)), n, float(weight), float(weight), _p(loss_out), _p(da), _p(scratch),
                       parts, _stream()), "loss_l1")
    _count(2)
    return da


def loss_kl_normal(mu, logvar, gscale, loss_out, want_grad=True):
    """loss_out[0] = -0.5 * sum(1 + logvar - mu^2 - exp(logvar)); -> (gscale * dloss/dmu, gscale * dloss/dlogvar)"""
    n = mu.numel()
    dmu = _T.empty_like(mu) if want_grad else None
    dlv = _T.empty_like(logvar) if want_grad else None
    check(_lib.lib().mt_loss_kl_normal(_p(_f32c(mu)), _p(_f32c(logvar)), n, float(gscale), _p(loss_out), _p(dmu), _p(dlv),
                                       _stream()), "loss_kl_normal")
    _count()
    return dmu, dlv
