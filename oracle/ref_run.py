"""TEST INFRASTRUCTURE — run the reference's OWN modules (oracle/ref_import.py: /root/reference/src, or
its byte-identical copy oracle/_ref/src on the GPU box) as the timed baselines of bench.py:

  * cpu_train_iterations : AdaINModel / BaseModel .optimize_parameters on the host cores (fp32),
  * cpu_config1          : BASELINE config 1 exactly — BaseModel, default flags, content_encoder -> decoder
                           forward on one 256x256 image, every net in .eval(),
  * gpu_train_iterations : the same reference modules on cuda:0 in fp32 / TF32 / bf16-autocast +
                           channels_last with cudnn.benchmark (SURVEY.md 8d "GPU baseline in the same run").

Nothing here is imported by the product path.
"""
import contextlib
import os
import time

import torch

from . import ref_import


def _batch(B, size, nd, seed):
    g = torch.Generator().manual_seed(seed)
    x1 = torch.rand((B, 3, size, size), generator=g) * 2 - 1
    x2 = torch.rand((B, 3, size, size), generator=g) * 2 - 1
    d1 = torch.randint(0, nd, (B,), generator=g)
    d2 = (d1 + 1 + torch.randint(0, nd - 1, (B,), generator=g)) % nd
    return {"x1": x1, "x2": x2, "y1": torch.eye(nd)[d1], "y2": torch.eye(nd)[d2]}


def make_model(model="AdaINModel", B=1, size=256, nd=4, device="cpu", mode="train", **flags):
    models = ref_import.import_reference()
    args = ref_import.default_args(dim=64, crop_size=size, batch_size=B, num_domains=nd, init_type="normal",
                                   mode=mode, **flags)
    torch.manual_seed(1234)
    m = getattr(models, model)(args)
    m.device = device
    if "train" in mode:          # the losses were moved to the constructor-time device
        for name in ("gan_loss", "classification_loss", "l1_loss"):
            if hasattr(m, name):
                setattr(m, name, getattr(m, name).to(device))
    m.initialize()
    return m


def cpu_train_iterations(model="AdaINModel", B=1, size=256, nd=4, steps=1, warmup=1, threads=None, **flags):
    """-> (images/s, seconds per iteration, threads) of the reference's optimize_parameters on the CPU."""
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    m = make_model(model, B, size, nd, "cpu", **flags)
    b = _batch(B, size, nd, 100)
    ts = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        m.set_inputs(b)
        m.optimize_parameters(i)
        ts.append(time.perf_counter() - t0)
    ts = ts[warmup:]
    sec = sum(ts) / len(ts)
    return 2 * B / sec, sec, threads


def cpu_config1(size=256, nd=4, warmup=2, reps=5, threads=None):
    """BASELINE config 1: BaseModel (default flags) content_encoder -> decoder forward, batch 1, .eval().
    -> dict(best_ms, median_ms, images_per_sec (from the best), threads)."""
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    m = make_model("BaseModel", 1, size, nd, "cpu", mode="test")
    for n in m.model:
        m.model[n].eval()
    g = torch.Generator().manual_seed(0)
    img = torch.rand((1, 3, size, size), generator=g) * 2 - 1
    c = torch.eye(nd)[:1]
    z = torch.randn((1, 8), generator=g)
    ts = []
    with torch.no_grad():
        for i in range(warmup + reps):
            t0 = time.perf_counter()
            m.model.decoder(m.model.content_encoder(img), z, c)
            ts.append(time.perf_counter() - t0)
    ts = sorted(ts[warmup:])
    return {"best_ms": ts[0] * 1e3, "median_ms": ts[len(ts) // 2] * 1e3, "images_per_sec": 1.0 / ts[0],
            "threads": threads}


def gpu_train_iterations(variant, model="AdaINModel", B=8, size=256, nd=4, steps=3, warmup=2, **flags):
    """The reference's modules on cuda:0.  variant: 'fp32' (TF32 off), 'tf32', 'bf16_autocast_channels_last'.
    -> (images/s, ms per iteration).  Device-timed with CUDA events around `steps` iterations."""
    assert torch.cuda.is_available()
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.benchmark)
    tf32 = variant != "fp32"
    torch.backends.cudnn.allow_tf32 = tf32
    torch.backends.cuda.matmul.allow_tf32 = tf32
    torch.backends.cudnn.benchmark = True
    try:
        m = make_model(model, B, size, nd, "cuda", **flags)
        amp = contextlib.nullcontext
        if variant == "bf16_autocast_channels_last":
            for n in m.model:
                m.model[n] = m.model[n].to(memory_format=torch.channels_last)
            amp = lambda: torch.autocast("cuda", dtype=torch.bfloat16)  # noqa: E731
        b = {k: v.cuda() for k, v in _batch(B, size, nd, 100).items()}
        if variant == "bf16_autocast_channels_last":
            b = {k: (v.contiguous(memory_format=torch.channels_last) if v.dim() == 4 else v) for k, v in b.items()}
        for i in range(warmup):
            m.set_inputs(b)
            with amp():
                m.optimize_parameters(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            m.set_inputs(b)
            with amp():
                m.optimize_parameters(warmup + i)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        finite = all(v == v and abs(v) < 1e9 for v in m.loss.values())
        del m
        torch.cuda.empty_cache()
        return 2 * B / (ms * 1e-3), ms, finite
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.benchmark = old
