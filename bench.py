"""Benchmark of the hot path: Model.optimize_parameters (full GAN+VAE training iteration, reference
src/models/adain_model.py:421-430 / base_model.py) on synthetic images.

    python bench.py --gpus N --steps K --warmup W [--config 3|4|5-adain|5-base]   # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W                          # the reference's own modules, CPU

Prints ONE JSON line (rank 0).  `value` = images/s of the whole job with the batch already resident in
HBM; `e2e` = the same through the public API with pinned HOST buffers (H2D of the batch and D2H of the
loss scalars inside the timed region); `roofline` = the dominant implicit-GEMM conv kernel's algorithmic
FLOP/s against the measured bf16 peak (+ `roofline_hbm` for the norm kernels against the measured copy
bandwidth); `cpu_baseline` = the reference's own modules on the host cores on a bounded sample;
`gpu_baseline` = the reference's own modules on the same GPU (fp32 / TF32 / bf16-autocast + channels_last).

--config selects the workload (BASELINE.json configs): 3 (default, the configuration the metric is quoted on:
AdaINModel, batch 8, 256x256), 4 (batch 32 per GPU), 5-adain / 5-base (512x512, batch 16 per GPU).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "train_images_per_sec"
UNIT = "images/s"
ND, LATENT = 4, 8
JSON_OUT = sys.stdout               # re-pointed in __main__ so that stdout carries only the JSON line

# BASELINE.json configs -> (model, pairs per GPU, crop size, algorithmic GFLOP per unit batch: SURVEY.md 8d /
# BASELINE.md 3, x4 at 512x512)
CONFIGS = {
    "3": dict(model="AdaINModel", B=8, size=256, gflop_per_b=2414.0,
              name="config 3: AdaINModel.optimize_parameters, batch 8, 256x256"),
    "4": dict(model="AdaINModel", B=32, size=256, gflop_per_b=2414.0,
              name="config 4: AdaINModel.optimize_parameters, batch 32 per GPU, 256x256, 4 domains, random style codes"),
    "5-adain": dict(model="AdaINModel", B=16, size=512, gflop_per_b=4 * 2414.0,
                    name="config 5: AdaINModel.optimize_parameters, batch 16 per GPU, 512x512"),
    "5-base": dict(model="BaseModel", B=16, size=512, gflop_per_b=4 * 3217.0,
                   name="config 5: BaseModel.optimize_parameters (default flags), batch 16 per GPU, 512x512"),
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"bf16_tflops": d.get("bf16_tflops_sustained", d.get("bf16_tflops")),
                "bf16_tflops_burst": d.get("bf16_tflops"), "hbm_gbs": d.get("hbm_gbs"),
                "source": "MEASURED_PEAKS.json (sustained bf16: the kernel is timed inside a long step)"}
    return {"bf16_tflops": 1400.0, "bf16_tflops_burst": 1590.0, "hbm_gbs": 6650.0,
            "source": "fallback (B200_PROFILING.md)"}


def make_args(B, mode="train", size=256):
    from masterthesis_b200.arguments import default_args
    return default_args(dim=64, crop_size=size, batch_size=B, num_domains=ND, latent_dim=LATENT,
                        mode=mode, init_type="normal")


def synth_batch(B, seed, pinned, size=256):
    g = torch.Generator().manual_seed(seed)
    x1 = torch.rand((B, 3, size, size), generator=g) * 2 - 1
    x2 = torch.rand((B, 3, size, size), generator=g) * 2 - 1
    d1 = torch.randint(0, ND, (B,), generator=g)
    d2 = (d1 + 1 + torch.randint(0, ND - 1, (B,), generator=g)) % ND      # a different domain
    b = {"x1": x1, "x2": x2, "y1": torch.eye(ND)[d1], "y2": torch.eye(ND)[d2]}
    if pinned:
        b = {k: v.pin_memory() for k, v in b.items()}
    return b


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index
        self.t0 = None          # start of the timed region (rows before it are warm-up samples)

    def start(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def mark(self):
        """the timed region starts now (the sampler itself is started before the warm-up: nvidia-smi needs a
        few hundred ms to deliver its first row, longer than a short timed region)"""
        self.t0 = time.perf_counter()

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        timed = [r for t, r in self.rows if self.t0 is None or t >= self.t0]
        # a timed region shorter than the sampling period: fall back to the last rows under load (warm-up
        # iterations of the same workload immediately before it)
        note = None
        if len(timed) < 2:
            timed, note = [r for _, r in self.rows[-5:]], "timed region shorter than the sampling period: last warm-up samples"
        self.rows = timed
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        out = {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
               "reasons": reasons, "samples": len(sm)}
        if note:
            out["note"] = note
        return out


# ------------------------------------------------------------------------------- CPU / GPU baselines
def reference_available():
    from oracle import ref_import
    return ref_import.available()


def cpu_iteration_port(steps, warmup, threads, B=1):
    """Fallback CPU leg when oracle/_ref is absent: oracle/step.py (fp32 restatement, pinned against the
    reference in tests/test_oracle_golden.py) — one batch-B iteration (2B images) per step."""
    from oracle import step as ostep, synth
    torch.set_num_threads(threads)
    import numpy as np
    meta = json.loads(str(np.load(os.path.join(ROOT, "tests", "golden", "nets_dim64_256.npz"))["meta"]))
    sh = meta["shapes"]                     # state_dict shapes of the dim=64 / 256x256 architecture
    sh = {"content_encoder": sh["content_encoder"], "style_encoder": sh["style_encoder"],
          "decoder": sh["decoder"], "discriminator1": sh["discriminator"], "discriminator2": sh["discriminator"]}
    params = {n: synth.synth_params(s_, 0, n + ".") for n, s_ in sh.items()}
    st = ostep.AdaINStep(params, B)
    b = synth_batch(B, 0, False)
    img, c = torch.cat((b["x1"], b["x2"])), torch.cat((b["y1"], b["y2"]))
    draws = synth.Draws(0)
    ts = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        st.optimize_parameters(img, c, draws)
        ts.append(time.perf_counter() - t0)
    ts = ts[warmup:]
    sec = sum(ts) / len(ts)
    return 2 * B / sec, sec


def cpu_reference(steps, warmup, budget_s=None):
    """The reference's own AdaINModel.optimize_parameters on the host cores, batch 1 (2 images) per step
    (a bounded sample of the workload: same networks, image size and schedule), all host threads.
    -> (images/s, s/iteration, threads, kind, steps actually timed)."""
    threads = os.cpu_count()
    if reference_available():
        from oracle import ref_run
        torch.set_num_threads(threads)
        m = ref_run.make_model("AdaINModel", 1, 256, ND, "cpu")
        b = ref_run._batch(1, 256, ND, 100)
        ts, t_start = [], time.perf_counter()
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            m.set_inputs(b)
            m.optimize_parameters(i)
            ts.append(time.perf_counter() - t0)
            if budget_s and i >= warmup and time.perf_counter() - t_start > budget_s:
                break
        ts = ts[warmup:]
        sec = sum(ts) / len(ts)
        return 2 / sec, sec, threads, "reference", len(ts)
    ips, sec = cpu_iteration_port(steps, warmup, threads)
    return ips, sec, threads, "port", steps


def sampling_throughput(iters=20, size=256):
    """Config 2: AdaINModel.forward_reference at N = 8, 256x256, one GPU — "translated
    images/s" (98.3 GFLOP per image).  Inputs resident; device-timed over `iters` calls."""
    from masterthesis_b200 import models
    N = 8
    m = models.AdaINModel(make_args(N, mode="test", size=size))
    m.initialize()
    g = torch.Generator().manual_seed(7)
    src = (torch.rand((N, 3, size, size), generator=g) * 2 - 1).cuda()
    ref = (torch.rand((N, 3, size, size), generator=g) * 2 - 1).cuda()
    c = torch.eye(ND)[torch.randint(0, ND, (N,), generator=g)].cuda()
    with torch.no_grad():
        for _ in range(3):
            m.forward_reference(src, ref, c)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            m.forward_reference(src, ref, c)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    return {"metric": "translated_images_per_sec", "value": N / (ms * 1e-3), "ms_per_call": ms,
            "config": "AdaINModel.forward_reference, N = 8, 256x256 (SURVEY config 2), launch sequence replayed as a CUDA graph (test mode); the "
                      "reference API returns (image, seconds, GiB) per call",
            "model_tflops": 98.3e-3 * N / (ms * 1e-3)}


def run_reference(a):
    """--impl reference: the reference's own CPU implementation of the path (its unmodified modules from
    oracle/_ref; the oracle port only if that copy is absent), all host threads, K timed steps after W
    warm-up steps, each step one batch-1 iteration (2 images) of the same networks at 256x256."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, a.steps), max(0, a.warmup)
    ips, sec, threads, kind, done = cpu_reference(steps, warmup, budget_s=420.0)
    what = "the reference's own AdaINModel.optimize_parameters (oracle/_ref)" if kind == "reference" else "oracle/step.py port"
    sample = (f"{done} timed iteration(s) after {warmup} warm-up, each at batch 1 (2 images), 256x256, fp32 torch CPU, {what}"
              + ("" if done == steps else f"; stopped after the 7-minute budget ({steps} requested)"))
    line = {"impl": "reference", "metric": METRIC, "value": ips, "unit": UNIT, "n_gpus": a.gpus,
            "steps": done, "warmup": warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": CONFIGS["3"]["name"] + " (CPU sample: batch 1 = 2 images per step)",
                       "cpu_processes": 1, "gpus_used": 0,
                       "note": "one CPU process whatever --gpus says: the reference's CPU path does not shard"},
            "cpu_baseline": {"value": ips, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": ips, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), file=JSON_OUT, flush=True)


# ------------------------------------------------------------------------------- rooflines
def kernel_rooflines(model, batch, pk, cfg):
    """One instrumented iteration: CUDA events around every implicit-GEMM launch and every norm kernel on
    the launching stream; aggregate algorithmic FLOPs (bytes) / device time per kernel class."""
    from masterthesis_b200 import ops
    ops.PROFILE = []
    graph_mode, model.use_graph = model.use_graph, False     # events between launches: eager iteration
    model.set_inputs(batch)
    model.optimize_parameters(0)
    torch.cuda.synchronize()
    model.use_graph = graph_mode
    recs, ops.PROFILE = ops.PROFILE, None
    by_kind, by_label, hbm = {}, {}, {}
    for kind, label, work, e0, e1 in recs:
        ms = e0.elapsed_time(e1)
        if kind.startswith("hbm."):
            h = hbm.setdefault((kind, label), [0.0, 0.0, 0])
            h[0] += work; h[1] += ms; h[2] += 1
            continue
        k = by_kind.setdefault(kind, [0.0, 0.0, 0])
        k[0] += work; k[1] += ms; k[2] += 1
        l = by_label.setdefault((kind, label), [0.0, 0.0, 0])
        l[0] += work; l[1] += ms; l[2] += 1
    tot_f = sum(v[0] for v in by_kind.values())
    tot_ms = sum(v[1] for v in by_kind.values())
    peak = pk["bf16_tflops"]
    top = sorted(by_label.items(), key=lambda kv: -kv[1][1])[:8]
    if os.environ.get('MT_BENCH_DUMP'):
        with open(os.environ['MT_BENCH_DUMP'], 'w') as f:
            for k, v in sorted(by_label.items(), key=lambda kv: -kv[1][1]):
                f.write(f"{k[0]:6s} {k[1]:40s} launches {v[2]:3d}  ms {v[1]:8.3f}  TFLOP/s {v[0] / (v[1] * 1e-3) / 1e12:8.1f}\n")
            for k, v in sorted(hbm.items(), key=lambda kv: -kv[1][1]):
                f.write(f"{k[0]:22s} {k[1]:24s} launches {v[2]:3d}  ms {v[1]:8.3f}  GB/s {v[0] / (v[1] * 1e-3) / 1e9:8.1f}\n")
    # Dominant kernel: igemm_conv_kernel<2,256,6> on the 3x3 256->256 conv of the residual blocks (forward
    # and data-gradient launches of ResnetBlock / AdaINResnetBlock: ~25 % of the step).
    hw = cfg["size"] // 4
    HOT = f"conv3x3s1_256->256@{hw}x{hw}"
    hot = [v for k, v in by_label.items() if k[1] == HOT and k[0] in ("fwd", "dgrad")]
    hot_f, hot_ms, hot_n = (sum(v[i] for v in hot) for i in range(3))
    ach = hot_f / (hot_ms * 1e-3) / 1e12 if hot_ms else None
    all_ach = tot_f / (tot_ms * 1e-3) / 1e12 if tot_ms else None
    # DRAM traffic of one launch of the dominant kernel: read from the committed ncu export (raw CSV of
    # `ncu --set full`, profiles/r2_ncu_conv_hot.csv: dram__bytes_read.sum + dram__bytes_write.sum)
    traffic, traffic_src = None, None
    for name in ("r2_ncu_conv_hot_raw.csv",):
        path = os.path.join(ROOT, "profiles", name)
        if os.path.exists(path):
            traffic, traffic_src = _ncu_dram_bytes(path), f"profiles/{name}"
    roof = {"bound": "tensor", "kernel": f"igemm_conv_kernel<2,256,6> on {HOT} (fwd + dgrad launches)",
            "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak if ach else None,
            "frac_of_burst_peak": ach / pk["bf16_tflops_burst"] if (ach and pk.get("bf16_tflops_burst")) else None,
            "traffic": traffic,
            "traffic_note": (f"dram__bytes_read.sum + dram__bytes_write.sum of one 16-image launch, read from {traffic_src}; "
                             "algorithmic bytes of that launch: 35.7 MB in + 1.2 MB weights + 33.5 MB out = 70.4 MB "
                             "(the output stays in L2)") if traffic else "no committed ncu export found",
            "peak_source": pk["source"], "launches": hot_n,
            "flop_per_launch_n16": 2.0 * 16 * hw * hw * 256 * 256 * 9,
            "ms_per_iter": hot_ms,
            "all_convs": {"kernel": "igemm_conv_kernel + igemm_wgrad_kernel (every conv launch of one iteration)",
                          "achieved": all_ach, "frac": all_ach / peak if all_ach else None,
                          "frac_of_burst_peak": all_ach / pk["bf16_tflops_burst"] if (all_ach and pk.get("bf16_tflops_burst")) else None,
                          "launches": sum(v[2] for v in by_kind.values()), "conv_ms_per_iter": tot_ms,
                          "by_kind": {k: {"tflops": v[0] / (v[1] * 1e-3) / 1e12, "ms": v[1], "launches": v[2]}
                                      for k, v in by_kind.items()}},
            "top_classes": [{"kind": k[0], "shape": k[1], "tflops": v[0] / (v[1] * 1e-3) / 1e12, "ms": v[1], "launches": v[2]}
                            for k, v in top]}
    # HBM side: the norm / AdaIN kernels (north_star: >= 70 % of HBM bandwidth), by kernel and tensor shape
    hpk = pk["hbm_gbs"]
    tot_b = sum(v[0] for v in hbm.values())
    tot_hms = sum(v[1] for v in hbm.values())
    per_kernel = {}
    for (kind, label), v in hbm.items():
        d = per_kernel.setdefault(kind, [0.0, 0.0, 0])
        d[0] += v[0]; d[1] += v[1]; d[2] += v[2]
    worst = min(hbm.items(), key=lambda kv: kv[1][0] / kv[1][1]) if hbm else None
    topb = sorted(hbm.items(), key=lambda kv: -kv[1][1])[:6]
    roof_hbm = {"bound": "hbm", "kernel": "affine_act_fwd / norm_bwd_reduce / norm_bwd_apply (InstanceNorm, AdaIN, LayerNorm "
                                          "apply passes, forward and backward, all launches of one iteration)",
                "achieved": tot_b / (tot_hms * 1e-3) / 1e9 if tot_hms else None, "peak": hpk, "unit": "GB/s",
                "frac": tot_b / (tot_hms * 1e-3) / 1e9 / hpk if tot_hms else None,
                "bytes": "algorithmic: fwd numel*(2+2) [+2 residual, +4 fp32 noise], bwd reduce numel*(2+2), bwd apply numel*(2+2+2)",
                "ms_per_iter": tot_hms, "launches": sum(v[2] for v in hbm.values()),
                "by_kernel": {k[4:]: {"gbs": v[0] / (v[1] * 1e-3) / 1e9, "frac": v[0] / (v[1] * 1e-3) / 1e9 / hpk,
                                      "ms": v[1], "launches": v[2]} for k, v in per_kernel.items()},
                "slowest_class": ({"kernel": worst[0][0][4:], "shape_nhwc": worst[0][1],
                                   "gbs": worst[1][0] / (worst[1][1] * 1e-3) / 1e9} if worst else None),
                "top_classes": [{"kernel": k[0][4:], "shape_nhwc": k[1], "gbs": v[0] / (v[1] * 1e-3) / 1e9,
                                 "ms": v[1], "launches": v[2]} for k, v in topb],
                "note": "tensors of the 64x64 layers (34 MB at 16 images) are partly L2-resident between producer and "
                        "consumer, so a class can read above the DRAM copy peak; launched back to back on one stream here"}
    return roof, roof_hbm


def _ncu_dram_bytes(path):
    """dram__bytes_read.sum + dram__bytes_write.sum of the first kernel row of an `ncu --page raw --csv` export"""
    import csv
    try:
        with open(path) as f:
            rows = list(csv.reader(l for l in f if not l.startswith("==")))
        hdr, units = rows[0], rows[1]
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        tot = 0.0
        for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = hdr.index(name)
            tot += float(rows[2][i].replace(",", "")) * scale.get(units[i], 1.0)
        return tot
    except Exception:
        return None


def gpu_baselines(cfg):
    """The reference's own modules on this GPU (same config), three precisions.  Context for `value`."""
    from oracle import ref_run
    out = {"config": cfg["name"], "how": "the reference's unmodified modules (oracle/_ref) on cuda:0, eager PyTorch, "
                                         "cudnn.benchmark on, 2 warm-up + 3 timed iterations, CUDA events"}
    for variant in ("fp32", "tf32", "bf16_autocast_channels_last"):
        try:
            ips, ms, finite = ref_run.gpu_train_iterations(variant, cfg["model"], cfg["B"], cfg["size"], ND)
            out[variant] = {"value": ips, "unit": UNIT, "ms_per_step": ms, "losses_finite": finite}
        except Exception as ex:
            out[variant] = {"failed": f"{type(ex).__name__}: {str(ex)[:200]}"}
            torch.cuda.empty_cache()
    return out


def run_ours(a):
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the CUDA path has no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    from masterthesis_b200 import models, ops
    from masterthesis_b200.parallel import DataParallelGroup
    cfg = CONFIGS[a.config]
    dp = DataParallelGroup.from_env()
    world = 1 if dp is None else dp.world
    rank = 0 if dp is None else dp.rank
    if dp is None:
        torch.cuda.set_device(0)
    dev_index = torch.cuda.current_device()
    B, size = cfg["B"], cfg["size"]
    torch.manual_seed(1234)
    model = getattr(models, cfg["model"])(make_args(B, size=size))
    model.initialize()
    model.dp = dp
    if dp is not None:
        dp.broadcast_parameters([model.model[n] for n in model.model])
    host = synth_batch(B, 100 + rank, pinned=True, size=size)
    resident = {k: v.cuda() for k, v in host.items()}

    def sync():
        if dp is not None:
            dp.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput
    clocks = ClockSampler(dev_index)
    clocks.start()
    model.set_inputs(resident)
    for _ in range(a.warmup):
        model.optimize_parameters(0)
    sync()
    clocks.mark()
    l0 = ops.launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.steps):
        model.optimize_parameters(i)
    e1.record()
    sync()
    ms = e0.elapsed_time(e1)
    launches = ops.launches() - l0
    clk = clocks.stop()
    if dp is not None:
        ms = dp.max_over_ranks(ms)
    ms_per_step = ms / a.steps
    value = world * 2 * B / (ms_per_step * 1e-3)

    # ---- end to end through the public API: pinned host batch -> set_inputs (H2D) -> iteration -> losses (D2H)
    for _ in range(min(2, a.warmup)):
        model.set_inputs(host); model.optimize_parameters(0)
    sync()
    e0.record()
    for i in range(a.steps):
        model.set_inputs(host)
        model.optimize_parameters(i)
        _ = dict(model.loss)
    e1.record()
    sync()
    ms2 = e0.elapsed_time(e1)
    if dp is not None:
        ms2 = dp.max_over_ranks(ms2)
    e2e = world * 2 * B / (ms2 / a.steps * 1e-3)
    h2d = sum(v.numel() * v.element_size() for v in host.values())
    d2h = len(model.loss) * 4

    # ---- data-parallel self check (N > 1): after the all-reduce every rank holds the same gradient bucket,
    # hence the same weights: checksum of every network's parameters, max relative spread over ranks
    dp_check = None
    if dp is not None:
        dp_check = dp.parameter_spread([model.model[n] for n in model.model])

    if rank != 0:
        return
    pk = peaks()
    # rank 0 alone from here on: the roofline leg must not wait in an all-reduce for ranks that left
    model.dp = None
    for opt in model.optimizer.values():
        opt.grad_scale = 1.0
    roof, roof_hbm = kernel_rooflines(model, resident, pk, cfg)
    extra = {}
    if world == 1 and not a.no_baselines:
        try:
            ips, sec, threads, kind, done = cpu_reference(steps=2, warmup=1)
            cpu = {"value": ips, "unit": UNIT, "cores": threads, "kind": kind,
                   "sample": f"{done} timed iterations after 1 warm-up at batch 1 (2 images), 256x256, fp32 torch CPU, {sec:.1f} s per iteration, "
                             + ("the reference's own AdaINModel.optimize_parameters (oracle/_ref)" if kind == "reference" else "oracle/step.py port")}
            if kind == "reference":
                from oracle import ref_run
                c1 = ref_run.cpu_config1()
                cpu["config1"] = dict(c1, what="BASELINE config 1 exactly: the reference's BaseModel (default flags), content_encoder -> "
                                               "decoder forward, batch 1, 256x256, every net in .eval(); 2 warm-up, best / median of 5")
        except Exception as ex:  # the GPU numbers must still be printed
            cpu = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "reference", "sample": f"failed: {ex}"}
        try:
            sampling = sampling_throughput()
        except Exception as ex:  # secondary metric: never lose the headline line over it
            sampling = {"failed": str(ex)}
        extra = {"cpu_baseline": cpu, "sampling": sampling}
        if reference_available():
            del model
            torch.cuda.empty_cache()
            try:
                extra["gpu_baseline"] = gpu_baselines(cfg)
            except Exception as ex:
                extra["gpu_baseline"] = {"failed": str(ex)[:300]}
        else:
            extra["gpu_baseline"] = {"unavailable": "oracle/_ref absent (python -m oracle.make_ref in the build container)"}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "warmup_requested": a.warmup_requested, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"{cfg['name']} (full GAN+VAE iteration, default flags), "
                                   f"batch {B} per GPU = {2 * B} images/iteration/GPU, {size}x{size}, 4 domains",
                       "bench_config": a.config,
                       "iters_per_sec": 1e3 / ms_per_step,
                       "l2_policy": "per-iteration working set (GBs of activations) exceeds the 126 MB L2",
                       "execution": "whole iteration replayed as one CUDA graph",
                       "warmup_note": "at least 3 warm-up iterations are always run (2 eager + the graph capture)",
                       "algorithmic_tflop_per_iter_per_gpu": cfg["gflop_per_b"] * B / 1e3,
                       "model_tflops_per_gpu": cfg["gflop_per_b"] * B / 1e3 / (ms_per_step * 1e-3)},
            "clocks": clk, "gpu_launches": launches,
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "roofline": roof, "roofline_hbm": roof_hbm}
    if dp_check is not None:
        line["dp_check"] = dp_check
    line.update(extra)
    print(json.dumps(line), file=JSON_OUT, flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="3", choices=sorted(CONFIGS))
    ap.add_argument("--no-baselines", dest="no_baselines", action="store_true",
                    help="skip the cpu_baseline / gpu_baseline / sampling legs (profiling runs)")
    a = ap.parse_args()
    # stdout carries exactly ONE line (the JSON record); everything else the libraries print (the model
    # mirrors the reference's "initialize network with ..." messages) goes to stderr
    # (fd 1 itself is re-pointed: NCCL prints its version banner from C)
    sys.stdout.flush()
    JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = sys.stderr
    a.warmup_requested = a.warmup
    a.warmup = max(a.warmup, 3) if a.impl == "ours" else a.warmup
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
    if torch.distributed.is_available() and torch.distributed.is_initialized():
        torch.distributed.destroy_process_group()
