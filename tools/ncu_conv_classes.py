"""One launch of every convolution class of the training iteration (batch 8: 16 images per call), in a fixed
order, for an `ncu` capture of the tensor-pipe / L2 / DRAM counters per class:

  python tools/ncu_conv_classes.py                       # plain run: prints the launch order + CUDA-event times
  ncu --metrics <list> --clock-control none -k regex:igemm --csv --log-file out.csv python tools/ncu_conv_classes.py
  python tools/ncu_conv_classes.py --join out.csv         # table: class, kernel, time, TFLOP/s, tensor-pipe %, L2->SM, DRAM

Each class is launched after a warm-up launch of the same class (not captured: `cudaProfilerStart` brackets only
the measured launches when ncu runs with --profile-from-start off).
"""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

N = 16
# (label, kind, Cin, Cout, H, W (input interior), k, stride, reflect pad, stem, transposed)
CLASSES = [
    ("res 3x3 256->256 @64", "fwd", 256, 256, 64, 64, 3, 1, 1, False, False),
    ("res 3x3 256->256 @64", "dgrad", 256, 256, 64, 64, 3, 1, 1, False, False),
    ("res 3x3 256->256 @64", "wgrad", 256, 256, 64, 64, 3, 1, 1, False, False),
    ("convT 3x3 s2 256->128 @64->128", "fwd", 256, 128, 64, 64, 3, 2, 0, False, True),
    ("convT 3x3 s2 256->128 @64->128", "dgrad", 256, 128, 64, 64, 3, 2, 0, False, True),
    ("convT 3x3 s2 256->128 @64->128", "wgrad", 256, 128, 64, 64, 3, 2, 0, False, True),
    ("convT 3x3 s2 128->64 @128->256", "fwd", 128, 64, 128, 128, 3, 2, 0, False, True),
    ("convT 3x3 s2 128->64 @128->256", "dgrad", 128, 64, 128, 128, 3, 2, 0, False, True),
    ("convT 3x3 s2 128->64 @128->256", "wgrad", 128, 64, 128, 128, 3, 2, 0, False, True),
    ("down 3x3 s2 64->128 @256->128", "fwd", 64, 128, 256, 256, 3, 2, 1, False, False),
    ("down 3x3 s2 64->128 @256->128", "dgrad", 64, 128, 256, 256, 3, 2, 1, False, False),
    ("down 3x3 s2 64->128 @256->128", "wgrad", 64, 128, 256, 256, 3, 2, 1, False, False),
    ("down 3x3 s2 128->256 @128->64", "fwd", 128, 256, 128, 128, 3, 2, 1, False, False),
    ("down 3x3 s2 128->256 @128->64", "dgrad", 128, 256, 128, 128, 3, 2, 1, False, False),
    ("down 3x3 s2 128->256 @128->64", "wgrad", 128, 256, 128, 128, 3, 2, 1, False, False),
    ("Es 3x3 64->64 @128", "fwd", 64, 64, 128, 128, 3, 1, 1, False, False),
    ("Es 3x3 64->64 @128", "dgrad", 64, 64, 128, 128, 3, 1, 1, False, False),
    ("Es 3x3 128->128 @64", "fwd", 128, 128, 64, 64, 3, 1, 1, False, False),
    ("Es 3x3 128->128 @64", "dgrad", 128, 128, 64, 64, 3, 1, 1, False, False),
    ("stem 7x7 3->64 @256", "fwd", 3, 64, 256, 256, 7, 1, 3, True, False),
    ("stem 7x7 3->64 @256", "dgrad", 3, 64, 256, 256, 7, 1, 3, True, False),
    ("stem 7x7 3->64 @256", "wgrad", 3, 64, 256, 256, 7, 1, 3, True, False),
    ("Es stem 4x4 s2 7->64 @256->128", "fwd", 7, 64, 256, 256, 4, 2, 1, True, False),
    ("Es stem 4x4 s2 7->64 @256->128", "wgrad", 7, 64, 256, 256, 4, 2, 1, True, False),
    ("Es 3x3 64->128 @128", "fwd", 64, 128, 128, 128, 3, 1, 1, False, False),
    ("Es 3x3 64->128 @128", "dgrad", 64, 128, 128, 128, 3, 1, 1, False, False),
    ("Es 3x3 64->128 @128", "wgrad", 64, 128, 128, 128, 3, 1, 1, False, False),
    ("D 3x3 s2 512->1024 @16->8", "fwd", 512, 1024, 16, 16, 3, 2, 1, False, False),
    ("D 3x3 s2 512->1024 @16->8", "dgrad", 512, 1024, 16, 16, 3, 2, 1, False, False),
    ("D 3x3 s2 512->1024 @16->8", "wgrad", 512, 1024, 16, 16, 3, 2, 1, False, False),
    ("D 3x3 s2 1024->1024 @8->4", "fwd", 1024, 1024, 8, 8, 3, 2, 1, False, False),
    ("D 3x3 s2 1024->1024 @8->4", "wgrad", 1024, 1024, 8, 8, 3, 2, 1, False, False),
]


def flops(c):
    _, kind, Cin, Cout, H, W, k, stride, rp, stem, tr = c
    if tr:
        px = H * W
    else:
        px = ((H + 2 * rp - k) // stride + 1) * ((W + 2 * rp - k) // stride + 1)
    return 2.0 * N * px * Cin * Cout * k * k


def run():
    import torch
    from masterthesis_b200 import ops
    from masterthesis_b200.ops import ConvGeom
    dev = 'cuda'
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
    rows = []
    prepared = []
    for c in CLASSES:
        label, kind, Cin, Cout, H, W, k, stride, rp, stem, tr = c
        g = ConvGeom(Cin, Cout, k, k, stride, 1 if tr else 0, tr, stem)
        if stem:
            x = ops.image_to_nhwc(torch.rand(N, 3, H, W, device=dev) * 2 - 1,
                                  torch.eye(4, device=dev)[torch.arange(N) % 4] if Cin == 7 else None, rp)
        else:
            x = torch.randn(N, H + 2 * rp, W + 2 * rp, g.Cin_s, device=dev).to(torch.bfloat16)
        w = torch.randn(*((Cin, Cout) if tr else (Cout, Cin)), k, k, device=dev) * 0.05
        wf, wb = ops.pack_weight(w, g)
        Ho, Wo = g.out_hw(x.shape[1], x.shape[2])
        dy = torch.randn(N, Ho, Wo, g.Cout_s, device=dev).to(torch.bfloat16)
        prepared.append((c, g, x, wf, wb, dy))

    def launch(c, g, x, wf, wb, dy):
        kind = c[1]
        if kind == 'fwd':
            ops.conv_fwd(x, wf, None, g)
        elif kind == 'dgrad':
            ops.conv_dgrad(dy, wb, g, x.shape[1], x.shape[2])
        else:
            ops.conv_wgrad(x, dy, g)
    for p in prepared:                 # warm-up (module load, attribute set-up, first-launch checks)
        launch(*p)
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    for p in prepared:
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); launch(*p); e1.record(); torch.cuda.synchronize()
        rows.append((p[0][0], p[0][1], e0.elapsed_time(e1) * 1e3, flops(p[0])))
    torch.cuda.profiler.stop()
    for label, kind, us, fl in rows:
        print(f"{label:34s} {kind:5s} {us:8.1f} us (events, incl. unpack / phase launches)  {fl / us / 1e6:8.1f} TFLOP/s")


def join(path):
    """group the ncu rows (launch order) back into classes: a class launch = its igemm kernels in order"""
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    by_id = {}
    for row in csv.DictReader(lines):
        i = int(row["ID"])
        d = by_id.setdefault(i, {"kernel": re.sub(r"\(.*", "", row["Kernel Name"])})
        try:
            d[row["Metric Name"]] = float(row["Metric Value"].replace(",", ""))
        except ValueError:
            pass
        d[row["Metric Name"] + ".unit"] = row["Metric Unit"]
    launches = [by_id[i] for i in sorted(by_id) if "igemm" in by_id[i]["kernel"]]
    # how many igemm launches each class issues (merged-parity: 2, sub-pixel phases: 4, else 1)
    per = []
    for c in CLASSES:
        _, kind, Cin, Cout, H, W, k, stride, rp, stem, tr = c
        n = 1
        if (tr and kind == 'fwd') or (not tr and stride == 2 and kind == 'dgrad'):
            merged = (tr and k == 3) or (not tr and (k == 4 or (k == 3 and Cin <= 128)))
            n = 2 if merged else 4
        per.append(n)
    assert sum(per) == len(launches), (sum(per), len(launches))
    print(f"{'class':34s} {'kind':5s} {'kernel':32s} {'us':>8s} {'TFLOP/s':>8s} {'tensor%':>8s} {'L2->SM MB':>10s} {'DRAM MB':>8s}")
    i = 0
    for c, n in zip(CLASSES, per):
        ls = launches[i:i + n]; i += n
        ns = sum(l["gpu__time_duration.sum"] * {"ns": 1, "us": 1e3, "ms": 1e6}.get(l["gpu__time_duration.sum.unit"], 1) for l in ls)
        tp = sum(l.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 0.0) * l["gpu__time_duration.sum"] for l in ls) / \
            max(sum(l["gpu__time_duration.sum"] for l in ls), 1e-9)
        def mb(key):
            tot = 0.0
            for l in ls:
                tot += l.get(key, 0.0) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(l.get(key + ".unit"), 1)
            return tot / 1e6
        print(f"{c[0]:34s} {c[1]:5s} {ls[0]['kernel'][9:41]:32s} {ns / 1e3:8.1f} {flops(c) / ns / 1e3:8.1f} {tp:8.1f} "
              f"{mb('l1tex__m_xbar2l1tex_read_bytes.sum'):10.1f} {mb('dram__bytes_read.sum') + mb('dram__bytes_write.sum'):8.1f}")


if __name__ == '__main__':
    if len(sys.argv) > 2 and sys.argv[1] == '--join':
        join(sys.argv[2])
    else:
        run()
