"""Run every op-level parity case in its own process (a trapped kernel poisons the CUDA context)
and print the error table.  Usage: python tools/gpu_probe.py [case ...]"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

if len(sys.argv) >= 3 and sys.argv[1] == "--one":
    import torch
    from tests.kernel_cases import CASES
    from tests.model_cases import CASES as MCASES
    CASES = dict(CASES); CASES.update(MCASES)
    res = CASES[sys.argv[2]]()
    torch.cuda.synchronize()
    print("RESULT " + json.dumps(res, default=str))
    sys.exit(0)

from tests.kernel_cases import CASES  # noqa: E402
from tests.model_cases import CASES as MCASES  # noqa: E402
CASES = dict(CASES); CASES.update(MCASES)

names = sys.argv[1:] or list(CASES)
for n in names:
    try:
        p = subprocess.run([sys.executable, __file__, "--one", n], capture_output=True, text=True,
                           timeout=1200, cwd=ROOT)
        line = [l for l in p.stdout.splitlines() if l.startswith("RESULT ")]
        if line:
            print(f"{n:24s} {line[0][7:]}", flush=True)
        else:
            tail = (p.stderr or p.stdout).strip().splitlines()[-14:]
            print(f"{n:24s} FAILED rc={p.returncode}\n    " + "\n    ".join(tail), flush=True)
    except subprocess.TimeoutExpired:
        print(f"{n:24s} TIMEOUT", flush=True)
