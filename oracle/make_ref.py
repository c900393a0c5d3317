"""TEST INFRASTRUCTURE — put the reference's own Python sources where the GPU box can import them.

    python -m oracle.make_ref          # only where /root/reference exists (the build container)

/root/reference does not exist on the GPU box, and a Python reference has no binary to build: this recipe
copies the reference's `src/` tree (pure Python, ~3k lines: models/, utils.py, arguments.py, dataset.py,
train.py, sample.py, ...) UNMODIFIED into oracle/_ref/src/.  oracle/_ref/ is git-ignored (no reference
source enters the history) but travels with the gpurun snapshot, like the built .so files.  It is used,
through oracle/ref_import.py, only by
  * bench.py --impl reference and bench.py's `cpu_baseline` / `gpu_baseline` legs (the reference's own
    modules timed on the host cores / on cuda:0 next to the CUDA path),
  * the -m gpu drop-in tests that drive the reference's own train.py / sample.py over this repo's models
    and exchange checkpoints with the reference's modules (tests/test_dropin_gpu.py).
The product (masterthesis_b200/) never imports it.  __graft_entry__.build() runs this recipe when
/root/reference is present.
"""
import filecmp
import os
import shutil
import sys

SRC = "/root/reference/src"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "src")


def make(verbose=True) -> bool:
    if not os.path.isdir(SRC):
        if verbose:
            print(f"oracle/make_ref: {SRC} not present; keeping {DST} as it is "
                  f"({'present' if os.path.isdir(DST) else 'absent'})")
        return os.path.isdir(DST)
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    shutil.copytree(SRC, DST, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    # the copy must be the reference, byte for byte
    cmp = filecmp.dircmp(SRC, DST, ignore=["__pycache__"])
    stack, bad = [cmp], []
    while stack:
        c = stack.pop()
        bad += c.diff_files + c.left_only + c.right_only
        stack += list(c.subdirs.values())
    if bad:
        raise RuntimeError(f"oracle/_ref differs from the reference: {bad}")
    if verbose:
        n = sum(len(f) for _, _, f in os.walk(DST))
        print(f"oracle/make_ref: copied {n} files from {SRC} to {DST} (unmodified)")
    return True


if __name__ == "__main__":
    sys.exit(0 if make() else 1)
