"""TEST INFRASTRUCTURE — import the UNMODIFIED reference as a library: from /root/reference/src where
that exists (the build container: oracle/make_golden.py, CPU tests that pin the oracle against the
reference directly), else from the byte-identical copy oracle/_ref/src that oracle/make_ref.py makes
(git-ignored; it travels to the GPU box, where bench.py's reference legs and the -m gpu drop-in tests
use it).  The product never imports this.  Two modules the reference imports but that are absent
here and never touch arithmetic are stubbed (tensorboardX: model.py:13, matplotlib.cm:
utils.py:7), and the two CPU-incompatible `.get_device()` lines (misc.py:25, loss.py:60-62)
are patched harness-side to `.device` so the training step can run on CPU.
"""
import os
import sys
import types

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference/src" if os.path.isdir("/root/reference/src") else os.path.join(_HERE, "_ref", "src")


def available() -> bool:
    return os.path.isdir(os.path.join(REF_SRC, "models"))


def import_reference():
    if not available():
        raise RuntimeError("reference sources not present on this machine")
    if "tensorboardX" not in sys.modules:
        tb = types.ModuleType("tensorboardX")

        class SummaryWriter:  # noqa: D401 - stub
            def __init__(self, *a, **k):
                pass

            def add_scalar(self, *a, **k):
                pass
        tb.SummaryWriter = SummaryWriter
        sys.modules["tensorboardX"] = tb
    if "matplotlib" not in sys.modules:
        mpl = types.ModuleType("matplotlib")
        cm = types.ModuleType("matplotlib.cm")
        mpl.cm = cm
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.cm"] = cm
    # the reference's top-level module names (models, utils, dataset) must win the import
    for name in ("models", "utils", "dataset"):
        mod = sys.modules.get(name)
        if mod is not None and not (getattr(mod, "__file__", None) or "").startswith(REF_SRC):
            raise RuntimeError(f"module name clash: {name} already imported from {mod.__file__}")
    if REF_SRC not in sys.path:
        sys.path.insert(0, REF_SRC)
    import torch
    import models  # noqa: F401  (reference package)
    from models.core import misc, loss

    def noise_forward(self, x):
        if not self.training:
            return x
        return x + torch.randn(x.size()).to(x.device)

    def gan_forward(self, inp, trg_is_real, is_dis=None):
        if "wgangp" in self.loss_type:
            return -inp.mean() if trg_is_real else inp.mean()
        trg = (self.ones if trg_is_real else self.zeros).expand_as(inp).to(inp.device)
        return self.loss(inp, trg)

    misc.GaussianNoiseLayer.forward = noise_forward
    loss.GANLoss.forward = gan_forward
    return models


def default_args(**over):
    """argparse.Namespace with the reference's defaults (arguments.py:20-46,85-118): ONE table, kept in
    masterthesis_b200/arguments.py (the values are the reference's; the golden generator hands the same
    namespace to the unmodified reference models)."""
    from masterthesis_b200.arguments import default_args as _d
    return _d(**over)
