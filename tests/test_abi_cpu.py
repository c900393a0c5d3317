"""CPU-side checks: the C-ABI library loads and exports every symbol include/mt_b200.h declares,
argument validation works without a GPU, the module mirror keeps the reference's state_dict keys,
and the product fails loudly (no CPU fallback)."""
import ctypes
import json
import os

import numpy as np
import pytest
import torch

from masterthesis_b200 import _lib

GOLD = os.path.join(os.path.dirname(__file__), 'golden')


def test_library_exports_every_declared_symbol():
    protos = _lib.header_prototypes()
    assert len(protos) >= 30
    lib = _lib.lib()
    for name in protos:
        assert hasattr(lib, name), name
    assert lib.mt_version() >= 100
    assert lib.mt_error_string(-1003)


def test_pure_host_entry_points():
    lib = _lib.lib()
    assert lib.mt_stored_channels(3) == 8 and lib.mt_stored_channels(268) == 272
    assert lib.mt_conv_window_mode(8, 7) == 1 and lib.mt_conv_window_mode(8, 4) == 1
    assert lib.mt_conv_window_mode(32, 1) == 0 and lib.mt_conv_window_mode(256, 3) == 0
    assert lib.mt_packed_weight_elems(256, 256, 3, 3, 0, 0) == 9 * 256 * 256
    assert lib.mt_packed_weight_elems(64, 3, 7, 7, 1, 0) == 7 * 64 * 64
    assert lib.mt_packed_weight_elems(64, 3, 7, 7, 1, 1) == 49 * 3 * 64
    d = _lib.MtConv2d(1, 8, 8, 12, 12, 6, 6, 8, 8, 3, 3, 1, 0, 6, 6, 0, 0, 0.01, 0)   # Cin_s % 8 != 0
    assert lib.mt_conv2d_fwd(ctypes.byref(d), None, None, None, None, None) == -1003


def test_state_dict_keys_match_reference():
    from oracle import ref_import
    from masterthesis_b200 import models
    meta = json.loads(str(np.load(os.path.join(GOLD, 'step_dim8_64.npz'))['meta']))
    m = models.AdaINModel(ref_import.default_args(dim=8, crop_size=64, batch_size=1))
    for n in m.model:
        mine = {k: list(v.shape) for k, v in m.model[n].state_dict().items()}
        assert list(mine) == list(meta['shapes'][n]) and mine == meta['shapes'][n], n


def test_factories_mirror_reference_errors():
    from masterthesis_b200.models.core import functions as fn
    for f, bad in ((fn.get_norm_layer, 'group'), (fn.get_activation_layer, 'gelu'), (fn.get_padding_layer, 'zero')):
        with pytest.raises(NotImplementedError):
            f(bad)
    assert fn.get_norm_layer(None) is None and fn.get_activation_layer(None) is None
    with pytest.raises(ValueError):
        fn.get_norm_layer(int)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    from oracle import ref_import
    from masterthesis_b200 import models, ops
    m = models.AdaINModel(ref_import.default_args(dim=8, crop_size=64, batch_size=1))
    with pytest.raises(RuntimeError):
        m.initialize()
    with pytest.raises(RuntimeError):
        ops.image_to_nhwc(torch.zeros(1, 3, 8, 8), None, 1)
    with pytest.raises(RuntimeError):
        m.optimizer.decoder.zero_grad()
