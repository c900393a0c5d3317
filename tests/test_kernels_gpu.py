"""Op-level parity of every CUDA entry point against plain PyTorch fp32 (GPU only)."""
import pytest

pytestmark = pytest.mark.gpu


def _cases():
    from tests.kernel_cases import CASES
    return sorted(CASES)


@pytest.mark.parametrize("name", [
    "conv3x3_256", "conv3x3_s2", "conv3x3_s2_act_pad", "stem7x7", "stem4x4_s2", "stem3x3_s2",
    "conv1x1", "dis_tail", "odd_channels", "conv4x4_s2_zpad", "conv7x7_s2_256", "convT_256_128",
    "convT_128_64", "in_relu", "in_res_noise", "adain_relu", "adain_res", "ln_relu",
    "in_lrelu_c1024", "pad1", "pad3", "c3_generic", "small_1x1_c32", "small_3x3_c32", "fused_stats_hot",
    "fused_stats_overhang", "fused_stats_s2", "fused_stats_convT", "fused_stats_convT_overhang",
    "conv3x3_c8_pair", "conv1x1_c5_pair", "conv4x4_s2_refl", "conv3x3_s2_odd", "convT_odd", "convT_odd_wide", "upsample2", "mask_mul", "conv3x3_zpad1", "conv7x7_zpad3_rgb"])
def test_kernel_case(name):
    from tests.kernel_cases import CASES, TOL
    res = CASES[name]()
    bad = {k: v for k, v in res.items() if not (v <= TOL)}
    assert not bad, f"{name}: relative errors above {TOL}: {bad} (all: {res})"
