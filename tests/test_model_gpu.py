"""Network / training-step / sampling parity of the CUDA path (GPU only).
Oracles: the reference's recorded outputs (tests/golden) and oracle/ run in fp32."""
import pytest

pytestmark = pytest.mark.gpu


def _check(res, rel_tol, cos_tol):
    bad = {}
    for k, v in res.items():
        if k.startswith('cos'):
            if not v >= cos_tol:
                bad[k] = v
        elif not v <= rel_tol:
            bad[k] = v
    assert not bad, f"out of tolerance (rel<={rel_tol}, cos>={cos_tol}): {bad}\nall: {res}"


def test_nets_small_vs_reference_golden():
    from tests.model_cases import nets_vs_golden
    # bf16 storage vs fp32 reference through ~25 conv layers: outputs 2e-2 of range, grads cos 0.995
    _check(nets_vs_golden('nets_dim16_64'), 2e-2, 0.995)


def test_nets_full_size_vs_reference_golden():
    from tests.model_cases import nets_vs_golden
    _check(nets_vs_golden('nets_dim64_256'), 2e-2, 0.995)


def test_training_step_vs_reference_golden():
    from tests.model_cases import step_vs_golden
    # 13 losses of two iterations within 2e-2 (relative, floor 1.0); per-phase gradient cosine
    _check(step_vs_golden(), 2e-2, 0.99)


def test_sampling_vs_oracle():
    from tests.model_cases import sampling_vs_oracle
    _check(sampling_vs_oracle(), 2e-2, 0.99)
