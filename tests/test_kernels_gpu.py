"""Op-level parity of every CUDA entry point against plain PyTorch fp32 (GPU only).  Every case of
tests/kernel_cases.py runs (the list is the dictionary itself, so a case cannot be defined and left out)."""
import pytest

pytestmark = pytest.mark.gpu


def _cases():
    from tests.kernel_cases import CASES
    return sorted(CASES)


@pytest.mark.parametrize("name", _cases())
def test_kernel_case(name):
    from tests.kernel_cases import CASES, TOL
    res = CASES[name]()
    bad = {k: v for k, v in res.items() if not (v <= TOL)}
    assert not bad, f"{name}: relative errors above {TOL}: {bad} (all: {res})"
