"""CPU: the drop-in modules' HOST logic (weight hoisting, PNA backward coefficients, permutations,
state_dict keys) against the golden fixtures, with the C-ABI kernels replaced by the plain-PyTorch
emulation in tests/emulator.py.  The real kernels are checked by the -m gpu tests."""
import pytest
import torch

from tests.emulator import emulated_kernels
from tests.parity import check_fixture

NAMES = ["gine_eu", "gine_eu_rev", "gine_noeu", "pnas_eu", "pnas_eu_rev", "pnas_h64", "tabgnn", "interleaved", "fused_mcm", "fused_lp", "cpna", "cpnatab"]


@pytest.fixture(autouse=True)
def _single_thread():
    n = torch.get_num_threads()
    torch.set_num_threads(1)
    yield
    torch.set_num_threads(n)


@pytest.mark.parametrize("name", NAMES)
def test_dropin_host_logic(name):
    with emulated_kernels():
        worst = check_fixture(name, "cpu")
    assert worst
