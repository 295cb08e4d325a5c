"""CPU tier: the kernel logic (host emulation of the same source) against the reference oracle's golden dumps."""
import pytest

import parity


@pytest.mark.parametrize("case", parity.CASES)
def test_kernel_logic_matches_reference(case, emu_lib):
    problems, _ = parity.run_case(case, lib_path=emu_lib)
    assert not problems, "\n".join(problems)
