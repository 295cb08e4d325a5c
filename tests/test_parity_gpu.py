"""GPU tier: libacclimate_b200.so through its C ABI against the reference oracle's golden dumps."""
import pytest

import parity


@pytest.mark.gpu
@pytest.mark.parametrize("case", parity.CASES)
def test_cuda_engine_matches_reference(case, cuda_lib):
    problems, launches = parity.run_case(case, lib_path=cuda_lib)
    assert not problems, "\n".join(problems)
    assert launches > 0
