"""GPU tier: libacclimate_b200.so through its C ABI against the reference oracle's golden dumps."""
import pytest

import parity


@pytest.mark.gpu
@pytest.mark.parametrize("case", parity.CASES)
def test_cuda_engine_matches_reference(case, cuda_lib):
    problems, launches = parity.run_case(case, lib_path=cuda_lib)
    assert not problems, "\n".join(problems)
    assert launches > 0


@pytest.mark.gpu
def test_full_size_flood_matches_reference(cuda_lib):
    """BASELINE.json configs[2] at FULL size (187 x 26, 24.5 M connections) through the shock onset: every array the oracle
    dumps, per connection ones included, as position-weighted checksums (int64 exact) of the reference's own serial run
    (tests/golden/make_golden.py eora_flood4; the 7 GB dump itself is not kept)."""
    import os

    if not os.path.exists(parity.fixture_path("eora_flood4")):
        pytest.fail("tests/golden/eora_flood4.shim.acc.gz is missing: generate it with `python tests/golden/make_golden.py eora_flood4`")
    problems, launches = parity.run_case("eora_flood4", lib_path=cuda_lib)
    assert not problems, "\n".join(problems)
    assert launches > 0
