"""The C-ABI library loads without a GPU and exports every symbol include/acclimate_b200.h declares."""
import ctypes
import os
import re

import __graft_entry__ as g


def declared_symbols():
    text = open(os.path.join(g.ROOT, "include", "acclimate_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(acc_[a-z_0-9]+)\s*\(", text)))


def test_header_symbols_exported(cuda_lib):
    lib = ctypes.CDLL(cuda_lib)
    names = declared_symbols()
    assert "acc_create" in names and "acc_step" in names and len(names) >= 15
    for n in names:
        assert hasattr(lib, n), n


def test_create_fails_loudly_without_gpu(cuda_lib):
    import torch

    if torch.cuda.is_available():
        return
    import pytest

    from acclimate_b200 import engine, synth

    net = synth.make_network(regions=3, sectors=2, seed=1, const_delay=1)
    with pytest.raises(engine.EngineError, match="no CUDA device"):
        engine.Engine(net, lib_path=cuda_lib)
