"""The conveyor's ring groups: connections of equal delay share a group of at most 2^ACC_RING_RANK_BITS connections;
larger ones are split.  A build with groups of 128 connections runs golden cases -- a shocked run, passage forcing over multi-hop
routes, the output surface -- with every delay split into several groups and must still equal the reference's run bit for bit."""
import os
import subprocess

import pytest

import parity

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "acclimate_b200", "csrc")


@pytest.fixture(scope="module")
def split_lib():
    out = os.path.join(ROOT, "tests", "emu", "_build", "libacc_emu_rings7.so")
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    if not os.path.exists(out) or any(os.path.getmtime(f) > os.path.getmtime(out) for f in srcs):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.run(["/usr/bin/g++", "-x", "c++", "-std=c++17", "-O2", "-DACC_HOST_EMU", "-DACC_RING_RANK_BITS=7", "-fPIC", "-shared", "-fno-math-errno",
                        "-ffp-contract=off", os.path.join(CSRC, "engine.cu"), "-o", out], check=True, cwd=CSRC)
    return out


@pytest.mark.parametrize("case", ["small_shock12", "passage25", "routes25"])
def test_split_ring_groups_match_the_reference(case, split_lib):
    problems, launches = parity.run_case(case, lib_path=split_lib)
    assert not problems, "\n".join(problems[:10])
    assert launches > 0
