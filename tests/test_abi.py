"""CPU tests: the C-ABI library loads and exports every symbol include/smcsmc_b200.h declares; without a GPU every
compute entry point fails loudly (no CPU fallback); the .seg reader mirrors the reference's validation."""
import ctypes
import os
import re

import numpy as np
import pytest

import smcsmc_b200 as S
from smcsmc_b200 import _abi
from smcsmc_b200.segdata import read_seg, InvalidSeg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_header_symbols_exported():
    hdr = open(os.path.join(ROOT, "include", "smcsmc_b200.h")).read()
    names = set(re.findall(r"\b(smc_[A-Za-z_0-9]+)\s*\(", hdr))
    names -= {"smc_engine", "smc_status", "smc_config"}
    assert len(names) >= 20
    L = ctypes.CDLL(S.lib_path())
    for name in sorted(names):
        assert hasattr(L, name), "libsmcgpu.so does not export " + name
    assert names == set(_abi.SYMBOLS), names ^ set(_abi.SYMBOLS)
    assert b"sm_100a" in S.lib().smc_version()


def test_no_cpu_fallback():
    if _have_gpu():
        pytest.skip("GPU present")
    with pytest.raises(S.SmcError) as ei:
        S.fastexp(np.array([0.5]))
    assert ei.value.code == 2 and "no CPU fallback" in str(ei.value)
    with pytest.raises(S.SmcError):
        S.ParticleFilter(4, 10, [0.0], [1 / 2e4], 2.5e-8, 1e-8, 1e5, [1e4])


def test_product_does_not_touch_oracle():
    """nothing under smcsmc_b200/ may import, link or name the oracle"""
    for dp, _, fs in os.walk(os.path.join(ROOT, "smcsmc_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")) or f == "Makefile":
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "liboracle" not in txt and "orc_" not in txt and "_ref/" not in txt, os.path.join(dp, f)


def _write(tmp_path, text, name="t.seg"):
    p = tmp_path / name
    p.write_text(text)
    return str(p)


def test_read_seg_basic(tmp_path):
    p = _write(tmp_path, "#c\n1\t10\tT\tF\t1\t0101\n11\t5\tT\tF\t1\t.1/0\n".replace(".1/0", "..//") + "16\t84.0\tT\tF\t1\t....\n")
    rows = read_seg(p, 4, 100)
    assert list(rows.x_end) == [10.0, 15.0, 99.0]
    assert rows.alleles.tolist() == [[0, 1, 0, 1], [-1, -1, 2, 2], [-1, -1, -1, -1]]
    assert list(rows.is_site) == [1, 1, 1]
    assert rows.dist_to_mutation[0] == 0 and rows.dist_to_mutation[2] == 0.0


def test_read_seg_three_columns_and_split(tmp_path):
    p = _write(tmp_path, "1\t25\t01\n26\t5\t11\n")
    rows = read_seg(p, 2, 30, max_segment_length=10)
    # the 25-bp row is fed in pieces of 10 (INVARIANT_PARTIAL) and the remainder carries the site (segdata.cpp:125-144)
    assert list(rows.x_end) == [10.0, 20.0, 25.0, 30.0]
    assert list(rows.is_site) == [0, 0, 1, 1]


def test_read_seg_errors(tmp_path):
    with pytest.raises(InvalidSeg, match="not consecutive"):
        read_seg(_write(tmp_path, "1\t10\t01\n12\t5\t11\n"), 2, 100)
    with pytest.raises(InvalidSeg, match="Unknown character"):
        read_seg(_write(tmp_path, "1\t10\t0x\n"), 2, 100)
    with pytest.raises(InvalidSeg, match="inconsistent"):
        read_seg(_write(tmp_path, "1\t10\t0.\n"), 2, 100)
    with pytest.raises(InvalidSeg, match="Require 3"):
        read_seg(_write(tmp_path, "1\t10\n"), 2, 100)
    with pytest.raises(InvalidSeg, match="No data"):
        read_seg(_write(tmp_path, "500\t10\t01\n"), 2, 100)


def test_read_seg_windowing_and_missing_distance(tmp_path):
    txt = "1\t100\t01\n101\t400\t..\n501\t300\t..\n801\t100\t10\n901\t100\t..\n"
    rows = read_seg(_write(tmp_path, txt), 2, 1000)
    assert list(rows.dist_to_mutation) == [0.0, 0.0, 400.0, 0.0, 0.0]
    assert list(rows.max_epoch_to_record([2000.0, 600.0, 100.0])) == [2, 2, 0, 2, 2]
    rows = read_seg(_write(tmp_path, txt), 2, 500, data_start=201)
    assert rows.x_end[0] == 300.0 and rows.x_end[-1] == 500.0
