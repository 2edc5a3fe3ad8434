"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol
include/tasktable.h declares, and fails loudly (no CPU fallback) when no GPU is present."""
import ctypes as C
import importlib
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pkg = importlib.import_module("aca-dotnet-workshop_b200")
tt = pkg.tasktable


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "tasktable.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tt_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = tt.lib()
    names = declared_symbols()
    assert len(names) >= 34
    for name in names:
        assert hasattr(L, name), f"{name} declared in include/tasktable.h but not exported"
    assert set(names) == set(tt.EXPORTS)


def test_struct_layouts_match_the_header():
    assert C.sizeof(tt.SynthCfg) == 56
    assert C.sizeof(tt.SweepStats) == 48
    assert C.sizeof(tt.SoaView) == 80
    assert C.sizeof(tt.DeviceView) == 72
    assert C.sizeof(tt.OpStats) == 32


def test_no_cpu_fallback_without_gpu():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(tt.TaskTableError) as e:
        tt.TaskTable(0)
    assert e.value.code == tt.TT_ERR_CUDA


def test_product_does_not_reference_the_oracle():
    # the oracle is test infrastructure: nothing under the package or include/ may name it
    bad = []
    for base in ("aca-dotnet-workshop_b200", "include"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".inc", ".h", ".hpp", ".cpp")):
                    s = open(os.path.join(dp, f), errors="replace").read()
                    if re.search(r"liboracle|oracle\.py|from oracle|import oracle|oracle/", s):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_bit_tricks_on_simulated_warp():
    """The nibble transpose + lane permutation of sweep.cuh, simulated over 32 lanes."""
    rng = np.random.default_rng(1)

    def step(x, y, lane, d):
        m = {4: 0x0000FFFF, 2: 0x00FF00FF, 1: 0x0F0F0F0F}[d]
        sh = 4 * d
        if lane & d:
            return (x & ~m & 0xFFFFFFFF) | ((y & ~m & 0xFFFFFFFF) >> sh)
        return (x & m) | (((y & m) << sh) & 0xFFFFFFFF)

    for _ in range(20):
        hit = rng.integers(0, 2, size=1024).astype(np.uint8)  # predicate per row of a warp's 1024 rows
        packed = [0] * 32
        for lane in range(32):
            for j in range(8):
                for k in range(4):
                    packed[lane] |= int(hit[128 * j + 4 * lane + k]) << (4 * j + k)
        for d in (4, 2, 1):
            packed = [step(packed[l], packed[l ^ d], l, d) for l in range(32)]
        words = [packed[8 * (t & 3) + (t >> 2)] for t in range(32)]
        want = [int(sum(int(hit[32 * t + b]) << b for b in range(32))) for t in range(32)]
        assert words == want


def test_due_code_packing_matches_reference_semantics():
    """code & mask == yesterday  <=>  the reference's equality (FAKE: ticks, STORE: JSON string)."""
    TPD = tt.TICKS_PER_DAY
    yd = tt.DEFAULT_TODAY // TPD - 1

    def pack(raw):
        ticks = raw & ((1 << 62) - 1)
        kind = raw >> 62
        day, tod = divmod(ticks, TPD)
        return day | (0x400000 if tod else 0) | (kind << 23)

    for kind in (0, 1, 2):
        for tod in (0, 1, TPD - 1):
            for day in (yd - 1, yd, yd + 1):
                raw = (day * TPD + tod) | (kind << 62)
                code = pack(raw)
                fake = (code & 0xFE7FFFFF) == yd
                store = (code & 0xFFFFFFFF) == yd
                assert fake == (day == yd and tod == 0)
                assert store == (day == yd and tod == 0 and kind == 0)
    assert (0xFFFFFFFF & 0xFE7FFFFF) != yd  # padding rows never match
