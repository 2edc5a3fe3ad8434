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


M64 = (1 << 64) - 1
ONES, HIGHS = 0x0101010101010101, 0x8080808080808080


def test_word_wise_tricks_of_the_parser_and_the_json_writer():
    """The byte-parallel arithmetic of csrc/ingest.cuh (hex8, the fixed-form date check) and csrc/json.cuh (hex4, dec2,
    JsonPacker's funnel shifts), restated on Python integers and checked against the obvious byte loops."""
    rng = np.random.default_rng(3)

    def hex8(x):  # eight ASCII hex digits -> four bytes, or None
        k7f = ONES * 0x7F
        d = x ^ (ONES * 0x30)
        l = (x | (ONES * 0x20)) ^ (ONES * 0x60)
        dig = ~((((d & k7f) + ONES * 0x76) & M64) | d) & HIGHS
        let = ~((((l & k7f) + ONES * 0x79) & M64) | l) & (((l & k7f) + k7f) & M64) & HIGHS
        if (dig | let) != HIGHS:
            return None
        nib = ((x & (ONES * 0x0F)) + ((let >> 7) * 9)) & M64
        y = ((nib << 4) | (nib >> 8)) & M64
        lo, hi = y & 0xFFFFFFFF, y >> 32
        return bytes([lo & 0xFF, (lo >> 16) & 0xFF, hi & 0xFF, (hi >> 16) & 0xFF])  # __byte_perm(lo, hi, 0x6420)

    for _ in range(3000):
        raw = bytes(rng.integers(0, 256, 4, dtype=np.uint8))
        text = raw.hex() if rng.integers(0, 2) else raw.hex().upper()
        assert hex8(int.from_bytes(text.encode(), "little")) == raw
    for c in range(256):  # one bad character in any position is refused; every hex digit is accepted
        ok = chr(c) in "0123456789abcdefABCDEF"
        for pos in range(8):
            t = bytearray(b"0a1B2c3D")
            t[pos] = c
            assert (hex8(int.from_bytes(bytes(t), "little")) is not None) == ok, (c, pos)

    def hex4(v):  # two bytes -> four lower-case hex characters
        t = ((v & 0xF0) >> 4) | ((v & 0x0F) << 8) | ((v & 0xF000) << 4) | ((v & 0x0F00) << 16)
        return (t + 0x30303030 + ((((t + 0x06060606) >> 4) & 0x01010101) * 39)) & 0xFFFFFFFF

    for v in range(0, 65536, 7):
        assert hex4(v).to_bytes(4, "little") == bytes([v & 0xFF, v >> 8]).hex().encode()
    for v in range(100):
        t = v // 10
        assert ((0x3030 + t) | ((v - 10 * t) << 8)).to_bytes(2, "little") == b"%02d" % v  # dec2

    # the date pattern: digits and separators checked per word, two-digit numbers by one multiplication
    p0, p1 = 0x2D30302D30303030, 0x30303A3030543030
    for _ in range(2000):
        s = bytearray(b"%04d-%02d-%02dT%02d:%02d:%02d" % tuple(int(x) for x in rng.integers(0, 100, 6) * [99, 1, 1, 1, 1, 1]))
        if rng.integers(0, 3) == 0:
            s[int(rng.integers(0, 16))] = int(rng.integers(32, 127))
        x0 = int.from_bytes(s[0:8], "little") ^ p0
        x1 = int.from_bytes(s[8:16], "little") ^ p1
        k7f = ONES * 0x7F
        bad = (((((x0 & k7f) + ONES * 0x76) & M64) | x0) | ((((x1 & k7f) + ONES * 0x76) & M64) | x1)) & HIGHS
        sep = (x0 & 0xFF0000FF00000000) | (x1 & 0x0000FF0000FF0000)
        want_ok = all(chr(s[i]).isdigit() for i in (0, 1, 2, 3, 5, 6, 8, 9, 11, 12, 14, 15)) and bytes([s[4], s[7], s[10], s[13]]) == b"--T:"
        assert (not (bad | sep)) == want_ok, bytes(s)
        if want_ok:
            t0 = (x0 * 10 + (x0 >> 8)) & M64
            t1 = (x1 * 10 + (x1 >> 8)) & M64
            got = (t0 & 0xFF, (t0 >> 16) & 0xFF, (t0 >> 40) & 0xFF, t1 & 0xFF, (t1 >> 24) & 0xFF, (t1 >> 48) & 0xFF)
            assert got == (int(s[0:2]), int(s[2:4]), int(s[5:7]), int(s[8:10]), int(s[11:13]), int(s[14:16]))

    # JsonPacker: words and 1..3-byte parts appended at any starting phase give the concatenation of the pieces
    def fsl(lo, hi, s):
        return ((hi << s) | (lo >> (32 - s))) & 0xFFFFFFFF if s else hi

    def fsr(lo, hi, s):
        return ((lo >> s) | (hi << (32 - s))) & 0xFFFFFFFF

    for _ in range(300):
        phase = int(rng.integers(0, 4))
        pieces = [bytes(rng.integers(1, 256, 4, dtype=np.uint8))] + [bytes(rng.integers(1, 256, int(rng.integers(1, 5)), dtype=np.uint8))
                                                                     for _ in range(int(rng.integers(0, 40)))]
        out = bytearray(b"\0" * (phase + sum(map(len, pieces)) + 8))
        first = int.from_bytes(pieces[0], "little")
        w0 = (first << (8 * phase)) & 0xFFFFFFFF
        for b in range(phase, 4):
            out[b] = (w0 >> (8 * b)) & 0xFF
        wp, prev, s = 4, first, 8 * phase
        for pc in pieces[1:]:
            v = int.from_bytes(pc, "little")
            if len(pc) == 4:
                out[wp:wp + 4] = fsl(prev, v, s).to_bytes(4, "little")
                wp, prev = wp + 4, v
            else:
                o = fsl(prev, v, s)
                prev = fsr(prev, v, 8 * len(pc))
                s += 8 * len(pc)
                if s >= 32:
                    out[wp:wp + 4] = o.to_bytes(4, "little")
                    wp, s = wp + 4, s - 32
        tail = prev >> (32 - s) if s else 0
        for b in range(s // 8):
            out[wp + b] = (tail >> (8 * b)) & 0xFF
        total = b"".join(pieces)
        assert bytes(out[phase:phase + len(total)]) == total
