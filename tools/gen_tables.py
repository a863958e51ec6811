#!/usr/bin/env python3
"""Transcribe the read-only CELT mode tables (pure data) from the reference tree.

Runs only in the build container (needs /root/reference); its two outputs are
committed so nothing at test/bench time reads the reference:

  oracle/celt_tables_data.h            (CPU oracle, test infrastructure)
  mousiki_b200/csrc/celt_tables_data.h (product: host arrays uploaded to the GPU)

Float tables are emitted as IEEE-754 bit patterns (uint32 hex) so that the
values are the *correctly rounded* f32 of the decimal literals in
  src/celt/window_48000_960.rs, mdct_twiddles_48000_960.rs, fft_twiddles_48000_960.rs
exactly as rustc would parse them.
"""
import re
import struct
import sys
from fractions import Fraction
from pathlib import Path

REF = Path("/root/reference/src/celt")
ROOT = Path(__file__).resolve().parent.parent


def f32_bits_exact(lit: str) -> int:
    """Correctly rounded decimal -> binary32 (round-half-even), no double rounding."""
    lit = lit.replace("_f32", "").replace("_", "")
    exact = Fraction(lit)
    d = float(lit)
    b = struct.unpack("<I", struct.pack("<f", d))[0]
    best = None
    for cand in (b - 1, b, b + 1):
        if cand < 0:
            continue
        cand &= 0xFFFFFFFF
        v = struct.unpack("<f", struct.pack("<I", cand))[0]
        if v != v or v in (float("inf"), float("-inf")):
            continue
        err = abs(Fraction(v) - exact)
        key = (err, cand & 1)
        if best is None or key < best[0]:
            best = (key, cand)
    return best[1]


def parse_floats(path: Path):
    txt = path.read_text()
    body = txt[txt.index("= [") :]
    return [f32_bits_exact(m) for m in re.findall(r"(-?\d[\d_]*\.?[\d_]*(?:e-?\d+)?)_f32", body)]


def parse_ints(path: Path, name: str):
    txt = path.read_text()
    m = re.search(name + r"[^=]*=\s*\[(.*?)\];", txt, re.S)
    assert m, name
    return [int(x) for x in re.findall(r"-?\d+", m.group(1))]


def emit_u32(name, vals, per=6):
    out = [f"static const uint32_t {name}[{len(vals)}] = {{"]
    for i in range(0, len(vals), per):
        out.append("  " + ", ".join(f"0x{v:08x}u" for v in vals[i : i + per]) + ",")
    out.append("};")
    return "\n".join(out)


def emit_int(ctype, name, vals, per=16):
    out = [f"static const {ctype} {name}[{len(vals)}] = {{"]
    for i in range(0, len(vals), per):
        out.append("  " + ", ".join(str(v) for v in vals[i : i + per]) + ",")
    out.append("};")
    return "\n".join(out)


def derived_tables():
    """PVQ codebook sizes U(n,k) (recurrence of src/celt/cwrs_pvq.rs:19-52, rows 0..14 packed) and
    LCG jump-ahead coefficients for celt_lcg_rand (seed_j = A_j*seed + C_j mod 2^32)."""
    row_len = [209, 208, 207, 206, 205, 204, 104, 54, 33, 21, 15, 10, 7, 4, 1]
    t = [[0] * 209 for _ in range(15)]
    t[0][0] = 1
    for n in range(1, 15):
        for k in range(1, 209):
            t[n][k] = min(t[n - 1][k] + t[n][k - 1] + t[n - 1][k - 1], (1 << 64) - 1)
    u = []
    for row in range(15):
        for col in range(row, row + row_len[row]):
            u.append(t[row][col] & 0xFFFFFFFF)
    assert len(u) == 1488
    a, c = 1, 0
    la, lc = [], []
    for _ in range(257):
        la.append(a); lc.append(c)
        a, c = (a * 1664525) & 0xFFFFFFFF, (c * 1664525 + 1013904223) & 0xFFFFFFFF
    def emit(name, vals, per=6):
        out = [f"MB_TBL uint32_t {name}[{len(vals)}] = {{"]
        for i in range(0, len(vals), per):
            out.append("  " + ", ".join(f"0x{v:08x}u" for v in vals[i : i + per]) + ",")
        out.append("};")
        return "\n".join(out)
    return ["// derived: U(n,k) rows 0..14 (cwrs_pvq.rs recurrence)", emit("kPvqU", u),
            "// derived: LCG jump-ahead, seed after j steps = kLcgA[j]*seed + kLcgC[j]", emit("kLcgA", la), emit("kLcgC", lc)]


def main():
    window = parse_floats(REF / "window_48000_960.rs")
    mdct_tw = parse_floats(REF / "mdct_twiddles_48000_960.rs")
    fft_tw = parse_floats(REF / "fft_twiddles_48000_960.rs")
    assert len(window) == 120 and len(mdct_tw) == 1800 and len(fft_tw) == 960
    sm = REF / "static_mode_48000_960.rs"
    log_n = parse_ints(sm, "LOG_N_400")
    cidx = parse_ints(sm, "CACHE_INDEX_50")
    cbits = parse_ints(sm, "CACHE_BITS_50")
    ccaps = parse_ints(sm, "CACHE_CAPS_50")
    br480 = parse_ints(sm, "FFT_BITREV_480_USIZE")
    br240 = parse_ints(sm, "FFT_BITREV_240_USIZE")
    br120 = parse_ints(sm, "FFT_BITREV_120_USIZE")
    br60 = parse_ints(sm, "FFT_BITREV_60_USIZE")
    # parse_ints picks up the array length too ("[usize; 480]"): drop leading dims
    def strip(v, n):
        return v[-n:]
    log_n, cidx, cbits, ccaps = strip(log_n, 21), strip(cidx, 105), strip(cbits, 392), strip(ccaps, 168)
    br480, br240, br120, br60 = strip(br480, 480), strip(br240, 240), strip(br120, 120), strip(br60, 60)
    modes = REF / "modes.rs"
    ebands = strip(parse_ints(modes, "EBAND_5MS"), 22)
    alloc = strip(parse_ints(modes, "BAND_ALLOCATION"), 231)
    qb = REF / "quant_bands.rs"
    txt = qb.read_text()
    m = re.search(r"E_PROB_MODEL[^=]*=\s*\[(.*?)\n\];", txt, re.S)
    eprob = [int(x) for x in re.findall(r"\d+", m.group(1))]
    assert len(eprob) == 4 * 2 * 42, len(eprob)
    assert sorted(br480) == list(range(480))

    hdr = [
        "// GENERATED by tools/gen_tables.py from the reference's static CELT mode data",
        "// (src/celt/{modes,static_mode_48000_960,window_48000_960,mdct_twiddles_48000_960,",
        "//  fft_twiddles_48000_960,quant_bands}.rs). Pure data; float tables are f32 bit patterns.",
        "#pragma once",
        "#include <stdint.h>",
        "",
        emit_int("int16_t", "kEBands5ms", ebands),
        emit_int("uint8_t", "kBandAllocation", alloc, 21),
        emit_int("int16_t", "kLogN400", log_n),
        emit_int("int16_t", "kCacheIndex50", cidx, 21),
        emit_int("uint8_t", "kCacheBits50", cbits, 20),
        emit_int("uint8_t", "kCacheCaps50", ccaps, 21),
        emit_int("uint8_t", "kEProbModel", eprob, 21),
        emit_int("int16_t", "kFftBitrev480", br480, 20),
        emit_int("int16_t", "kFftBitrev240", br240, 20),
        emit_int("int16_t", "kFftBitrev120", br120, 20),
        emit_int("int16_t", "kFftBitrev60", br60, 20),
        emit_u32("kWindow120Bits", window),
        emit_u32("kMdctTwiddles960Bits", mdct_tw),
        "// interleaved (re, im) pairs, 480 complex values",
        emit_u32("kFftTwiddles480Bits", fft_tw),
        "",
    ]
    text = "\n".join(hdr)
    # Product copy: same data behind the MB_TBL qualifier (device-resident arrays when compiled by
    # nvcc, plain host arrays in the CPU emulation build) plus two derived, reference-free tables.
    prod = text.replace("static const", "MB_TBL").replace("#include <stdint.h>", "#include <stdint.h>\n#ifndef MB_TBL\n#define MB_TBL static const\n#endif")
    prod += "\n".join(derived_tables()) + "\n"
    for out, t in ((ROOT / "oracle" / "celt_tables_data.h", text), (ROOT / "mousiki_b200" / "csrc" / "celt_tables_data.h", prod)):
        out.parent.mkdir(parents=True, exist_ok=True)
        out.write_text(t)
        print("wrote", out, len(t), "bytes")


if __name__ == "__main__":
    sys.exit(main())
