#!/usr/bin/env python3
"""Copies the known-answer vectors the reference holds for its lightweight decoder (row A2) into
tests/golden/lite_vectors.json: src/decoder.rs:275-283 (packet + first 12 samples), src/silk/decoder.rs:1805-2476 (Go
fixtures: two consecutive unvoiced frames -> 320 samples each, NLSF / LPC unit vectors, pitch-lag and LCG frames),
:2806-2857 (excitation). Run in the build container (reads /root/reference)."""
import json
import re
from pathlib import Path

REF = Path("/root/reference/src")
OUT = Path(__file__).resolve().parent.parent / "tests" / "golden" / "lite_vectors.json"


def const(src, name, kind=float):
    m = re.search(r"const " + name + r"\s*:[^=]*=\s*&?\[(.*?)\];", src, re.S)
    assert m, name
    body = re.sub(r"//[^\n]*", "", m.group(1))
    out = []
    for tok in body.replace("\n", " ").split(","):
        tok = tok.strip().replace("_", "")
        if not tok:
            continue
        tok = re.sub(r"(f32|i32|i16|u8)$", "", tok)
        out.append(int(tok, 0) if kind is int else float(tok))
    return out


silk = (REF / "silk" / "decoder.rs").read_text()
top = (REF / "decoder.rs").read_text()
v = {
    "packet": const(top, "TEST_PACKET", int),
    "packet_prefix": const(top, "EXPECTED_FLOATS_PREFIX"),
    "silk_frame": const(silk, "TEST_SILK_FRAME", int),
    "silk_frame_second": const(silk, "TEST_SILK_FRAME_SECOND", int),
    "decode_out_0": const(silk, "EXPECTED_DECODE_OUT_0"),
    "decode_out_1": const(silk, "EXPECTED_DECODE_OUT_1"),
    "res_q10": const(silk, "TEST_RES_Q_10", int),
    "nlsf_q15": const(silk, "TEST_NLSF_Q_15", int),
    "convert_nlsf_q15": const(silk, "TEST_CONVERT_NLSF_Q15", int),
    "a32_q17": const(silk, "EXPECTED_A32_Q17", int),
    "aq12": const(silk, "EXPECTED_AQ12"),
    "pitch_lag_frame": const(silk, "TEST_PITCH_LAG_FRAME", int),
    "lcg_frame": const(silk, "TEST_LCG_FRAME", int),
    "gains_q16": [210944.0, 112640.0, 96256.0, 96256.0],           # :2526
    "pitch_lags": [206, 206, 206, 206],                            # :2625
    "ltp_taps": [[1, 1, 8, 1, 1], [2, 0, 77, 11, 9], [1, 1, 8, 1, 1], [-1, 36, 64, 27, -6]],   # :2645-2650
    "stab_in": [[856, 2310, 3452, 4865, 4852, 7547, 9662, 11512, 13884, 15919, 18467, 20487, 23559, 25900, 28222, 30700],
                [1533, 1674, 2506, 4374, 6630, 9867, 10260, 10691, 14397, 16969, 19355, 21645, 25228, 26972, 30514, 30208]],
    "stab_out": [[856, 2310, 3452, 4858, 4861, 7547, 9662, 11512, 13884, 15919, 18467, 20487, 23559, 25900, 28222, 30700],
                 [1533, 1674, 2506, 4374, 6630, 9867, 10260, 10691, 14397, 16969, 19355, 21645, 25228, 26972, 30360, 30363]],
}
m = re.search(r"fn decode_excitation_matches_go_fixture\(\) \{\s*const EXPECTED: &\[i32\] = &\[(.*?)\];", silk, re.S)
v["excitation"] = [int(x) for x in m.group(1).replace("\n", " ").split(",") if x.strip()]
assert len(v["decode_out_0"]) == 320 and len(v["decode_out_1"]) == 320 and len(v["excitation"]) == 320
OUT.write_text(json.dumps(v))
print("wrote", OUT, OUT.stat().st_size, "bytes")
