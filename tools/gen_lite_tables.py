#!/usr/bin/env python3
"""Transcribes the constant tables of the reference's lightweight SILK decoder (src/silk/icdf.rs, src/silk/codebook.rs)
into oracle/lite_tables_data.h for the oracle's restatement of src/silk/decoder.rs (row A2). Data only, verbatim; run in
the build container (reads /root/reference).   python tools/gen_lite_tables.py"""
import re
import sys
from pathlib import Path

REF = Path("/root/reference/src/silk")
OUT = Path(__file__).resolve().parent.parent / "oracle" / "lite_tables_data.h"
OUT_DEV = Path(__file__).resolve().parent.parent / "mousiki_b200" / "csrc" / "lite_tables_data.h"

ICDF_SINGLE = [
    "FRAME_TYPE_VAD_INACTIVE", "FRAME_TYPE_VAD_ACTIVE", "INDEPENDENT_QUANTIZATION_GAIN_MSB_INACTIVE",
    "INDEPENDENT_QUANTIZATION_GAIN_MSB_UNVOICED", "INDEPENDENT_QUANTIZATION_GAIN_MSB_VOICED",
    "INDEPENDENT_QUANTIZATION_GAIN_LSB", "DELTA_QUANTIZATION_GAIN",
    "NORMALIZED_LSF_STAGE_1_INDEX_NARROWBAND_OR_MEDIUMBAND_UNVOICED", "NORMALIZED_LSF_STAGE_1_INDEX_NARROWBAND_OR_MEDIUMBAND_VOICED",
    "NORMALIZED_LSF_STAGE_1_INDEX_WIDEBAND_UNVOICED", "NORMALIZED_LSF_STAGE_1_INDEX_WIDEBAND_VOICED",
    "NORMALIZED_LSF_STAGE_2_INDEX_EXTENSION", "NORMALIZED_LSF_INTERPOLATION_INDEX", "LINEAR_CONGRUENTIAL_GENERATOR_SEED",
    "RATE_LEVEL_UNVOICED", "RATE_LEVEL_VOICED", "EXCITATION_LSB", "PRIMARY_PITCH_LAG_HIGH_PART",
    "PRIMARY_PITCH_LAG_LOW_PART_NARROWBAND", "PRIMARY_PITCH_LAG_LOW_PART_MEDIUMBAND", "PRIMARY_PITCH_LAG_LOW_PART_WIDEBAND",
    "SUBFRAME_PITCH_CONTOUR_NARROWBAND20_MS", "SUBFRAME_PITCH_CONTOUR_MEDIUMBAND_OR_WIDEBAND20_MS", "PERIODICITY_INDEX",
    "LTP_FILTER_INDEX0", "LTP_FILTER_INDEX1", "LTP_FILTER_INDEX2", "LTP_SCALING_PARAMETER",
]
ICDF_ARRAY = ["NORMALIZED_LSF_STAGE_2_INDEX", "PULSE_COUNT", "PULSE_COUNT_SPLIT16_SAMPLE_PARTITIONS",
              "PULSE_COUNT_SPLIT8_SAMPLE_PARTITIONS", "PULSE_COUNT_SPLIT4_SAMPLE_PARTITIONS", "PULSE_COUNT_SPLIT2_SAMPLE_PARTITIONS"]
CODEBOOK_2D = [
    "NORMALIZED_LSF_STAGE_TWO_INDEX_NARROWBAND_OR_MEDIUMBAND", "NORMALIZED_LSF_STAGE_TWO_INDEX_WIDEBAND",
    "PREDICTION_WEIGHT_FOR_NARROWBAND_AND_MEDIUMBAND_NORMALIZED_LSF", "PREDICTION_WEIGHT_FOR_WIDEBAND_NORMALIZED_LSF",
    "PREDICTION_WEIGHT_SELECTION_FOR_NARROWBAND_AND_MEDIUMBAND_NORMALIZED_LSF", "PREDICTION_WEIGHT_SELECTION_FOR_WIDEBAND_NORMALIZED_LSF",
    "NORMALIZED_LSF_STAGE_ONE_NARROWBAND_OR_MEDIUMBAND", "NORMALIZED_LSF_STAGE_ONE_WIDEBAND",
    "CODEBOOK_LTP_FILTER_PERIODICITY_INDEX_0", "CODEBOOK_LTP_FILTER_PERIODICITY_INDEX_1", "CODEBOOK_LTP_FILTER_PERIODICITY_INDEX_2",
]
CODEBOOK_1D = [
    "LSF_ORDERING_FOR_POLYNOMIAL_EVALUATION_NARROWBAND_AND_MEDIUMBAND", "LSF_ORDERING_FOR_POLYNOMIAL_EVALUATION_WIDEBAND",
    "Q12_COSINE_TABLE_FOR_LSFCONVERION", "MINIMUM_SPACING_FOR_NORMALIZED_LSCOEFFICIENTS_NARROWBAND_AND_MEDIUMBAND",
    "MINIMUM_SPACING_FOR_NORMALIZED_LSCOEFFICIENTS_WIDEBAND",
]


def strip_comments(src):
    return re.sub(r"//[^\n]*", "", src)


def const_body(src, name):
    m = re.search(r"pub const " + name + r"\s*:[^=]*=\s*", src)
    if not m:
        raise SystemExit(f"constant {name} not found")
    i = m.end()
    depth = 0
    j = i
    while True:
        ch = src[j]
        if ch in "[{(":
            depth += 1
        elif ch in "]})":
            depth -= 1
        elif ch == ";" and depth == 0:
            break
        j += 1
    return src[i:j]


def icdf_contexts(body):
    out = []
    for m in re.finditer(r"icdf!\s*\{\s*(\d+)\s*;([^}]*)\}", body):
        out.append((int(m.group(1)), [int(x.replace("_", "")) for x in m.group(2).replace("\n", " ").split(",") if x.strip()]))
    return out


def nested_rows(body):
    body = body.strip()
    if body.startswith("&"):
        body = body[1:].strip()
    rows = re.findall(r"\[([^\[\]]*)\]", body[1:-1] if body.count("[") > 1 else body)
    return [[int(x.replace("_", "")) for x in r.replace("\n", " ").split(",") if x.strip()] for r in rows]


def main():
    icdf = strip_comments((REF / "icdf.rs").read_text())
    cb = strip_comments((REF / "codebook.rs").read_text())
    L = ["// Constant tables of mousiki's lightweight SILK decoder (src/silk/icdf.rs, src/silk/codebook.rs), transcribed verbatim by",
         "// tools/gen_lite_tables.py. Data only. Cumulative distributions as the reference stores them: {total, n, c[0..n)}.",
         "#pragma once", "#include <stdint.h>", "namespace lite_tab {", "struct Icdf { int total; int n; const uint16_t* c; };"]
    for name in ICDF_SINGLE:
        (total, vals), = icdf_contexts(const_body(icdf, name))
        L.append(f"static const uint16_t {name}_C[{len(vals)}] = {{{', '.join(map(str, vals))}}};")
        L.append(f"static const Icdf {name} = {{{total}, {len(vals)}, {name}_C}};")
    for name in ICDF_ARRAY:
        ctxs = icdf_contexts(const_body(icdf, name))
        for k, (total, vals) in enumerate(ctxs):
            L.append(f"static const uint16_t {name}_C{k}[{len(vals)}] = {{{', '.join(map(str, vals))}}};")
        L.append(f"static const Icdf {name}[{len(ctxs)}] = {{" + ", ".join(f"{{{t}, {len(v)}, {name}_C{k}}}" for k, (t, v) in enumerate(ctxs)) + "};")
    for name in CODEBOOK_2D:
        rows = nested_rows(const_body(cb, name))
        w = max(len(r) for r in rows)
        assert all(len(r) == w for r in rows), name
        L.append(f"static const int16_t {name}[{len(rows)}][{w}] = {{")
        for r in rows:
            L.append("  {" + ", ".join(map(str, r)) + "},")
        L.append("};")
    for name in CODEBOOK_1D:
        rows = nested_rows(const_body(cb, name))
        vals = [x for r in rows for x in r]
        L.append(f"static const int32_t {name}[{len(vals)}] = {{{', '.join(map(str, vals))}}};")
    L.append("}  // namespace lite_tab")
    OUT.write_text("\n".join(L) + "\n")
    print("wrote", OUT, len(L), "lines")
    # The product's copy: the same distributions in the inverse-cumulative byte form its range decoder reads
    # (icdf[i] = 256 - c[i], every total is 256), contexts of an array back to back with an offset table.
    D = ["// Constant tables of mousiki's lightweight SILK decoder (src/silk/icdf.rs, src/silk/codebook.rs) for the k_lite kernel,",
         "// written by tools/gen_lite_tables.py: distributions as inverse cumulative bytes (256 - c[i]), codebooks verbatim.",
         "#pragma once", "#include <stdint.h>", "namespace mb { namespace lite_tab {"]
    for name in ICDF_SINGLE:
        (total, vals), = icdf_contexts(const_body(icdf, name))
        assert total == 256 and vals[-1] == 256
        D.append(f"MB_TBL uint8_t {name}[{len(vals)}] = {{{', '.join(str(256 - v) for v in vals)}}};")
    for name in ICDF_ARRAY:
        ctxs = icdf_contexts(const_body(icdf, name))
        flat, offs = [], []
        for total, vals in ctxs:
            assert total == 256 and vals[-1] == 256
            offs.append(len(flat))
            flat += [256 - v for v in vals]
        D.append(f"MB_TBL uint8_t {name}[{len(flat)}] = {{{', '.join(map(str, flat))}}};")
        D.append(f"MB_TBL uint16_t {name}_OFF[{len(offs)}] = {{{', '.join(map(str, offs))}}};")
    for name in CODEBOOK_2D:
        rows = nested_rows(const_body(cb, name))
        w = len(rows[0])
        ty = "int8_t" if name.startswith("CODEBOOK_LTP") else "uint8_t"
        D.append(f"MB_TBL {ty} {name}[{len(rows)}][{w}] = {{")
        for r in rows:
            D.append("  {" + ", ".join(map(str, r)) + "},")
        D.append("};")
    for name in CODEBOOK_1D:
        rows = nested_rows(const_body(cb, name))
        vals = [x for r in rows for x in r]
        ty = "int16_t" if max(vals) < 32768 else "int32_t"
        D.append(f"MB_TBL {ty} {name}[{len(vals)}] = {{{', '.join(map(str, vals))}}};")
    D.append("} }  // namespace mb::lite_tab")
    OUT_DEV.write_text("\n".join(D) + "\n")
    print("wrote", OUT_DEV, len(D), "lines")


if __name__ == "__main__":
    sys.exit(main())
