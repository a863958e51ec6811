#!/usr/bin/env python3
"""Transcribe the read-only SILK decoder tables (pure data) from the reference tree.

Runs only in the build container (needs /root/reference); its two outputs are committed so nothing at test or
bench time reads the reference:

  oracle/silk_tables_data.h            (CPU oracle, test infrastructure)
  mousiki_b200/csrc/silk_tables_data.h (product: constant tables of the SILK kernels)

Multi-dimensional Rust arrays are flattened row-major.
"""
import re
import sys
from pathlib import Path

REF = Path("/root/reference/src/silk")
ROOT = Path(__file__).resolve().parent.parent

# (file, rust name, C name, C type)
TABLES = [
    ("tables_other.rs", "SILK_STEREO_PRED_QUANT_Q13", "stereo_pred_quant_Q13", "int16_t"),
    ("tables_other.rs", "SILK_STEREO_PRED_JOINT_ICDF", "stereo_pred_joint_iCDF", "uint8_t"),
    ("tables_other.rs", "SILK_STEREO_ONLY_CODE_MID_ICDF", "stereo_only_code_mid_iCDF", "uint8_t"),
    ("tables_other.rs", "SILK_LBRR_FLAGS_2_ICDF", "LBRR_flags_2_iCDF", "uint8_t"),
    ("tables_other.rs", "SILK_LBRR_FLAGS_3_ICDF", "LBRR_flags_3_iCDF", "uint8_t"),
    ("tables_other.rs", "SILK_LSB_ICDF", "lsb_iCDF", "uint8_t"),
    ("tables_other.rs", "SILK_LTPSCALE_ICDF", "LTPscale_iCDF", "uint8_t"),
    ("tables_other.rs", "SILK_TYPE_OFFSET_VAD_ICDF", "type_offset_VAD_iCDF", "uint8_t"),
    ("tables_other.rs", "SILK_TYPE_OFFSET_NO_VAD_ICDF", "type_offset_no_VAD_iCDF", "uint8_t"),
    ("tables_other.rs", "SILK_NLSF_INTERPOLATION_FACTOR_ICDF", "NLSF_interpolation_factor_iCDF", "uint8_t"),
    ("tables_other.rs", "SILK_LTPSCALES_TABLE_Q14", "LTPScales_table_Q14", "int16_t"),
    ("tables_other.rs", "SILK_UNIFORM3_ICDF", "uniform3_iCDF", "uint8_t"),
    ("tables_other.rs", "SILK_UNIFORM4_ICDF", "uniform4_iCDF", "uint8_t"),
    ("tables_other.rs", "SILK_UNIFORM5_ICDF", "uniform5_iCDF", "uint8_t"),
    ("tables_other.rs", "SILK_UNIFORM6_ICDF", "uniform6_iCDF", "uint8_t"),
    ("tables_other.rs", "SILK_UNIFORM8_ICDF", "uniform8_iCDF", "uint8_t"),
    ("tables_other.rs", "SILK_NLSF_EXT_ICDF", "NLSF_EXT_iCDF", "uint8_t"),
    ("tables_ltp.rs", "SILK_LTP_PER_INDEX_ICDF", "LTP_per_index_iCDF", "uint8_t"),
    ("tables_ltp.rs", "SILK_LTP_GAIN_ICDF_0", "LTP_gain_iCDF_0", "uint8_t"),
    ("tables_ltp.rs", "SILK_LTP_GAIN_ICDF_1", "LTP_gain_iCDF_1", "uint8_t"),
    ("tables_ltp.rs", "SILK_LTP_GAIN_ICDF_2", "LTP_gain_iCDF_2", "uint8_t"),
    ("tables_ltp.rs", "SILK_LTP_GAIN_VQ_0", "LTP_gain_vq_0", "int8_t"),
    ("tables_ltp.rs", "SILK_LTP_GAIN_VQ_1", "LTP_gain_vq_1", "int8_t"),
    ("tables_ltp.rs", "SILK_LTP_GAIN_VQ_2", "LTP_gain_vq_2", "int8_t"),
    ("tables_gain.rs", "SILK_GAIN_ICDF", "gain_iCDF", "uint8_t"),
    ("tables_gain.rs", "SILK_DELTA_GAIN_ICDF", "delta_gain_iCDF", "uint8_t"),
    ("tables_pitch_lag.rs", "PITCH_LAG_ICDF", "pitch_lag_iCDF", "uint8_t"),
    ("tables_pitch_lag.rs", "PITCH_DELTA_ICDF", "pitch_delta_iCDF", "uint8_t"),
    ("tables_pitch_lag.rs", "PITCH_CONTOUR_ICDF", "pitch_contour_iCDF", "uint8_t"),
    ("tables_pitch_lag.rs", "PITCH_CONTOUR_NB_ICDF", "pitch_contour_NB_iCDF", "uint8_t"),
    ("tables_pitch_lag.rs", "PITCH_CONTOUR_10_MS_ICDF", "pitch_contour_10_ms_iCDF", "uint8_t"),
    ("tables_pitch_lag.rs", "PITCH_CONTOUR_10_MS_NB_ICDF", "pitch_contour_10_ms_NB_iCDF", "uint8_t"),
    ("pitch_est_tables.rs", "SILK_CB_LAGS_STAGE2_10_MS", "CB_lags_stage2_10_ms", "int8_t"),
    ("pitch_est_tables.rs", "SILK_CB_LAGS_STAGE3_10_MS", "CB_lags_stage3_10_ms", "int8_t"),
    ("pitch_est_tables.rs", "SILK_CB_LAGS_STAGE2", "CB_lags_stage2", "int8_t"),
    ("pitch_est_tables.rs", "SILK_CB_LAGS_STAGE3", "CB_lags_stage3", "int8_t"),
    ("table_lsf_cos.rs", "SILK_LSF_COS_TAB_FIX_Q12", "LSFCosTab_FIX_Q12", "int16_t"),
    ("resampler_rom.rs", "SILK_RESAMPLER_UP2_HQ_0", "resampler_up2_hq_0", "int16_t"),
    ("resampler_rom.rs", "SILK_RESAMPLER_UP2_HQ_1", "resampler_up2_hq_1", "int16_t"),
    ("resampler_rom.rs", "SILK_RESAMPLER_FRAC_FIR_12", "resampler_frac_FIR_12", "int16_t"),
    ("resampler_rom.rs", "SILK_RESAMPLER_3_4_COEFS", "Resampler_3_4_COEFS", "int16_t"),
    ("resampler_rom.rs", "SILK_RESAMPLER_2_3_COEFS", "Resampler_2_3_COEFS", "int16_t"),
    ("resampler_rom.rs", "SILK_RESAMPLER_1_2_COEFS", "Resampler_1_2_COEFS", "int16_t"),
    ("resampler_rom.rs", "SILK_RESAMPLER_1_3_COEFS", "Resampler_1_3_COEFS", "int16_t"),
    ("resampler_rom.rs", "SILK_RESAMPLER_1_4_COEFS", "Resampler_1_4_COEFS", "int16_t"),
    ("resampler_rom.rs", "SILK_RESAMPLER_1_6_COEFS", "Resampler_1_6_COEFS", "int16_t"),
    ("tables_pulses_per_block.rs", "SILK_MAX_PULSES_TABLE", "max_pulses_table", "uint8_t"),
    ("tables_pulses_per_block.rs", "SILK_PULSES_PER_BLOCK_ICDF", "pulses_per_block_iCDF", "uint8_t"),
    ("tables_pulses_per_block.rs", "SILK_RATE_LEVELS_ICDF", "rate_levels_iCDF", "uint8_t"),
    ("tables_pulses_per_block.rs", "SILK_SHELL_CODE_TABLE0", "shell_code_table0", "uint8_t"),
    ("tables_pulses_per_block.rs", "SILK_SHELL_CODE_TABLE1", "shell_code_table1", "uint8_t"),
    ("tables_pulses_per_block.rs", "SILK_SHELL_CODE_TABLE2", "shell_code_table2", "uint8_t"),
    ("tables_pulses_per_block.rs", "SILK_SHELL_CODE_TABLE3", "shell_code_table3", "uint8_t"),
    ("tables_pulses_per_block.rs", "SILK_SHELL_CODE_TABLE_OFFSETS", "shell_code_table_offsets", "uint8_t"),
    ("tables_pulses_per_block.rs", "SILK_SIGN_ICDF", "sign_iCDF", "uint8_t"),
    ("tables_nlsf_cb_wb.rs", "SILK_NLSF_CB1_WB_Q8", "NLSF_CB1_WB_Q8", "uint8_t"),
    ("tables_nlsf_cb_wb.rs", "SILK_NLSF_CB1_WB_WGHT_Q9", "NLSF_CB1_WB_Wght_Q9", "int16_t"),
    ("tables_nlsf_cb_wb.rs", "SILK_NLSF_CB1_ICDF_WB", "NLSF_CB1_iCDF_WB", "uint8_t"),
    ("tables_nlsf_cb_wb.rs", "SILK_NLSF_CB2_SELECT_WB", "NLSF_CB2_SELECT_WB", "uint8_t"),
    ("tables_nlsf_cb_wb.rs", "SILK_NLSF_CB2_ICDF_WB", "NLSF_CB2_iCDF_WB", "uint8_t"),
    ("tables_nlsf_cb_wb.rs", "SILK_NLSF_PRED_WB_Q8", "NLSF_PRED_WB_Q8", "uint8_t"),
    ("tables_nlsf_cb_wb.rs", "SILK_NLSF_DELTA_MIN_WB_Q15", "NLSF_DELTA_MIN_WB_Q15", "int16_t"),
    ("tables_nlsf_cb_nb_mb.rs", "SILK_NLSF_CB1_NB_MB_Q8", "NLSF_CB1_NB_MB_Q8", "uint8_t"),
    ("tables_nlsf_cb_nb_mb.rs", "SILK_NLSF_CB1_NB_MB_WGHT_Q9", "NLSF_CB1_NB_MB_Wght_Q9", "int16_t"),
    ("tables_nlsf_cb_nb_mb.rs", "SILK_NLSF_CB1_ICDF_NB_MB", "NLSF_CB1_iCDF_NB_MB", "uint8_t"),
    ("tables_nlsf_cb_nb_mb.rs", "SILK_NLSF_CB2_SELECT_NB_MB", "NLSF_CB2_SELECT_NB_MB", "uint8_t"),
    ("tables_nlsf_cb_nb_mb.rs", "SILK_NLSF_CB2_ICDF_NB_MB", "NLSF_CB2_iCDF_NB_MB", "uint8_t"),
    ("tables_nlsf_cb_nb_mb.rs", "SILK_NLSF_PRED_NB_MB_Q8", "NLSF_PRED_NB_MB_Q8", "uint8_t"),
    ("tables_nlsf_cb_nb_mb.rs", "SILK_NLSF_DELTA_MIN_NB_MB_Q15", "NLSF_DELTA_MIN_NB_MB_Q15", "int16_t"),
]


def strip_comments(txt):
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return re.sub(r"//[^\n]*", "", txt)


def parse(path, name):
    txt = strip_comments((REF / path).read_text())
    m = re.search(r"\b(?:const|static)\s+" + name + r"\s*:\s*([^=]*)=\s*", txt)
    assert m, (path, name)
    i = m.end()
    assert txt[i] == "[", (path, name, txt[i : i + 20])
    depth, j = 0, i
    while True:
        if txt[j] == "[":
            depth += 1
        elif txt[j] == "]":
            depth -= 1
            if depth == 0:
                break
        j += 1
    body = txt[i : j + 1]
    vals = []
    for tok in re.findall(r"-?[A-Za-z_0-9]+", body):
        t = tok.replace("_", "") if re.fullmatch(r"-?[\d_]+", tok) else tok
        if re.fullmatch(r"-?\d+", t):
            vals.append(int(t))
        else:
            raise SystemExit(f"{path}:{name}: non-literal token {tok!r}")
    return vals, m.group(1).strip()


def main():
    out = ["// GENERATED by tools/gen_silk_tables.py from the reference's SILK table files (data only). Do not edit.",
           "#pragma once", "#include <stdint.h>", ""]
    for path, rname, cname, ctype in TABLES:
        vals, decl = parse(path, rname)
        out.append(f"// src/silk/{path} {rname}: {decl}")
        out.append(f"SILK_TABLE {ctype} silk_{cname}[{len(vals)}] = {{")
        for k in range(0, len(vals), 16):
            out.append("  " + ", ".join(str(v) for v in vals[k : k + 16]) + ",")
        out.append("};")
    out.append("// src/silk/tables_other.rs SILK_QUANTIZATION_OFFSETS_Q10 = [[UVL, UVH], [VL, VH]]")
    out.append("SILK_TABLE int16_t silk_Quantization_Offsets_Q10[4] = {100, 240, 32, 100};")
    out.append("// src/silk/resampler.rs:228 DELAY_MATRIX_DEC [in 8,12,16][out 8,12,16,24,48]")
    out.append("SILK_TABLE int8_t silk_delay_matrix_dec[15] = {4, 0, 2, 0, 0, 0, 9, 4, 7, 4, 0, 3, 12, 7, 7};")
    text = "\n".join(out) + "\n"
    for dst in (ROOT / "oracle" / "silk_tables_data.h", ROOT / "mousiki_b200" / "csrc" / "silk_tables_data.h"):
        dst.write_text(text)
        print("wrote", dst, len(text), "bytes")


if __name__ == "__main__":
    sys.exit(main())
