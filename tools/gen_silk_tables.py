#!/usr/bin/env python3
"""Generate concentus_b200/csrc/silk_tables.h: the SILK decoder's static tables (normative data of
RFC 6716 as shipped in libopus 1.1.2), read out of the reference tree (read-only) and re-emitted in
our own layout -- one flat array per table, 2-D tables flattened row-major.  Run in the build
container only (needs /root/reference); the emitted header is committed.

Reference locations: silk/tables_other.c, tables_gain.c, tables_pitch_lag.c, pitch_est_tables.c,
tables_LTP.c, tables_pulses_per_block.c, tables_NLSF_CB_WB.c, tables_NLSF_CB_NB_MB.c,
table_LSF_cos.c, resampler_rom.c.
"""
import os, re, sys

REF = "/root/reference/opus-fix/silk"


def grab(path, name):
    src = open(os.path.join(REF, path)).read()
    m = re.search(r"\b" + re.escape(name) + r"\s*(\[[^\]]*\]\s*)*=\s*\{", src)
    assert m, name
    i = m.end()
    depth, j = 1, i
    while depth:
        c = src[j]
        depth += (c == "{") - (c == "}")
        j += 1
    body = src[i:j - 1]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    body = re.sub(r"#.*", "", body)
    # macro-valued entries (define.h:125-128)
    for k, v in (("OFFSET_VL_Q10", "32"), ("OFFSET_VH_Q10", "100"), ("OFFSET_UVL_Q10", "100"), ("OFFSET_UVH_Q10", "240")):
        body = body.replace(k, v)
    assert not re.search(r"[A-Za-z_]", body), (name, body[:200])
    return [int(t) for t in re.findall(r"-?\d+", body)]


def fmt(vals, per=16):
    return "\n".join("  " + ", ".join(str(v) for v in vals[i:i + per]) + "," for i in range(0, len(vals), per))


TABLES = [
    # (our name, ctype, file, reference name, expected length)
    ("stereo_pred_quant_Q13", "int16_t", "tables_other.c", "silk_stereo_pred_quant_Q13", 16),
    ("stereo_pred_joint_iCDF", "uint8_t", "tables_other.c", "silk_stereo_pred_joint_iCDF", 25),
    ("LBRR_flags_2_iCDF", "uint8_t", "tables_other.c", "silk_LBRR_flags_2_iCDF", 3),
    ("LBRR_flags_3_iCDF", "uint8_t", "tables_other.c", "silk_LBRR_flags_3_iCDF", 7),
    ("type_offset_VAD_iCDF", "uint8_t", "tables_other.c", "silk_type_offset_VAD_iCDF", 4),
    ("type_offset_no_VAD_iCDF", "uint8_t", "tables_other.c", "silk_type_offset_no_VAD_iCDF", 2),
    ("NLSF_interp_iCDF", "uint8_t", "tables_other.c", "silk_NLSF_interpolation_factor_iCDF", 5),
    ("quant_offsets_Q10", "int16_t", "tables_other.c", "silk_Quantization_Offsets_Q10", 4),
    ("LTPScales_Q14", "int16_t", "tables_other.c", "silk_LTPScales_table_Q14", 3),
    ("NLSF_EXT_iCDF", "uint8_t", "tables_other.c", "silk_NLSF_EXT_iCDF", 7),
    ("gain_iCDF", "uint8_t", "tables_gain.c", "silk_gain_iCDF", 24),
    ("delta_gain_iCDF", "uint8_t", "tables_gain.c", "silk_delta_gain_iCDF", 41),
    ("pitch_lag_iCDF", "uint8_t", "tables_pitch_lag.c", "silk_pitch_lag_iCDF", 32),
    ("pitch_delta_iCDF", "uint8_t", "tables_pitch_lag.c", "silk_pitch_delta_iCDF", 21),
    ("pitch_contour_iCDF", "uint8_t", "tables_pitch_lag.c", "silk_pitch_contour_iCDF", 34),
    ("pitch_contour_NB_iCDF", "uint8_t", "tables_pitch_lag.c", "silk_pitch_contour_NB_iCDF", 11),
    ("pitch_contour_10ms_iCDF", "uint8_t", "tables_pitch_lag.c", "silk_pitch_contour_10_ms_iCDF", 12),
    ("pitch_contour_10ms_NB_iCDF", "uint8_t", "tables_pitch_lag.c", "silk_pitch_contour_10_ms_NB_iCDF", 3),
    ("CB_lags_stage2_10ms", "int8_t", "pitch_est_tables.c", "silk_CB_lags_stage2_10_ms", 2 * 3),
    ("CB_lags_stage3_10ms", "int8_t", "pitch_est_tables.c", "silk_CB_lags_stage3_10_ms", 2 * 12),
    ("CB_lags_stage2", "int8_t", "pitch_est_tables.c", "silk_CB_lags_stage2", 4 * 11),
    ("CB_lags_stage3", "int8_t", "pitch_est_tables.c", "silk_CB_lags_stage3", 4 * 34),
    ("LTP_per_index_iCDF", "uint8_t", "tables_LTP.c", "silk_LTP_per_index_iCDF", 3),
    ("LTP_gain_iCDF_0", "uint8_t", "tables_LTP.c", "silk_LTP_gain_iCDF_0", 8),
    ("LTP_gain_iCDF_1", "uint8_t", "tables_LTP.c", "silk_LTP_gain_iCDF_1", 16),
    ("LTP_gain_iCDF_2", "uint8_t", "tables_LTP.c", "silk_LTP_gain_iCDF_2", 32),
    ("LTP_gain_vq_0", "int8_t", "tables_LTP.c", "silk_LTP_gain_vq_0", 40),
    ("LTP_gain_vq_1", "int8_t", "tables_LTP.c", "silk_LTP_gain_vq_1", 80),
    ("LTP_gain_vq_2", "int8_t", "tables_LTP.c", "silk_LTP_gain_vq_2", 160),
    ("pulses_per_block_iCDF", "uint8_t", "tables_pulses_per_block.c", "silk_pulses_per_block_iCDF", 180),
    ("rate_levels_iCDF", "uint8_t", "tables_pulses_per_block.c", "silk_rate_levels_iCDF", 18),
    ("shell_code_table0", "uint8_t", "tables_pulses_per_block.c", "silk_shell_code_table0", 152),
    ("shell_code_table1", "uint8_t", "tables_pulses_per_block.c", "silk_shell_code_table1", 152),
    ("shell_code_table2", "uint8_t", "tables_pulses_per_block.c", "silk_shell_code_table2", 152),
    ("shell_code_table3", "uint8_t", "tables_pulses_per_block.c", "silk_shell_code_table3", 152),
    ("shell_code_table_offsets", "uint8_t", "tables_pulses_per_block.c", "silk_shell_code_table_offsets", 17),
    ("sign_iCDF", "uint8_t", "tables_pulses_per_block.c", "silk_sign_iCDF", 42),
    ("NLSF_CB1_WB_Q8", "uint8_t", "tables_NLSF_CB_WB.c", "silk_NLSF_CB1_WB_Q8", 512),
    ("NLSF_CB1_iCDF_WB", "uint8_t", "tables_NLSF_CB_WB.c", "silk_NLSF_CB1_iCDF_WB", 64),
    ("NLSF_CB2_SELECT_WB", "uint8_t", "tables_NLSF_CB_WB.c", "silk_NLSF_CB2_SELECT_WB", 256),
    ("NLSF_CB2_iCDF_WB", "uint8_t", "tables_NLSF_CB_WB.c", "silk_NLSF_CB2_iCDF_WB", 72),
    ("NLSF_PRED_WB_Q8", "uint8_t", "tables_NLSF_CB_WB.c", "silk_NLSF_PRED_WB_Q8", 30),
    ("NLSF_DELTA_MIN_WB_Q15", "int16_t", "tables_NLSF_CB_WB.c", "silk_NLSF_DELTA_MIN_WB_Q15", 17),
    ("NLSF_CB1_NB_MB_Q8", "uint8_t", "tables_NLSF_CB_NB_MB.c", "silk_NLSF_CB1_NB_MB_Q8", 320),
    ("NLSF_CB1_iCDF_NB_MB", "uint8_t", "tables_NLSF_CB_NB_MB.c", "silk_NLSF_CB1_iCDF_NB_MB", 64),
    ("NLSF_CB2_SELECT_NB_MB", "uint8_t", "tables_NLSF_CB_NB_MB.c", "silk_NLSF_CB2_SELECT_NB_MB", 160),
    ("NLSF_CB2_iCDF_NB_MB", "uint8_t", "tables_NLSF_CB_NB_MB.c", "silk_NLSF_CB2_iCDF_NB_MB", 72),
    ("NLSF_PRED_NB_MB_Q8", "uint8_t", "tables_NLSF_CB_NB_MB.c", "silk_NLSF_PRED_NB_MB_Q8", 18),
    ("NLSF_DELTA_MIN_NB_MB_Q15", "int16_t", "tables_NLSF_CB_NB_MB.c", "silk_NLSF_DELTA_MIN_NB_MB_Q15", 11),
    ("LSFCosTab_Q12", "int16_t", "table_LSF_cos.c", "silk_LSFCosTab_FIX_Q12", 129),
    ("Resampler_3_4_COEFS", "int16_t", "resampler_rom.c", "silk_Resampler_3_4_COEFS", 29),
    ("Resampler_2_3_COEFS", "int16_t", "resampler_rom.c", "silk_Resampler_2_3_COEFS", 20),
    ("Resampler_1_2_COEFS", "int16_t", "resampler_rom.c", "silk_Resampler_1_2_COEFS", 14),
    ("resampler_frac_FIR_12", "int16_t", "resampler_rom.c", "silk_resampler_frac_FIR_12", 48),
]


def main():
    out = ["/* GENERATED by tools/gen_silk_tables.py -- do not edit.\n"
           " * Static tables of the SILK decoder (normative RFC 6716 data as shipped in libopus 1.1.2;\n"
           " * see the generator for the reference file of each table).  One SILK stream is decoded by\n"
           " * one thread, so every table is indexed per lane: all are CB_GTABLE (global, L1-cached). */\n"
           "#pragma once\n#include <stdint.h>\n"]
    for ours, ty, path, name, n in TABLES:
        vals = grab(path, name)
        assert len(vals) == n, (name, len(vals), n)
        # byte tables carry 4 zero bytes of padding: rd_icdf_g reads four entries ahead of its linear search
        pad = 4 if ty == "uint8_t" else 0
        out.append("CB_GTABLE %s st_%s[%d] = {\n%s\n};\n" % (ty, ours, n + pad, fmt(vals + [0] * pad)))
    # the two up2_hq coefficient triplets are written in the header as expressions; restate them
    hdr = open(os.path.join(REF, "resampler_rom.h")).read()
    m0 = re.search(r"silk_resampler_up2_hq_0\[ 3 \] = \{([^}]*)\}", hdr)
    m1 = re.search(r"silk_resampler_up2_hq_1\[ 3 \] = \{([^}]*)\}", hdr)
    assert m0 and m1
    def ev(s):
        return [int(eval(t.replace("(opus_int16)", ""))) for t in s.split(",")]
    v0, v1 = ev(m0.group(1)), ev(m1.group(1))
    v0 = [((v + 32768) % 65536) - 32768 for v in v0]
    v1 = [((v + 32768) % 65536) - 32768 for v in v1]
    out.append("CB_GTABLE int16_t st_resampler_up2_hq_0[3] = {%s};\n" % ", ".join(map(str, v0)))
    out.append("CB_GTABLE int16_t st_resampler_up2_hq_1[3] = {%s};\n" % ", ".join(map(str, v1)))
    here = os.path.dirname(os.path.abspath(__file__))
    p = os.path.join(here, "..", "concentus_b200", "csrc", "silk_tables.h")
    open(p, "w").write("\n".join(out))
    print("wrote", os.path.normpath(p), sum(len(x) for x in out), "bytes; up2_hq", v0, v1)


if __name__ == "__main__":
    main()
