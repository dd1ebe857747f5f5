"""The same robustness scenarios through the CUDA library's batch entry point: every stream of the
batch follows its own script (losses, DTX, FEC requests, multi-frame and damaged packets), all
streams are decoded by ONE opus_decode_batch call per step, and each must match the compiled
reference decoder fed the same calls (return value, PCM, final range)."""
import ctypes

import numpy as np
import pytest

import concentus_b200 as cb
from conftest import RefDecoder, needs_ref
from scenarios import fuzz_scenario, loss_scenario

pytestmark = pytest.mark.gpu

LOSS = ["c2_96k_cbr_st", "vbr_128k_st", "f10_64k_st", "f2_5_96k_st", "mono_12k", "low_24k_st",
        "st_to_mono_out", "mono_to_st_out", "st_to_24k_out", "mono_to_16k_out", "nb_20k_mono"]
FUZZ = ["c2_96k_cbr_st", "vbr_128k_st", "f5_64k_st", "f2_5_96k_st", "mono_64k", "low_24k_st",
        "st_to_24k_out", "mono_to_st_out"]


def _run_batch(scripts):
    lib = cb.load()
    n = len(scripts)
    metas = [m for m, _ in scripts]
    decs = [cb.OpusDecoder(m["fs"], m["channels"]) for m in metas]
    refs = [RefDecoder(m["fs"], m["channels"]) for m in metas]
    handles = (ctypes.c_void_p * n)(*[d._h for d in decs])
    steps = max(len(c) for _, c in scripts)
    try:
        for t in range(steps):
            # streams whose script is over, and streams with another frame_size, sit this call out:
            # one batch call has one frame_size, so group the streams by it
            by_fsz = {}
            for i, (_, calls) in enumerate(scripts):
                if t < len(calls):
                    by_fsz.setdefault(calls[t][1], []).append(i)
            for fsz, idx in sorted(by_fsz.items()):
                k = len(idx)
                bufs = []
                for i in idx:
                    d = scripts[i][1][t][0]
                    bufs.append(ctypes.create_string_buffer(d, len(d)) if d else None)
                data = (ctypes.c_void_p * k)(*[ctypes.cast(b, ctypes.c_void_p) if b is not None else None for b in bufs])
                lens = np.array([len(scripts[i][1][t][0]) if scripts[i][1][t][0] else 0 for i in idx], dtype=np.int32)
                fec = np.array([scripts[i][1][t][2] for i in idx], dtype=np.int32)
                outs = [np.zeros((fsz, metas[i]["channels"]), dtype=np.int16) for i in idx]
                pcm = (ctypes.c_void_p * k)(*[o.ctypes.data for o in outs])
                sub = (ctypes.c_void_p * k)(*[decs[i]._h for i in idx])
                ret = np.zeros(k, dtype=np.int32)
                rc = lib.opus_decode_batch(sub, data, lens.ctypes.data, pcm, fsz, fec.ctypes.data, k, ret.ctypes.data)
                assert rc == 0
                fr = np.zeros(k, dtype=np.uint32)
                assert lib.opus_decoder_get_final_range_batch(sub, k, fr.ctypes.data) == 0
                for j, i in enumerate(idx):
                    d, _, f = scripts[i][1][t]
                    ra, pa = refs[i].decode_raw(d, fsz, f)
                    assert ra == ret[j], "stream %d step %d: ret %d vs reference %d" % (i, t, ret[j], ra)
                    if ra > 0:
                        assert np.array_equal(pa[:ra], outs[j][:ra]), "stream %d step %d: pcm" % (i, t)
                    assert refs[i].final_range == int(fr[j]), "stream %d step %d: final range" % (i, t)
    finally:
        for d in decs:
            d.close()


@needs_ref
def test_gpu_loss_concealment_batch():
    scripts = []
    for rep in range(3):
        for name in LOSS:
            scripts.append(loss_scenario(name, seed=len(name) * 7 + 1 + 100 * rep))
    _run_batch(scripts)


@needs_ref
def test_gpu_multiframe_and_corruption_batch():
    scripts = []
    for rep in range(3):
        for name in FUZZ:
            scripts.append(fuzz_scenario(name, seed=len(name) + 3 + 100 * rep))
    _run_batch(scripts)
