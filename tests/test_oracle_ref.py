"""Pins the oracle: the compiled reference (oracle/_ref, libopus 1.1.2 FIXED_POINT built from the
sources under /root/reference) must reproduce the committed golden fixtures -- i.e. the fixtures
really are the reference's output -- and its decoder final range must equal the range the
reference ENCODER stored per packet (the check opus_demo itself makes, opus_demo.c:820-833)."""
import numpy as np
import pytest

from conftest import RefDecoder, crc32, golden_case, golden_index, needs_ref

CASES = sorted(golden_index().keys())


@needs_ref
@pytest.mark.parametrize("name", CASES)
def test_ref_reproduces_golden(name):
    meta, pk = golden_case(name)
    dec = RefDecoder(meta["fs"], meta["channels"])
    spf = meta["samples_per_frame"]
    for n, (p, rng) in enumerate(pk):
        ret, pcm = dec.decode_raw(p, spf)
        assert ret == spf
        assert dec.final_range == rng
        assert crc32(pcm[:ret]) == meta["pcm_crc32"][n]
