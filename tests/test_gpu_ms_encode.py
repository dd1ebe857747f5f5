"""Multistream encoder on the GPU against the compiled reference: same packets (bytes) and final range for
explicit layouts (coupled + mono streams), surround families 0 / 255, VBR and CBR, generous and tight output
buffers (the byte budget of a stream depends on the streams before it), plus the creation error codes."""
import ctypes

import numpy as np
import pytest

import concentus_b200 as cb
import enc_cases
from conftest import REF_SO, needs_ref

pytestmark = pytest.mark.gpu
LOWDELAY = enc_cases.LOWDELAY


def _ref():
    lib = ctypes.CDLL(REF_SO)
    vp = ctypes.c_void_p
    lib.opus_multistream_encoder_create.restype = vp
    lib.opus_multistream_encoder_create.argtypes = [ctypes.c_int32, ctypes.c_int, ctypes.c_int, ctypes.c_int, vp, ctypes.c_int, vp]
    lib.opus_multistream_surround_encoder_create.restype = vp
    lib.opus_multistream_surround_encoder_create.argtypes = [ctypes.c_int32, ctypes.c_int, ctypes.c_int, vp, vp, vp, ctypes.c_int, vp]
    lib.opus_multistream_encode.argtypes = [vp, vp, ctypes.c_int, vp, ctypes.c_int32]
    lib.opus_multistream_encoder_destroy.argtypes = [vp]
    return lib


def pcm_for(channels, frames, frame):
    cols = [enc_cases.signal(enc_cases.KINDS[(3 * c) % len(enc_cases.KINDS)], frame * frames, 1, 500 + c) for c in range(channels)]
    return np.ascontiguousarray(np.stack(cols, axis=1))          # [samples, channels] int16


def ctl_both(libs, handles, req, val):
    for lib, h in zip(libs, handles):
        assert lib.opus_multistream_encoder_ctl(ctypes.c_void_p(h), req, ctypes.c_int32(val)) == 0


def final_range(lib, h):
    v = ctypes.c_uint32(0)
    assert lib.opus_multistream_encoder_ctl(ctypes.c_void_p(h), 4031, ctypes.byref(v)) == 0
    return v.value


LAYOUTS = [  # channels, streams, coupled, mapping
    (4, 2, 2, [0, 1, 2, 3]),
    (3, 2, 1, [0, 1, 2]),
    (5, 3, 2, [0, 4, 1, 2, 3]),
    (2, 2, 0, [1, 0]),
]


@needs_ref
@pytest.mark.parametrize("layout", LAYOUTS)
@pytest.mark.parametrize("vbr,maxbytes,frame", [(1, 4000, 960), (0, 1500, 960), (1, 260, 480), (0, 400, 240), (1, 4000, 2880), (0, 2000, 1920)])
def test_explicit_layouts_match_reference(layout, vbr, maxbytes, frame):
    ch, streams, coupled, mapping = layout
    ours, ref = cb.load(), _ref()
    m = (ctypes.c_ubyte * ch)(*mapping)
    e1, e2 = ctypes.c_int(0), ctypes.c_int(0)
    a = ours.opus_multistream_encoder_create(48000, ch, streams, coupled, m, LOWDELAY, ctypes.byref(e1))
    b = ref.opus_multistream_encoder_create(48000, ch, streams, coupled, m, LOWDELAY, ctypes.byref(e2))
    assert a and b and e1.value == 0 and e2.value == 0
    try:
        ctl_both((ours, ref), (a, b), enc_cases.SET_BITRATE, 48000 * ch)
        ctl_both((ours, ref), (a, b), enc_cases.SET_VBR, vbr)
        ctl_both((ours, ref), (a, b), enc_cases.SET_COMPLEXITY, 7)
        nframes = 12
        pcm = pcm_for(ch, nframes, frame)
        for t in range(nframes):
            x = np.ascontiguousarray(pcm[t * frame:(t + 1) * frame])
            o1 = ctypes.create_string_buffer(maxbytes)
            o2 = ctypes.create_string_buffer(maxbytes)
            n1 = ours.opus_multistream_encode(a, x.ctypes.data, frame, o1, maxbytes)
            n2 = ref.opus_multistream_encode(b, x.ctypes.data, frame, o2, maxbytes)
            assert n1 == n2, (layout, vbr, maxbytes, t, n1, n2)
            assert n1 > 0 and o1.raw[:n1] == o2.raw[:n2], (layout, vbr, maxbytes, t)
            assert final_range(ours, a) == final_range(ref, b), (layout, t)
        # the packets are decodable multistream packets of the right duration
        assert cb.load().opus_packet_get_nb_frames(o1, n1) >= 1
    finally:
        ours.opus_multistream_encoder_destroy(a)
        ref.opus_multistream_encoder_destroy(b)


@needs_ref
@pytest.mark.parametrize("family,ch", [(0, 1), (0, 2), (255, 3), (1, 2)])
def test_surround_create_families_without_analysis(family, ch):
    ours, ref = cb.load(), _ref()
    res = []
    for lib in (ours, ref):
        streams, coupled, err = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_int(0)
        mapping = (ctypes.c_ubyte * 8)()
        h = lib.opus_multistream_surround_encoder_create(48000, ch, family, ctypes.byref(streams), ctypes.byref(coupled),
                                                         mapping, LOWDELAY, ctypes.byref(err))
        assert h and err.value == 0
        res.append((h, streams.value, coupled.value, list(mapping)[:ch]))
    assert res[0][1:] == res[1][1:]
    a, b = res[0][0], res[1][0]
    try:
        ctl_both((ours, ref), (a, b), enc_cases.SET_BITRATE, 64000 * ch)
        pcm = pcm_for(ch, 6, 960)
        for t in range(6):
            x = np.ascontiguousarray(pcm[t * 960:(t + 1) * 960])
            o1, o2 = ctypes.create_string_buffer(4000), ctypes.create_string_buffer(4000)
            n1 = ours.opus_multistream_encode(a, x.ctypes.data, 960, o1, 4000)
            n2 = ref.opus_multistream_encode(b, x.ctypes.data, 960, o2, 4000)
            assert n1 == n2 and n1 > 0 and o1.raw[:n1] == o2.raw[:n2], (family, ch, t)
    finally:
        ours.opus_multistream_encoder_destroy(a)
        ref.opus_multistream_encoder_destroy(b)


def test_creation_errors():
    ours = cb.load()
    err = ctypes.c_int(0)
    streams, coupled = ctypes.c_int(0), ctypes.c_int(0)
    mapping = (ctypes.c_ubyte * 8)()
    # 5.1 surround needs the masking analysis and the LFE stream: refused, not approximated
    assert not ours.opus_multistream_surround_encoder_create(48000, 6, 1, ctypes.byref(streams), ctypes.byref(coupled),
                                                             mapping, LOWDELAY, ctypes.byref(err))
    assert err.value == cb.OPUS_UNIMPLEMENTED
    bad = (ctypes.c_ubyte * 3)(0, 1, 1)                 # stream 1 (mono) has no channel: mapping never names index 2
    assert not ours.opus_multistream_encoder_create(48000, 3, 2, 1, bad, LOWDELAY, ctypes.byref(err))
    assert err.value == cb.OPUS_BAD_ARG
    assert not ours.opus_multistream_encoder_create(48000, 2, 1, 2, bad, LOWDELAY, ctypes.byref(err))
    assert err.value == cb.OPUS_BAD_ARG
