"""Multistream encoder on the GPU against the compiled reference: same packets (bytes) and final range for
explicit layouts (coupled + mono streams), surround families 0 / 255 / 1 (3..8 channels: masking analysis on the
device, LFE stream), VBR and CBR, generous and tight output buffers (the byte budget of a stream depends on the
streams before it), plus the creation error codes."""
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


def surround_pcm(channels, frames, frame, fs):
    """Channels of different material and level (some silent for a while), so that the masks move."""
    cols = []
    for c in range(channels):
        x = enc_cases.signal(enc_cases.KINDS[(2 * c + 1) % len(enc_cases.KINDS)], frame * frames, 1, 900 + c).astype(np.float64)
        g = np.ones(frame * frames)
        for t in range(frames):
            g[t * frame:(t + 1) * frame] = [1.0, 0.5, 0.1, 0.02, 0.0, 1.0][(t + 2 * c) % 6]
        cols.append((x * g).astype(np.int16))
    return np.ascontiguousarray(np.stack(cols, axis=1))


@needs_ref
@pytest.mark.parametrize("ch", [3, 4, 5, 6, 7, 8])
@pytest.mark.parametrize("app,fs,vbr,cvbr,rate_per_ch,maxbytes,frame_ms", [
    (LOWDELAY, 48000, 1, 1, 42000, 4000, 20), (2049, 48000, 1, 1, 48000, 4000, 20), (2049, 48000, 0, 0, 40000, 1500, 20),
    (LOWDELAY, 48000, 1, 0, 24000, 600, 10), (LOWDELAY, 48000, 1, 1, 64000, 4000, 40), (2049, 24000, 1, 1, 56000, 4000, 20),
    (LOWDELAY, 48000, 0, 0, 30000, 900, 5), (LOWDELAY, 48000, 1, 1, 6500, 4000, 20), (2049, 48000, 1, 1, 96000, 4000, 60)])
def test_surround_family1_matches_reference(ch, app, fs, vbr, cvbr, rate_per_ch, maxbytes, frame_ms):
    """opus_multistream_surround_encoder_create(family 1, 3..8 channels): vorbis channel order, surround masking
    analysis (surround_analysis, on the device) feeding OPUS_SET_ENERGY_MASK of every stream, the LFE stream of
    5.1 / 6.1 / 7.1, forced stereo CELT on coupled streams, bandwidth from the total rate.  Packets byte-identical
    with the reference, final range too; a reset in the middle."""
    ours, ref = cb.load(), _ref()
    res = []
    for lib in (ours, ref):
        streams, coupled, err = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_int(0)
        mapping = (ctypes.c_ubyte * 8)()
        h = lib.opus_multistream_surround_encoder_create(fs, ch, 1, ctypes.byref(streams), ctypes.byref(coupled),
                                                         mapping, app, ctypes.byref(err))
        assert h and err.value == 0
        res.append((h, streams.value, coupled.value, list(mapping)[:ch]))
    assert res[0][1:] == res[1][1:]
    a, b = res[0][0], res[1][0]
    try:
        ctl_both((ours, ref), (a, b), enc_cases.SET_BITRATE, rate_per_ch * ch)
        ctl_both((ours, ref), (a, b), enc_cases.SET_VBR, vbr)
        ctl_both((ours, ref), (a, b), 4020, cvbr)
        frame = fs * frame_ms // 1000
        nframes = 10
        pcm = surround_pcm(ch, nframes, frame, fs)
        for t in range(nframes):
            if t == 6:
                for lib, h in ((ours, a), (ref, b)):
                    assert lib.opus_multistream_encoder_ctl(ctypes.c_void_p(h), 4028) == 0      # OPUS_RESET_STATE
            x = np.ascontiguousarray(pcm[t * frame:(t + 1) * frame])
            o1, o2 = ctypes.create_string_buffer(maxbytes), ctypes.create_string_buffer(maxbytes)
            n1 = ours.opus_multistream_encode(a, x.ctypes.data, frame, o1, maxbytes)
            n2 = ref.opus_multistream_encode(b, x.ctypes.data, frame, o2, maxbytes)
            assert n1 == n2, (ch, app, t, n1, n2)
            assert n1 > 0 and o1.raw[:n1] == o2.raw[:n2], (ch, app, t)
            assert final_range(ours, a) == final_range(ref, b), (ch, t)
    finally:
        ours.opus_multistream_encoder_destroy(a)
        ref.opus_multistream_encoder_destroy(b)


def test_creation_errors():
    ours = cb.load()
    err = ctypes.c_int(0)
    streams, coupled = ctypes.c_int(0), ctypes.c_int(0)
    mapping = (ctypes.c_ubyte * 8)()
    # families other than 0 / 1 / 255, family 1 above 8 channels, family 0 above stereo (:488-537)
    for fam, ch in ((2, 6), (1, 9), (0, 3)):
        big = (ctypes.c_ubyte * 16)()
        assert not ours.opus_multistream_surround_encoder_create(48000, ch, fam, ctypes.byref(streams), ctypes.byref(coupled),
                                                                 big, LOWDELAY, ctypes.byref(err))
        assert err.value == cb.OPUS_UNIMPLEMENTED
    bad = (ctypes.c_ubyte * 3)(0, 1, 1)                 # stream 1 (mono) has no channel: mapping never names index 2
    assert not ours.opus_multistream_encoder_create(48000, 3, 2, 1, bad, LOWDELAY, ctypes.byref(err))
    assert err.value == cb.OPUS_BAD_ARG
    assert not ours.opus_multistream_encoder_create(48000, 2, 1, 2, bad, LOWDELAY, ctypes.byref(err))
    assert err.value == cb.OPUS_BAD_ARG


def test_surround_float_entry_equals_int16_entry():
    """opus_multistream_encode_float on a FIXED_POINT build is FLOAT2INT16 followed by the int16 path
    (opus_multistream_encoder.c:919-936, 1006-1022): two 5.1 surround encoders fed the same audio through the two
    entry points produce the same packets."""
    ours = cb.load()
    hs = []
    for _ in range(2):
        streams, coupled, err = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_int(0)
        mapping = (ctypes.c_ubyte * 8)()
        h = ours.opus_multistream_surround_encoder_create(48000, 6, 1, ctypes.byref(streams), ctypes.byref(coupled),
                                                          mapping, LOWDELAY, ctypes.byref(err))
        assert h and err.value == 0 and (streams.value, coupled.value) == (4, 2)
        hs.append(h)
    try:
        for h in hs:
            assert ours.opus_multistream_encoder_ctl(ctypes.c_void_p(h), enc_cases.SET_BITRATE, ctypes.c_int32(256000)) == 0
        pcm = surround_pcm(6, 6, 960, 48000)
        for t in range(6):
            x = np.ascontiguousarray(pcm[t * 960:(t + 1) * 960])
            xf = np.ascontiguousarray(x.astype(np.float32) / 32768.0)
            o1, o2 = ctypes.create_string_buffer(4000), ctypes.create_string_buffer(4000)
            n1 = ours.opus_multistream_encode(hs[0], x.ctypes.data, 960, o1, 4000)
            n2 = ours.opus_multistream_encode_float(hs[1], xf.ctypes.data, 960, o2, 4000)
            assert n1 == n2 and n1 > 0 and o1.raw[:n1] == o2.raw[:n2], t
    finally:
        for h in hs:
            ours.opus_multistream_encoder_destroy(h)
