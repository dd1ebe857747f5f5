"""Repacketizer / padding entry points (host-side, no GPU needed): byte-identical with the compiled reference
on packets from the committed fixtures -- merged, split, padded and unpadded -- including the error codes.
Follows the checks of the reference's own tests/test_opus_padding.c and the repacketizer part of
tests/test_opus_api.c (toc mismatch, 120 ms limit, buffer too small, out_range bounds)."""
import ctypes
import os

import numpy as np
import pytest

import concentus_b200 as cb
from conftest import GOLDEN, REF_SO, golden_case, needs_ref, read_bit, silk_golden_index
from test_multistream import ms_case


def _proto(lib):
    vp = ctypes.c_void_p
    lib.opus_repacketizer_create.restype = vp
    lib.opus_repacketizer_init.restype = vp
    lib.opus_repacketizer_init.argtypes = [vp]
    lib.opus_repacketizer_destroy.argtypes = [vp]
    lib.opus_repacketizer_cat.argtypes = [vp, vp, ctypes.c_int32]
    lib.opus_repacketizer_get_nb_frames.argtypes = [vp]
    lib.opus_repacketizer_out.argtypes = [vp, vp, ctypes.c_int32]
    lib.opus_repacketizer_out_range.argtypes = [vp, ctypes.c_int, ctypes.c_int, vp, ctypes.c_int32]
    lib.opus_packet_pad.argtypes = [vp, ctypes.c_int32, ctypes.c_int32]
    lib.opus_packet_unpad.argtypes = [vp, ctypes.c_int32]
    lib.opus_multistream_packet_pad.argtypes = [vp, ctypes.c_int32, ctypes.c_int32, ctypes.c_int]
    lib.opus_multistream_packet_unpad.argtypes = [vp, ctypes.c_int32, ctypes.c_int]
    return lib


@pytest.fixture(scope="module")
def libs():
    ours = cb.load()
    ref = _proto(ctypes.CDLL(REF_SO)) if os.path.exists(REF_SO) else None
    return ours, ref


def streams():
    """(name, packets) of different configurations: CELT 20 ms CBR, CELT 2.5 ms, SILK 20 ms, hybrid, SILK 60 ms."""
    out = []
    for name in ("c2_96k_cbr_st",):
        out.append((name, [p for p, _ in golden_case(name)[1]]))
    idx = silk_golden_index()
    for name in ("silk_wb_mono_32k", "hyb_fb_mono_32k", "silk_wb_mono_60ms", "silk_wb_mono_10ms", "trans_all_st_to24"):
        assert name in idx
        out.append((name, [p for p, _ in read_bit(os.path.join(GOLDEN, name + ".bit"))]))
    return out


class Rp(object):
    def __init__(self, lib):
        self.lib, self.h = lib, lib.opus_repacketizer_create()
        self.keep = []

    def cat(self, pkt):
        buf = ctypes.create_string_buffer(pkt, len(pkt))
        self.keep.append(buf)                      # the repacketizer keeps pointers into the caller's packets
        return self.lib.opus_repacketizer_cat(self.h, buf, len(pkt))

    def out(self, maxlen, rng=None):
        o = ctypes.create_string_buffer(max(1, maxlen))
        if rng is None:
            n = self.lib.opus_repacketizer_out(self.h, o, maxlen)
        else:
            n = self.lib.opus_repacketizer_out_range(self.h, rng[0], rng[1], o, maxlen)
        return n, o.raw[:max(n, 0)]

    def reset(self):
        self.lib.opus_repacketizer_init(self.h)
        self.keep = []

    def close(self):
        self.lib.opus_repacketizer_destroy(self.h)


@needs_ref
def test_merge_and_split_match_reference(libs):
    ours, ref = libs
    rng = np.random.default_rng(3)
    for name, pk in streams():
        a, b = Rp(ours), Rp(ref)
        try:
            for trial in range(60):
                a.reset(); b.reset()
                k = int(rng.integers(1, 9))
                start = int(rng.integers(0, len(pk) - k))
                rets = []
                for p in pk[start:start + k]:
                    r1, r2 = a.cat(p), b.cat(p)
                    assert r1 == r2, (name, trial)
                    rets.append(r1)
                n1 = ours.opus_repacketizer_get_nb_frames(a.h)
                assert n1 == ref.opus_repacketizer_get_nb_frames(b.h)
                if n1 == 0:
                    continue
                for maxlen in (4000, 1277, int(rng.integers(1, 400))):
                    assert a.out(maxlen) == b.out(maxlen), (name, trial, maxlen)
                for _ in range(4):
                    lo = int(rng.integers(-1, n1 + 1))
                    hi = int(rng.integers(-1, n1 + 2))
                    assert a.out(3000, (lo, hi)) == b.out(3000, (lo, hi)), (name, trial, lo, hi)
        finally:
            a.close(); b.close()


@needs_ref
def test_mismatched_configurations_and_limits(libs):
    ours, ref = libs
    s = dict(streams())
    celt, silk = s["c2_96k_cbr_st"], s["silk_wb_mono_32k"]
    for lib in (ours, ref):
        r = Rp(lib)
        assert r.cat(celt[0]) == 0
        assert r.cat(silk[0]) == cb.OPUS_INVALID_PACKET          # different TOC configuration
        assert r.cat(b"") == cb.OPUS_INVALID_PACKET
        for i in range(1, 6):
            assert r.cat(celt[i]) == 0
        assert r.cat(celt[7]) == cb.OPUS_INVALID_PACKET          # 7 x 20 ms > 120 ms
        assert lib.opus_repacketizer_get_nb_frames(r.h) == 6
        assert r.out(10)[0] == cb.OPUS_BUFFER_TOO_SMALL
        assert r.out(3000, (2, 2))[0] == cb.OPUS_BAD_ARG
        assert r.out(3000, (0, 7))[0] == cb.OPUS_BAD_ARG
        r.close()


@needs_ref
def test_pad_and_unpad_match_reference(libs):
    ours, ref = libs
    rng = np.random.default_rng(5)
    for name, pk in streams():
        for trial in range(40):
            p = pk[int(rng.integers(0, len(pk)))]
            for extra in (0, 1, 2, 3, 254, 255, 256, 257, 511, int(rng.integers(1, 900))):
                new_len = len(p) + extra
                b1 = ctypes.create_string_buffer(p + bytes(extra), new_len)
                b2 = ctypes.create_string_buffer(p + bytes(extra), new_len)
                r1 = ours.opus_packet_pad(b1, len(p), new_len)
                r2 = ref.opus_packet_pad(b2, len(p), new_len)
                assert r1 == r2 == 0 and b1.raw == b2.raw, (name, trial, extra)
                u1 = ours.opus_packet_unpad(b1, new_len)
                u2 = ref.opus_packet_unpad(b2, new_len)
                assert u1 == u2 and b1.raw[:u1] == b2.raw[:u2], (name, trial, extra)
                assert u1 <= len(p)
    for lib in (ours, ref):
        b = ctypes.create_string_buffer(b"\x00" * 8, 8)
        assert lib.opus_packet_pad(b, 0, 8) == cb.OPUS_BAD_ARG
        assert lib.opus_packet_pad(b, 8, 4) == cb.OPUS_BAD_ARG
        assert lib.opus_packet_unpad(b, 0) == cb.OPUS_BAD_ARG


@needs_ref
def test_multistream_pad_and_unpad_match_reference(libs):
    ours, ref = libs
    for case in ("ms51_256k", "ms51_80k_voip"):
        meta, pk = ms_case(case)
        for p, _ in pk[:60]:
            if not p:
                continue
            for extra in (1, 2, 255, 300):
                new_len = len(p) + extra
                b1 = ctypes.create_string_buffer(p + bytes(extra), new_len)
                b2 = ctypes.create_string_buffer(p + bytes(extra), new_len)
                r1 = ours.opus_multistream_packet_pad(b1, len(p), new_len, meta["streams"])
                r2 = ref.opus_multistream_packet_pad(b2, len(p), new_len, meta["streams"])
                assert r1 == r2 == 0 and b1.raw == b2.raw, (case, extra)
                u1 = ours.opus_multistream_packet_unpad(b1, new_len, meta["streams"])
                u2 = ref.opus_multistream_packet_unpad(b2, new_len, meta["streams"])
                assert u1 == u2 and b1.raw[:u1] == b2.raw[:u2], (case, extra)


def test_pad_then_unpad_is_identity_on_code0_packets(libs):
    """Property that needs no reference: a one-frame packet survives pad -> unpad unchanged, and the padded
    packet still parses to the same single frame."""
    ours, _ = libs
    pk = [p for p, _ in golden_case("c2_96k_cbr_st")[1]][:50]
    for p in pk:
        assert (p[0] & 3) == 0
        for extra in (1, 7, 300):
            n = len(p) + extra
            b = ctypes.create_string_buffer(p + bytes(extra), n)
            assert ours.opus_packet_pad(b, len(p), n) == 0
            assert ours.opus_packet_get_nb_frames(b, n) == 1
            assert ours.opus_packet_unpad(b, n) == len(p)
            assert b.raw[:len(p)] == p


def test_bit_container_index(libs):
    """opus_b200_bit_index against the Python reader of the same container, incl. a fixture with lost packets."""
    ours, _ = libs
    for name in ("c2_96k_cbr_st.bit", "ms51_256k.bit", "silk_wb_mono_32k.bit"):
        path = os.path.join(GOLDEN, name)
        raw = open(path, "rb").read()
        pk = read_bit(path)
        n = len(pk)
        offs = np.zeros(n, dtype=np.int32)
        lens = np.zeros(n, dtype=np.int32)
        rngs = np.zeros(n, dtype=np.uint32)
        assert ours.opus_b200_bit_index(raw, len(raw), offs.ctypes.data, lens.ctypes.data, rngs.ctypes.data, n) == n
        for i, (p, r) in enumerate(pk):
            assert lens[i] == len(p) and int(rngs[i]) == r
            assert (offs[i] == -1) if not p else raw[offs[i]:offs[i] + lens[i]] == p
        assert ours.opus_b200_bit_index(raw, len(raw) - 3, None, None, None, 0) == cb.OPUS_INVALID_PACKET
        assert ours.opus_b200_bit_index(raw, len(raw), None, None, None, 0) == n
