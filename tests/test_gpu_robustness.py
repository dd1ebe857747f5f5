"""The same robustness scenarios through the CUDA library's batch entry point: every stream of the
batch follows its own script (losses, DTX, FEC requests, multi-frame and damaged packets), all
streams are decoded by ONE opus_decode_batch call per step, and each must match the compiled
reference decoder fed the same calls (return value, PCM, final range)."""
import ctypes
import os

import numpy as np
import pytest

import concentus_b200 as cb
from conftest import RefDecoder, needs_ref
from scenarios import fuzz_scenario, loss_scenario, random_payload_scenario

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


@needs_ref
def test_gpu_random_celt_payloads_batch():
    """48 streams fed random payloads behind CELT-only TOCs (every frame size, mono / stereo packets into mono /
    stereo decoders at 48 / 24 / 16 kHz): each must follow the reference."""
    scripts = []
    for rep in range(12):
        for fs, ch in ((48000, 2), (48000, 1), (24000, 2), (16000, 1)):
            scripts.append(random_payload_scenario(fs, ch, seed=1000 + 17 * rep + ch + fs // 8000, n=60))
    _run_batch(scripts)


@needs_ref
@pytest.mark.parametrize("fsz", [40320, 46080])
def test_gpu_long_concealment(fsz):
    """A lost call asking for most of a second (42 / 48 pieces of 20 ms): PCM offsets beyond 32767 samples."""
    from conftest import golden_case
    meta, pk = golden_case("c2_96k_cbr_st")
    dec, ref = cb.OpusDecoder(48000, 2), RefDecoder(48000, 2)
    try:
        calls = [(pk[t][0], 960) for t in range(6)] + [(None, fsz), (pk[6][0], 960), (pk[7][0], 960)]
        for n, (d, f) in enumerate(calls):
            ra, pa = ref.decode_raw(d, f)
            pb = dec.decode(d, f)
            assert ra == pb.shape[0], "call %d" % n
            assert np.array_equal(pa[:ra], pb), "pcm, call %d" % n
            assert ref.final_range == dec.final_range
    finally:
        dec.close()


@needs_ref
def test_gpu_device_entry_limits():
    """Device-resident entry: a stream needing more passes than max_frames fails alone with OPUS_BAD_ARG and keeps
    its state; a packet longer than max_packet_bytes is decoded all the same (read in place instead of staged);
    a pcm_stride shorter than a row is refused."""
    import torch
    from conftest import golden_case
    lib = cb.load()
    meta10, pk10 = golden_case("f10_64k_st")
    metab, pkb = golden_case("hi_510k_st")               # 20 ms packets of ~1275 bytes
    n, spf, ch = 8, 960, 2
    grp = cb.OpusBatchDecoder(48000, ch, n)
    ref10, refb = RefDecoder(48000, 2), RefDecoder(48000, 2)
    stream = torch.cuda.Stream()
    sp = ctypes.c_void_p(stream.cuda_stream)

    def call(packets, max_frames, max_bytes, stride=spf * ch):
        lens_h = np.array([len(p) if p else 0 for p in packets], dtype=np.int32)
        offs_h = np.concatenate([[0], np.cumsum(lens_h)[:-1]]).astype(np.int32)
        offs_h[lens_h == 0] = -1
        blob = b"".join(p for p in packets if p) or b"\0"
        arena = torch.from_numpy(np.frombuffer(blob, dtype=np.uint8).copy()).cuda()
        offs, lens = torch.from_numpy(offs_h).cuda(), torch.from_numpy(lens_h).cuda()
        pcm = torch.zeros(n * spf * ch, dtype=torch.int16, device="cuda")
        ret = torch.zeros(n, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        rc = lib.opus_decode_group_device(grp._g, arena.data_ptr(), offs.data_ptr(), lens.data_ptr(), pcm.data_ptr(),
                                          stride, spf, None, ret.data_ptr(), max_frames, max_bytes, sp)
        stream.synchronize()
        return rc, ret.cpu().numpy(), pcm.cpu().numpy().reshape(n, spf, ch)

    try:
        # streams 0..3 play the 10 ms fixture, 4..7 the 510 kbit/s one; max_packet_bytes understated on purpose
        for t in range(4):
            rc, ret, pcm = call([pk10[t][0]] * 4 + [pkb[t][0]] * 4, 1, 64)
            assert rc == 0
            ra, pa = ref10.decode_raw(pk10[t][0], spf)
            rb, pb = refb.decode_raw(pkb[t][0], spf)
            assert (ret[:4] == ra).all() and (ret[4:] == rb).all()
            for i in range(4):
                assert np.array_equal(pcm[i, :ra], pa[:ra])
                assert np.array_equal(pcm[4 + i, :rb], pb[:rb])
        # a 20 ms loss after 10 ms frames takes two pieces: with max_frames = 1 those streams fail, the others decode
        rc, ret, pcm = call([None] * 4 + [pkb[4][0]] * 4, 1, 1275)
        assert rc == 0
        assert (ret[:4] == cb.OPUS_BAD_ARG).all()
        rb, pb = refb.decode_raw(pkb[4][0], spf)
        assert (ret[4:] == rb).all() and np.array_equal(pcm[5, :rb], pb[:rb])
        # the refused streams are where they were: the next packets decode as if nothing had happened
        rc, ret, pcm = call([pk10[4][0]] * 4 + [pkb[5][0]] * 4, 1, 1275)
        ra, pa = ref10.decode_raw(pk10[4][0], spf)
        assert (ret[:4] == ra).all() and np.array_equal(pcm[2, :ra], pa[:ra])
        # and with two passes the loss is concealed like the reference does
        rc, ret, pcm = call([None] * 4 + [None] * 4, 2, 1275)
        ra, pa = ref10.decode_raw(None, spf)
        assert (ret[:4] == ra).all() and np.array_equal(pcm[1, :ra], pa[:ra])
        # rows must not overlap
        rc, _, _ = call([pk10[5][0]] * 8, 1, 1275, stride=spf * ch - 2)
        assert rc == cb.OPUS_BAD_ARG
    finally:
        grp.close()


@needs_ref
def test_gpu_pipelined_device_calls_follow_reference():
    """Device-resident calls queued back to back on the library's own streams (cuda_stream = NULL): the part of call
    k+1 that reads packets (plan, routing, SILK layer, parse) runs beside the rebuild / synthesis kernels of call k.
    A batch that mixes CELT, SILK, hybrid and mode-switching streams, with lost packets, must still do exactly what
    the reference does, call for call.  Sixteen calls are queued without any synchronisation in between; every call
    has its own PCM and return-value buffers, which are only read at the end."""
    import torch
    from conftest import GOLDEN, read_bit, golden_case
    lib = cb.load()
    names = ["c2_96k_cbr_st", "vbr_128k_st", "hyb_swb_st_48k", "silk_wb_st_40k", "trans_hyb_celt_st", "f10_64k_st",
             "trans_all_st_to24", "hi_510k_st"]
    packs = [[p for p, _ in read_bit(os.path.join(GOLDEN, nm + ".bit"))] for nm in names]
    n, spf, ch, calls = 48, 960, 2, 16
    rng = np.random.default_rng(21)
    grp = cb.OpusBatchDecoder(48000, ch, n)
    refs = [RefDecoder(48000, ch) for _ in range(n)]
    script = []          # per call: list of packets (None = lost)
    for t in range(calls):
        row = []
        for i in range(n):
            pk = packs[i % len(packs)]
            p = pk[(3 * i + t) % len(pk)]
            row.append(None if (t > 1 and rng.random() < 0.08) else p)
        script.append(row)
    keep = []
    try:
        for t in range(calls):
            row = script[t]
            lens_h = np.array([len(p) if p else 0 for p in row], dtype=np.int32)
            offs_h = np.concatenate([[0], np.cumsum(lens_h)[:-1]]).astype(np.int32)
            offs_h[lens_h == 0] = -1
            arena = torch.from_numpy(np.frombuffer(b"".join(p for p in row if p) or b"\0", dtype=np.uint8).copy()).cuda()
            offs, lens = torch.from_numpy(offs_h).cuda(), torch.from_numpy(lens_h).cuda()
            pcm = torch.zeros(n * spf * ch, dtype=torch.int16, device="cuda")
            ret = torch.zeros(n, dtype=torch.int32, device="cuda")
            keep.append((arena, offs, lens, pcm, ret))
        torch.cuda.synchronize()
        for arena, offs, lens, pcm, ret in keep:                     # queue everything, no waiting in between
            rc = lib.opus_decode_group_device(grp._g, arena.data_ptr(), offs.data_ptr(), lens.data_ptr(), pcm.data_ptr(),
                                              spf * ch, spf, None, ret.data_ptr(), 8, 1275, None)
            assert rc == 0
        assert lib.opus_b200_synchronize() == 0
        for t in range(calls):
            pcm = keep[t][3].cpu().numpy().reshape(n, spf, ch)
            ret = keep[t][4].cpu().numpy()
            for i in range(n):
                ra, pa = refs[i].decode_raw(script[t][i], spf)
                assert ret[i] == ra, (t, i, ret[i], ra)
                if ra > 0:
                    assert np.array_equal(pcm[i, :ra], pa[:ra]), (t, i)
    finally:
        grp.close()


@needs_ref
def test_gpu_entry_points_interleave_on_one_group():
    """One group driven through every decode entry point in a random order -- the synchronous host call, the pipelined
    host pair, device-resident calls on the library's streams and on a stream of the caller, the static-scratch call,
    final-range queries in between -- has to stay in step with the reference: each form queues its work differently
    (run-ahead or in order), and all of them must order themselves behind whatever the one before left in flight."""
    import torch
    from conftest import GOLDEN, read_bit
    lib = cb.load()
    names = ["c2_96k_cbr_st", "hyb_swb_st_48k", "silk_wb_st_40k", "vbr_128k_st", "trans_hyb_celt_st"]
    packs = [[p for p, _ in read_bit(os.path.join(GOLDEN, nm + ".bit"))] for nm in names]
    n, spf, ch, steps = 40, 960, 2, 36
    rng = np.random.default_rng(77)
    grp = cb.OpusBatchDecoder(48000, ch, n)
    handles = (ctypes.c_void_p * n)(*[d._h for d in grp.decoders])
    refs = [RefDecoder(48000, ch) for _ in range(n)]
    stream = torch.cuda.Stream()
    pending = []                      # (step, pcm tensor/array, ret tensor/array, packets) whose check is deferred
    keep = []

    def check(t, pcm, ret, row):
        pcm = np.asarray(pcm.cpu() if hasattr(pcm, "cpu") else pcm).reshape(n, spf, ch)
        ret = np.asarray(ret.cpu() if hasattr(ret, "cpu") else ret)
        for i in range(n):
            ra, pa = refs[i].decode_raw(row[i], spf)
            assert ret[i] == ra, (t, i, ret[i], ra)
            if ra > 0:
                assert np.array_equal(pcm[i, :ra], pa[:ra]), (t, i)

    def drain():
        assert lib.opus_decode_group_wait(grp._g) == 0
        assert lib.opus_b200_synchronize() == 0
        torch.cuda.synchronize()
        for item in pending:
            check(*item)
        del pending[:]

    try:
        for t in range(steps):
            row = []
            for i in range(n):
                pk = packs[i % len(packs)]
                row.append(None if (t > 1 and rng.random() < 0.06) else pk[(5 * i + t) % len(pk)])
            lens_h = np.array([len(p) if p else 0 for p in row], dtype=np.int32)
            offs_h = np.concatenate([[0], np.cumsum(lens_h)[:-1]]).astype(np.int32)
            offs_h[lens_h == 0] = -1
            blob = np.frombuffer(b"".join(p for p in row if p) or b"\0", dtype=np.uint8).copy()
            mode = int(rng.integers(0, 5))
            if mode == 0:                                     # synchronous host call
                drain()
                ret, pcm = grp.decode_flat(blob, len(blob), offs_h, lens_h, spf)
                check(t, pcm, ret, row)
            elif mode == 1:                                   # pipelined host pair
                pcm = torch.zeros(n * spf * ch, dtype=torch.int16).pin_memory()
                ret = torch.zeros(n, dtype=torch.int32).pin_memory()
                a = torch.from_numpy(blob).pin_memory()
                o, l = torch.from_numpy(offs_h).pin_memory(), torch.from_numpy(lens_h).pin_memory()
                keep.append((a, o, l))
                assert lib.opus_decode_group_submit(grp._g, a.data_ptr(), a.numel(), o.data_ptr(), l.data_ptr(), pcm.data_ptr(),
                                                    spf * ch, spf, None, ret.data_ptr()) == 0
                pending.append((t, pcm, ret, row))
            else:                                             # device-resident forms
                a, o, l = torch.from_numpy(blob).cuda(), torch.from_numpy(offs_h).cuda(), torch.from_numpy(lens_h).cuda()
                pcm = torch.zeros(n * spf * ch, dtype=torch.int16, device="cuda")
                ret = torch.zeros(n, dtype=torch.int32, device="cuda")
                keep.append((a, o, l))
                torch.cuda.synchronize()
                if mode == 2:                                 # library streams (run-ahead)
                    rc = lib.opus_decode_group_device(grp._g, a.data_ptr(), o.data_ptr(), l.data_ptr(), pcm.data_ptr(), spf * ch,
                                                      spf, None, ret.data_ptr(), 8, 1275, None)
                elif mode == 3:                               # the caller's stream: order it behind the library's work first
                    drain()
                    rc = lib.opus_decode_group_device(grp._g, a.data_ptr(), o.data_ptr(), l.data_ptr(), pcm.data_ptr(), spf * ch,
                                                      spf, None, ret.data_ptr(), 8, 1275, ctypes.c_void_p(stream.cuda_stream))
                    stream.synchronize()
                else:                                         # static scratch, library stream
                    rc = lib.opus_decode_batch_device_static(handles, n, a.data_ptr(), o.data_ptr(), l.data_ptr(), pcm.data_ptr(),
                                                             spf * ch, spf, ret.data_ptr(), 8, 1275, None)
                assert rc == 0
                pending.append((t, pcm, ret, row))
            if rng.random() < 0.2:
                drain()
                fr = grp.final_ranges()
                assert [int(x) for x in fr] == [r.final_range for r in refs]
        drain()
    finally:
        grp.close()
