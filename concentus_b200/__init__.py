"""concentus_b200 -- Python face of libopus_b200.so (the B200-native batched Opus engine).

The classes mirror the reference's native plugin wrappers so that tests read like the
reference's own (CSharp/Concentus/Native/NativeOpusDecoder.cs:74-214):

    OpusDecoder(Fs, channels).decode(packet_bytes_or_None, frame_size, fec=False) -> int16 ndarray
    OpusDecoder.final_range / bandwidth / last_packet_duration / reset_state()
    OpusBatchDecoder(Fs, channels, n).decode(...)      # n independent streams per call

Everything is computed by the CUDA library; there is no Python or CPU fallback.  Importing this
package does not need a GPU, creating a decoder does.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libopus_b200.so")

OPUS_OK = 0
OPUS_BAD_ARG = -1
OPUS_BUFFER_TOO_SMALL = -2
OPUS_INTERNAL_ERROR = -3
OPUS_INVALID_PACKET = -4
OPUS_UNIMPLEMENTED = -5
OPUS_INVALID_STATE = -6
OPUS_ALLOC_FAIL = -7

OPUS_GET_BANDWIDTH_REQUEST = 4009
OPUS_RESET_STATE = 4028
OPUS_GET_SAMPLE_RATE_REQUEST = 4029
OPUS_GET_FINAL_RANGE_REQUEST = 4031
OPUS_GET_PITCH_REQUEST = 4033
OPUS_SET_GAIN_REQUEST = 4034
OPUS_GET_LAST_PACKET_DURATION_REQUEST = 4039
OPUS_GET_GAIN_REQUEST = 4045

# every symbol include/opus_b200.h declares
EXPORTS = [
    "opus_get_version_string", "opus_strerror", "opus_decoder_get_size", "opus_decoder_create",
    "opus_decode", "opus_decode_float", "opus_decoder_ctl", "opus_decoder_destroy",
    "opus_packet_get_bandwidth", "opus_packet_get_nb_channels", "opus_packet_get_nb_frames",
    "opus_packet_get_samples_per_frame", "opus_packet_get_nb_samples", "opus_decoder_get_nb_samples",
    "opus_packet_parse", "opus_repacketizer_get_size", "opus_repacketizer_init", "opus_repacketizer_create",
    "opus_repacketizer_destroy", "opus_repacketizer_cat", "opus_repacketizer_out_range", "opus_repacketizer_get_nb_frames",
    "opus_repacketizer_out", "opus_packet_pad", "opus_packet_unpad", "opus_multistream_packet_pad",
    "opus_multistream_packet_unpad", "opus_b200_bit_index", "opus_multistream_encoder_get_size",
    "opus_multistream_surround_encoder_get_size", "opus_multistream_encoder_create",
    "opus_multistream_surround_encoder_create", "opus_multistream_encode", "opus_multistream_encode_float",
    "opus_multistream_encoder_ctl", "opus_multistream_encoder_destroy", "opus_decode_batch", "opus_decode_batch_flat", "opus_decoder_group_create",
    "opus_decoder_group_destroy", "opus_decode_group_flat", "opus_decode_group_device",
    "opus_decode_group_submit", "opus_decode_group_wait", "opus_encode_group_submit", "opus_encode_group_wait",
    "opus_decoder_get_final_range_batch", "opus_b200_synchronize", "opus_b200_launch_count",
    "opus_b200_profile_begin", "opus_b200_profile_end", "opus_b200_timer_mark", "opus_b200_timer_elapsed",
    "opus_b200_trace_begin", "opus_b200_trace_end", "opus_b200_selftest_cwrs", "opus_b200_selftest_isqrt",
    "opus_decode_batch_device_static",
    "opus_multistream_decoder_get_size", "opus_multistream_decoder_create", "opus_multistream_decode",
    "opus_multistream_decode_float", "opus_multistream_decoder_ctl", "opus_multistream_decoder_destroy",
    "opus_multistream_decode_batch", "opus_multistream_decoder_group_create",
    "opus_multistream_decoder_group_destroy", "opus_multistream_decode_group_flat",
    "opus_multistream_decode_group_submit", "opus_multistream_decode_group_wait",
    "opus_encoder_get_size", "opus_encoder_create", "opus_encode", "opus_encode_float", "opus_encoder_ctl",
    "opus_encoder_destroy", "opus_encode_batch", "opus_encoder_group_create", "opus_encoder_group_destroy",
    "opus_encode_group_flat", "opus_encode_group_device", "opus_encoder_get_final_range_batch",
    "opus_b200_encode_profile_begin", "opus_b200_encode_profile_end",
]

_lib = None


class OpusError(RuntimeError):
    def __init__(self, code, what=""):
        self.code = code
        msg = load().opus_strerror(code).decode() if _lib is not None else str(code)
        super().__init__("%s: %s (%d)" % (what, msg, code))


def load():
    """dlopen libopus_b200.so and declare prototypes.  Raises if the library was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "libopus_b200.so is missing (%s). Build it with `python -m concentus_b200.build`; "
            "there is no CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    c_int_p = ctypes.POINTER(ctypes.c_int)
    vp = ctypes.c_void_p
    lib.opus_get_version_string.restype = ctypes.c_char_p
    lib.opus_strerror.restype = ctypes.c_char_p
    lib.opus_strerror.argtypes = [ctypes.c_int]
    lib.opus_decoder_create.restype = vp
    lib.opus_decoder_create.argtypes = [ctypes.c_int32, ctypes.c_int, c_int_p]
    lib.opus_decoder_destroy.argtypes = [vp]
    lib.opus_decoder_destroy.restype = None
    lib.opus_decode.argtypes = [vp, vp, ctypes.c_int32, vp, ctypes.c_int, ctypes.c_int]
    lib.opus_decode_float.argtypes = [vp, vp, ctypes.c_int32, vp, ctypes.c_int, ctypes.c_int]
    lib.opus_decode_batch.argtypes = [vp, vp, vp, vp, ctypes.c_int, vp, ctypes.c_int, vp]
    lib.opus_decode_batch_flat.argtypes = [vp, ctypes.c_int, vp, ctypes.c_int32, vp, vp, vp, ctypes.c_int32,
                                           ctypes.c_int, vp, vp]
    lib.opus_decoder_group_create.restype = vp
    lib.opus_decoder_group_create.argtypes = [vp, ctypes.c_int, c_int_p]
    lib.opus_decoder_group_destroy.argtypes = [vp]
    lib.opus_decoder_group_destroy.restype = None
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
    lib.opus_multistream_encoder_create.restype = vp
    lib.opus_multistream_encoder_create.argtypes = [ctypes.c_int32, ctypes.c_int, ctypes.c_int, ctypes.c_int, vp, ctypes.c_int, vp]
    lib.opus_multistream_surround_encoder_create.restype = vp
    lib.opus_multistream_surround_encoder_create.argtypes = [ctypes.c_int32, ctypes.c_int, ctypes.c_int, vp, vp, vp, ctypes.c_int, vp]
    lib.opus_multistream_encode.argtypes = [vp, vp, ctypes.c_int, vp, ctypes.c_int32]
    lib.opus_multistream_encode_float.argtypes = [vp, vp, ctypes.c_int, vp, ctypes.c_int32]
    lib.opus_multistream_encoder_destroy.argtypes = [vp]
    lib.opus_b200_bit_index.argtypes = [vp, ctypes.c_int32, vp, vp, vp, ctypes.c_int32]
    lib.opus_decode_group_flat.argtypes = [vp, vp, ctypes.c_int32, vp, vp, vp, ctypes.c_int32, ctypes.c_int, vp, vp]
    lib.opus_decode_group_submit.argtypes = [vp, vp, ctypes.c_int32, vp, vp, vp, ctypes.c_int32, ctypes.c_int, vp, vp]
    lib.opus_decode_group_wait.argtypes = [vp]
    lib.opus_decode_group_device.argtypes = [vp, vp, vp, vp, vp, ctypes.c_int32, ctypes.c_int, vp, vp,
                                             ctypes.c_int, ctypes.c_int, vp]
    lib.opus_decode_batch_device_static.argtypes = [vp, ctypes.c_int, vp, vp, vp, vp, ctypes.c_int32, ctypes.c_int,
                                                    vp, ctypes.c_int, ctypes.c_int, vp]
    lib.opus_multistream_decoder_create.restype = vp
    lib.opus_multistream_decoder_create.argtypes = [ctypes.c_int32, ctypes.c_int, ctypes.c_int, ctypes.c_int, vp, c_int_p]
    lib.opus_multistream_decoder_destroy.argtypes = [vp]
    lib.opus_multistream_decoder_destroy.restype = None
    lib.opus_multistream_decode.argtypes = [vp, vp, ctypes.c_int32, vp, ctypes.c_int, ctypes.c_int]
    lib.opus_multistream_decode_float.argtypes = [vp, vp, ctypes.c_int32, vp, ctypes.c_int, ctypes.c_int]
    lib.opus_multistream_decode_batch.argtypes = [vp, vp, vp, vp, ctypes.c_int, vp, ctypes.c_int, vp]
    lib.opus_multistream_decoder_group_create.restype = vp
    lib.opus_multistream_decoder_group_create.argtypes = [vp, ctypes.c_int, c_int_p]
    lib.opus_multistream_decoder_group_destroy.argtypes = [vp]
    lib.opus_multistream_decoder_group_destroy.restype = None
    lib.opus_multistream_decode_group_flat.argtypes = [vp, vp, ctypes.c_int32, vp, vp, vp, ctypes.c_int32,
                                                       ctypes.c_int, vp, vp]
    lib.opus_multistream_decode_group_submit.argtypes = [vp, vp, ctypes.c_int32, vp, vp, vp, ctypes.c_int32,
                                                       ctypes.c_int, vp, vp]
    lib.opus_multistream_decode_group_wait.argtypes = [vp]
    lib.opus_decoder_get_final_range_batch.argtypes = [vp, ctypes.c_int, vp]
    lib.opus_encoder_create.restype = vp
    lib.opus_encoder_create.argtypes = [ctypes.c_int32, ctypes.c_int, ctypes.c_int, c_int_p]
    lib.opus_encoder_destroy.argtypes = [vp]
    lib.opus_encoder_destroy.restype = None
    lib.opus_encode.argtypes = [vp, vp, ctypes.c_int, vp, ctypes.c_int32]
    lib.opus_encode_float.argtypes = [vp, vp, ctypes.c_int, vp, ctypes.c_int32]
    lib.opus_encode_batch.argtypes = [vp, ctypes.c_int, vp, ctypes.c_int, vp, ctypes.c_int32, vp]
    lib.opus_encoder_group_create.restype = vp
    lib.opus_encoder_group_create.argtypes = [vp, ctypes.c_int, c_int_p]
    lib.opus_encoder_group_destroy.argtypes = [vp]
    lib.opus_encoder_group_destroy.restype = None
    lib.opus_encode_group_flat.argtypes = [vp, vp, ctypes.c_int, vp, ctypes.c_int32, vp]
    lib.opus_encode_group_submit.argtypes = [vp, vp, ctypes.c_int, vp, ctypes.c_int32, vp]
    lib.opus_encode_group_wait.argtypes = [vp]
    lib.opus_encode_group_device.argtypes = [vp, vp, ctypes.c_int, vp, ctypes.c_int32, vp]
    lib.opus_encoder_get_final_range_batch.argtypes = [vp, ctypes.c_int, vp]
    lib.opus_b200_encode_profile_end.argtypes = [vp, vp]
    lib.opus_b200_launch_count.restype = ctypes.c_longlong
    lib.opus_packet_get_nb_samples.argtypes = [vp, ctypes.c_int32, ctypes.c_int32]
    lib.opus_packet_get_nb_frames.argtypes = [vp, ctypes.c_int32]
    lib.opus_packet_get_samples_per_frame.argtypes = [vp, ctypes.c_int32]
    lib.opus_packet_get_bandwidth.argtypes = [vp]
    lib.opus_packet_get_nb_channels.argtypes = [vp]
    _lib = lib
    return lib


def _ptr(a):
    return a.ctypes.data if a is not None else None


class OpusDecoder(object):
    """One stream.  Mirrors NativeOpusDecoder (reference Native/NativeOpusDecoder.cs)."""

    def __init__(self, sample_rate, channels):
        lib = load()
        err = ctypes.c_int(0)
        self._h = lib.opus_decoder_create(sample_rate, channels, ctypes.byref(err))
        if not self._h:
            raise OpusError(err.value, "opus_decoder_create")
        self.sample_rate = sample_rate
        self.channels = channels

    def close(self):
        if getattr(self, "_h", None):
            load().opus_decoder_destroy(self._h)
            self._h = None

    __del__ = close

    def decode_raw(self, data, frame_size, fec=False):
        """Returns (ret, pcm[frame_size, channels]); ret is what opus_decode returned."""
        pcm = np.zeros((frame_size if frame_size > 0 else 1, self.channels), dtype=np.int16)
        if data is None:
            ret = load().opus_decode(self._h, None, 0, pcm.ctypes.data, frame_size, int(fec))
        else:
            buf = ctypes.create_string_buffer(bytes(data), len(data))
            ret = load().opus_decode(self._h, ctypes.cast(buf, ctypes.c_void_p), len(data), pcm.ctypes.data,
                                     frame_size, int(fec))
        return ret, pcm

    def decode(self, data, frame_size, fec=False):
        ret, pcm = self.decode_raw(data, frame_size, fec)
        if ret < 0:
            raise OpusError(ret, "opus_decode")
        return pcm[:ret]

    def _get(self, request):
        v = ctypes.c_int32(0)
        rc = load().opus_decoder_ctl(self._h, request, ctypes.byref(v))
        if rc != OPUS_OK:
            raise OpusError(rc, "opus_decoder_ctl(%d)" % request)
        return v.value

    @property
    def final_range(self):
        return self._get(OPUS_GET_FINAL_RANGE_REQUEST) & 0xFFFFFFFF

    @property
    def bandwidth(self):
        return self._get(OPUS_GET_BANDWIDTH_REQUEST)

    @property
    def last_packet_duration(self):
        return self._get(OPUS_GET_LAST_PACKET_DURATION_REQUEST)

    @property
    def pitch(self):
        return self._get(OPUS_GET_PITCH_REQUEST)

    @property
    def gain(self):
        return self._get(OPUS_GET_GAIN_REQUEST)

    @gain.setter
    def gain(self, value):
        rc = load().opus_decoder_ctl(self._h, OPUS_SET_GAIN_REQUEST, ctypes.c_int32(int(value)))
        if rc != OPUS_OK:
            raise OpusError(rc, "OPUS_SET_GAIN")

    def reset_state(self):
        rc = load().opus_decoder_ctl(self._h, OPUS_RESET_STATE)
        if rc != OPUS_OK:
            raise OpusError(rc, "OPUS_RESET_STATE")


class OpusBatchDecoder(object):
    """n independent streams of the same format, decoded by one call (opus_decode_group_flat)."""

    def __init__(self, sample_rate, channels, n):
        lib = load()
        self.n = n
        self.sample_rate = sample_rate
        self.channels = channels
        self.decoders = [OpusDecoder(sample_rate, channels) for _ in range(n)]
        self._handles = (ctypes.c_void_p * n)(*[d._h for d in self.decoders])
        err = ctypes.c_int(0)
        self._g = lib.opus_decoder_group_create(self._handles, n, ctypes.byref(err))
        if not self._g:
            raise OpusError(err.value, "opus_decoder_group_create")

    def close(self):
        if getattr(self, "_g", None):
            load().opus_decoder_group_destroy(self._g)
            self._g = None
        for d in getattr(self, "decoders", []):
            d.close()
        self.decoders = []

    __del__ = close

    def decode(self, packets, frame_size, fec=None):
        """packets: list of n bytes objects (None/b'' = lost).  Returns (ret[n], pcm[n, frame_size, ch])."""
        assert len(packets) == self.n
        lens = np.array([len(p) if p else 0 for p in packets], dtype=np.int32)
        offs = np.zeros(self.n, dtype=np.int32)
        o = 0
        for i, p in enumerate(packets):
            if p:
                offs[i] = o
                o += len(p)
            else:
                offs[i] = -1
        arena = np.frombuffer(b"".join(p for p in packets if p) + b"\0", dtype=np.uint8).copy()
        return self.decode_flat(arena, o, offs, lens, frame_size, fec)

    def decode_flat(self, arena, nbytes, offs, lens, frame_size, fec=None, pcm=None):
        stride = frame_size * self.channels
        if pcm is None:
            pcm = np.zeros((self.n, frame_size, self.channels), dtype=np.int16)
        ret = np.zeros(self.n, dtype=np.int32)
        fec_a = None if fec is None else np.ascontiguousarray(fec, dtype=np.int32)
        rc = load().opus_decode_group_flat(self._g, _ptr(arena), nbytes, _ptr(offs), _ptr(lens), _ptr(pcm), stride,
                                           frame_size, _ptr(fec_a), _ptr(ret))
        if rc != OPUS_OK:
            raise OpusError(rc, "opus_decode_group_flat")
        return ret, pcm

    def submit_flat(self, arena, nbytes, offs, lens, frame_size, pcm, ret, fec=None):
        """Pipelined form (opus_decode_group_submit): queues the call and returns.  The caller owns every
        array and keeps it alive and unchanged until wait(); pcm / ret are filled by then."""
        fec_a = None if fec is None else np.ascontiguousarray(fec, dtype=np.int32)
        rc = load().opus_decode_group_submit(self._g, _ptr(arena), nbytes, _ptr(offs), _ptr(lens), _ptr(pcm),
                                             frame_size * self.channels, frame_size, _ptr(fec_a), _ptr(ret))
        if rc != OPUS_OK:
            raise OpusError(rc, "opus_decode_group_submit")
        self._keep = (getattr(self, "_keep", ()) + ((arena, offs, lens, fec_a, pcm, ret),))[-3:]

    def wait(self):
        rc = load().opus_decode_group_wait(self._g)
        if rc != OPUS_OK:
            raise OpusError(rc, "opus_decode_group_wait")
        self._keep = ()

    def final_ranges(self):
        out = np.zeros(self.n, dtype=np.uint32)
        rc = load().opus_decoder_get_final_range_batch(self._handles, self.n, _ptr(out))
        if rc != OPUS_OK:
            raise OpusError(rc, "opus_decoder_get_final_range_batch")
        return out


class OpusMSDecoder(object):
    """One multistream (surround) decoder: mirrors NativeOpusMultistreamDecoder of the reference."""

    def __init__(self, sample_rate, channels, streams, coupled, mapping):
        lib = load()
        err = ctypes.c_int(0)
        m = (ctypes.c_ubyte * len(mapping))(*mapping)
        self._h = lib.opus_multistream_decoder_create(sample_rate, channels, streams, coupled, m, ctypes.byref(err))
        if not self._h:
            raise OpusError(err.value, "opus_multistream_decoder_create")
        self.channels = channels

    def close(self):
        if getattr(self, "_h", None):
            load().opus_multistream_decoder_destroy(self._h)
            self._h = None

    __del__ = close

    def decode_raw(self, data, frame_size, fec=False):
        pcm = np.zeros((max(frame_size, 1), self.channels), dtype=np.int16)
        if not data:
            ret = load().opus_multistream_decode(self._h, None, 0, pcm.ctypes.data, frame_size, int(fec))
        else:
            buf = ctypes.create_string_buffer(bytes(data), len(data))
            ret = load().opus_multistream_decode(self._h, ctypes.cast(buf, ctypes.c_void_p), len(data), pcm.ctypes.data,
                                                 frame_size, int(fec))
        return ret, pcm

    @property
    def final_range(self):
        v = ctypes.c_uint32(0)
        rc = load().opus_multistream_decoder_ctl(self._h, OPUS_GET_FINAL_RANGE_REQUEST, ctypes.byref(v))
        if rc != OPUS_OK:
            raise OpusError(rc, "opus_multistream_decoder_ctl")
        return v.value


class OpusMSBatchDecoder(object):
    """n multistream decoders of the same layout decoded by one call (opus_multistream_decode_group_flat)."""

    def __init__(self, sample_rate, channels, streams, coupled, mapping, n):
        lib = load()
        self.n, self.channels = n, channels
        self.decoders = [OpusMSDecoder(sample_rate, channels, streams, coupled, mapping) for _ in range(n)]
        self._handles = (ctypes.c_void_p * n)(*[d._h for d in self.decoders])
        err = ctypes.c_int(0)
        self._g = lib.opus_multistream_decoder_group_create(self._handles, n, ctypes.byref(err))
        if not self._g:
            raise OpusError(err.value, "opus_multistream_decoder_group_create")

    def close(self):
        if getattr(self, "_g", None):
            load().opus_multistream_decoder_group_destroy(self._g)
            self._g = None
        for d in getattr(self, "decoders", []):
            d.close()
        self.decoders = []

    __del__ = close

    def decode(self, packets, frame_size):
        lens = np.array([len(p) if p else 0 for p in packets], dtype=np.int32)
        offs = np.zeros(self.n, dtype=np.int32)
        o = 0
        for i, p in enumerate(packets):
            if p:
                offs[i] = o
                o += len(p)
            else:
                offs[i] = -1
        arena = np.frombuffer(b"".join(p for p in packets if p) + b"\0", dtype=np.uint8).copy()
        pcm = np.zeros((self.n, frame_size, self.channels), dtype=np.int16)
        ret = np.zeros(self.n, dtype=np.int32)
        rc = load().opus_multistream_decode_group_flat(self._g, _ptr(arena), o, _ptr(offs), _ptr(lens), _ptr(pcm),
                                                       frame_size * self.channels, frame_size, None, _ptr(ret))
        if rc != OPUS_OK:
            raise OpusError(rc, "opus_multistream_decode_group_flat")
        return ret, pcm


# ---- encoder (reference: CSharp/Concentus/Native/NativeOpusEncoder.cs) ---------------------------
OPUS_APPLICATION_VOIP = 2048
OPUS_APPLICATION_AUDIO = 2049
OPUS_APPLICATION_RESTRICTED_LOWDELAY = 2051
OPUS_SET_BITRATE_REQUEST = 4002
OPUS_SET_VBR_REQUEST = 4006
OPUS_SET_COMPLEXITY_REQUEST = 4010
OPUS_SET_VBR_CONSTRAINT_REQUEST = 4020
OPUS_SET_FORCE_MODE_REQUEST = 11002
OPUS_GET_FINAL_RANGE_REQUEST_ENC = 4031


class OpusEncoder(object):
    """One stream.  Mirrors NativeOpusEncoder: encode(pcm int16 [frame*channels]) -> bytes."""

    def __init__(self, sample_rate, channels, application=OPUS_APPLICATION_AUDIO):
        lib = load()
        err = ctypes.c_int(0)
        self._h = lib.opus_encoder_create(sample_rate, channels, application, ctypes.byref(err))
        if not self._h:
            raise OpusError(err.value, "opus_encoder_create")
        self.sample_rate = sample_rate
        self.channels = channels

    def close(self):
        if getattr(self, "_h", None):
            load().opus_encoder_destroy(self._h)
            self._h = None

    __del__ = close

    def ctl(self, request, value):
        rc = load().opus_encoder_ctl(ctypes.c_void_p(self._h), ctypes.c_int(request), ctypes.c_int(value))
        if rc < 0:
            raise OpusError(rc, "opus_encoder_ctl(%d)" % request)

    def get(self, request):
        out = ctypes.c_int32(0)
        rc = load().opus_encoder_ctl(ctypes.c_void_p(self._h), ctypes.c_int(request), ctypes.byref(out))
        if rc < 0:
            raise OpusError(rc, "opus_encoder_ctl(%d)" % request)
        return out.value

    @property
    def final_range(self):
        return self.get(OPUS_GET_FINAL_RANGE_REQUEST_ENC) & 0xFFFFFFFF

    def encode(self, pcm, frame_size, max_data_bytes=1275):
        """Returns the packet bytes; raises OpusError on a negative return code."""
        pcm = np.ascontiguousarray(pcm, dtype=np.int16)
        buf = np.zeros(max_data_bytes, dtype=np.uint8)
        n = load().opus_encode(self._h, _ptr(pcm), frame_size, _ptr(buf), max_data_bytes)
        if n < 0:
            raise OpusError(n, "opus_encode")
        return buf[:n].tobytes()


class OpusBatchEncoder(object):
    """n independent encoders driven by one kernel launch per frame."""

    def __init__(self, n, sample_rate, channels, application=OPUS_APPLICATION_AUDIO):
        self.encoders = [OpusEncoder(sample_rate, channels, application) for _ in range(n)]
        self.n = n
        self.channels = channels
        lib = load()
        arr = (ctypes.c_void_p * n)(*[e._h for e in self.encoders])
        err = ctypes.c_int(0)
        self._g = lib.opus_encoder_group_create(arr, n, ctypes.byref(err))
        if not self._g:
            raise OpusError(err.value, "opus_encoder_group_create")

    def close(self):
        if getattr(self, "_g", None):
            load().opus_encoder_group_destroy(self._g)
            self._g = None
        for e in getattr(self, "encoders", []):
            e.close()

    __del__ = close

    def ctl(self, request, value):
        for e in self.encoders:
            e.ctl(request, value)

    def encode(self, pcm, frame_size, max_data_bytes=1275):
        """pcm: int16 [n, frame_size*channels].  Returns (data uint8 [n, max_data_bytes], lens int32 [n])."""
        pcm = np.ascontiguousarray(pcm, dtype=np.int16)
        assert pcm.shape == (self.n, frame_size * self.channels)
        data = np.zeros((self.n, max_data_bytes), dtype=np.uint8)
        lens = np.zeros(self.n, dtype=np.int32)
        rc = load().opus_encode_group_flat(self._g, _ptr(pcm), frame_size, _ptr(data), max_data_bytes, _ptr(lens))
        if rc < 0:
            raise OpusError(rc, "opus_encode_group_flat")
        return data, lens

    def submit(self, pcm, frame_size, data, lens):
        """Pipelined form (opus_encode_group_submit): pcm int16 [n, frame_size*channels], data uint8
        [n, max_data_bytes], lens int32 [n], all caller-owned and kept alive until wait()."""
        assert pcm.dtype == np.int16 and pcm.flags.c_contiguous and pcm.shape == (self.n, frame_size * self.channels)
        assert data.dtype == np.uint8 and data.flags.c_contiguous and data.shape[0] == self.n
        rc = load().opus_encode_group_submit(self._g, _ptr(pcm), frame_size, _ptr(data), data.shape[1], _ptr(lens))
        if rc < 0:
            raise OpusError(rc, "opus_encode_group_submit")
        self._keep = (getattr(self, "_keep", ()) + ((pcm, data, lens),))[-2:]

    def wait(self):
        rc = load().opus_encode_group_wait(self._g)
        if rc < 0:
            raise OpusError(rc, "opus_encode_group_wait")
        self._keep = ()

    def encode_device(self, d_pcm_ptr, frame_size, d_data_ptr, max_data_bytes, d_len_ptr):
        """Raw device pointers (e.g. torch tensors' data_ptr()); asynchronous."""
        rc = load().opus_encode_group_device(self._g, d_pcm_ptr, frame_size, d_data_ptr, max_data_bytes, d_len_ptr)
        if rc < 0:
            raise OpusError(rc, "opus_encode_group_device")

    def final_ranges(self):
        arr = (ctypes.c_void_p * self.n)(*[e._h for e in self.encoders])
        out = np.zeros(self.n, dtype=np.uint32)
        rc = load().opus_encoder_get_final_range_batch(arr, self.n, _ptr(out))
        if rc < 0:
            raise OpusError(rc, "opus_encoder_get_final_range_batch")
        return out
