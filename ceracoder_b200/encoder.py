"""ctypes mirror of include/ceracuda.h, shaped like the GStreamer element the reference instantiates.

`CudaH264Enc` mirrors what ceracoder sees of its encoder element
(/root/reference/pipeline/generic/x264_veryfast_v4l_mjpeg_1080p30:6, /root/reference/src/gst/encoder_control.c:45-57):
construct with the pipeline-line properties (`speed_preset`, `key_int_max`, `bitrate` in kbit/s, `name`), then
`set_property("bitrate"|"bps", v)` at any time from any thread, and `encode(frame)` once per buffer.
"""
import ctypes as C
import os
import numpy as np

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libceracuda.so")
CC_FORCE_IDR, CC_FRAME_DEVICE = 1, 2
CC_FMT_I420, CC_FMT_NV12, CC_FMT_YUY2, CC_FMT_BGRX = 0, 1, 2, 3


class CeraCudaError(RuntimeError):
    pass


class Params(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("device", "width", "height", "fps_n", "fps_d", "format", "bitrate_bps", "key_int_max",
                                       "rows_per_slice", "qp_fixed", "qp_min", "qp_max", "partitions", "threads", "host_entropy")]


class Frame(C.Structure):
    _fields_ = [("plane", C.c_void_p * 3), ("stride", C.c_int * 3), ("pts", C.c_int64)]


class Packet(C.Structure):
    _fields_ = [("data", C.POINTER(C.c_uint8)), ("size", C.c_int), ("is_idr", C.c_int), ("qp", C.c_int), ("pts", C.c_int64),
                ("n_skip", C.c_int), ("n_inter", C.c_int), ("n_intra", C.c_int), ("sym_bytes", C.c_int), ("gpu_ms", C.c_float), ("host_ms", C.c_float),
                ("n_intra_analysed", C.c_int), ("n_skip_probed", C.c_int), ("submit_ms", C.c_float), ("wait_ms", C.c_float)]


class SeqStats(C.Structure):
    _fields_ = [(n, C.c_long) for n in ("frames", "au_bytes", "sym_bytes", "n_skip", "n_inter", "n_intra", "n_intra_analysed", "n_skip_probed")] + \
               [(n, C.c_double) for n in ("submit_ms", "wait_ms", "host_ms", "gpu_ms")] + [("last_qp", C.c_int), ("n_idr", C.c_int)]


_libs = {}


def load_library(path=None):
    """Load libceracuda.so (built in-tree by `make lib` / __graft_entry__.build()). Fails loudly if missing."""
    path = path or os.environ.get("CERACUDA_LIB") or LIB_PATH   # CERACUDA_LIB: an alternative build of the same library (kernel experiments)
    if path in _libs:
        return _libs[path]
    if not os.path.exists(path):
        raise CeraCudaError(f"{path} not found: build it with `make lib` (nvcc, sm_100a). There is no CPU fallback.")
    L = C.CDLL(path)
    vp, ci = C.c_void_p, C.c_int
    L.cc_version.restype = C.c_char_p
    L.cc_strerror.restype = C.c_char_p
    L.cc_strerror.argtypes = [ci]
    L.cc_enc_params_default.argtypes = [C.POINTER(Params)]
    L.cc_enc_open.argtypes = [C.POINTER(vp), C.POINTER(Params)]
    L.cc_enc_close.argtypes = [vp]
    L.cc_enc_set_bitrate.argtypes = [vp, ci]
    L.cc_enc_get_bitrate.argtypes = [vp]
    L.cc_enc_encode.argtypes = [vp, C.POINTER(Frame), C.c_uint, C.POINTER(Packet)]
    L.cc_enc_submit.argtypes = [vp, C.POINTER(Frame), C.c_uint]
    L.cc_enc_collect.argtypes = [vp, C.POINTER(Packet)]
    L.cc_enc_encode_sequence.argtypes = [vp, C.POINTER(Frame), C.POINTER(ci), ci, C.c_uint, C.POINTER(SeqStats), C.POINTER(Packet)]
    L.cc_enc_set_next_qp.argtypes = [vp, ci]
    L.cc_enc_get_recon.argtypes = [vp, vp, vp, vp]
    L.cc_enc_mb_w.argtypes = [vp]
    L.cc_enc_mb_h.argtypes = [vp]
    L.cc_enc_get_me.argtypes = [vp, vp, vp]
    L.cc_enc_get_me8.argtypes = [vp, vp, vp, vp]
    L.cc_enc_get_mbinfo.argtypes = [vp, vp]
    L.cc_enc_device_bytes.argtypes = [vp]
    L.cc_enc_device_bytes.restype = C.c_size_t
    L.cc_enc_kernel_launches.argtypes = [vp]
    L.cc_enc_kernel_launches.restype = C.c_long
    L.cc_enc_staged_frames.argtypes = [vp]
    L.cc_enc_staged_frames.restype = C.c_long
    L.cc_enc_overlapped_frames.argtypes = [vp]
    L.cc_enc_overlapped_frames.restype = C.c_long
    L.cc_enc_set_profile.argtypes = [vp, ci]
    L.cc_enc_get_stage_ms.argtypes = [vp, vp, vp]
    L.cc_host_alloc.restype = vp
    L.cc_host_alloc.argtypes = [C.c_size_t]
    L.cc_host_free.argtypes = [vp]
    L.cc_host_register.argtypes = [vp, C.c_size_t]
    L.cc_host_unregister.argtypes = [vp]
    L.cc_dev_alloc.restype = vp
    L.cc_dev_alloc.argtypes = [C.c_size_t]
    L.cc_dev_free.argtypes = [vp]
    L.cc_dev_upload.argtypes = [vp, vp, C.c_size_t]
    L.cc_prim_sad.argtypes = [vp, vp, ci, ci, ci, vp]
    L.cc_prim_satd4x4.argtypes = [vp, vp, ci, vp]
    L.cc_prim_satd16x16.argtypes = [vp, vp, ci, vp]
    L.cc_prim_tq4x4.argtypes = [vp, vp, ci, ci, ci, vp, vp]
    L.cc_prim_dct4x4.argtypes = [vp, vp, ci, vp]
    L.cc_prim_tq8x8.argtypes = [vp, vp, ci, ci, vp, vp, vp]
    L.cc_prim_pred4x4.argtypes = [vp, ci, ci, vp]
    L.cc_prim_mc_luma.argtypes = [vp, ci, ci, vp, vp, ci, vp]
    L.cc_bench_int_peak.argtypes = [ci, C.POINTER(C.c_double)]
    _libs[path] = L
    return L


def _check(L, rc, what):
    if rc != 0:
        raise CeraCudaError(f"{what}: {L.cc_strerror(rc).decode()} ({rc})")


# x264enc's speed-preset enum (accepted for pipeline-text compatibility; this encoder has one speed)
_SPEED_PRESETS = {"ultrafast": 1, "superfast": 2, "veryfast": 3, "faster": 4, "fast": 5, "medium": 6, "slow": 7, "slower": 8, "veryslow": 9, "placebo": 10}


class CudaH264Enc:
    """The `cudah264enc` element as a Python object: same properties as the pipeline line that names x264enc."""

    def __init__(self, width, height, fps=(30, 1), bitrate=6000, key_int_max=60, speed_preset=3, tune=0, name="venc_kbps",
                 fmt="I420", device=0, rows_per_slice=8, qp_fixed=0, lib=None, byte_stream=True, threads=0, partitions=True, host_entropy=False):
        if isinstance(speed_preset, str):
            if speed_preset not in _SPEED_PRESETS:
                raise ValueError(f"unknown speed-preset {speed_preset!r}")
            speed_preset = _SPEED_PRESETS[speed_preset]
        if not 0 <= int(speed_preset) <= 10:
            raise ValueError("speed-preset out of range")
        if name not in ("venc_kbps", "venc_bps"):
            # /root/reference/src/gst/encoder_control.c:29-32 looks the element up by one of these two names
            raise ValueError("ceracoder finds its encoder by name: venc_kbps or venc_bps")
        self.L = lib or load_library()
        self.name = name
        self.width, self.height = int(width), int(height)
        p = Params()
        self.L.cc_enc_params_default(C.byref(p))
        p.device, p.width, p.height = device, self.width, self.height
        p.fps_n, p.fps_d = (fps if isinstance(fps, tuple) else (int(fps), 1))
        p.format = {"I420": CC_FMT_I420, "NV12": CC_FMT_NV12, "YUY2": CC_FMT_YUY2, "BGRx": CC_FMT_BGRX}[fmt]  # packed: one plane, converted on the device
        p.bitrate_bps = int(bitrate) * 1000
        p.key_int_max, p.rows_per_slice, p.qp_fixed = key_int_max, rows_per_slice, qp_fixed
        p.partitions = int(bool(partitions))
        p.threads = int(threads)  # x264enc's `threads`: host threads writing the slices of one access unit
        p.host_entropy = int(bool(host_entropy))  # True: the host C CAVLC writer instead of the device kernels (same bytes)
        self.fmt = fmt
        self._h = C.c_void_p()
        _check(self.L, self.L.cc_enc_open(C.byref(self._h), C.byref(p)), "cc_enc_open")
        self.mb_w, self.mb_h = self.L.cc_enc_mb_w(self._h), self.L.cc_enc_mb_h(self._h)
        self._pkt = Packet()

    # -- GObject-property face (what encoder_control.c:53 and the docs' "bitrate" write) --
    def set_property(self, prop, value):
        value = int(value)
        if prop == "bitrate":      # x264enc-compatible: kbit/s
            bps = value * 1000
        elif prop == "bps":        # what the reference actually writes: bitrate_bps / div (div 1000 for venc_kbps)
            bps = value * 1000 if (self.name == "venc_kbps" and value <= 30000) else value
        else:
            raise AttributeError(f"cudah264enc has no writable property {prop!r}")
        _check(self.L, self.L.cc_enc_set_bitrate(self._h, bps), "cc_enc_set_bitrate")

    def get_property(self, prop):
        bps = self.L.cc_enc_get_bitrate(self._h)
        if prop == "bitrate":
            return bps // 1000
        if prop == "bps":
            return bps // 1000 if self.name == "venc_kbps" else bps
        raise AttributeError(prop)

    # -- data path --
    def _frame(self, y, u=None, v=None, pts=0):
        f = Frame()
        planes = [y] + ([u] if u is not None else []) + ([v] if v is not None else [])
        for i, pl in enumerate(planes):
            if isinstance(pl, tuple):  # (pointer, stride): device or pinned memory handed over raw
                f.plane[i], f.stride[i] = pl
            else:
                assert pl.dtype == np.uint8 and pl.strides[1] == 1
                f.plane[i], f.stride[i] = pl.ctypes.data, pl.strides[0]
        f.pts = pts
        return f

    def encode(self, y, u=None, v=None, pts=0, force_idr=False, qp=None, device_frame=False):
        """One frame in, one Annex-B access unit out (bytes). Frame planes: numpy uint8 arrays (any row stride); packed formats
        (YUY2, BGRx) pass their one plane as `y` (rows of 2 w or 4 w bytes)."""
        if qp is not None:
            _check(self.L, self.L.cc_enc_set_next_qp(self._h, int(qp)), "cc_enc_set_next_qp")
        f = self._frame(y, u, v, pts)
        flags = (CC_FORCE_IDR if force_idr else 0) | (CC_FRAME_DEVICE if device_frame else 0)
        _check(self.L, self.L.cc_enc_encode(self._h, C.byref(f), flags, C.byref(self._pkt)), "cc_enc_encode")
        return C.string_at(self._pkt.data, self._pkt.size)

    def submit(self, y, u=None, v=None, pts=0, force_idr=False, device_frame=False):
        f = self._frame(y, u, v, pts)
        flags = (CC_FORCE_IDR if force_idr else 0) | (CC_FRAME_DEVICE if device_frame else 0)
        _check(self.L, self.L.cc_enc_submit(self._h, C.byref(f), flags), "cc_enc_submit")

    def prepare_sequence(self, frames, bitrates_kbps):
        """Frame descriptors (cc_frame: plane pointers, strides, pts) and the per-frame bitrates of a sequence as C arrays.
        Building them is interpreter work (it holds the GIL); callers that time encode_sequence() build them beforehand."""
        n = len(frames)
        arr = (Frame * n)()
        for i, (y, u, v) in enumerate(frames):
            f = self._frame(y, u, v, i)
            arr[i].plane, arr[i].stride, arr[i].pts = f.plane, f.stride, f.pts
        br = (C.c_int * n)(*[int(b) * 1000 for b in bitrates_kbps])
        return arr, br, n, frames  # `frames` keeps the source arrays alive

    def encode_sequence(self, frames, bitrates_kbps=None, device_frame=False, stats=None):
        """Consecutive frames in one C call (cc_enc_encode_sequence): frames = [(y, u, v)] as for encode(), or the result of
        prepare_sequence(); bitrates in kbit/s (the `bitrate` property, written only when it changes).
        Returns the accumulated SeqStats."""
        arr, br, n, _keep = frames if bitrates_kbps is None else self.prepare_sequence(frames, bitrates_kbps)
        stats = stats if stats is not None else SeqStats()
        _check(self.L, self.L.cc_enc_encode_sequence(self._h, arr, br, n, CC_FRAME_DEVICE if device_frame else 0, C.byref(stats), C.byref(self._pkt)),
               "cc_enc_encode_sequence")
        return stats

    def collect(self, copy=True):
        _check(self.L, self.L.cc_enc_collect(self._h, C.byref(self._pkt)), "cc_enc_collect")
        return C.string_at(self._pkt.data, self._pkt.size) if copy else self._pkt.size

    @property
    def last(self):
        """Statistics of the last access unit (size, is_idr, qp, MB counts, gpu_ms, host_ms)."""
        return self._pkt

    # -- test hooks --
    def recon(self):
        W, H = self.mb_w * 16, self.mb_h * 16
        y = np.zeros((H, W), np.uint8); u = np.zeros((H // 2, W // 2), np.uint8); v = np.zeros_like(u)
        _check(self.L, self.L.cc_enc_get_recon(self._h, y.ctypes.data, u.ctypes.data, v.ctypes.data), "cc_enc_get_recon")
        return y, u, v

    def motion(self):
        n = self.mb_w * self.mb_h
        mv = np.zeros((n, 4), np.int16); satd = np.zeros(n, np.int32)
        _check(self.L, self.L.cc_enc_get_me(self._h, mv.ctypes.data, satd.ctypes.data), "cc_enc_get_me")
        return mv, satd

    def motion8(self):
        """8x8-partition search results of the last P frame: (mv[n,4,2] quarter-pel, satd[n,4], part[n])."""
        n = self.mb_w * self.mb_h
        mv = np.zeros((n, 4, 2), np.int16); satd = np.zeros((n, 4), np.int32); part = np.zeros(n, np.uint8)
        _check(self.L, self.L.cc_enc_get_me8(self._h, mv.ctypes.data, satd.ctypes.data, part.ctypes.data), "cc_enc_get_me8")
        return mv, satd, part

    def mbinfo(self):
        a = np.zeros((self.mb_w * self.mb_h, 8), np.uint8)
        _check(self.L, self.L.cc_enc_get_mbinfo(self._h, a.ctypes.data), "cc_enc_get_mbinfo")
        return a

    STAGES = ("upload", "ingest", "me_coarse", "me_fine", "p_code", "mb_encode", "pack", "deblock", "border", "hpel", "l1ref", "wait_ref")

    def set_profile(self, on=True):
        _check(self.L, self.L.cc_enc_set_profile(self._h, int(on)), "cc_enc_set_profile")

    def stage_ms(self):
        """{stage: (total_ms, launches)} measured with CUDA events on the encoder's stream since set_profile()."""
        ms = (C.c_double * len(self.STAGES))(); n = (C.c_long * len(self.STAGES))()
        _check(self.L, self.L.cc_enc_get_stage_ms(self._h, ms, n), "cc_enc_get_stage_ms")
        return {s: (ms[i], n[i]) for i, s in enumerate(self.STAGES)}

    def staged_frames(self):
        """frames that arrived in pageable host memory and went through the handle's pinned staging frame"""
        return self.L.cc_enc_staged_frames(self._h)

    def overlapped_frames(self):
        """Frames whose in-loop filter was launched next to the macroblock kernel (cc_enc_overlapped_frames, include/ceracuda.h)."""
        return self.L.cc_enc_overlapped_frames(self._h)

    def kernel_launches(self):
        return self.L.cc_enc_kernel_launches(self._h)

    def close(self):
        if getattr(self, "_h", None):
            self.L.cc_enc_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
