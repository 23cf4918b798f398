"""ctypes binding of the CPU oracle (oracle/libceraoracle.so). Test infrastructure only."""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(ROOT, "oracle", "libceraoracle.so")


class MbInfo(C.Structure):
    _fields_ = [("type", C.c_uint8), ("i16_mode", C.c_uint8), ("chroma_mode", C.c_uint8), ("cbp", C.c_uint8),
                ("i4_mode", C.c_uint8 * 16), ("mv", C.c_int16 * 2), ("mvd", C.c_int16 * 2), ("nnz", C.c_uint8 * 28),
                ("me_mv", C.c_int16 * 2), ("me_satd", C.c_int32), ("l1_mv", C.c_int16 * 2), ("bits", C.c_int32),
                ("t8", C.c_uint8), ("part", C.c_uint8), ("me_part", C.c_uint8), ("pad8", C.c_uint8),
                ("mv4", C.c_int16 * 8), ("mvd4", C.c_int16 * 8), ("me_mv8", C.c_int16 * 8), ("me_satd8", C.c_int32 * 4),
                ("me_cost16", C.c_int32), ("me_cost8", C.c_int32),
                ("dbg_i4_blk", C.c_uint16 * 16), ("dbg_i4_blk_exact", C.c_int32 * 16), ("dbg_est_intra", C.c_int32), ("dbg_cost_inter", C.c_int32), ("dbg_cost_i16", C.c_int32), ("dbg_cost_i4", C.c_int32)]


def build():
    srcs = [os.path.join(ROOT, "oracle", f) for f in os.listdir(os.path.join(ROOT, "oracle")) if f.endswith((".c", ".h"))]
    if os.path.exists(_SO) and all(os.path.getmtime(_SO) >= os.path.getmtime(s) for s in srcs):
        return _SO
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        L = _lib
        L.ref_enc_open.restype = C.c_void_p
        L.ref_enc_open.argtypes = [C.c_int] * 5
        L.ref_enc_close.argtypes = [C.c_void_p]
        L.ref_enc_set_partitions.argtypes = [C.c_void_p, C.c_int]
        L.ref_set_threads.argtypes = [C.c_int]
        L.ref_enc_encode.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.ref_enc_get_recon.argtypes = [C.c_void_p] * 4
        L.ref_enc_mb_info.restype = C.POINTER(MbInfo)
        L.ref_enc_mb_info.argtypes = [C.c_void_p]
        L.ref_enc_mb_w.argtypes = [C.c_void_p]
        L.ref_enc_mb_h.argtypes = [C.c_void_p]
    return _lib


class OracleEncoder:
    def __init__(self, w, h, fps=30, rows_per_slice=4, partitions=True):
        self.L = lib()
        self.h_ = self.L.ref_enc_open(w, h, fps, 1, rows_per_slice)
        assert self.h_
        self.L.ref_enc_set_partitions(self.h_, int(partitions))
        self.w, self.h = w, h
        self.mb_w, self.mb_h = self.L.ref_enc_mb_w(self.h_), self.L.ref_enc_mb_h(self.h_)
        self.out = np.zeros(self.mb_w * self.mb_h * 1024 + 4096, np.uint8)

    def encode(self, y, u, v, idr, qp):
        y, u, v = (np.ascontiguousarray(p) for p in (y, u, v))
        n = self.L.ref_enc_encode(self.h_, y.ctypes.data, y.strides[0], u.ctypes.data, u.strides[0], v.ctypes.data, v.strides[0],
                                  int(idr), int(qp), self.out.ctypes.data, self.out.size)
        assert n > 0, n
        return self.out[:n].tobytes()

    def recon(self):
        W, H = self.mb_w * 16, self.mb_h * 16
        y = np.zeros((H, W), np.uint8); u = np.zeros((H // 2, W // 2), np.uint8); v = np.zeros_like(u)
        self.L.ref_enc_get_recon(self.h_, y.ctypes.data, u.ctypes.data, v.ctypes.data)
        return y, u, v

    def mb_info(self):
        p = self.L.ref_enc_mb_info(self.h_)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(self.mb_w * self.mb_h, C.sizeof(MbInfo))).copy()

    def mb_struct(self, i):
        return self.L.ref_enc_mb_info(self.h_)[i]

    def close(self):
        if self.h_:
            self.L.ref_enc_close(self.h_)
            self.h_ = None

    def __del__(self):
        self.close()
