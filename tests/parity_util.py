"""Shared helpers for encoder parity tests (emulated on CPU, or the real library on a B200)."""
import numpy as np
from ceracoder_b200 import CudaH264Enc
from ceracoder_b200.synth import make_source
from h264_decode import decode_aus
from oracle_lib import OracleEncoder


def oracle_mb_table(orc):
    n = orc.mb_w * orc.mb_h
    rows = []
    for j in range(n):
        o = orc.mb_struct(j)
        rows.append((o.type, o.i16_mode, o.chroma_mode, o.cbp, o.l1_mv[0], o.l1_mv[1], o.me_mv[0], o.me_mv[1], o.me_satd, o.me_part, o.part, o.t8))
    return np.array(rows, dtype=np.int64)


def oracle_me8_table(orc):
    n = orc.mb_w * orc.mb_h
    mv = np.zeros((n, 4, 2), np.int64); satd = np.zeros((n, 4), np.int64)
    for j in range(n):
        o = orc.mb_struct(j)
        mv[j] = np.array(o.me_mv8[:]).reshape(4, 2); satd[j] = o.me_satd8[:]
    return mv, satd


def run_parity(lib, source, w, h, qps, rps=4, gop=60, fmt="I420", check_decode=True, frames=None, partitions=True):
    """Encode the same frames at the same per-frame QP with libceracuda and with the CPU oracle; assert that the
    access units are byte-identical, the reconstructions identical, and (optionally) that libavcodec decodes the
    stream to exactly that reconstruction. Returns the list of access units."""
    src = make_source(source, w, h) if isinstance(source, str) else source
    enc = CudaH264Enc(w, h, fps=(30, 1), rows_per_slice=rps, key_int_max=gop, lib=lib, fmt=fmt, partitions=partitions)
    orc = OracleEncoder(w, h, 30, rps, partitions=partitions)
    aus, recs = [], []
    try:
        for i, qp in enumerate(qps):
            y, u, v = src.frame(i) if frames is None else frames[i]
            if fmt == "NV12":
                uv = np.empty((h // 2, w), np.uint8); uv[:, 0::2] = u; uv[:, 1::2] = v
                a = enc.encode(y, uv, None, qp=qp)
            else:
                a = enc.encode(y, u, v, qp=qp)
            b = orc.encode(y, u, v, i % gop == 0, qp)
            ry, ru, rv = enc.recon()
            oy, ou, ov = orc.recon()
            if i % gop:
                mv, satd = enc.motion()
                t = oracle_mb_table(orc)
                assert (mv[:, :2] == t[:, 4:6]).all(), f"frame {i}: coarse motion vectors differ"
                assert (mv[:, 2:] == t[:, 6:8]).all(), f"frame {i}: refined motion vectors differ"
                assert (satd == t[:, 8]).all(), f"frame {i}: motion SATD differs"
                mv8, satd8, part = enc.motion8()
                assert (part == t[:, 9]).all(), f"frame {i}: partition choice of the search differs at {np.argwhere(part != t[:, 9])[:4].ravel()}"
                if partitions:
                    omv8, osatd8 = oracle_me8_table(orc)
                    assert (mv8 == omv8).all(), f"frame {i}: 8x8 search vectors differ at {np.argwhere(mv8 != omv8)[:3]}"
                    assert (satd8 == osatd8).all(), f"frame {i}: 8x8 search SATD differs"
            info = enc.mbinfo()
            t = oracle_mb_table(orc)
            assert (info[:, 0] == t[:, 0]).all(), f"frame {i}: macroblock types differ at {np.argwhere(info[:, 0] != t[:, 0])[:4].ravel()}"
            assert (info[:, 2] == t[:, 3]).all(), f"frame {i}: cbp differs"
            assert (info[:, 3] == (t[:, 11] | ((t[:, 10] & (t[:, 0] == 1)) << 1))).all(), f"frame {i}: transform size / P_8x8 flags differ"
            assert (ry == oy).all(), f"frame {i}: luma reconstruction differs at {np.argwhere(ry != oy)[:3]}"
            assert (ru == ou).all() and (rv == ov).all(), f"frame {i}: chroma reconstruction differs"
            assert a == b, f"frame {i}: access units differ ({len(a)} vs {len(b)} bytes)"
            assert enc.last.qp == qp and enc.last.is_idr == (i % gop == 0)
            aus.append(a); recs.append((ry, ru, rv))
        if check_decode:
            dec = decode_aus(aus)
            assert len(dec) == len(aus)
            for (y, u, v), (ry, ru, rv) in zip(dec, recs):
                assert (y == ry[:h, :w]).all() and (u == ru[:h // 2, :w // 2]).all() and (v == rv[:h // 2, :w // 2]).all()
    finally:
        enc.close(); orc.close()
    return aus


def psnr(a, b):
    mse = ((a.astype(np.float64) - b.astype(np.float64)) ** 2).mean()
    return 99.0 if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)


def run_rc_parity(lib, src, w, h, fps, n, kbps_of, gop=60, rps=4, recon_every=10, host_entropy=False):
    """Rate control ON: the CUDA encoder runs its own per-frame QP decisions under the bitrate schedule kbps_of(i); the oracle is
    then driven with exactly the frame types and QPs the encoder chose (the QP trace), and every access unit must be
    byte-identical — the whole stream, across IDR restarts and bitrate steps.  Returns (access units, qp trace)."""
    enc = CudaH264Enc(w, h, fps=(fps, 1), bitrate=kbps_of(0), rows_per_slice=rps, key_int_max=gop, lib=lib, host_entropy=host_entropy)
    orc = OracleEncoder(w, h, fps, rps)
    aus, trace = [], []
    try:
        for i in range(n):
            y, u, v = src.frame(i)
            enc.set_property("bitrate", kbps_of(i))
            a = enc.encode(y, u, v)
            qp, idr = enc.last.qp, bool(enc.last.is_idr)
            b = orc.encode(y, u, v, idr, qp)
            assert a == b, f"frame {i} (qp {qp}, idr {idr}): access units differ ({len(a)} vs {len(b)} bytes)"
            if i % recon_every == recon_every - 1 or i == n - 1:
                for p, q in zip(enc.recon(), orc.recon()):
                    assert (p == q).all(), f"frame {i}: reconstruction differs"
            aus.append(a); trace.append((int(idr), qp, len(a)))
    finally:
        enc.close(); orc.close()
    return aus, trace
