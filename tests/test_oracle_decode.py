"""Oracle pin: every stream of the CPU oracle decodes with libavcodec's h264 decoder (the decoder avdec_h264
wraps) and the decoded Y,U,V equal the oracle's reconstruction exactly, every frame."""
import numpy as np
import pytest
from ceracoder_b200.synth import make_source
from h264_decode import decode_aus, split_aus
from oracle_lib import OracleEncoder


def _run(src, w, h, qps, rps, gop=60, partitions=True):
    enc = OracleEncoder(w, h, 30, rps, partitions=partitions)
    aus, recs = [], []
    for i, qp in enumerate(qps):
        y, u, v = src.frame(i)
        aus.append(enc.encode(y, u, v, i % gop == 0, qp))
        recs.append(enc.recon())
    return aus, recs


@pytest.mark.parametrize("name,w,h,rps,qps", [
    ("pan", 64, 48, 3, [30, 30, 24, 36]),
    ("pan", 176, 144, 4, [26, 28, 30, 32, 22, 40]),
    ("pan", 200, 120, 2, [20, 33, 33, 45, 51, 12]),   # width/height not multiples of 16 -> cropping
    ("testsrc", 160, 96, 6, [28, 28, 28]),
    ("ball", 128, 80, 1, [35, 30, 25, 20]),
])
@pytest.mark.parametrize("partitions", [False, True])
def test_decoder_equals_recon(name, w, h, rps, qps, partitions):
    src = make_source(name, w, h)
    aus, recs = _run(src, w, h, qps, rps, gop=4, partitions=partitions)
    frames = decode_aus(aus)
    assert len(frames) == len(aus)
    for (y, u, v), (ry, ru, rv) in zip(frames, recs):
        assert y.shape == (h, w)
        assert (y == ry[:h, :w]).all()
        assert (u == ru[:h // 2, :w // 2]).all() and (v == rv[:h // 2, :w // 2]).all()
    assert [len(a) for a in split_aus(b"".join(aus))] == [len(a) for a in aus]


def test_random_noise_high_and_low_qp():
    rng = np.random.default_rng(11)
    w, h = 96, 64
    enc = OracleEncoder(w, h, 30, 2)
    aus, recs = [], []
    for i, qp in enumerate([10, 10, 51, 51, 18]):
        y = rng.integers(0, 256, (h, w), dtype=np.uint8); u = rng.integers(0, 256, (h // 2, w // 2), dtype=np.uint8); v = u[::-1].copy()
        aus.append(enc.encode(y, u, v, i == 0, qp)); recs.append(enc.recon())
    for (y, u, v), (ry, ru, rv) in zip(decode_aus(aus), recs):
        assert (y == ry[:h, :w]).all() and (u == ru[:h // 2, :w // 2]).all() and (v == rv[:h // 2, :w // 2]).all()


def test_psnr_sane():
    w, h = 176, 144
    src = make_source("pan", w, h)
    aus, recs = _run(src, w, h, [26] * 4, 3)
    for i, (ry, _, _) in enumerate(recs):
        y = src.frame(i)[0].astype(float)
        mse = ((y - ry[:h, :w]) ** 2).mean()
        assert 10 * np.log10(255 ** 2 / mse) > 34
