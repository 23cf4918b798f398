"""The QP trace behind `bench.py --impl reference`: the CPU arm replays the frame types and QPs the CUDA arm's
rate control chose for stream 0 of the bench workload, so the two arms run at equal bitrate with byte-identical output."""
import os
import zlib
import pytest
from oracle_lib import ROOT

TRACE = os.path.join(ROOT, "tests", "golden", "qp_trace_1080p60.csv")


def _bench():
    import importlib
    import sys
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    return importlib.import_module("bench")


def test_trace_is_committed_and_well_formed():
    b = _bench()
    t = b.load_trace()
    assert len(t) >= 600, "the reference arm needs (steps + warmup) * 12 frames: 300 for the driver's --steps 20 --warmup 5"
    assert [i for i, r in enumerate(t[:200]) if r[0]] == [0, 60, 120, 180]          # GOP 60
    qps = [r[1] for r in t]
    assert 10 <= min(qps) and max(qps) <= 51 and max(qps) - min(qps) >= 8          # the 12 -> 2 Mbit/s drop moves the QP
    # the schedule the trace was recorded under: 12 Mbit/s for the first second, then the drop
    assert b.schedule_kbps(0) == 12000 and b.schedule_kbps(59) == 12000 and b.schedule_kbps(60) == 2000
    hi = sum(r[2] for r in t[:60]) * 8 / 1e6
    lo = sum(r[2] for r in t[120:180]) * 8 / 1e6
    assert 9.0 < hi < 15.0 and lo < 4.0, f"bits of the first second {hi:.1f} Mbit and of the third {lo:.1f} Mbit do not look like the schedule"


def test_both_arms_describe_the_same_config():
    """the driver compares the two arms' `config` objects: they come from one function"""
    import argparse
    b = _bench()
    a = argparse.Namespace(streams=8, input_format="I420", frames_per_step=120, rows_per_slice=4, partitions=1, host_entropy=False, pool=900)
    assert b.workload_config(a, 1) == b.workload_config(a, 1)
    assert "frames_per_step" in b.workload_config(a, 1) and b.workload_config(a, 2)["parallelism"] == "replicas x2"


@pytest.mark.gpu
def test_committed_trace_is_what_the_encoder_does_today():
    """stale-trace guard: stream 0 of the bench workload, coded now, reproduces the committed (frame type, QP, size, CRC) lines"""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import record_qp_trace
    b = _bench()
    t = b.load_trace()
    rows = record_qp_trace.record(150)  # on the full frame store, which is what the trace was recorded on
    for r, g in zip(rows, t):
        assert (r[1], r[2], r[3], r[4]) == g, f"frame {r[0]}: encoder ({r[1:]}) != committed trace {g}: re-record with tools/record_qp_trace.py"


@pytest.mark.gpu
def test_oracle_replay_of_the_trace_is_byte_identical():
    """what `bench.py --impl reference` does, on the first 24 frames: same sizes, same CRCs"""
    b = _bench()
    store = b.FrameStore(b.POOL, 0)
    fps, timed, same = b.replay_oracle(store, b.load_trace(), 24, 12, 0, os.cpu_count())
    assert same and fps > 0
