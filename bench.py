#!/usr/bin/env python
"""bench.py — 1080p H.264 encode throughput of libceracuda on B200 (BASELINE.json metric), one JSON line.

Workload (config.workload): BASELINE configs[1] in the shape the target is defined on — 1080p60 `cudah264enc` streams under a
balancer-shaped 12 -> 2 -> 12 Mbit/s bitrate schedule, 8 concurrent streams per GPU (north_star: "8 concurrent 1080p60 streams
per GPU"), synthetic S2 pan-sprites content as SURVEY.md §8d specifies it: 900 distinct frames = three scenes of 300, played in
a loop, so every stream meets a scene cut every 300 frames.  A *step* = every stream of this rank encodes FRAMES_PER_STEP
frames.
  value         frames/s with the frames already resident in HBM (CC_FRAME_DEVICE)
  e2e           the same through the public C-ABI call with page-locked HOST frames (H2D inside) and the access unit in host memory
  e2e_pageable  the same with pageable host frames (what a mapped GstBuffer is): the library's pinned staging copy is inside
Multi-GPU = independent replicas (one process per GPU, no collective on the data path): scaling "weak".

`--impl reference`: the CPU arm.  x264enc / libx264 do not exist on this image, so it is the oracle port (oracle/, kind "port"),
all host threads, on a bounded sample of the SAME workload at EQUAL BITRATE: it encodes stream 0's frames with the frame types and
QPs the CUDA arm's rate control chose for them (tests/golden/qp_trace_1080p60.csv, recorded by tools/record_qp_trace.py), which
makes its access units byte-identical to the CUDA arm's (checked frame by frame against the sizes and CRCs in the trace).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl cuda|reference]
"""
import argparse
import csv
import ctypes as C
import json
import mmap
import os
import subprocess
import sys
import tempfile
import threading
import time
import zlib

import numpy as np

# more than 8 streams per process need more hardware queues than the default 8, or they serialise pairwise
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, FPS = 1920, 1080, 60
FSZ = W * H * 3 // 2
STREAMS = 8
FRAMES_PER_STEP = 200           # per stream; the driver's --steps 20 gives 4000 timed frames per stream (> 5 s at 6000 frames/s)
POOL = 900                      # S2 frames 0..899: scenes of 300, looped -> a cut every 300 frames
STREAM_SHIFT = 113              # stream k plays the loop 113 frames ahead of stream k - 1
GOP_PHASE = 7                   # and sits 7 frames further into its GOP (live streams do not emit their IDR frames in the same instant)
ALG_BYTES_PER_FRAME = 4.5 * W * H   # read cur + read ref + write recon (+ bitstream)
METRIC = "1080p H.264 encode fps per B200 at equal bitrate"
TRACE = os.path.join(ROOT, "tests", "golden", "qp_trace_1080p60.csv")
REF_FRAMES_PER_STEP = 12        # reference arm: frames of stream 0 per step (bounded sample; the oracle runs at tens of frames/s)


def schedule_kbps(frame_idx):
    """Balancer-shaped target for config 2 (the reference's adaptive ladder, /root/reference/src/core/bitrate_control.c:176-206):
    12 Mbit/s, emergency drop to 2 Mbit/s after 1 s, then +100..300 kbit/s every 520 ms back to 12; period 30 s."""
    t = (frame_idx / FPS + 4.0) % 30.0
    if t < 5.0:
        return 12000
    kb = 2000
    for _ in range(int((t - 5.0) / 0.52)):
        kb = min(12000, kb + 30 + kb // 30)
        kb = kb // 100 * 100 if kb % 100 == 0 else (kb // 100 + 1) * 100
    return kb


def pingpong(i, n):
    p = i % (2 * n - 2)
    return p if p < n else 2 * n - 2 - p


# ---- the frame store: S2 frames as tight I420, in a shared-memory file all ranks (and both arms) map ---------------------------
def _gen_chunk(job):
    path, a, b, n = job
    from ceracoder_b200.synth import make_source
    arr = np.memmap(path, dtype=np.uint8, mode="r+", shape=(n, FSZ))
    src = make_source("pan", W, H)
    for i in range(a, b):
        y, u, v = src.frame(i)
        arr[i, :W * H] = y.ravel(); arr[i, W * H:W * H * 5 // 4] = u.ravel(); arr[i, W * H * 5 // 4:] = v.ravel()
    arr.flush()
    return b - a


def _store_dir(nbytes):
    for d in ("/dev/shm", tempfile.gettempdir()):
        try:
            st = os.statvfs(d)
            if os.access(d, os.W_OK) and st.f_bavail * st.f_frsize > nbytes * 1.1:
                return d
        except OSError:
            pass
    return tempfile.gettempdir()


class FrameStore:
    """n S2 frames.  n = POOL: the specified sequence, looped (index = sequence position mod n); smaller n (--pool, for ncu
    runs): the first n frames played ping-pong.  One generator per box (local rank 0, one process per host core); everybody maps
    the same file twice: one mapping gets page-locked (cc_host_register), the other stays pageable."""

    def __init__(self, n, local_rank=0):
        self.n = n
        nbytes = n * FSZ
        done = None
        for d in ("/dev/shm", tempfile.gettempdir()):  # a store an earlier run on this box left behind
            p = os.path.join(d, f"ceracuda_s2_{W}x{H}_{n}_v2.i420")
            if os.path.exists(p + ".done") and os.path.getsize(p) == nbytes:
                self.path, done = p, p + ".done"
                break
        t0 = time.perf_counter()
        if done is None:
            self.path = os.path.join(_store_dir(nbytes), f"ceracuda_s2_{W}x{H}_{n}_v2.i420")
            done = self.path + ".done"
            if local_rank == 0:
                with open(self.path, "wb") as f:
                    f.truncate(nbytes)
                import multiprocessing as mp
                procs = max(1, min(os.cpu_count() or 1, 32, n // 8 or 1))
                step = (n + procs - 1) // procs
                jobs = [(self.path, a, min(a + step, n), n) for a in range(0, n, step)]
                # fork: the children only draw frames with numpy (no CUDA, no threads of the parent are touched); spawn would need an
                # importable __main__, which a harness that feeds this module from stdin does not have
                with mp.get_context("fork").Pool(procs) as pool:
                    pool.map(_gen_chunk, jobs)
                with open(done + ".tmp", "w") as f:
                    f.write("ok\n")
                os.replace(done + ".tmp", done)
            else:
                while not os.path.exists(done):
                    time.sleep(0.2)
                    if time.perf_counter() - t0 > 900:
                        raise RuntimeError("frame store was not generated by local rank 0")
        self.gen_s = time.perf_counter() - t0
        self._f = open(self.path, "r+b")
        self._m1 = mmap.mmap(self._f.fileno(), nbytes)            # to be page-locked
        self._m2 = mmap.mmap(self._f.fileno(), nbytes)            # stays pageable
        self.arr = np.frombuffer(self._m1, dtype=np.uint8).reshape(n, FSZ)
        self.arr_pageable = np.frombuffer(self._m2, dtype=np.uint8).reshape(n, FSZ)

    def index(self, k, i):
        """pool frame of stream k at sequence position i"""
        return (i + STREAM_SHIFT * k) % self.n if self.n >= POOL else pingpong(i + 5 * k, self.n)

    def planes(self, k, i, pageable=False):
        f = (self.arr_pageable if pageable else self.arr)[self.index(k, i)]
        return f[:W * H].reshape(H, W), f[W * H:W * H * 5 // 4].reshape(H // 2, W // 2), f[W * H * 5 // 4:].reshape(H // 2, W // 2)


def load_trace():
    """[(is_idr, qp, au_bytes, crc32)] of stream 0 of this workload, as the CUDA arm codes it (tools/record_qp_trace.py)"""
    with open(TRACE) as f:
        rows = [r for r in csv.reader(f) if r and not r[0].startswith("#")]
    return [(int(r[1]), int(r[2]), int(r[3]), int(r[4], 16)) for r in rows[1:]]


def dist_info():
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def aggregate(local_frames, local_ms, world, backend="nccl", device=None):
    """whole-job throughput: sum of frames over ranks / max of device times over ranks (driver contract)."""
    if world == 1:
        return local_frames, local_ms
    import torch
    import torch.distributed as dist
    t = torch.tensor([float(local_frames), float(local_ms)], dtype=torch.float64, device=device or "cpu")
    frames = t[:1].clone(); ms = t[1:].clone()
    dist.all_reduce(frames, op=dist.ReduceOp.SUM)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return frames.item(), ms.item()


class ClockSampler:
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu):
        self.gpu, self.p = gpu, None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200", "-i", str(self.gpu)],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.p = None

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        out = self.p.communicate()[0]
        sm, mx, reasons = [], [], set()
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}


def prepare_streams(encs, frames_of, n_frames, start_idx, warm=False):
    """Per stream the C arrays of frame descriptors (pointers into the resident pool, strides, pts) and bitrates.  Interpreter
    work under the GIL (8 threads would serialise on it), so it is done before the timed region; the frames themselves are
    copied / encoded inside it.  Stream k runs GOP_PHASE * k frames ahead of stream 0 (its warm-up is that much longer)."""
    out = []
    for k, e in enumerate(encs):
        n = n_frames + (GOP_PHASE * k if warm else 0)
        first = start_idx + (0 if warm else GOP_PHASE * k)
        out.append(e.prepare_sequence([frames_of(k, first + i) for i in range(n)], [schedule_kbps(first + i + 37 * k) for i in range(n)]))
    return out


STAT_KEYS = ("frames", "au_bytes", "sym_bytes", "n_skip", "n_inter", "n_intra", "n_intra_analysed", "n_skip_probed", "submit_ms", "wait_ms", "host_ms", "gpu_ms", "n_idr")


def run_streams(encs, prepared, device_frames):
    """Each encoder on its own host thread / CUDA stream; one C call per stream (cc_enc_encode_sequence: the frame loop the element's
    streaming thread runs, off the interpreter lock).  Returns the summed statistics."""
    tot = dict.fromkeys(STAT_KEYS, 0)
    lock = threading.Lock()
    errs = []

    def work(k, e):
        try:
            s = e.encode_sequence(prepared[k], device_frame=device_frames)
            with lock:
                for key in STAT_KEYS:
                    tot[key] += getattr(s, key)
        except Exception as ex:  # noqa: BLE001
            errs.append(ex)
    th = [threading.Thread(target=work, args=(k, e)) for k, e in enumerate(encs)]
    [t.start() for t in th]; [t.join() for t in th]
    if errs:
        raise errs[0]
    return tot


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)["hbm_gbs"], "measured (MEASURED_PEAKS.json)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed `ncu --set full` capture
    (profiles/traffic.json, written by tools/summarize_profiles.py); None if no capture is committed."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)[kernel]
        return t["dram_bytes_read"] + t["dram_bytes_write"]
    except Exception:  # noqa: BLE001
        return None


def int_peaks(L):
    out = {}
    for kind, name in enumerate(("vabsdiff4_acc", "iadd", "lop3", "idp4a")):
        g = C.c_double()
        if L.cc_bench_int_peak(kind, C.byref(g)) == 0:
            out[name] = g.value
    return out


def workload_config(args, world):
    """the `config` object of the JSON line — the same for both arms (the reference arm runs a bounded sample of this workload)"""
    S = args.streams
    return {"workload": f"{S} concurrent 1080p60 cudah264enc streams per GPU, balancer-shaped 12->2->12 Mbps schedule (configs[1] x{S}, north_star target shape)",
            "width": W, "height": H, "fps": FPS, "input_format": args.input_format, "streams_per_gpu": S, "frames_per_step": args.frames_per_step * S,
            "frames_per_stream_per_step": args.frames_per_step, "rows_per_slice": args.rows_per_slice, "partitions": "16x16+8x8" if args.partitions else "16x16",
            "key_int_max": 60, "entropy_coder": "host C (cc_cavlc.c)" if args.host_entropy else "device (k_cavlc_size + k_cavlc_write)",
            "content": (f"S2 pan-sprites (SURVEY.md 8d): {POOL} distinct frames = 3 scenes of 300, looped: a scene cut every 300 frames; stream k offset by {STREAM_SHIFT} k frames"
                        if args.pool >= POOL else f"S2 pan-sprites, first {args.pool} frames played ping-pong (--pool: short runs under ncu only)"),
            "bitrate_schedule": "12 Mbit/s for 1 s, drop to 2 Mbit/s, +100..300 kbit/s every 520 ms back to 12 (period 30 s); per-frame QP by the encoder's rate control",
            "l2": "inputs (2.8 GB frame store) + 8 streams' planes exceed the 126 MB L2; no explicit flush", "parallelism": f"replicas x{world}"}


def replay_oracle(store, trace, n_frames, chunk, warm_chunks, threads):
    """The CPU arm: stream 0's frames through the oracle with the (frame type, QP) trace of the CUDA arm.  Returns
    (frames/s over the timed chunks, per-chunk seconds, every access unit matched the trace's size and CRC)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_lib import OracleEncoder, lib as oracle_lib_handle
    oracle_lib_handle().ref_set_threads(threads)  # torchrun exports OMP_NUM_THREADS=1; the baseline uses every host thread
    enc = OracleEncoder(W, H, FPS, 8)  # the library default rows_per_slice (what the trace was recorded with)
    secs, same = [], True
    for c in range(0, n_frames, chunk):
        t0 = time.perf_counter()
        for i in range(c, min(c + chunk, n_frames)):
            idr, qp, size, crc = trace[i]
            y, u, v = store.planes(0, i)
            au = enc.encode(y, u, v, bool(idr), qp)
            same = same and len(au) == size and (zlib.crc32(au) & 0xffffffff) == crc
        secs.append(time.perf_counter() - t0)
    enc.close()
    timed = secs[warm_chunks:]
    frames = n_frames - warm_chunks * chunk
    return frames / max(sum(timed), 1e-9), timed, same


def main_reference(args, rank, world, local):
    if rank != 0:
        return 0
    store = FrameStore(args.pool, 0)
    trace = load_trace()
    chunk = REF_FRAMES_PER_STEP
    n = min((args.steps + args.warmup) * chunk, len(trace))
    cores = os.cpu_count()
    t0 = time.perf_counter()
    fps, timed, same = replay_oracle(store, trace, n, chunk, args.warmup, cores)
    kbps = sum(t[2] for t in trace[args.warmup * chunk:n]) * 8 / max((n - args.warmup * chunk) / FPS, 1e-9) / 1e3
    line = {"impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * float(np.mean(timed)) if timed else None, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32",
            "data": "synthetic S2 pan-sprites 1080p (ceracoder_b200/synth.py), the CUDA arm's frame store",
            "config": workload_config(args, args.gpus),
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port",
                             "sample": f"frames 0..{n - 1} of stream 0 of the workload ({chunk} per step, the first {args.warmup} steps untimed, IDR frames included), coded by the "
                                       f"oracle (OpenMP, {cores} host threads) with the frame types and QPs of the CUDA arm's rate control (tests/golden/qp_trace_1080p60.csv): "
                                       f"equal bitrate, {kbps:.0f} kbit/s over the timed frames",
                             "note": "x264enc / libx264 are not installed on this box: this is the repo's CPU port of the same encoder, not x264"},
            "same_output_as_cuda_arm": bool(same), "achieved_kbps": kbps, "wall_s": time.perf_counter() - t0,
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--streams", type=int, default=STREAMS)
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the CPU-baseline and quality legs (kernel experiments, ncu runs)")
    ap.add_argument("--rows-per-slice", type=int, default=8, help="macroblock rows per slice (experiments; the reported workload uses the library default, 8)")
    ap.add_argument("--partitions", type=int, default=1, help="0: 16x16 partitions only (experiments)")
    ap.add_argument("--input-format", default="I420", choices=["I420", "YUY2", "BGRx"],
                    help="frame format handed to the encoder; packed formats are converted inside the ingest kernel (SURVEY 8f N2); needs --pool <= 100")
    ap.add_argument("--host-entropy", action="store_true", help="CAVLC by the host C writer instead of the device kernels (round-1 path; comparison runs)")
    ap.add_argument("--frames-per-step", type=int, default=FRAMES_PER_STEP, help="frames each stream encodes per step (shorten for ncu runs)")
    ap.add_argument("--pool", type=int, default=POOL, help=f"distinct frames; {POOL} = the specified S2 sequence looped; smaller = ping-pong over the first frames (ncu runs only)")
    args = ap.parse_args()
    rank, world, local = dist_info()
    if args.impl == "reference":
        return main_reference(args, rank, world, local)

    store = FrameStore(args.pool, local)  # before anything CUDA in this process
    import torch
    import torch.distributed as dist
    assert torch.cuda.is_available(), "bench.py --impl cuda needs a CUDA device (there is no CPU fallback)"
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from ceracoder_b200 import CudaH264Enc, load_library
    L = load_library()
    S, FMT, n_pool, fps_step = args.streams, args.input_format, store.n, args.frames_per_step
    fsz = W * H * {"I420": 3, "YUY2": 4, "BGRx": 8}[FMT] // 2
    if FMT == "I420":
        harr, parr = store.arr, store.arr_pageable
    else:  # packed copies of the same pictures (the encoder converts back with its integer matrix)
        assert n_pool <= 100, "--input-format YUY2/BGRx: use --pool <= 100"
        harr = np.empty((n_pool, fsz), np.uint8)
        for i in range(n_pool):
            fr = store.arr[i]
            y, u, v = fr[:W * H].reshape(H, W), fr[W * H:W * H * 5 // 4].reshape(H // 2, W // 2), fr[W * H * 5 // 4:].reshape(H // 2, W // 2)
            if FMT == "YUY2":
                f = harr[i].reshape(H, 2 * W)
                f[:, 0::2] = y; f[:, 1::4] = np.repeat(u, 2, axis=0); f[:, 3::4] = np.repeat(v, 2, axis=0)
            else:
                yy = (y.astype(np.float32) - 16) * (255 / 219)
                uu = np.repeat(np.repeat(u, 2, 0), 2, 1).astype(np.float32) - 128; vv = np.repeat(np.repeat(v, 2, 0), 2, 1).astype(np.float32) - 128
                f = harr[i].reshape(H, W, 4)
                f[..., 2] = np.clip(yy + 1.7927 * vv, 0, 255); f[..., 1] = np.clip(yy - 0.2132 * uu - 0.5329 * vv, 0, 255)
                f[..., 0] = np.clip(yy + 2.1124 * uu, 0, 255); f[..., 3] = 255
        parr = harr.copy()
    hbase, pbase = harr.ctypes.data, parr.ctypes.data
    assert L.cc_host_register(hbase, harr.nbytes) == 0, "cudaHostRegister of the frame store failed"
    dbase = L.cc_dev_alloc(harr.nbytes)
    assert dbase and L.cc_dev_upload(dbase, hbase, harr.nbytes) == 0

    def planes(base, k, fi):
        p = base + store.index(k, fi) * fsz
        if FMT != "I420":
            return (p, W * (2 if FMT == "YUY2" else 4)), None, None
        return (p, W), (p + W * H, W // 2), (p + W * H * 5 // 4, W // 2)

    def make_encs():
        return [CudaH264Enc(W, H, fps=(FPS, 1), bitrate=12000, key_int_max=60, speed_preset="veryfast", device=local, lib=L, fmt=FMT,
                            rows_per_slice=args.rows_per_slice, partitions=bool(args.partitions), host_entropy=args.host_entropy) for _ in range(S)]

    def timed(base, device_frames, steps, warmup, profile=False):
        encs = make_encs()
        fo = lambda k, fi: planes(base, k, fi)  # noqa: E731
        run_streams(encs, prepare_streams(encs, fo, fps_step * max(warmup, 1), 0, warm=True), device_frames)  # warm-up steps (IDR + RC settle), GOP phases spread
        prepared = prepare_streams(encs, fo, fps_step * steps, fps_step * max(warmup, 1))
        if profile:
            [e.set_profile(True) for e in encs]
        launches0 = sum(e.kernel_launches() for e in encs)
        staged0 = sum(e.staged_frames() for e in encs)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); ev0.record()
        tot = run_streams(encs, prepared, device_frames)
        torch.cuda.synchronize(); ev1.record(); torch.cuda.synchronize()
        wall_ms = 1e3 * (time.perf_counter() - t0)
        ms = max(ev0.elapsed_time(ev1), 1e-3)
        if world > 1:
            dist.barrier()
        tot.update(ms=ms, wall_ms=wall_ms, launches=sum(e.kernel_launches() for e in encs) - launches0,
                   staged=sum(e.staged_frames() for e in encs) - staged0, qps=[e.last.qp for e in encs], stages={},
                   overlapped=sum(e.overlapped_frames() for e in encs))
        if profile:
            for e in encs:
                for s, (tms, n) in e.stage_ms().items():
                    a = tot["stages"].setdefault(s, [0.0, 0]); a[0] += tms; a[1] += n
        [e.close() for e in encs]
        return tot

    sampler = ClockSampler(local)
    sampler.start()
    r_dev = timed(dbase, True, args.steps, args.warmup)
    clocks = sampler.stop()
    # per-stage CUDA events in a pass of their own: with them on a handle launches the in-loop filter AFTER the wavefront kernel (so that a
    # stage's time is that stage's), while the product path measured above launches it next to it (cc_api.c, CC_OVERLAP)
    r_prof = timed(dbase, True, max(args.steps // 2, 1), 1, profile=True)
    r_e2e = timed(hbase, False, args.steps, args.warmup)
    r_page = timed(pbase, False, max(args.steps // 4, 1), 1)

    def quality_leg(n_frames=600):
        """Stream 0 of the workload alone, frame by frame through the public call (pageable frames, as the element hands them over):
        its (frame type, QP, size, CRC) per frame against the committed trace the CPU arm replays, the bitrate error over 1 s
        windows and the Y-PSNR of the reconstruction against the source (the decoder reproduces the reconstruction
        bit for bit — parity tests).  Untimed."""
        enc = CudaH264Enc(W, H, fps=(FPS, 1), bitrate=12000, key_int_max=60, device=local, lib=L, partitions=bool(args.partitions),
                          rows_per_slice=args.rows_per_slice, host_entropy=args.host_entropy)
        try:
            trace = load_trace()
        except Exception:  # noqa: BLE001
            trace = []
        bits, targ, psnr, match = [], [], [], bool(trace)
        for i in range(n_frames):
            kb = schedule_kbps(i)
            enc.set_property("bitrate", kb)
            y, u, v = store.planes(0, i, pageable=True)
            au = enc.encode(y, u, v)
            bits.append(8 * len(au)); targ.append(kb * 1000.0 / FPS)
            if i < len(trace):
                match = match and trace[i] == (int(enc.last.is_idr), int(enc.last.qp), len(au), zlib.crc32(au) & 0xffffffff)
            if i % 12 == 11:
                ry = enc.recon()[0][:H, :W]
                mse = float(((ry.astype(np.float32) - y.astype(np.float32)) ** 2).mean())
                psnr.append((kb, 10 * np.log10(255.0 ** 2 / max(mse, 1e-9))))
        enc.close()
        b, t = np.array(bits, float), np.array(targ, float)
        err = [abs(b[i:i + FPS].sum() - t[i:i + FPS].sum()) / t[i:i + FPS].sum() for i in range(FPS, n_frames - FPS + 1, FPS // 4)]
        hi = [p for k, p in psnr if k >= 10000]; lo = [p for k, p in psnr if k <= 3000]
        return {"frames": n_frames, "schedule": "stream 0 of the workload: 12 Mbit/s for 1 s, drop to 2, ladder back up; scene cut at frame 300",
                "achieved_kbps": float(b.sum() / (n_frames / FPS) / 1e3), "target_kbps": float(t.sum() / (n_frames / FPS) / 1e3),
                "bitrate_error_1s_windows": {"mean": float(np.mean(err)), "p95": float(np.percentile(err, 95)), "note": "first second left out"},
                "y_psnr_db": {"at_10_to_12_mbps": float(np.mean(hi)) if hi else None, "at_2_to_3_mbps": float(np.mean(lo)) if lo else None},
                "qp_trace_match": match, "qp_trace_frames_checked": min(n_frames, len(trace)),
                "vs_x264enc": None, "note": "x264enc / libx264 are not installed on this box: the PSNR delta against x264 veryfast was not measured"}

    quality = quality_leg() if rank == 0 and not args.no_cpu_baseline else None

    dev = torch.device("cuda", local)
    f_dev, ms_dev = aggregate(r_dev["frames"], r_dev["ms"], world, device=dev)
    f_e2e, ms_e2e = aggregate(r_e2e["frames"], r_e2e["ms"], world, device=dev)
    f_pg, ms_pg = aggregate(r_page["frames"], r_page["ms"], world, device=dev)
    launches, _ = aggregate(r_dev["launches"], 0.0, world, device=dev)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    value = f_dev / (ms_dev / 1e3)
    e2e = f_e2e / (ms_e2e / 1e3)
    e2e_pg = f_pg / (ms_pg / 1e3)
    peak, peak_src = peaks()
    st = r_prof["stages"]
    names = {"mb_encode": "k_mb_encode", "me_coarse": "k_me_l1", "me_fine": "k_me_l0", "p_code": "k_p_code", "hpel": "k_hpel", "deblock": "k_deblock",
             "pack": "k_pack" if args.host_entropy else "k_cavlc_size + k_cavlc_write"}
    dom = max((s for s in st if s != "upload"), key=lambda s: st[s][0]) if st else None
    roof = None
    if dom:
        avg_ms = st[dom][0] / max(st[dom][1], 1)
        ach = ALG_BYTES_PER_FRAME / (avg_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": names.get(dom, dom), "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": ncu_traffic(names.get(dom, dom)),
                "avg_launch_ms": avg_ms, "launches_timed": st[dom][1], "peak_source": peak_src,
                "timed_by": "CUDA events around every stage, on the streams the kernels are launched on, in a second pass of the same workload "
                            f"({max(args.steps // 2, 1)} step(s)): with the events on, the in-loop filter is launched after k_mb_encode instead of next to it",
                "algorithmic_bytes_per_launch": ALG_BYTES_PER_FRAME,
                "note": "one launch = one 1080p frame of one stream; 8 streams' launches overlap on the GPU, so the per-launch duration includes sharing; "
                        "aggregate_achieved = value * 4.5*W*H.  The kernel is bound by the macroblock dependency chain, not by HBM (DESIGN.md 5)",
                "aggregate_achieved": value * ALG_BYTES_PER_FRAME / 1e9,
                "stage_ms_per_frame": {s: st[s][0] / max(st[s][1], 1) for s in st}}
    # integer roofline: algorithmic 4-pixel abs-diff-accumulate ops of the exhaustive coarse search
    # (33*33 candidates x 16 ops per macroblock) over k_me_l1's CUDA-event time, against the VABSDIFF4 peak measured here
    introof = None
    if st and "me_coarse" in st and st["me_coarse"][1]:
        pk = int_peaks(L)
        ops = (W // 16) * ((H + 15) // 16) * 33 * 33 * 16
        ms_l1 = st["me_coarse"][0] / st["me_coarse"][1]
        a = ops / (ms_l1 * 1e-3) / 1e9
        introof = {"kernel": "k_me_l1", "op": "4-pixel abs-diff-accumulate (VABSDIFF4.U8.ACC)", "algorithmic_ops_per_launch": ops, "avg_launch_ms": ms_l1,
                   "achieved": a, "peak": pk.get("vabsdiff4_acc"), "unit": "Gop/s (thread-level)", "frac": a / pk["vabsdiff4_acc"] if pk.get("vabsdiff4_acc") else None,
                   "measured_peaks_gops": pk, "note": "launches of 8 streams overlap on the GPU; the kernel alone (ncu, profiles/) runs faster"}
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        try:
            trace = load_trace()
            n = min(240, len(trace))  # ~10 s of host work
            t0 = time.perf_counter()
            fps_cpu, _, same = replay_oracle(store, trace, n, 12, 1, os.cpu_count())
            cpu = {"value": fps_cpu, "unit": "frames/s", "cores": os.cpu_count(), "kind": "port",
                   "sample": f"frames 0..{n - 1} of stream 0 of the workload (the first 12 untimed) coded by the oracle (OpenMP, all host threads) with the frame types and "
                             f"QPs of this arm's rate control (tests/golden/qp_trace_1080p60.csv): equal bitrate, {time.perf_counter() - t0:.1f} s",
                   "same_output_as_cuda_arm": bool(same)}
        except Exception as ex:  # noqa: BLE001
            cpu = {"value": None, "unit": "frames/s", "cores": os.cpu_count(), "kind": "port", "sample": f"not run: {ex}"}
    nf = max(r_e2e["frames"], 1)
    line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32",
            "data": f"synthetic S2 pan-sprites 1080p (SURVEY.md 8d), {n_pool}-frame store" + (", looped" if n_pool >= POOL else ", ping-pong"),
            "config": workload_config(args, world),
            "e2e": {"value": e2e, "unit": "frames/s", "h2d_bytes_per_step": fps_step * S * fsz, "d2h_bytes_per_step": int(r_e2e["sym_bytes"] / args.steps),
                    "au_bytes_per_step": int(r_e2e["au_bytes"] / args.steps), "ms_per_step": ms_e2e / args.steps, "host_frames": "page-locked (cc_host_register)"},
            "e2e_pageable": {"value": e2e_pg, "unit": "frames/s", "frames": int(f_pg), "staged_frames": int(r_page["staged"]),
                             "note": "pageable host frames, as gst_video_frame_map hands them to the element: each frame is copied by the calling thread into the handle's "
                                     "pinned staging frame, then uploaded asynchronously; inside cc_enc_encode_sequence that copy (and the upload) of frame i + 1 runs while the device codes frame i"},
            "timed_region": {"frames_per_stream": fps_step * args.steps, "frames": int(f_dev), "idr_frames_rank0": int(r_dev["n_idr"]), "frames_rank0": int(r_dev["frames"]),
                             "seconds": ms_dev / 1e3, "scene_cuts_per_stream": (fps_step * args.steps) // 300 if n_pool >= POOL else 0},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "int_roofline": introof, "cpu_baseline": cpu, "quality": quality,
            "achieved_kbps_per_stream": r_dev["au_bytes"] * 8 / (r_dev["frames"] / FPS) / 1e3, "last_qp": r_dev["qps"],
            "mb_per_frame": {k: round(r_dev["n_" + k] / max(r_dev["frames"], 1), 1) for k in ("skip", "inter", "intra", "intra_analysed", "skip_probed")},
            "host_ms_per_frame": {"submit_api": round(r_e2e["submit_ms"] / nf, 4), "wait_device": round(r_e2e["wait_ms"] / nf, 4),
                                  "cavlc_nal": round(r_e2e["host_ms"] / nf, 4), "device_frame_latency": round(r_e2e["gpu_ms"] / nf, 4),
                                  "submit_api_pageable": round(r_page["submit_ms"] / max(r_page["frames"], 1), 4)},
            "frame_store": {"path": store.path, "generate_s": round(store.gen_s, 1)},
            "launch_order": {"filter_next_to_mb_encode_frames": {"value_leg": r_dev["overlapped"], "e2e_leg": r_e2e["overlapped"], "stage_timing_pass": r_prof["overlapped"]},
                             "note": "frames since open (warm-up included) whose k_deblock was launched behind k_mb_encode by programmatic stream serialisation; DESIGN.md 5"},
            "wall_ms": {"value_leg": r_dev["wall_ms"], "e2e_leg": r_e2e["wall_ms"], "e2e_pageable_leg": r_page["wall_ms"]}}
    print(json.dumps(line))
    L.cc_host_unregister(hbase)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
