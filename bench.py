#!/usr/bin/env python
"""bench.py — 1080p H.264 encode throughput of libceracuda on B200 (BASELINE.json metric), one JSON line.

Workload (config.workload): BASELINE configs[1] as the target defines it — 1080p60 `cudah264enc` streams under a
2<->12 Mbps balancer-style bitrate schedule, 8 concurrent streams per GPU (north_star: "8 concurrent 1080p60 streams
per GPU"), synthetic S2 pan-sprites content.  A *step* = every stream of this rank encodes
FRAMES_PER_STEP frames.  `value` = frames/s with the frames already resident in HBM (CC_FRAME_DEVICE), `e2e` = the
same through the public C-ABI call with pinned HOST frames (H2D inside) and the access unit in host memory.
Multi-GPU = independent replicas (one process per GPU, no collective on the data path): scaling "weak".

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl cuda|reference]
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

# more than 8 streams per process need more hardware queues than the default 8, or they serialise pairwise
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, FPS = 1920, 1080, 60
STREAMS = 8
FRAMES_PER_STEP = 30
POOL = 40                       # distinct synthetic frames, played ping-pong; 40 * 3.1 MB = 124 MB host, same in HBM
ALG_BYTES_PER_FRAME = 4.5 * W * H   # read cur + read ref + write recon (+ bitstream)
METRIC = "1080p H.264 encode fps per B200 at equal bitrate"


def schedule_kbps(frame_idx):
    """Balancer-shaped target for config 2: 12 Mbps, emergency drop to 2 Mbps, then +100..300 kbps every 520 ms
    (the reference's adaptive ladder, /root/reference/src/core/bitrate_control.c:176-206), period 30 s."""
    t = (frame_idx / FPS) % 30.0
    if t < 5.0:
        return 12000
    steps = int((t - 5.0) / 0.52)
    kb = 2000
    for _ in range(steps):
        kb = min(12000, kb + 30 + kb // 30)
        kb = kb // 100 * 100 if kb % 100 == 0 else (kb // 100 + 1) * 100
    return kb


def pingpong(i, n):
    p = i % (2 * n - 2)
    return p if p < n else 2 * n - 2 - p


def make_pool():
    from ceracoder_b200.synth import make_source
    src = make_source("pan", W, H)
    return [src.frame(i) for i in range(POOL)]


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


def prepare_streams(encs, frames_of, n_frames, start_idx, phase=0, warm=False):
    """Per stream the C arrays of frame descriptors (pointers into the resident pool, strides, pts) and bitrates.  Interpreter
    work under the GIL (8 threads would serialise on it), so it is done before the timed region; the frames themselves are
    copied / encoded inside it."""
    out = []
    for k, e in enumerate(encs):
        n = n_frames + (phase * k if warm else 0)
        first = start_idx + (0 if warm else phase * k)
        out.append(e.prepare_sequence([frames_of(k, first + i) for i in range(n)], [schedule_kbps(first + i + 37 * k) for i in range(n)]))
    return out


def run_streams(encs, prepared, device_frames):
    """Each encoder on its own host thread / CUDA stream; returns (frames, au_bytes, sym_bytes, ...).
    phase: stream k runs phase*k frames ahead of stream 0 (its warm-up is that much longer), so that independent streams do
    not sit on the same GOP phase (eight live streams do not emit their IDR frames in the same instant)."""
    tot = [0, 0, 0, 0, 0, 0, 0, 0, 0.0, 0.0, 0.0, 0.0]
    lock = threading.Lock()
    errs = []

    def work(k, e):
        try:  # one C call per stream per run_streams(): the frame loop itself runs in C (cc_enc_encode_sequence), off the interpreter lock
            s = e.encode_sequence(prepared[k], device_frame=device_frames)
            with lock:
                tot[0] += s.frames; tot[1] += s.au_bytes; tot[2] += s.sym_bytes
                for j, v in enumerate((s.n_skip, s.n_inter, s.n_intra, s.n_intra_analysed, s.n_skip_probed, s.submit_ms, s.wait_ms, s.host_ms, s.gpu_ms)):
                    tot[3 + j] += v
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


def cpu_baseline(pool, n_frames, rps=4):
    """The oracle (CPU port of the same algorithm, OpenMP over slices/MBs) timed on this box's host cores."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_lib import OracleEncoder, lib as oracle_lib_handle
    oracle_lib_handle().ref_set_threads(os.cpu_count())  # torchrun exports OMP_NUM_THREADS=1; the baseline uses every host thread
    enc = OracleEncoder(W, H, FPS, rps)
    y, u, v = pool[0]
    enc.encode(y, u, v, True, 30)  # IDR outside the timed sample, like a warm stream
    t0 = time.perf_counter()
    for i in range(1, n_frames + 1):
        y, u, v = pool[pingpong(i, len(pool))]
        enc.encode(y, u, v, False, 30)
    dt = time.perf_counter() - t0
    enc.close()
    return n_frames / dt


def main_reference(args, rank, world):
    if rank != 0:
        return 0
    from ceracoder_b200.synth import make_source
    src = make_source("pan", W, H)
    pool = [src.frame(i) for i in range(8)]
    sample = FRAMES_PER_STEP  # P frames per step: one stream's share of a step of the CUDA arm
    for _ in range(args.warmup):
        cpu_baseline(pool, 1)
    t0 = time.perf_counter()
    fps_list = [cpu_baseline(pool, sample) for _ in range(args.steps)]
    dt = time.perf_counter() - t0
    fps = float(np.mean(fps_list))
    cores = os.cpu_count()
    line = {"impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * dt / max(args.steps, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32",
            "data": "synthetic S2 pan-sprites 1080p (tests inputs: ceracoder_b200/synth.py)",
            "config": {"workload": "1080p60 H.264 encode, CPU reference encoder (oracle port; x264enc/libx264 is not installed on this image)",
                       "width": W, "height": H, "fps": FPS, "qp": 30},
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port",
                             "sample": f"{sample} P frames of one 1080p stream per step at QP 30 (IDR untimed), OpenMP on all {cores} host threads"},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


def main():
    global FRAMES_PER_STEP
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--streams", type=int, default=STREAMS)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--rows-per-slice", type=int, default=4, help="macroblock rows per slice (experiments; the reported workload uses 4)")
    ap.add_argument("--partitions", type=int, default=1, help="0: 16x16 partitions only (experiments)")
    ap.add_argument("--input-format", default="I420", choices=["I420", "YUY2", "BGRx"],
                    help="frame format handed to the encoder; packed formats are converted inside the ingest kernel (SURVEY 8f N2). "
                         "The YUY2 pool carries exactly the I420 pool's pictures (chroma rows doubled), so the encoded workload is the same")
    ap.add_argument("--host-entropy", action="store_true", help="CAVLC by the host C writer instead of the device kernels (round-1 path; comparison runs)")
    ap.add_argument("--frames-per-step", type=int, default=FRAMES_PER_STEP, help="frames each stream encodes per step (shorten for ncu runs)")
    args = ap.parse_args()
    FRAMES_PER_STEP = args.frames_per_step
    rank, world, local = dist_info()
    if args.impl == "reference":
        return main_reference(args, rank, world)

    import torch
    import torch.distributed as dist
    assert torch.cuda.is_available(), "bench.py --impl cuda needs a CUDA device (there is no CPU fallback)"
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from ceracoder_b200 import CudaH264Enc, load_library
    L = load_library()
    S = args.streams
    pool = make_pool()
    FMT = args.input_format
    fsz = W * H * {"I420": 3, "YUY2": 4, "BGRx": 8}[FMT] // 2
    # pinned host pool (e2e leg) and HBM-resident pool (value leg); both > 126 MB L2 together with the encoders' planes
    hbase = L.cc_host_alloc(fsz * POOL); dbase = L.cc_dev_alloc(fsz * POOL)
    assert hbase and dbase
    harr = np.ctypeslib.as_array(C.cast(hbase, C.POINTER(C.c_uint8)), shape=(POOL, fsz))
    for i, (y, u, v) in enumerate(pool):
        if FMT == "I420":
            harr[i, :W * H] = y.ravel(); harr[i, W * H:W * H * 5 // 4] = u.ravel(); harr[i, W * H * 5 // 4:] = v.ravel()
        elif FMT == "YUY2":
            f = harr[i].reshape(H, 2 * W)
            f[:, 0::2] = y; f[:, 1::4] = np.repeat(u, 2, axis=0); f[:, 3::4] = np.repeat(v, 2, axis=0)
        else:  # BGRx from the same pictures (BT.709 limited range, float; the encoder converts back with its integer matrix)
            yy = (y.astype(np.float32) - 16) * (255 / 219)
            uu = np.repeat(np.repeat(u, 2, 0), 2, 1).astype(np.float32) - 128; vv = np.repeat(np.repeat(v, 2, 0), 2, 1).astype(np.float32) - 128
            f = harr[i].reshape(H, W, 4)
            f[..., 2] = np.clip(yy + 1.7927 * vv, 0, 255); f[..., 1] = np.clip(yy - 0.2132 * uu - 0.5329 * vv, 0, 255)
            f[..., 0] = np.clip(yy + 2.1124 * uu, 0, 255); f[..., 3] = 255
    assert L.cc_dev_upload(dbase, hbase, fsz * POOL) == 0

    def planes(base, k, fi):
        p = base + pingpong(fi + 5 * k, POOL) * fsz
        if FMT != "I420":
            return (p, W * (2 if FMT == "YUY2" else 4)), None, None
        return (p, W), (p + W * H, W // 2), (p + W * H * 5 // 4, W // 2)

    def make_encs():
        return [CudaH264Enc(W, H, fps=(FPS, 1), bitrate=12000, key_int_max=60, speed_preset="veryfast", device=local, lib=L, fmt=FMT,
                            rows_per_slice=args.rows_per_slice, partitions=bool(args.partitions), host_entropy=args.host_entropy) for _ in range(S)]

    def timed(device_frames, steps, warmup, profile=False):
        encs = make_encs()
        base = dbase if device_frames else hbase
        fo = lambda k, fi: planes(base, k, fi)  # noqa: E731
        run_streams(encs, prepare_streams(encs, fo, FRAMES_PER_STEP * max(warmup, 1), 0, phase=7, warm=True), device_frames)  # warm-up steps (IDR + RC settle), GOP phases spread
        prepared = prepare_streams(encs, fo, FRAMES_PER_STEP * steps, FRAMES_PER_STEP * max(warmup, 1), phase=7)
        if profile:
            [e.set_profile(True) for e in encs]
        launches0 = sum(e.kernel_launches() for e in encs)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); ev0.record()
        frames, au_bytes, sym_bytes, *mbstats = run_streams(encs, prepared, device_frames)
        torch.cuda.synchronize(); ev1.record(); torch.cuda.synchronize()
        wall_ms = 1e3 * (time.perf_counter() - t0)
        ms = max(ev0.elapsed_time(ev1), 1e-3)
        if world > 1:
            dist.barrier()
        launches = sum(e.kernel_launches() for e in encs) - launches0
        stages = {}
        if profile:
            for e in encs:
                for s, (tms, n) in e.stage_ms().items():
                    a = stages.setdefault(s, [0.0, 0]); a[0] += tms; a[1] += n
        qps = [e.last.qp for e in encs]
        [e.close() for e in encs]
        return dict(frames=frames, ms=ms, wall_ms=wall_ms, au_bytes=au_bytes, sym_bytes=sym_bytes, launches=launches, stages=stages, qps=qps, mbstats=mbstats)

    sampler = ClockSampler(local)
    sampler.start()
    r_dev = timed(True, args.steps, args.warmup, profile=True)
    clocks = sampler.stop()
    r_e2e = timed(False, args.steps, args.warmup)

    def quality_leg(n_frames=360):
        """One 1080p60 stream, frame by frame through the public call, under a 12 -> 2 -> 12 Mbps square wave (2 s per level):
        Y-PSNR of the reconstruction against the source (the decoder reproduces the reconstruction bit-exactly — parity
        tests) and the bitrate error over 1 s windows.  Untimed."""
        enc = CudaH264Enc(W, H, fps=(FPS, 1), bitrate=12000, key_int_max=60, device=local, lib=L, partitions=bool(args.partitions))
        bits, targ, psnr = [], [], []
        for i in range(n_frames):
            kb = 12000 if (i // (2 * FPS)) % 2 == 0 else 2000
            enc.set_property("bitrate", kb)
            y, u, v = pool[pingpong(i, POOL)]
            au = enc.encode(y, u, v)
            bits.append(8 * len(au)); targ.append(kb * 1000.0 / FPS)
            if i % 12 == 11:
                ry = enc.recon()[0][:H, :W]
                mse = float(((ry.astype(np.float32) - y.astype(np.float32)) ** 2).mean())
                psnr.append((kb, 10 * np.log10(255.0 ** 2 / max(mse, 1e-9))))
        enc.close()
        b, t = np.array(bits, float), np.array(targ, float)
        err = [abs(b[i:i + FPS].sum() - t[i:i + FPS].sum()) / t[i:i + FPS].sum() for i in range(0, n_frames - FPS + 1, FPS // 4)]
        return {"frames": n_frames, "schedule": "square wave 12 <-> 2 Mbps, 2 s per level, 1080p60, one stream",
                "achieved_kbps": float(b.sum() / (n_frames / FPS) / 1e3), "target_kbps": float(t.sum() / (n_frames / FPS) / 1e3),
                "bitrate_error_1s_windows": {"mean": float(np.mean(err)), "p95": float(np.percentile(err, 95))},
                "y_psnr_db": {"at_12mbps": float(np.mean([p for k, p in psnr if k == 12000])), "at_2mbps": float(np.mean([p for k, p in psnr if k == 2000]))},
                "vs_x264enc": None, "note": "x264enc / libx264 are not installed on this box: the PSNR delta against x264 veryfast was not measured"}

    quality = quality_leg() if rank == 0 and not args.no_cpu_baseline else None

    dev = torch.device("cuda", local)
    f_dev, ms_dev = aggregate(r_dev["frames"], r_dev["ms"], world, device=dev)
    f_e2e, ms_e2e = aggregate(r_e2e["frames"], r_e2e["ms"], world, device=dev)
    launches, _ = aggregate(r_dev["launches"], 0.0, world, device=dev)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    value = f_dev / (ms_dev / 1e3)
    e2e = f_e2e / (ms_e2e / 1e3)
    peak, peak_src = peaks()
    st = r_dev["stages"]
    dom = max((s for s in st if s != "upload"), key=lambda s: st[s][0]) if st else None
    roof = None
    if dom:
        avg_ms = st[dom][0] / max(st[dom][1], 1)
        ach = ALG_BYTES_PER_FRAME / (avg_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": {"mb_encode": "k_mb_encode", "me_coarse": "k_me_l1", "me_fine": "k_me_l0", "hpel": "k_hpel"}.get(dom, dom),
                "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": ncu_traffic({"mb_encode": "k_mb_encode", "me_coarse": "k_me_l1", "me_fine": "k_me_l0", "p_code": "k_p_code"}.get(dom, dom)),
                "avg_launch_ms": avg_ms, "launches_timed": st[dom][1], "peak_source": peak_src,
                "algorithmic_bytes_per_launch": ALG_BYTES_PER_FRAME,
                "note": "one launch = one 1080p frame of one stream; 8 streams' launches overlap on the GPU, so the per-launch duration includes sharing; "
                        "aggregate_achieved = value * 4.5*W*H",
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
    if not args.no_cpu_baseline:
        t0 = time.perf_counter()
        n = 160  # ~10 s of host work
        fps_cpu = cpu_baseline(pool, n)
        cpu = {"value": fps_cpu, "unit": "frames/s", "cores": os.cpu_count(), "kind": "port",
               "sample": f"{n} P frames of one 1080p stream at QP 30 with the oracle (OpenMP, all host threads), {time.perf_counter() - t0:.1f} s"}
    steps_frames = FRAMES_PER_STEP * S
    line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32",
            "data": "synthetic S2 pan-sprites 1080p (SURVEY.md §8d), 40-frame pool played ping-pong",
            "config": {"workload": f"{S} concurrent 1080p60 cudah264enc streams per GPU, balancer-shaped 12->2->12 Mbps schedule (configs[1] x8, north_star target shape)",
                       "width": W, "height": H, "fps": FPS, "input_format": FMT, "streams_per_gpu": S, "frames_per_step": steps_frames, "rows_per_slice": args.rows_per_slice, "partitions": "16x16+8x8" if args.partitions else "16x16", "key_int_max": 60, "entropy_coder": "host C (cc_cavlc.c)" if args.host_entropy else "device (k_cavlc_size + k_cavlc_write)",
                       "l2": "inputs (124 MB pool) + 8 streams' planes exceed the 126 MB L2; no explicit flush", "parallelism": f"replicas x{world}"},
            "e2e": {"value": e2e, "unit": "frames/s", "h2d_bytes_per_step": steps_frames * fsz, "d2h_bytes_per_step": int(r_e2e["sym_bytes"] / args.steps),
                    "au_bytes_per_step": int(r_e2e["au_bytes"] / args.steps), "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "int_roofline": introof, "cpu_baseline": cpu, "quality": quality,
            "achieved_kbps_per_stream": r_dev["au_bytes"] * 8 / (r_dev["frames"] / FPS) / 1e3, "last_qp": r_dev["qps"],
            "mb_per_frame": dict(zip(("skip", "inter", "intra", "intra_analysed", "skip_probed"), [round(v / max(r_dev["frames"], 1), 1) for v in r_dev["mbstats"][:5]])),
            "host_ms_per_frame": dict(zip(("submit_api", "wait_device", "cavlc_nal", "device_frame_latency"), [round(v / max(r_e2e["frames"], 1), 4) for v in r_e2e["mbstats"][5:9]])),
            "wall_ms": {"value_leg": r_dev["wall_ms"], "e2e_leg": r_e2e["wall_ms"]}}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
