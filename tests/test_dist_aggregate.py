"""Multi-GPU = independent replicas: the only cross-rank step is the reduction of the timing numbers
(sum of frames over ranks / max of device time over ranks).  Covered here with world_size-2 gloo on CPU."""
import os
import socket
import sys
import torch.multiprocessing as mp
from oracle_lib import ROOT


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import bench
    assert bench.dist_info() == (rank, world, rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    frames, ms = bench.aggregate(720 + rank, 100.0 + 50.0 * rank, world, device="cpu")
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, frames, ms))


def test_aggregate_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    res = sorted(q.get(timeout=120) for _ in ps)
    [p.join(60) for p in ps]
    assert all(p.exitcode == 0 for p in ps)
    for _, frames, ms in res:
        assert frames == 720 + 721      # whole-job frames
        assert ms == 150.0              # max over ranks


def test_reference_arm_runs_on_rank0_only():
    """`bench.py --impl reference` under torchrun: ranks != 0 exit 0 without work (driver contract)."""
    import subprocess
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, env=env, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""
