"""Multi-GPU host logic on CPU (gloo, world_size 2): ranks own contiguous world ranges, seeds use
the GLOBAL world index, no data-path collective -- the union of the ranks' results equals a
single-process run. The per-world work is done by the oracle port here (no GPU on this machine);
the same first_world plumbing drives EcsExecutor / PhysExecutor on the GPU."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world_size, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    from gpu_ecs_madrona_b200 import ecs_apps as A
    from tests import ecs_util as U
    total = 8
    per = total // world_size
    o = U.make_oracle("port", A.APP_FVS, per, [6, 20], first_world=rank * per)
    dist.barrier()
    o.step(50)
    rows = torch.from_numpy(o.row_counts().astype(np.int64))
    # the only cross-rank traffic: episode statistics (here total units alive), outside the step
    stats = rows.sum(dim=0)
    dist.all_reduce(stats)
    hp = np.concatenate([o.read_column(w, 1, 2).view(np.int32).ravel() for w in range(per)])
    np.save(os.path.join(out_dir, f"hp_{rank}.npy"), hp)
    if rank == 0:
        np.save(os.path.join(out_dir, "stats.npy"), stats.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_equal_one_process(tmp_path):
    import subprocess
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "libmadrona_ecs_port.so"], check=True,
                   capture_output=True)
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    from gpu_ecs_madrona_b200 import ecs_apps as A
    from tests import ecs_util as U
    o = U.make_oracle("port", A.APP_FVS, 8, [6, 20])
    o.step(50)
    want = np.concatenate([o.read_column(w, 1, 2).view(np.int32).ravel() for w in range(8)])
    got = np.concatenate([np.load(tmp_path / f"hp_{r}.npy") for r in range(2)])
    assert np.array_equal(want, got)
    assert np.array_equal(np.load(tmp_path / "stats.npy"), o.row_counts().sum(axis=0))
