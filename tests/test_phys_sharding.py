"""Multi-GPU host logic of the physics step on CPU (gloo, world_size 2): contiguous world ranges, global
world index as the scene seed, no data-path collective; the union of the ranks' states equals the
single-process run bit for bit, and the optional episode-statistics all-gather (sharding.EpisodeStats)
adds up. The per-world step is done by the device source compiled for the host (tests/emu, a test tool);
on the GPU box the same plumbing drives the CUDA executor (bench.py --gpus N)."""
import ctypes
import os
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOTAL, STEPS = 6, 12


def _worker(rank, world_size, port, emu_path, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    from gpu_ecs_madrona_b200 import phys, scenes, sharding
    lo, hi = sharding.world_range(rank, world_size, TOTAL)
    sc = scenes.collisions64(hi - lo, world_offset=lo, num_cubes=16)
    ex = phys.PhysExecutor(sc, lib=phys.bind(ctypes.CDLL(emu_path)))
    stats = sharding.EpisodeStats(every=6)
    tables = []
    dist.barrier()
    for step in range(1, STEPS + 1):
        ex.run()
        if stats.due(step):
            stats.launch(sharding.local_statistics(ex, step, ex.state()[0]))
            tables.append(stats.result())
    pos, rot, vel = ex.state()
    np.savez(os.path.join(out_dir, f"state_{rank}.npz"), pos=pos, rot=rot, vel=vel, tables=np.stack(tables))
    dist.barrier()
    dist.destroy_process_group()


def test_world_range_partitions():
    from gpu_ecs_madrona_b200 import sharding
    for total in (0, 1, 7, 16384, 1 << 20):
        for g in (1, 2, 3, 8):
            r = [sharding.world_range(k, g, total) for k in range(g)]
            assert r[0][0] == 0 and r[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1


def test_two_ranks_equal_one_process(tmp_path, emu_lib):
    from gpu_ecs_madrona_b200 import phys, scenes, sharding
    emu_path = emu_lib._name
    port = 31500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, emu_path, str(tmp_path)), nprocs=2, join=True)
    ex = phys.PhysExecutor(scenes.collisions64(TOTAL, num_cubes=16), lib=emu_lib)
    tables = []
    for step in range(1, STEPS + 1):
        ex.run()
        if step % 6 == 0:
            tables.append(sharding.local_statistics(ex, step, ex.state()[0]))
    got = [np.load(tmp_path / f"state_{r}.npz") for r in range(2)]
    for k, want in zip(("pos", "rot", "vel"), ex.state()):
        both = np.concatenate([g[k] for g in got])
        assert np.array_equal(both.view(np.uint32), want.view(np.uint32)), k
    # every rank holds the same gathered table; its totals equal the single-process statistics
    assert np.array_equal(got[0]["tables"], got[1]["tables"])
    for table, want in zip(got[0]["tables"], tables):
        assert table.shape == (2, len(sharding.STAT_NAMES))
        tot = sharding.EpisodeStats.totals(table)
        for name, w in zip(sharding.STAT_NAMES, want):
            assert tot[name] == w, name
