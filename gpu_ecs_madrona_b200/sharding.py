"""Multi-GPU plumbing of the world step: worlds shard, nothing else does.

One process per GPU (`torch.distributed`), each stepping its own contiguous range of worlds with
NO per-step communication: no node of the path reads another world (only the replicated < 1 KB asset
tables). World-indexed seeds use the GLOBAL world index (`scenes.collisions64(world_offset=...)`,
`EcsExecutor(first_world=...)`), so results do not depend on the number of GPUs.

The one collective the path has is optional and off the step: every K steps a small per-GPU vector of
episode statistics is all-gathered (NCCL over NVLink on the GPU box, gloo in the CPU tests) on a side
stream, so the step stream never waits for it.
"""
import numpy as np
import torch
import torch.distributed as dist


def world_range(rank, world_size, total_worlds):
    """Contiguous partition: GPU g of G owns worlds [g*W/G, (g+1)*W/G)."""
    if not 0 <= rank < world_size:
        raise ValueError("rank outside the group")
    return rank * total_worlds // world_size, (rank + 1) * total_worlds // world_size


STAT_NAMES = ("worlds", "world_steps", "pairs", "contacts", "max_contacts", "bodies_below_ground")


def local_statistics(executor, steps_done, position=None):
    """Per-GPU statistics vector of a PhysExecutor (float64[len(STAT_NAMES)]): counts from the step
    kernel's per-world counters (`mdr_phys_stats`), optionally a scan of the Position column."""
    pairs, contacts = executor.stats()
    below = 0
    if position is not None:
        below = int((position[..., 2] < -0.5).sum())
    return np.array([executor.W, executor.W * steps_done, pairs.sum(), contacts.sum(),
                     contacts.max() if len(contacts) else 0, below], np.float64)


class EpisodeStats:
    """All-gather of the per-GPU statistics vector, every `every` steps, outside the step."""

    def __init__(self, every=100, device=None, group=None):
        self.every, self.group = every, group
        self.device = torch.device(device) if device is not None else torch.device("cpu")
        self.side = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        self.pending = None

    def due(self, step):
        return self.every > 0 and step % self.every == 0

    def launch(self, local_vector):
        """Starts the all-gather (side stream on CUDA) and returns at once."""
        n = dist.get_world_size(self.group) if dist.is_initialized() else 1
        src = torch.as_tensor(np.asarray(local_vector, np.float64)).to(self.device, non_blocking=True)
        # flat output (concatenation along dim 0): the form both NCCL and gloo accept
        out = torch.empty(n * src.numel(), dtype=torch.float64, device=self.device)
        if n == 1:
            out.copy_(src)
            self.pending = (out, None, src)
            return
        if self.side is not None:
            self.side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self.side):
                work = dist.all_gather_into_tensor(out, src, group=self.group, async_op=True)
        else:
            work = dist.all_gather_into_tensor(out, src, group=self.group, async_op=True)
        self.pending = (out, work, src)

    def result(self):
        """[world_size, len(vector)] float64 on the host; waits for the gather started by launch()."""
        out, work, src = self.pending
        if work is not None:
            work.wait()
        if self.side is not None:
            self.side.synchronize()
        self.pending = None
        return out.view(-1, src.numel()).cpu().numpy()

    @staticmethod
    def totals(table):
        """Whole-job figures from the gathered table: sums, except the maximum column."""
        t = table.sum(axis=0)
        t[STAT_NAMES.index("max_contacts")] = table[:, STAT_NAMES.index("max_contacts")].max()
        return dict(zip(STAT_NAMES, t))
