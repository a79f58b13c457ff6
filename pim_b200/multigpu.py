"""Multi-GPU driver for the image path: one process per GPU, scene replicated, samples sharded.

The path tracer shards by sample index with no data-path exchange: rank r of W renders the
sample indices `sample_range(step, r, W, spp)` of every pixel into its own sum images, and the
only collective is one reduce(SUM) of those images per step. Lightmap bakes
shard by texel range instead (`texel_range`), because the spherical-gaussian fit is sequential
per texel.
"""
import torch
import torch.distributed as dist


def sample_range(step, rank, world, spp_per_step):
    """First sample index and count that `rank` renders in `step` (disjoint, gap free)."""
    return (step * world + rank) * spp_per_step, spp_per_step


def texel_range(texel_count, rank, world):
    """Contiguous texel slab of `rank` (all texels covered once; sizes differ by at most 1)."""
    base, rem = divmod(texel_count, world)
    begin = rank * base + min(rank, rem)
    return begin, base + (1 if rank < rem else 0)


def reduce_sums(tensors, dst=0):
    """One reduce(SUM) per image to `dst` (NCCL over NVLink on GPUs, gloo in the CPU tests)."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        for t in tensors:
            dist.reduce(t, dst=dst, op=dist.ReduceOp.SUM)


def resolve(sum_image, total_spp):
    """Running mean from the reduced sums."""
    return sum_image / float(total_spp)


def max_over_ranks(value, device):
    if dist.is_initialized() and dist.get_world_size() > 1:
        t = torch.tensor([value], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return float(value)


class LightCounterSync:
    """Keeps the adaptive light grid of every rank's replica learning from ALL ranks' paths: note the
    counters before a call, and after it all-reduce what the call added and write base + sum back. Integer adds, so the
    replicas hold exactly the counters one device tracing every rank's samples would hold."""

    def __init__(self, scene, device, stream=None):
        self.scene, self.stream = scene, stream
        n = scene.light_counters()
        self.base = torch.zeros(max(n, 1), dtype=torch.int32, device=device)
        self.live = torch.zeros_like(self.base)
        self.n = n

    def _s(self):
        return self.stream.cuda_stream if self.stream is not None else None

    def before(self):
        if self.n:
            self.scene.light_counters(self.base.data_ptr(), False, self._s())

    def after(self):
        if self.n and dist.is_initialized() and dist.get_world_size() > 1:
            self.scene.light_counters(self.live.data_ptr(), False, self._s())
            self.live -= self.base
            dist.all_reduce(self.live, op=dist.ReduceOp.SUM)
            self.live += self.base
            self.scene.light_counters(self.live.data_ptr(), True, self._s())
