"""Multi-GPU sharding of the path: one process per GPU (torchrun), the frame is split by
SAMPLE RANGE — rank r renders samples [begin_r, begin_r + n_r) of every pixel, RNG keyed by the
global sample index — and the float accumulation buffers (Film layout, W*H*3 fp32, film.cu:79-88)
are sum-reduced to rank 0 with ONE collective (NCCL over NVLink on GPUs, gloo in the CPU tests).
The reference has no multi-GPU code; scene and LBVH are replicated (each rank builds identical bits).
"""
import torch
import torch.distributed as dist


def sample_range(rank, world, spp_total, sample_begin=0):
    """Contiguous, balanced split of [sample_begin, sample_begin+spp_total) over `world` ranks.
    The first (spp_total % world) ranks get one extra sample. Returns (begin, count)."""
    base, extra = divmod(int(spp_total), int(world))
    count = base + (1 if rank < extra else 0)
    begin = sample_begin + rank * base + min(rank, extra)
    return begin, count


def pixel_range(rank, world, n_pixels, width=1):
    """TILE partition (SURVEY 8e alternative): contiguous, balanced split of the row-major pixel range over `world`
    ranks, cut at multiples of `width` (= whole rows -> horizontal bands) when width > 1. Returns (begin, count)."""
    units = int(n_pixels) // int(width)
    base, extra = divmod(units, int(world))
    count = base + (1 if rank < extra else 0)
    begin = rank * base + min(rank, extra)
    return begin * int(width), count * int(width)


def reduce_accum(accum, dst=0, all_ranks=False):
    """Sum the per-rank accumulation buffers in place; the result lands on `dst` (or everywhere)."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return accum
    if all_ranks:
        dist.all_reduce(accum, op=dist.ReduceOp.SUM)
    else:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
    return accum


def render_sharded(render_fn, accum, spp_total, sample_begin=0, dst=0, all_ranks=False):
    """render_fn(begin, count) must ADD samples [begin, begin+count) into `accum` (a torch tensor
    of W*H*3 floats on this rank's device). Returns (begin, count) rendered by this rank."""
    rank = dist.get_rank() if dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_initialized() else 1
    begin, count = sample_range(rank, world, spp_total, sample_begin)
    if count > 0:
        render_fn(begin, count)
    reduce_accum(accum, dst=dst, all_ranks=all_ranks)
    return begin, count


def render_tiled(render_fn, accum, n_pixels, width=1, dst=0, all_ranks=False):
    """render_fn(pixel_begin, pixel_count) must ADD every sample of pixels [pixel_begin, +pixel_count) into `accum`
    (zero elsewhere on this rank); the same sum-reduce then assembles the frame (disjoint supports: exact, no
    dependence on summation order). Returns (pixel_begin, pixel_count) rendered by this rank."""
    rank = dist.get_rank() if dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_initialized() else 1
    begin, count = pixel_range(rank, world, n_pixels, width)
    if count > 0:
        render_fn(begin, count)
    reduce_accum(accum, dst=dst, all_ranks=all_ranks)
    return begin, count


class ShardedRenderer:
    """B200Renderer on every rank + the reduce. ProgressiveRender(spp_total) renders spp_total new
    samples split over the ranks into a per-call buffer, sum-reduces it to rank 0 and adds it to
    rank 0's running total (the Film accumulator: sum of samples + weight)."""

    def __init__(self, renderer, film, device, partition="samples"):
        assert partition in ("samples", "tiles")
        self.r, self.film, self.partition = renderer, film, partition
        n = film.width * film.height * 3
        self.batch = torch.zeros(n, dtype=torch.float32, device=device)    # this call's samples (this rank)
        self.total = torch.zeros(n, dtype=torch.float32, device=device)    # rank 0: sum since Clear()
        renderer.ctx.set_accum_buffer(self.batch.data_ptr())
        self.next_sample = 0
        self.weight = 0

    def Clear(self):
        self.total.zero_()
        self.next_sample = self.weight = 0

    def ProgressiveRender(self, spp_total):
        self.batch.zero_()
        torch.cuda.synchronize()

        if self.partition == "samples":
            def fn(begin, count):
                self.r.ctx.clear_accum()
                self.r.ctx.render(count, self.r.m_Depth, self.r.m_Seed, begin)      # synchronous
            out = render_sharded(fn, self.batch, spp_total, self.next_sample)
        else:                                                                       # horizontal bands of whole rows
            def fn(pixel_begin, pixel_count):
                self.r.ctx.clear_accum()
                self.r.ctx.render_region(spp_total, self.r.m_Depth, self.r.m_Seed, self.next_sample, pixel_begin, pixel_count)
            out = render_tiled(fn, self.batch, self.film.width * self.film.height, self.film.width)
        self.total += self.batch
        self.next_sample += spp_total
        self.weight += spp_total
        return out
