"""Host-side logic of the two multi-GPU partitions of the path (the reference is
single-GPU, docs/technical-details.md:327).  One process per GPU, `torch.distributed` for the plumbing.

  spp split      every rank holds the whole scene, rebuilds the same LBVH locally and renders a
                 contiguous block of the frame's samples for every pixel; ONE exchange per frame:
                 a sum-reduce of the float4 accumulation buffer to rank 0 (NCCL over NVLink on the
                 GPU box, gloo in the CPU tests).  Sample s of pixel p uses the same XORWOW stream
                 (seed, p, s*64) whatever the partition, so the image does not depend on the world
                 size except for the fp32 re-association of the per-rank partial sums.
  frame parallel frame f of a time series goes to rank f mod world; no collective on the data path.

Nothing here computes pixels: rendering goes through the C ABI (Scene.render)."""
from __future__ import annotations

import os
from typing import List, Tuple


def env_world() -> Tuple[int, int, int]:
    """(rank, local_rank, world_size) as torchrun exports them; (0, 0, 1) for a plain launch."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def spp_partition(spp: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous block [offset, offset+count) of the frame's `spp` samples rendered by `rank`.
    The first spp % world ranks take one extra sample; count may be 0 when spp < world."""
    if spp < 0 or world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad partition request spp={spp} world={world} rank={rank}")
    base, extra = divmod(spp, world)
    count = base + (1 if rank < extra else 0)
    offset = rank * base + min(rank, extra)
    return offset, count


def frame_sample_offset(frame: int, spp: int, world: int, rank: int) -> Tuple[int, int]:
    """Sample block of `rank` for frame `frame` when the per-pixel streams continue across frames
    (the reference initialises its RNG states once and keeps consuming them frame after frame,
    src/Global/RendererTime.cu:395 + shader/Shader.cu:170): frame f owns samples [f*spp, (f+1)*spp)."""
    off, cnt = spp_partition(spp, world, rank)
    return frame * spp + off, cnt


def frame_partition(n_frames: int, world: int, rank: int) -> List[int]:
    """Frames of a time series rendered by `rank` (round robin, BASELINE config 5)."""
    if n_frames < 0 or world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad partition request frames={n_frames} world={world} rank={rank}")
    return list(range(rank, n_frames, world))


def reduce_accum(accum, dst: int = 0):
    """Sum the per-rank accumulation buffers (float4 per pixel: rgb sum, w = sample count) into
    rank `dst`.  One collective per frame; a no-op for a single process."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst, op=dist.ReduceOp.SUM)
    return accum


def render_spp_split(scene, cam, width, height, spp, max_depth, seed, accum, *, frame: int = 0,
                     background=(0.7, 0.8, 0.9), stream=None, rank: int | None = None,
                     world: int | None = None, dst: int = 0):
    """One frame of the spp-split path on this rank: render this rank's sample block into `accum`
    (a torch float32 tensor [w*h, 4] on the rank's GPU) and reduce to `dst`.  Returns the number of
    samples this rank rendered."""
    if rank is None or world is None:
        r, _, w = env_world()
        rank = r if rank is None else rank
        world = w if world is None else world
    off, cnt = frame_sample_offset(frame, spp, world, rank)
    if cnt:
        scene.render(cam, width, height, cnt, max_depth, seed, background, off, 0, accum.data_ptr(), None, None,
                     None, stream)
    else:
        accum.zero_()
    reduce_accum(accum, dst)
    return cnt
