"""Host-side logic of the two multi-GPU partitions of the path (the reference is
single-GPU, docs/technical-details.md:327).  One process per GPU, `torch.distributed` for the plumbing.

  spp split      every rank holds the whole scene, rebuilds the same LBVH locally and renders a
                 contiguous block of the frame's samples for every pixel; ONE exchange per frame:
                 a sum-reduce of the float4 accumulation buffer to rank 0 (NCCL over NVLink on the
                 GPU box, gloo in the CPU tests).  Sample s of pixel p uses the same XORWOW stream
                 (seed, p, s*32) whatever the partition, so the image does not depend on the world
                 size except for the fp32 re-association of the per-rank partial sums (layout "frame";
                 layout "rank" gives every rank its own contiguous region of the streams instead,
                 which spares the per-frame re-positioning of the RNG states: frame_sample_offset).
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


RANK_REGION = 1 << 20  # samples of a pixel's stream reserved per rank by the "rank" layout (2^25 draws of the 2^67 a sequence holds)


def frame_sample_offset(frame: int, spp: int, world: int, rank: int, layout: str = "frame") -> Tuple[int, int]:
    """Sample block of `rank` for frame `frame` when the per-pixel streams continue across frames
    (the reference initialises its RNG states once and keeps consuming them frame after frame,
    src/Global/RendererTime.cu:395 + shader/Shader.cu:170).
      layout "frame": frame f owns samples [f*spp, (f+1)*spp) and the ranks split that block — the frame is the one a
                      single GPU renders, whatever the world size (partial sums re-associated), but from one frame to
                      the next a rank's block is not contiguous: its per-pixel states are re-positioned every frame
                      (one GF(2) matrix-vector product per pixel, ~1 ms at 4K).
      layout "rank":  rank r owns the region [r*RANK_REGION, (r+1)*RANK_REGION) of every pixel's stream and walks it
                      contiguously, frame after frame: no re-positioning.  The frame then is the sum of `world`
                      single-GPU renders at those offsets (same estimator, disjoint samples)."""
    off, cnt = spp_partition(spp, world, rank)
    if layout == "rank":
        if (frame + 1) * cnt > RANK_REGION:
            raise ValueError("rank layout: a rank's region of the sample stream is exhausted")
        return rank * RANK_REGION + frame * cnt, cnt
    if layout != "frame":
        raise ValueError(f"unknown layout {layout!r}")
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
                     world: int | None = None, dst: int = 0, layout: str = "frame"):
    """One frame of the spp-split path on this rank: render this rank's sample block into `accum`
    (a torch float32 tensor [w*h, 4] on the rank's GPU) and reduce to `dst`.  Returns the number of
    samples this rank rendered."""
    if rank is None or world is None:
        r, _, w = env_world()
        rank = r if rank is None else rank
        world = w if world is None else world
    off, cnt = frame_sample_offset(frame, spp, world, rank, layout)
    if cnt:
        scene.render(cam, width, height, cnt, max_depth, seed, background, off, 0, accum.data_ptr(), None, None,
                     None, stream)
    else:
        accum.zero_()
    reduce_accum(accum, dst)
    return cnt


# ---------------------------------------------------------------------------------------------------
# Exchange step over peer memory (include/rtb200.h rt_peer_* / rt_resolve_parts): instead of an NCCL
# reduce of the whole accumulation buffer to one rank followed by a resolve there, every rank sums ITS
# slice of the pixels straight out of all ranks' accumulation buffers (mapped with CUDA IPC, i.e. NVLink
# loads on a multi-GPU box), resolves it and stores the RGBA8 pixels into the destination rank's frame
# buffer: reduce-scatter + resolve + gather in one kernel per rank, parts added in rank order.

def pixel_slice(n_pixels: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous pixel range [first, first+count) that `rank` sums and resolves."""
    if n_pixels < 0 or world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad slice request n_pixels={n_pixels} world={world} rank={rank}")
    base, extra = divmod(n_pixels, world)
    return rank * base + min(rank, extra), base + (1 if rank < extra else 0)


class _DevMem:
    """Minimal __cuda_array_interface__ carrier so torch can view memory the C ABI allocated."""
    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def tensor_view(ptr: int, shape, dtype: str, device):
    import torch
    return torch.as_tensor(_DevMem(ptr, shape, {"float32": "<f4", "uint8": "|u1"}[dtype]), device=device)


class PeerExchange:
    """The per-rank accumulation buffers (float4 per pixel) and frame buffers (RGBA8) of all ranks,
    mapped into this process.  Handles travel through torch.distributed (any backend)."""

    def __init__(self, ctx, n_pixels: int, rank: int, world: int, device):
        """Collective over all ranks.  Raises RuntimeError on EVERY rank if any rank could not allocate,
        export or map a buffer (the collectives below are executed by all ranks regardless, so nobody
        is left waiting), so that callers can fall back to the NCCL reduce together."""
        import torch
        import torch.distributed as dist
        self.ctx, self.n_pixels, self.rank, self.world, self.device = ctx, n_pixels, rank, world, device
        self.parts, self.frames, self._opened = [], [], []
        self.accum_ptr = self.frame_ptr = self.hits_ptr = None
        self.accum = self.frame = self.hits = None
        self.peer_hits = []
        nccl = dist.get_backend() == "nccl"
        self._nccl = nccl
        err = None
        h_acc = h_frm = h_hit = bytes(64)
        try:
            self.accum_ptr, h_acc = ctx.peer_alloc(n_pixels * 16)
            self.frame_ptr, h_frm = ctx.peer_alloc(n_pixels * 4)
            self.hits_ptr, h_hit = ctx.peer_alloc(n_pixels * 16)
            self.hits = tensor_view(self.hits_ptr, (n_pixels, 4), "float32", device)
            self.accum = tensor_view(self.accum_ptr, (n_pixels, 4), "float32", device)
            self.frame = tensor_view(self.frame_ptr, (n_pixels, 4), "uint8", device)
            self.accum.zero_()
            self.frame.zero_()
        except Exception as e:  # noqa: BLE001 - reported to all ranks below
            err = e
        mine = torch.frombuffer(bytearray(h_acc + h_frm + h_hit), dtype=torch.uint8)
        if nccl:
            mine = mine.to(device)
        allh = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allh, mine)
        if err is None:
            try:
                for r in range(world):
                    if r == rank:
                        self.parts.append(self.accum_ptr)
                        self.frames.append(self.frame_ptr)
                        self.peer_hits.append(self.hits)
                        continue
                    hb = bytes(allh[r].cpu().numpy().tobytes())
                    a, f, h = ctx.peer_open(hb[:64]), ctx.peer_open(hb[64:128]), ctx.peer_open(hb[128:192])
                    self.parts.append(a)
                    self.frames.append(f)
                    self.peer_hits.append(tensor_view(h, (n_pixels, 4), "float32", device))
                    self._opened += [a, f, h]
            except Exception as e:  # noqa: BLE001
                err = e
        ok = torch.tensor([0.0 if err else 1.0], device=device if nccl else "cpu")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        self._token = torch.zeros(1, device=device if nccl else "cpu")
        if float(ok.item()) < 1.0:
            self.close()
            raise RuntimeError(f"peer-memory exchange unavailable on at least one rank ({err!r} on rank {rank})")

    def _check_stream(self, stream):
        """fence() and the peer copies are enqueued on torch's CURRENT stream, the kernels on `stream`: they are
        ordered only when both are the same stream (rtb200.h, rt_resolve_parts).  Anything else would let a peer
        read a buffer before its producer has finished: refuse it."""
        import torch
        # None = the legacy default stream (handle 0) on the C-ABI side: ordered with the fences only while torch's
        # current stream is that stream too, not inside a torch.cuda.stream(side_stream) block
        handle = 0 if stream is None else int(stream)
        if handle != int(torch.cuda.current_stream().cuda_stream):
            raise ValueError("PeerExchange: pass torch's current stream (None only while torch's default stream is current)"
                             " - the NCCL fences run on it")

    def fence(self):
        """Every rank's work enqueued so far is complete before anything enqueued after this runs
        (NCCL: a stream-ordered 4-byte all-reduce; other backends: device sync + host barrier)."""
        import torch
        import torch.distributed as dist
        if self._nccl:
            dist.all_reduce(self._token)
        else:
            torch.cuda.synchronize()
            dist.barrier()

    def gather_primary(self, scene, cam, width, height, stream=None):
        """Primary visibility split over the ranks: this rank traces the camera rays of its pixel slice
        (rt_render_primary), then copies the other slices out of the peers' buffers (NVLink reads) and
        hands the complete frame of hit records to the scene (rt_scene_set_primary_hits), so that the
        following rt_render does not repeat a pass every rank would otherwise do identically."""
        self._check_stream(stream)
        first, cnt = pixel_slice(self.n_pixels, self.world, self.rank)
        scene.render_primary(cam, width, height, first, cnt, self.hits_ptr, stream)
        self.fence()
        for r in range(self.world):
            if r == self.rank:
                continue
            f, c = pixel_slice(self.n_pixels, self.world, r)
            if c:
                self.hits[f:f + c].copy_(self.peer_hits[r][f:f + c], non_blocking=True)
        scene.set_primary_hits(self.hits_ptr)

    def resolve_slice(self, dst: int = 0, stream=None, sum_out_ptr=None):
        """Sum + resolve this rank's pixel slice from all ranks' partial sums into rank `dst`'s frame."""
        self._check_stream(stream)
        first, cnt = pixel_slice(self.n_pixels, self.world, self.rank)
        self.fence()
        self.ctx.resolve_parts(self.parts, first, cnt, sum_out_ptr, None, self.frames[dst], stream)
        self.fence()

    def close(self):
        import torch
        torch.cuda.synchronize()
        for p in self._opened:
            try:
                self.ctx.peer_close(p)
            except Exception:  # noqa: BLE001 - best effort on the failure path
                pass
        self._opened = []
        self.accum = self.frame = self.hits = None
        self.peer_hits = []
        for p in (self.accum_ptr, self.frame_ptr, self.hits_ptr):
            if p:
                self.ctx.peer_free(p)
        self.accum_ptr = self.frame_ptr = self.hits_ptr = None


def render_spp_split_peer(scene, xch: PeerExchange, cam, width, height, spp, max_depth, seed, *, frame: int = 0,
                          background=(0.7, 0.8, 0.9), stream=None, dst: int = 0, split_primary: bool = False,
                          layout: str = "frame"):
    """One frame of the spp split with the peer-memory exchange: this rank's sample block goes into its
    own exchange buffer, then every rank sums and resolves its pixel slice into rank `dst`'s RGBA8 frame."""
    off, cnt = frame_sample_offset(frame, spp, xch.world, xch.rank, layout)
    split = split_primary and xch.world > 1
    if split:
        xch.gather_primary(scene, cam, width, height, stream)
    try:
        if cnt:
            scene.render(cam, width, height, cnt, max_depth, seed, background, off, 0, xch.accum_ptr, None, None, None, stream)
        else:
            xch.accum.zero_()
    finally:
        if split:
            scene.set_primary_hits(None)
    xch.resolve_slice(dst, stream)
    return cnt
