"""spp split on 2 ranks through the CUDA path == the 1-rank frame (SURVEY §4 "multi-GPU without a
cluster").  With >= 2 visible GPUs the ranks use cuda:0 / cuda:1 and NCCL; on a 1-GPU box both ranks
share cuda:0 and reduce host copies over gloo — the partition / stream mapping under test is the same."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
W, H, SPP, DEPTH, SEED = 160, 96, 6, 5, 21


def _scene(pkg, synth):
    import scenes
    return scenes.random_scene(pkg, synth, 3000, n_tris=60, seed=8, metal_every=5, fuzz=0.15)


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    import torch
    import torch.distributed as dist
    import scenes
    pkg = importlib.import_module("nvidia-optix-ray-tracer_b200")
    synth = importlib.import_module("nvidia-optix-ray-tracer_b200.synth")
    mg = importlib.import_module("nvidia-optix-ray-tracer_b200.multigpu")
    ngpu = torch.cuda.device_count()
    nccl = ngpu >= world
    dev = rank if nccl else 0
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl" if nccl else "gloo", rank=rank, world_size=world)
    gs = scenes.GpuScene(pkg, _scene(pkg, synth), device=f"cuda:{dev}")
    cam = pkg.camera_configure([2.2, 0.1, 0.2], [0, 0, 0], [0, 0, 1], False)
    for frame in range(2):
        accum = torch.zeros((W * H, 4), dtype=torch.float32, device=f"cuda:{dev}")
        if nccl:
            mg.render_spp_split(gs.scene, cam, W, H, SPP, DEPTH, SEED, accum, frame=frame, rank=rank, world=world,
                                stream=torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
            res = accum.cpu()
        else:
            off, cnt = mg.frame_sample_offset(frame, SPP, world, rank)
            gs.scene.render(cam, W, H, cnt, DEPTH, SEED, (0.7, 0.8, 0.9), off, 0, accum.data_ptr(), None, None, None,
                            torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
            res = accum.cpu()
            mg.reduce_accum(res, 0)
        if rank == 0:
            np.save(os.path.join(out_dir, f"split_{frame}.npy"), res.numpy())
    gs.close()
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_spp_split_matches_single_rank(tmp_path, pkg, synth):
    import torch.multiprocessing as mp
    import scenes
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    gs = scenes.GpuScene(pkg, _scene(pkg, synth))
    cam = pkg.camera_configure([2.2, 0.1, 0.2], [0, 0, 0], [0, 0, 1], False)
    for frame in range(2):
        full = gs.render(cam, W, H, SPP, DEPTH, SEED, sample_offset=frame * SPP, want=("accum",))["accum"]
        got = np.load(tmp_path / f"split_{frame}.npy")
        assert np.array_equal(got[:, 3], full[:, 3])
        assert np.abs(got[:, :3] - full[:, :3]).max() < 4e-6   # fp32 re-association of two partial sums
    gs.close()


# ---------------------------------------------------------------------------------------------------
# Peer-memory exchange (rt_peer_* + rt_resolve_parts): two processes map each other's accumulation
# buffers (NVLink peers on a multi-GPU box, CUDA IPC on the shared GPU otherwise); each sums and
# resolves its pixel slice into rank 0's RGBA8 frame.  Fixed part order => bit-exact against the same
# two partial sums added on the host and resolved by the reference's colour functions (oracle).
def _peer_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    import torch
    import torch.distributed as dist
    import scenes
    pkg = importlib.import_module("nvidia-optix-ray-tracer_b200")
    synth = importlib.import_module("nvidia-optix-ray-tracer_b200.synth")
    mg = importlib.import_module("nvidia-optix-ray-tracer_b200.multigpu")
    ngpu = torch.cuda.device_count()
    nccl = ngpu >= world
    dev = rank if nccl else 0
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl" if nccl else "gloo", rank=rank, world_size=world)
    gs = scenes.GpuScene(pkg, _scene(pkg, synth), device=f"cuda:{dev}")
    cam = pkg.camera_configure([2.2, 0.1, 0.2], [0, 0, 0], [0, 0, 1], False)
    xch = mg.PeerExchange(gs.ctx, W * H, rank, world, f"cuda:{dev}")
    stream = torch.cuda.current_stream().cuda_stream
    for frame in range(2):
        # frame 0 with every rank tracing all primary rays, frame 1 with the primary pass split by pixel slice
        mg.render_spp_split_peer(gs.scene, xch, cam, W, H, SPP, DEPTH, SEED, frame=frame, stream=stream,
                                 split_primary=frame == 1)
        torch.cuda.synchronize()
        np.save(os.path.join(out_dir, f"part_{frame}_{rank}.npy"), xch.accum.cpu().numpy())
        if rank == 0:
            np.save(os.path.join(out_dir, f"frame_{frame}.npy"), xch.frame.cpu().numpy())
        xch.fence()  # nobody starts the next frame before the copies above are done
    xch.close()
    gs.close()
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_peer_exchange_is_bit_exact(tmp_path, pkg, synth):
    import torch
    import torch.multiprocessing as mp
    import oracle_lib
    import scenes
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_peer_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    orc = oracle_lib.load()
    gs = scenes.GpuScene(pkg, _scene(pkg, synth))
    cam = pkg.camera_configure([2.2, 0.1, 0.2], [0, 0, 0], [0, 0, 1], False)
    mg = importlib.import_module("nvidia-optix-ray-tracer_b200.multigpu")
    for frame in range(2):
        parts = [np.load(tmp_path / f"part_{frame}_{r}.npy") for r in range(2)]
        # each rank's partial sum is exactly what one GPU renders for that sample block
        for r in range(2):
            off, cnt = mg.frame_sample_offset(frame, SPP, 2, r)
            solo = gs.render(cam, W, H, cnt, DEPTH, SEED, sample_offset=off, want=("accum",))["accum"]
            assert np.array_equal(parts[r].view(np.uint32), solo.view(np.uint32))
        total = (parts[0] + parts[1]).astype(np.float32)  # the kernel's order: part 0 + part 1
        got = np.load(tmp_path / f"frame_{frame}.npy")
        # (a) bit-exact against the single-GPU resolve kernel applied to the host-side sum
        t = torch.from_numpy(total).cuda()
        u8 = torch.zeros((W * H, 4), dtype=torch.uint8, device="cuda")
        gs.ctx.resolve(t.data_ptr(), W * H, None, u8.data_ptr(), torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert np.array_equal(got, u8.cpu().numpy())
        # (b) within one code value of the reference's colorToUchar4 (CUDA powf vs glibc powf differ by <= 2 ulp)
        idx = np.random.default_rng(frame).integers(0, W * H, 200)
        inv = np.float32(1.0) / total[idx, 3]
        wu = np.array([orc.color_to_uchar4(total[i, :3] * k) for i, k in zip(idx, inv)])
        assert np.abs(got[idx].astype(int) - wu.astype(int)).max() <= 1
    gs.close()


def test_peer_exchange_refuses_unordered_streams():
    """stream=None means the legacy default stream on the C-ABI side; the fences run on torch's current stream.  The two
    are ordered only while torch's default stream is current: inside a side-stream block None must be refused, the
    side stream's own handle accepted."""
    import torch
    mg = importlib.import_module("nvidia-optix-ray-tracer_b200.multigpu")
    x = mg.PeerExchange.__new__(mg.PeerExchange)   # _check_stream uses no state
    x._check_stream(None)
    x._check_stream(torch.cuda.current_stream().cuda_stream)
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        x._check_stream(side.cuda_stream)
        with pytest.raises(ValueError):
            x._check_stream(None)
    with pytest.raises(ValueError):
        x._check_stream(side.cuda_stream)
