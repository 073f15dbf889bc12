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
