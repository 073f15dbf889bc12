"""N > 1 host logic on CPU: world_size-2 `gloo` runs of the spp-split reduce and the frame-parallel
partition.  The pixels come from the CPU oracle here (test infrastructure); on the GPU
box the same partition / reduce code drives the CUDA path (tests/test_gpu_multigpu.py, bench.py)."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _mg():
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    return importlib.import_module("nvidia-optix-ray-tracer_b200.multigpu")


def test_spp_partition_covers_every_sample_once():
    mg = _mg()
    for spp in (0, 1, 3, 8, 64, 65):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for r in range(world):
                off, cnt = mg.spp_partition(spp, world, r)
                seen += list(range(off, off + cnt))
            assert seen == list(range(spp)), (spp, world)
    assert mg.frame_sample_offset(3, 64, 8, 2) == (3 * 64 + 16, 8)
    # "rank" layout: every rank walks its own region contiguously from frame to frame; all blocks disjoint
    seen = set()
    for frame in range(3):
        for r in range(4):
            off, cnt = mg.frame_sample_offset(frame, 10, 4, r, "rank")
            assert off == r * mg.RANK_REGION + frame * cnt and cnt == (3 if r < 2 else 2)
            blk = set(range(off, off + cnt))
            assert not (blk & seen)
            seen |= blk
    with pytest.raises(ValueError):
        mg.frame_sample_offset(0, 8, 2, 0, "tile")
    with pytest.raises(ValueError):
        mg.spp_partition(8, 2, 2)


def test_frame_partition_round_robin():
    mg = _mg()
    for n in (0, 1, 7, 256):
        for world in (1, 2, 8):
            allf = sorted(f for r in range(world) for f in mg.frame_partition(n, world, r))
            assert allf == list(range(n))
    assert mg.frame_partition(10, 4, 1) == [1, 5, 9]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank), OMP_NUM_THREADS="2")
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    import torch
    import torch.distributed as dist
    import oracle_lib
    import scenes
    pkg = importlib.import_module("nvidia-optix-ray-tracer_b200")
    synth = importlib.import_module("nvidia-optix-ray-tracer_b200.synth")
    mg = importlib.import_module("nvidia-optix-ray-tracer_b200.multigpu")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    orc = oracle_lib.load()
    d = scenes.random_scene(pkg, synth, 400, n_tris=20, seed=5, metal_every=6, fuzz=0.2)
    wd = scenes.flatten_world(d)
    cam = pkg.camera_configure([2.2, 0.1, 0.2], [0, 0, 0], [0, 0, 1], False)
    W, H, SPP, D, SEED = 96, 64, 7, 5, 3
    # ---- spp split: two frames with streams continuing across frames
    for frame in range(2):
        off, cnt = mg.frame_sample_offset(frame, SPP, world, rank)
        part = orc.render(wd, cam, W, H, cnt, D, SEED, sample_offset=off)["accum"] if cnt else np.zeros((W * H, 4), np.float32)
        acc = torch.from_numpy(part.copy())
        mg.reduce_accum(acc, 0)
        if rank == 0:
            np.save(os.path.join(out_dir, f"split_{frame}.npy"), acc.numpy())
    # ---- frame parallel: each rank renders its frames of a 5-frame "series" (seed = frame)
    for f in mg.frame_partition(5, world, rank):
        img = orc.render(wd, cam, 48, 32, 2, D, 100 + f)["accum"]
        np.save(os.path.join(out_dir, f"frame_{f}.npy"), img)
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_spp_split_and_frame_parallel(tmp_path, pkg, synth, oracle, built_lib):
    import torch.multiprocessing as mp
    import scenes
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    d = scenes.random_scene(pkg, synth, 400, n_tris=20, seed=5, metal_every=6, fuzz=0.2)
    wd = scenes.flatten_world(d)
    cam = pkg.camera_configure([2.2, 0.1, 0.2], [0, 0, 0], [0, 0, 1], False)
    for frame in range(2):
        full = oracle.render(wd, cam, 96, 64, 7, 5, 3, sample_offset=frame * 7)["accum"]
        got = np.load(tmp_path / f"split_{frame}.npy")
        assert np.array_equal(got[:, 3], full[:, 3])          # sample counts add up exactly
        # rgb: same samples, only the order of the final additions differs -> a few ulp of the sum
        assert np.allclose(got[:, :3], full[:, :3], rtol=0, atol=7 * 2.0 ** -22)
        assert np.abs(got[:, :3] - full[:, :3]).max() < 4e-6
    for f in range(5):
        want = oracle.render(wd, cam, 48, 32, 2, 5, 100 + f)["accum"]
        got = np.load(tmp_path / f"frame_{f}.npy")
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))   # byte-identical to the serial run


def test_pixel_slices_tile_the_frame():
    mg = _mg()
    for n in (0, 1, 7, 1920 * 1080, 3840 * 2160 + 3):
        for world in (1, 2, 3, 8):
            pos = 0
            for r in range(world):
                first, cnt = mg.pixel_slice(n, world, r)
                assert first == pos and cnt >= 0
                pos += cnt
            assert pos == n
    with pytest.raises(ValueError):
        mg.pixel_slice(10, 2, 2)
