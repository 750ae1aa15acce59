"""Multi-GPU host logic on CPU: strip partition, frame sharding, and a world_size-2 gloo gather of oracle-rendered
strips that must reproduce the whole-frame oracle render byte for byte."""
import os
import sys
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from b200rt import sharding
from conftest import load_golden


@pytest.mark.parametrize("W,world,strip", [(320, 2, 16), (7680, 8, 16), (100, 3, 16), (7, 4, 16), (33, 2, 5)])
def test_strips_partition_the_frame(W, world, strip):
    seen = np.zeros(W, dtype=int)
    for r in range(world):
        for a, b in sharding.strip_ranges(W, r, world, strip):
            assert 0 <= a < b <= W and b - a <= strip
            seen[a:b] += 1
        assert len(sharding.columns_of_rank(W, r, world, strip)) == sum(b - a for a, b in sharding.strip_ranges(W, r, world, strip))
    assert (seen == 1).all()
    counts = [len(sharding.columns_of_rank(W, r, world, strip)) for r in range(world)]
    assert max(counts) - min(counts) <= strip


def test_frames_of_rank():
    got = sorted(k for r in range(8) for k in sharding.frames_of_rank(64, r, 8))
    assert got == list(range(64))
    assert all(len(sharding.frames_of_rank(64, r, 8)) == 8 for r in range(8))


def test_pack_unpack_roundtrip():
    rng = np.random.default_rng(0)
    f = rng.integers(0, 256, (3, 77, 9), dtype=np.uint8)
    parts = [sharding.pack_strips(f, r, 3, 4) for r in range(3)]
    assert np.array_equal(sharding.unpack_strips(parts, 77, 4), f)


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle_f64 as O
        g = load_golden("rand_a")
        W, H = int(g["W"]), int(g["H"])
        args = (g["cam_origin"], g["cam_rotation"], g["spheres"], g["lights"], g["planes"], float(g["amb"]), float(g["lamb"]),
                float(g["refl"]), int(g["refl_depth"]), bool(g["aliasing"]))
        mine = np.zeros((3, W, H), np.uint8)
        for a, b in sharding.strip_ranges(W, rank, world, 4):            # this rank renders only its strips
            mine[:, a:b, :] = O.render(None, *args, grid=g["grid"], shape=(W, H), x_range=(a, b))[:, a:b, :]
        t = torch.from_numpy(mine)
        sharding.allgather_frame(t, rank, world, strip=4)
        ms = torch.tensor([float(rank + 1)])
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)                        # the benchmark's max-over-ranks reduction
        q.put((rank, t.numpy().copy(), float(ms.item())))
    finally:
        dist.destroy_process_group()


def test_gloo_world2_gather_equals_whole_frame(oracle):
    g = load_golden("rand_a")
    W, H = int(g["W"]), int(g["H"])
    args = (g["cam_origin"], g["cam_rotation"], g["spheres"], g["lights"], g["planes"], float(g["amb"]), float(g["lamb"]),
            float(g["refl"]), int(g["refl_depth"]), bool(g["aliasing"]))
    whole = oracle.render(None, *args, grid=g["grid"], shape=(W, H))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, frame, mx in res:
        assert mx == 2.0
        assert np.array_equal(frame, whole), f"rank {rank}: gathered frame differs from the whole-frame render"


@pytest.mark.gpu
def test_library_strips_match_sharding_module():
    """libb200rt's strip_width/strip_rank/strip_world select exactly sharding.strip_ranges, and the union of the
    ranks' strips is bit-identical to the whole frame (AA frame: strip-edge neighbours are index arithmetic)."""
    from b200rt import render_host
    for name, strip, world in (("rand_b", 5, 3), ("c4_small", 16, 2)):
        g = load_golden(name)
        W, H = int(g["W"]), int(g["H"])
        args = (g["cam_origin"], g["cam_rotation"], g["spheres"], g["lights"], g["planes"], float(g["amb"]), float(g["lamb"]),
                float(g["refl"]), int(g["refl_depth"]), bool(g["aliasing"]))
        whole = render_host(None, *args, grid=g["grid"], shape=(W, H))
        acc = np.zeros_like(whole)
        for r in range(world):
            part = render_host(None, *args, grid=g["grid"], shape=(W, H), strips=(strip, r, world))
            cols = sharding.columns_of_rank(W, r, world, strip)
            other = np.setdiff1d(np.arange(W), cols)
            assert np.array_equal(part[:, cols, :], whole[:, cols, :])
            assert not part[:, other, :].any(), "pixels outside the rank's strips must stay untouched"
            acc[:, cols, :] = part[:, cols, :]
        assert np.array_equal(acc, whole)


@pytest.mark.gpu
def test_multi_device_frame_equals_single_device():
    import torch
    from b200rt import render_frame
    from b200rt.scene import Scene, Camera
    scene = Scene.default_scene()
    cam = Camera(resolution=(200, 120), position=[-2, 0, 2.0], euler=[0, -30, 0])
    one = render_frame(scene, cam, 0.0, 0.6, 0.3, 2, True)
    n = torch.cuda.device_count()
    devs = list(range(n)) if n > 1 else [0, 0]           # one GPU: two host threads share the device (serialised by the library)
    assert np.array_equal(render_frame(scene, cam, 0.0, 0.6, 0.3, 2, True, devices=devs), one)


def test_shared_frame_assembles_strips_across_mappings(tmp_path):
    """SharedFrame without pinning (no GPU here): two mappings of one /dev/shm-style file, each 'rank' writes its strips,
    the creator sees the assembled frame (the multi-process direct device-to-host gather of bench.py --gpus N)."""
    from b200rt.sharding import SharedFrame, columns_of_rank
    W, H = 100, 24
    whole = np.random.default_rng(0).integers(0, 256, (3, W, H), dtype=np.uint8)
    a = SharedFrame("frame_test", W, H, create=True, pin=False, directory=str(tmp_path))
    b = SharedFrame("frame_test", W, H, create=False, pin=False, directory=str(tmp_path))
    for rank, fr in enumerate((a, b)):
        cols = columns_of_rank(W, rank, 2, 16)
        fr.array[:, cols, :] = whole[:, cols, :]
    assert np.array_equal(np.asarray(a.array), whole)
    b.close(); a.close()
    import os
    assert not os.path.exists(os.path.join(str(tmp_path), "frame_test"))


def _shared_worker(rank, world, port, q, name, directory):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle_f64 as O
        g = load_golden("rand_a")
        W, H = int(g["W"]), int(g["H"])
        args = (g["cam_origin"], g["cam_rotation"], g["spheres"], g["lights"], g["planes"], float(g["amb"]), float(g["lamb"]),
                float(g["refl"]), int(g["refl_depth"]), bool(g["aliasing"]))
        fr = None
        if rank == 0:                                                    # bench.py's sequence: rank 0 creates, barrier, the others attach
            fr = sharding.SharedFrame(name, W, H, create=True, pin=False, directory=directory)
        dist.barrier()
        if rank != 0:
            fr = sharding.SharedFrame(name, W, H, create=False, pin=False, directory=directory)
        for a, b in sharding.strip_ranges(W, rank, world, 4):            # every rank writes its strips straight into the one frame
            fr.array[:, a:b, :] = O.render(None, *args, grid=g["grid"], shape=(W, H), x_range=(a, b))[:, a:b, :]
        fr.array.flush()
        dist.barrier()
        if rank == 0:
            q.put(np.array(fr.array))                                    # the creating rank holds the assembled frame
        dist.barrier()
        fr.close()
    finally:
        dist.destroy_process_group()


def test_gloo_world2_shared_frame_is_assembled_on_rank0(oracle, tmp_path):
    """The multi-process direct gather of bench.py --gpus N (SharedFrame: one mapping, every rank writes its strips, no collective on the
    data path) with two gloo processes on CPU: rank 0 ends up holding the whole frame."""
    g = load_golden("rand_a")
    W, H = int(g["W"]), int(g["H"])
    args = (g["cam_origin"], g["cam_rotation"], g["spheres"], g["lights"], g["planes"], float(g["amb"]), float(g["lamb"]),
            float(g["refl"]), int(g["refl_depth"]), bool(g["aliasing"]))
    whole = oracle.render(None, *args, grid=g["grid"], shape=(W, H))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_shared_worker, args=(r, 2, port, q, "shared_frame_gloo", str(tmp_path))) for r in range(2)]
    for p in procs:
        p.start()
    frame = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert np.array_equal(frame, whole)
    assert not os.path.exists(os.path.join(str(tmp_path), "shared_frame_gloo"))
