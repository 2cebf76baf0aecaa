"""CPU tests of the multi-GPU host logic (world_size 2, gloo): the sample-slice partition and the single
reduce per progressive frame.  The per-rank renderer is the CPU oracle here (test-only stand-in for the CUDA
kernel, which has no CPU path)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sample_slice_tiles_the_range(gpt):
    from importlib import import_module
    mg = import_module("gpt_b200.multigpu")
    for total in [0, 1, 2, 7, 16, 17, 1024]:
        for world in [1, 2, 3, 4, 8]:
            seen = []
            for r in range(world):
                off, cnt = mg.sample_slice(total, world, r)
                assert cnt in (total // world, total // world + 1) or (total < world and cnt in (0, 1))
                seen += list(range(off, off + cnt))
            assert seen == list(range(total))
    with pytest.raises(ValueError):
        mg.sample_slice(4, 2, 2)


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import gpt_b200 as gpt
    from importlib import import_module
    mg = import_module("gpt_b200.multigpu")
    from support import scenes
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = scenes.load_scene("cornell")
    sph = gpt.default_spheres()
    info = gpt.default_kernel_info(48, 27)
    info.cam.dist = 12.0 * 27 / 720.0
    accum = torch.zeros(27, 48, 4)
    frame = torch.zeros(27, 48, 4)

    def render_fn(begin, count):
        frame.add_(torch.from_numpy(scenes.oracle_render(sc, sph, info, 9, begin, count, threads=1)))

    done = []
    for f in range(2):                                              # two progressive frames of 5 spp
        done.append(mg.render_progressive_frame(render_fn, frame, accum, f * 5, 5, dist=dist))
    if rank == 0:
        np.save(os.path.join(out_dir, "sum.npy"), accum.numpy())
    np.save(os.path.join(out_dir, "slices_%d.npy" % rank), np.array(done))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_frame_equals_single_rank(gpt, tmp_path):
    import torch.multiprocessing as mp
    from support import scenes
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = np.load(tmp_path / "sum.npy")
    s0, s1 = np.load(tmp_path / "slices_0.npy"), np.load(tmp_path / "slices_1.npy")
    assert s0.tolist() == [[0, 3], [0, 3]] and s1.tolist() == [[3, 2], [3, 2]]
    sc = scenes.load_scene("cornell")
    info = gpt.default_kernel_info(48, 27)
    info.cam.dist = 12.0 * 27 / 720.0
    want = scenes.oracle_render(sc, gpt.default_spheres(), info, 9, 0, 10, threads=2)
    np.testing.assert_array_equal(got[..., 3], want[..., 3])          # 10 samples everywhere
    np.testing.assert_allclose(got[..., :3], want[..., :3], rtol=1e-6, atol=1e-6)
