"""Multi-GPU frame reduce fused with accumulate + resolve (gpt_reduce_resolve) -- GPU tests.
The kernel arithmetic and the single-rank path run on one GPU; the IPC peer path needs two and is skipped otherwise
(the driver's scaling run and `gpurun --gpus 2` exercise it through bench.py)."""
import os
import sys

import numpy as np
import pytest

from support import scenes

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reduce_resolve_kernel_matches_sequential_sum(gpt, torch_cuda):
    torch = torch_cuda
    W, H = 97, 41                                                       # ragged
    g = torch.Generator(device="cuda").manual_seed(1)
    frames = [torch.rand(H, W, 4, device="cuda", generator=g) * 3 for _ in range(5)]
    for f in frames:
        f[..., 3] = 4.0
    total0 = torch.rand(H, W, 4, device="cuda", generator=g)
    total0[..., 3] = 8.0
    with gpt.Context(0) as ctx:
        for n in (1, 2, 5):
            total = total0.clone()
            rgba = torch.zeros(H, W, dtype=torch.int32, device="cuda")
            rgb = torch.zeros(H, W, 3, dtype=torch.float32, device="cuda")
            ctx.reduce_resolve_ptrs(frames[:n], W, H, total, rgba, rgb)
            s = frames[0].clone()
            for f in frames[1:n]:
                s += f                                                   # rank order
            want = total0 + s
            rgba2 = torch.zeros_like(rgba); rgb2 = torch.zeros_like(rgb)
            ctx.resolve(want, W, H, rgba2, rgb2)
            torch.cuda.synchronize()
            assert torch.equal(total, want) and torch.equal(rgba, rgba2) and torch.equal(rgb, rgb2)
        # no running sum: resolve of the reduced frame alone
        rgb = torch.zeros(H, W, 3, dtype=torch.float32, device="cuda")
        ctx.reduce_resolve_ptrs(frames[:2], W, H, None, None, rgb)
        rgb2 = torch.zeros_like(rgb)
        ctx.resolve(frames[0] + frames[1], W, H, None, rgb2)
        torch.cuda.synchronize()
        assert torch.equal(rgb, rgb2)


def test_peer_frames_single_rank_equals_render_plus_resolve(gpt, torch_cuda):
    from importlib import import_module
    mg = import_module("gpt_b200.multigpu")
    torch = torch_cuda
    sc = scenes.load_scene("cornell")
    info = gpt.default_kernel_info(96, 54)
    info.cam.dist = 0.9
    with gpt.Context(0) as ctx:
        ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], gpt.default_spheres())
        pf = mg.PeerFrames(ctx, 96, 54, dist=None, device="cuda")
        total = torch.zeros(54, 96, 4, dtype=torch.float32, device="cuda")
        rgba = torch.zeros(54, 96, dtype=torch.int32, device="cuda")
        for f in range(3):                                               # three progressive frames of 4 spp, alternating buffers
            pf.frame(lambda ptr, b, n: ctx.render(info, ptr, 5, b, n), f * 4, 4, total, rgba)
        want = torch.zeros_like(total)
        for f in range(3):
            fr = torch.zeros_like(total)
            ctx.render(info, fr, 5, f * 4, 4)
            want += fr
        rgba2 = torch.zeros_like(rgba)
        ctx.resolve(want, 96, 54, rgba2, None)
        torch.cuda.synchronize()
        assert torch.equal(total, want) and torch.equal(rgba, rgba2)
        pf.close()


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import gpt_b200 as gpt
    from importlib import import_module
    mg = import_module("gpt_b200.multigpu")
    from support import scenes as sc_mod
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    sc = sc_mod.load_scene("cornell")
    info = gpt.default_kernel_info(160, 90)
    info.cam.dist = 1.5
    dev = torch.device("cuda", rank)
    ctx = gpt.Context(rank)
    ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], gpt.default_spheres())
    st = torch.cuda.current_stream(dev).cuda_stream
    pf = mg.PeerFrames(ctx, 160, 90, dist=dist, device=dev)
    total = torch.zeros(90, 160, 4, dtype=torch.float32, device=dev)
    for f in range(4):
        pf.frame(lambda ptr, b, n: ctx.render(info, ptr, 11, b, n, st), f * 6, 6, total if rank == 0 else None, None, None, st)
    # the NCCL path on the same samples
    frame = torch.zeros(90, 160, 4, dtype=torch.float32, device=dev)
    total2 = torch.zeros_like(total) if rank == 0 else None
    for f in range(4):
        mg.render_progressive_frame(lambda b, n: ctx.render(info, frame, 11, b, n, st), frame, total2, f * 6, 6, dist=dist)
    torch.cuda.synchronize()
    if rank == 0:
        np.save(os.path.join(out_dir, "peer.npy"), total.cpu().numpy())
        np.save(os.path.join(out_dir, "nccl.npy"), total2.cpu().numpy())
    pf.close()
    ctx.close()
    dist.barrier()
    dist.destroy_process_group()


def test_two_gpu_peer_reduce_equals_nccl_reduce(gpt, torch_cuda, tmp_path):
    torch = torch_cuda
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp
    port = 29700 + (os.getpid() % 1000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    peer, nccl = np.load(tmp_path / "peer.npy"), np.load(tmp_path / "nccl.npy")
    assert (peer[..., 3] == 24).all()
    np.testing.assert_array_equal(peer, nccl)                            # two ranks: a+b is order independent
    sc = scenes.load_scene("cornell")
    info = gpt.default_kernel_info(160, 90)
    info.cam.dist = 1.5
    want = scenes.oracle_render(sc, gpt.default_spheres(), info, 11, 0, 24)
    np.testing.assert_allclose(peer[..., :3], want[..., :3], rtol=2e-6, atol=2e-6)
