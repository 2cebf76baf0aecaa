"""-m gpu: gpt_render_static, the render entry point that allocates nothing (include/gpt.h; no counterpart in the reference),
called through the C-ABI and compared with gpt_render on the same inputs.  Both run the same
kernels on the same samples, so the float sums must be BIT-IDENTICAL (the tolerance of tests/test_gpu_parity.py: zero differing
pixels).  Named test_zz_* so that pytest collects it last: it must pass alone in a fresh process and after every other test."""
import numpy as np
import pytest

from support import scenes

pytestmark = pytest.mark.gpu


def make_ctx(gpt, name, flags=0):
    ctx = gpt.Context(0, flags)
    sc = scenes.load_scene(name)
    ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], gpt.default_spheres())
    return ctx


def frame_info(gpt, w, h, depth):
    info = gpt.default_kernel_info(w, h)
    info.depth = depth
    info.cam.dist = 12.0 * h / 720.0
    return info


def test_static_render_equals_render_on_two_calls_in_a_row(gpt, torch_cuda):
    torch = torch_cuda
    ctx = make_ctx(gpt, "bunny_cornell")
    other = make_ctx(gpt, "cornell")          # a second context of the same device: it shares the static array
    try:
        info = frame_info(gpt, 200, 150, 6)
        big = frame_info(gpt, 512, 512, 4)      # exactly GPT_STATIC_MAX_PIXELS
        assert big.width * big.height == gpt.STATIC_MAX_PIXELS
        acc_s = torch.zeros(150, 200, 4, dtype=torch.float32, device="cuda")
        acc_r = torch.zeros_like(acc_s)
        big_s = torch.zeros(512, 512, 4, dtype=torch.float32, device="cuda")
        big_r = torch.zeros_like(big_s)
        free = []
        for rnd in range(2):                    # round 0 loads every kernel (lazy module loading takes device memory), round 1 is measured
            acc_s.zero_(); acc_r.zero_(); big_s.zero_(); big_r.zero_()
            torch.cuda.synchronize()
            free.append(torch.cuda.mem_get_info()[0])
            # call 1: 30 000 pixels x 4 samples = one wave, summed straight from the per-path slots
            ctx.render_static(info, acc_s, 5, 0, 4)
            # a call of another context and another frame shape in between: leaves the shared array full of its own state.
            # 512 x 512 x 3 samples = three waves of 262 144 paths
            other.render_static(big, big_s, 9, 0, 3)
            # call 2, in a row on the first context: 30 000 x 9 = 270 000 paths > GPT_STATIC_MAX_PATHS, two waves (8 + 1 samples)
            ctx.render_static(info, acc_s, 5, 4, 9)
            torch.cuda.synchronize()
            free.append(torch.cuda.mem_get_info()[0])
            ctx.render(info, acc_r, 5, 0, 4)
            other.render(big, big_r, 9, 0, 3)
            ctx.render(info, acc_r, 5, 4, 9)
            torch.cuda.synchronize()
            a, b = acc_s.cpu().numpy(), acc_r.cpu().numpy()
            assert np.isfinite(b).all() and (b[..., 3] == 13.0).all() and b[..., :3].max() > 0.0
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), "round %d: %d floats differ" % (rnd, (a != b).sum())
            a, b = big_s.cpu().numpy(), big_r.cpu().numpy()
            assert (b[..., 3] == 3.0).all() and b[..., :3].max() > 0.0
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), "round %d, 512 x 512: %d floats differ" % (rnd, (a != b).sum())
        # free device memory before and after the three static calls of the measured round: nothing was allocated
        assert free[3] == free[2], "device memory changed by %d bytes across gpt_render_static calls" % (free[2] - free[3])
    finally:
        ctx.close(); other.close()


def test_static_render_equals_the_megakernel(gpt, torch_cuda):
    """the same samples whatever the pipeline: the static entry against gpt_render forced onto the single-launch megakernel"""
    torch = torch_cuda
    ctx = make_ctx(gpt, "sphere_mesh")
    mega = make_ctx(gpt, "sphere_mesh", gpt.FLAG_MEGAKERNEL)
    try:
        info = frame_info(gpt, 96, 54, 8)
        acc_s = torch.zeros(54, 96, 4, dtype=torch.float32, device="cuda")
        acc_m = torch.zeros_like(acc_s)
        for call in range(2):
            ctx.render_static(info, acc_s, 77, 3 * call, 3)
            mega.render(info, acc_m, 77, 3 * call, 3)
            torch.cuda.synchronize()
            a, b = acc_s.cpu().numpy(), acc_m.cpu().numpy()
            assert (b[..., 3] == 3.0 * (call + 1)).all()
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), "call %d: %d floats differ" % (call, (a != b).sum())
    finally:
        ctx.close(); mega.close()


def test_static_render_rejects_what_does_not_fit(gpt, torch_cuda):
    torch = torch_cuda
    ctx = make_ctx(gpt, "cornell")
    empty = gpt.Context(0, 0)
    try:
        acc = torch.zeros(1024 * 512 * 4, dtype=torch.float32, device="cuda")
        with pytest.raises(gpt.GptError, match="GPT_STATIC_MAX_PIXELS"):
            ctx.render_static(frame_info(gpt, 1024, 512, 4), acc, 1, 0, 1)
        with pytest.raises(gpt.GptError, match="GPT_STATIC_MAX_PIXELS"):
            ctx.render_static(frame_info(gpt, 513, 512, 4), acc, 1, 0, 1)       # one column too many
        with pytest.raises(gpt.GptError, match="GPT_STATIC_MAX_DEPTH"):
            ctx.render_static(frame_info(gpt, 64, 64, gpt.STATIC_MAX_DEPTH + 1), acc, 1, 0, 1)
        with pytest.raises(gpt.GptError, match="no uploaded scene"):
            empty.render_static(frame_info(gpt, 64, 64, 4), acc, 1, 0, 1)
        torch.cuda.synchronize()
        assert float(acc.abs().max()) == 0.0                                     # a rejected call writes nothing
        # the limits themselves are inside: depth 64 on a small frame, and sample_count = 0 is a no-op
        ctx.render_static(frame_info(gpt, 64, 64, gpt.STATIC_MAX_DEPTH), acc, 1, 0, 1)
        ctx.render_static(frame_info(gpt, 64, 64, 4), acc, 1, 0, 0)
        torch.cuda.synchronize()
        assert float(acc[:64 * 64 * 4].view(-1, 4)[:, 3].min()) == 1.0
    finally:
        ctx.close(); empty.close()
