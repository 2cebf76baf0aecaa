#!/usr/bin/env python
"""bench.py -- Mrays/s (primary+secondary) and ms/frame of the progressive path-tracing hot path.

Workload (BASELINE.json configs[1]): Stanford bunny (Assets/bunny_low.obj, translated
(0,-15.2,-40)) inside Assets/cornell.obj, BVH built by the reference's SplitBVHBuilder/CudaBVH (committed
fixture tests/golden/scene_bunny_cornell.npz), the 8 default spheres of BasicScene.cpp:177-218 plus one METAL
sphere ("mixed diffuse/metal/specular"), 1920x1080, 16 spp per step per GPU, depth 8, default camera.

One step = one progressive frame: every rank traces 16 samples/pixel (disjoint GLOBAL sample indices), the
float4 sums are reduced onto rank 0 with one NCCL reduce, rank 0 resolves to RGBA8.  Weak scaling: the frame
carries 16*N spp.  `value` = rays traced by all ranks / device time (max over ranks).

  python bench.py [--gpus N --steps K --warmup W]        our CUDA path
  python bench.py --impl reference ...                   the reference's CPU tracer (oracle/_ref/libcpuref.so)
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOAD = "bunny_cornell_1080p_16spp_depth8"
METRIC = "Mrays/s (prim+sec) @1080p 16spp depth 8"
# Algorithmic bytes per ray of THIS workload, counted by the CPU oracle executing the reference algorithm on
# the reference BVH2 arrays (64*N_inner + 48*N_tri + 16*N_leaf + 4*[tri hit] + 44*N_sphere), full
# frame, 1 spp, seed 0 (profiles/r01_algorithmic_bytes.json; re-measured on a sample by the cpu_baseline leg).
_REAL_STDOUT = None
ALGO_BYTES_PER_RAY = None   # filled from profiles/r01_algorithmic_bytes.json
ACCUM_BYTES_PER_PIXEL_PER_LAUNCH = 32


def emit_line(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def load_algo_bytes():
    """-> (bytes per ray of the whole bounce-loop iteration, bytes per ray of the BVH traversal alone)"""
    p = os.path.join(ROOT, "profiles", "r01_algorithmic_bytes.json")
    with open(p) as fh:
        a = json.load(fh)[WORKLOAD]
    trav = 64 * a["inner_nodes_per_ray"] + 48 * a["tri_tests_per_ray"] + 16 * a["leaf_terminators_per_ray"] + 4 * a["tri_hits_per_ray"]
    return a["bytes_per_ray"], trav


def load_issue_profile():
    """warp instructions, lanes per instruction and issue-slot utilisation of one k_wf_extend_ww launch from the same committed
    ncu capture (profiles/r01_traffic.json), or None."""
    p = os.path.join(ROOT, "profiles", "r01_traffic.json")
    if not os.path.exists(p):
        return None
    with open(p) as fh:
        d = json.load(fh).get("k_wf_extend_ww", {})
    return d if "warp_instructions" in d else None


def load_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of one k_wf_extend_ww launch from the committed `ncu --set full`
    capture of this same command (profiles/r01_traffic.json), or None."""
    p = os.path.join(ROOT, "profiles", "r01_traffic.json")
    if not os.path.exists(p):
        return None
    with open(p) as fh:
        return json.load(fh).get("k_wf_extend_ww", {}).get("dram_bytes_per_launch")


def load_scene_arrays():
    z = np.load(os.path.join(ROOT, "tests", "golden", "scene_bunny_cornell.npz"))
    return {k: np.ascontiguousarray(z[k]) for k in z.files}


def bench_spheres(gpt):
    items = [(list(s.posRad), list(s.emi), list(s.col), s.mat) for s in gpt.default_spheres()]
    items.append(([-9.0, -9.0, -33.0, 5.0], [0.0, 0.0, 0.0], [0.9, 0.8, 0.7], gpt.METAL))
    return gpt.spheres_from_list(items)


def bench_info(gpt, args):
    info = gpt.default_kernel_info(args.width, args.height)
    info.depth = args.depth
    return info


# ---------------------------------------------------------------- clocks
class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) < 9:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[5:9]):
                if v.lower() == "active":
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.rows)}


# ---------------------------------------------------------------- CPU legs (the only places oracle/ is executed)
def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def write_cpuref_obj(sc, path):
    """The bench mesh, moved into the CpuRayTracer's own room (main.cpp:25-35: z-up, floor z=0, walls x=+-4,
    camera (-2,-5,2.5) looking at the origin): uniform scale to a 4-unit footprint, GPU -z (depth) -> +y."""
    v = sc["verts"].astype(np.float64)
    lo, hi = v.min(0), v.max(0)
    s = 4.0 / max(hi[0] - lo[0], hi[2] - lo[2])
    cx, cz = (lo[0] + hi[0]) / 2, (lo[2] + hi[2]) / 2
    out = np.stack([(v[:, 0] - cx) * s, -(v[:, 2] - cz) * s, (v[:, 1] - lo[1]) * s], axis=1)
    with open(path, "w") as fh:
        fh.write("o bench_mesh\n")
        for p in out:
            fh.write("v %.7f %.7f %.7f\n" % (p[0], p[1], p[2]))
        for t in sc["faces"]:
            fh.write("f %d %d %d\n" % (t[0] + 1, t[1] + 1, t[2] + 1))


_cpuref = None


def cpuref_lib(sc):
    """oracle/_ref/libcpuref.so = the reference's CpuRayTracer compiled in place (oracle/Makefile)."""
    global _cpuref
    if _cpuref is None:
        so = os.path.join(ROOT, "oracle", "_ref", "libcpuref.so")
        if not os.path.exists(so):
            return None
        L = C.CDLL(so)
        L.cpuref_render.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.POINTER(C.c_ulonglong), C.POINTER(C.c_double), C.c_void_p]
        tmp = os.path.join(tempfile.gettempdir(), "gpt_bench_mesh_%d.obj" % os.getpid())
        write_cpuref_obj(sc, tmp)
        devnull = os.open(os.devnull, os.O_WRONLY)
        saved = os.dup(1)
        os.dup2(devnull, 1)                      # the reference prints load banners
        try:
            rc = L.cpuref_load(tmp.encode())
        finally:
            os.dup2(saved, 1)
            os.close(saved)
            os.close(devnull)
            os.unlink(tmp)
        if rc != 0:
            return None
        _cpuref = L
    return _cpuref


def time_cpuref(sc, args, rows, threads):
    """Reference CPU tracer on `rows` rows centred in the frame, args.spp samples. -> (rays, seconds)"""
    L = cpuref_lib(sc)
    if L is None:
        return None
    y0 = max(0, args.height // 2 - rows // 2)
    rays, secs = C.c_ulonglong(0), C.c_double(0)
    rc = L.cpuref_render(args.width, args.height, args.spp, y0, min(args.height, y0 + rows), threads, C.byref(rays), C.byref(secs), None)
    if rc != 0:
        return None
    return rays.value, secs.value


def time_oracle_port(gpt, sc, args, rows, threads):
    """The scalar oracle (same algorithm as the GPU path) on a horizontal band of the same frame."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from support import reflib, scenes
    L = reflib.oracle()
    sph, ns = scenes.orc_spheres(bench_spheres(gpt))
    osc = scenes.orc_scene(sc, sph, ns)
    info = bench_info(gpt, args)
    p = scenes.orc_params(info, 0)
    y0 = max(0, args.height // 2 - rows // 2)
    cnt = reflib.OrcCounters()
    band = np.zeros((args.height, args.width, 4), np.float32)
    t0 = time.time()
    rc = L.orc_render_rows(C.byref(osc), C.byref(p), C.c_uint64(0), C.c_uint32(args.spp), C.c_int(y0), C.c_int(min(args.height, y0 + rows)),
                           band.ctypes.data_as(C.c_void_p), C.byref(cnt), C.c_int(threads))
    secs = time.time() - t0
    assert rc == 0
    return cnt.rays, secs, cnt.algorithmic_bytes() / max(1, cnt.rays)


def sized_rows(probe_rows, probe_secs, target_secs, height):
    """rows of the frame that should take about target_secs, from a short probe run"""
    rate = probe_rows / max(probe_secs, 1e-3)
    return int(max(probe_rows, min(height, rate * target_secs)))


def cpu_baseline(gpt, sc, args):
    """Bounded samples (~12 s each) of the same workload on all host cores: the reference CPU tracer
    (kind=reference) and the scalar oracle port of the GPU algorithm (kind=port)."""
    threads = os.cpu_count() or 1
    probe = max(2, min(args.height, 2 * threads))
    out = None
    r = time_cpuref(sc, args, probe, threads)
    if r is not None and r[1] > 0:
        rows = sized_rows(probe, r[1], 12.0, args.height)
        r = time_cpuref(sc, args, rows, threads)
        out = {"value": r[0] / r[1] / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "reference",
               "sample": "%d of %d rows at %dx%d, %d spp (reference CpuRayTracer: own scene/tree/RR, throughput only); %.1f s; cpu=%s"
                         % (rows, args.height, args.width, args.height, args.spp, r[1], cpu_model())}
    o = time_oracle_port(gpt, sc, args, probe, threads)
    rows = sized_rows(probe, o[1], 12.0, args.height)
    o = time_oracle_port(gpt, sc, args, rows, threads)
    port = {"value": o[0] / o[1] / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "port",
            "sample": "%d of %d rows, %d spp, scalar oracle with OpenMP over rows; %.1f s" % (rows, args.height, args.spp, o[1]),
            "bytes_per_ray_sample": o[2]}
    if out is None:
        return port, None
    return out, port


def reference_gpu_kernel(gpt, torch, sc, sph, args, rays_per_step):
    """The reference's OWN CUDA kernel (oracle/_ref/reftrace.cubin = GpuPathTracer/tracer.cu compiled in place) on
    this GPU, same scene / camera / depth, 16x16 blocks, args.spp one-sample frames per step exactly as
    BasicScene::run drives it (BasicScene.cpp:395-404).  Baseline leg only: nothing of it is on the product path."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from support import refgpu
    if not refgpu.available():
        return None
    k = refgpu.RefGpuKernel(torch.cuda.current_device())
    info = bench_info(gpt, args)
    rs = refgpu.RefScene(torch, info, sc, sph, torch.cuda.current_device())

    def step(i):
        for f in range(args.spp):
            info.hash = gpt.hash64(i * args.spp + f)
            info.constantPdf = f + 1
            k.launch(info)
    for i in range(2):
        step(i)
    torch.cuda.synchronize()
    n = 3
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for i in range(n):
        step(2 + i)
    t1.record()
    torch.cuda.synchronize()
    ms = t0.elapsed_time(t1) / n
    k.close()
    return {"ms_per_frame_16spp": ms, "value": rays_per_step / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "kind": "reference",
            "what": "reference __global__ trace (tracer.cu:343-400) built for sm_100a from /root/reference (PTX-level vote.sync fix), "
                    "%d launches of 1 spp per step, 16x16 blocks; rays per step taken from our counter (same algorithm, same "
                    "expected ray count; the reference kernel does not count rays)" % args.spp}


def run_reference_arm(args):
    """--impl reference: the reference's own CPU implementation on this box's host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import gpt_b200 as gpt      # host-only helpers (scene preset); no GPU used on this arm
    sc = load_scene_arrays()
    threads = os.cpu_count() or 1
    probe = max(2, min(args.height, int(2 * threads)))
    have_ref = cpuref_lib(sc) is not None
    pr = time_cpuref(sc, args, probe, threads) if have_ref else time_oracle_port(gpt, sc, args, probe, threads)
    # whole run (warmup + steps) bounded to about two minutes of CPU time
    rows = sized_rows(probe, pr[1], 120.0 / max(1, args.steps + args.warmup), args.height)
    if not have_ref:
        # oracle/_ref absent (reference never compiled on this snapshot): time the oracle port instead
        kind = "port"

        def step():
            r = time_oracle_port(gpt, sc, args, rows, threads)
            return r[0], r[1]
    else:
        kind = "reference"

        def step():
            return time_cpuref(sc, args, rows, threads)
    for _ in range(args.warmup):
        step()
    rays = secs = 0.0
    for _ in range(args.steps):
        r = step()
        rays += r[0]
        secs += r[1]
    val = rays / secs / 1e6
    sample = "%d of %d rows at %dx%d, %d spp per step, %s; cpu=%s" % (
        rows, args.height, args.width, args.height, args.spp,
        "CpuRayTracer compiled from /root/reference (own scene, tree, RNG, Russian roulette)" if kind == "reference" else "scalar oracle port",
        cpu_model())
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64" if kind == "reference" else "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "width": args.width, "height": args.height, "spp": args.spp, "depth": args.depth},
            "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit_line(line)
    return 0


# ---------------------------------------------------------------- our arm
def run_ours(args):
    import torch
    import gpt_b200 as gpt
    from importlib import import_module
    mg = import_module("gpt_b200.multigpu")

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    W, H, SPP = args.width, args.height, args.spp
    sc = load_scene_arrays()
    sph = bench_spheres(gpt)
    info = bench_info(gpt, args)
    flags = (gpt.FLAG_MEGAKERNEL if args.pipeline == "mega" else 0) | gpt.FLAG_TIME_KERNELS
    ctx = gpt.Context(local, flags)
    ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], sph)
    frame = torch.zeros(H, W, 4, dtype=torch.float32, device=dev)
    total = torch.zeros(H, W, 4, dtype=torch.float32, device=dev) if rank == 0 else None
    rgba = torch.zeros(H, W, dtype=torch.int32, device=dev)
    host_rgba = torch.zeros(H, W, dtype=torch.int32).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2
    stream = torch.cuda.current_stream(dev)
    k_events = []

    def render_fn(begin, count):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        ctx.render(info, frame, args.seed, begin, count, stream.cuda_stream)
        e1.record(stream)
        k_events.append((e0, e1))

    peer = None
    if world > 1 and args.reduce == "peer":
        peer = mg.PeerFrames(ctx, W, H, dist=dist, device=dev)

    def render_ptr(ptr, begin, count):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        ctx.render(info, ptr, args.seed, begin, count, stream.cuda_stream)
        e1.record(stream)
        k_events.append((e0, e1))

    def step(i, e2e=False):
        flush.zero_()                                                  # L2 flush between timed iterations
        # weak scaling: every rank renders SPP samples of a SPP*world-sample progressive frame
        if peer is not None:
            # fused path: root P2P-reads the peers' frames inside ONE reduce + accumulate + resolve kernel
            peer.frame(render_ptr, i * SPP * world, SPP * world, total, rgba if rank == 0 else None, None, stream.cuda_stream)
        else:
            mg.render_progressive_frame(render_fn, frame, total, i * SPP * world, SPP * world, dist=dist)
            if rank == 0:
                ctx.resolve(total, W, H, rgba, None, stream.cuda_stream)
        if rank == 0 and e2e:
            host_rgba.copy_(rgba, non_blocking=True)
            stream.synchronize()

    def sync_all():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def timed(n_steps, first, e2e):
        sync_all()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record(stream)
        for i in range(n_steps):
            step(first + i, e2e)
        t1.record(stream)
        sync_all()
        ms = torch.tensor([t0.elapsed_time(t1)], device=dev)
        if dist is not None:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for i in range(args.warmup):
        step(i)
    sync_all()
    ctx.reset_stats()
    k_events.clear()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = timed(args.steps, args.warmup, False)
    st = ctx.stats()
    ktimes = ctx.kernel_times()
    kernel_ms = [a.elapsed_time(b) for a, b in k_events]
    rays = torch.tensor([float(st.rays)], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(rays)
    total_rays = float(rays.item())
    launches = int(st.kernel_launches)
    # ---- e2e: same steps through the C-ABI with HOST output, D2H (+ the 296-B kernelInfo H2D) inside the timing ----
    if world == 1:
        def e2e_step(i):
            flush.zero_()
            ctx.render_frame_host(info, args.seed, i * SPP, SPP, host_rgba, None)     # gpt_render_frame_host: host buffers
        for i in range(2):
            e2e_step(i)
        sync_all()
        ctx.reset_stats()
        t0 = time.perf_counter()
        for i in range(args.steps):
            e2e_step(i)
        torch.cuda.synchronize(dev)
        e2e_ms = 1e3 * (time.perf_counter() - t0)
        e2e_rays = float(ctx.stats().rays)
    else:
        ctx.reset_stats()
        e2e_ms = timed(args.steps, args.warmup + args.steps, True)
        r2 = torch.tensor([float(ctx.stats().rays)], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(r2)
        e2e_rays = float(r2.item())
    sampler.stop_flag = True

    if rank == 0:
        peaks = {}
        pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
        peak_src = "fallback"
        hbm = 6650.0
        if os.path.exists(pk):
            peaks = json.load(open(pk))
            hbm = float(peaks.get("hbm_gbs", hbm))
            peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)"
        bpr, bpr_trav = load_algo_bytes()
        pass_avg_ms = float(np.mean(kernel_ms))
        pass_bytes = bpr * float(st.rays) / max(1, len(kernel_ms)) + ACCUM_BYTES_PER_PIXEL_PER_LAUNCH * W * H
        pass_achieved = pass_bytes / (pass_avg_ms * 1e-3) / 1e9
        if args.pipeline == "mega" or ktimes["extend"][1] == 0:
            kname, k_avg_ms, rays_per_launch, k_bytes_per_ray, k_total_ms = "k_render_mega<false>", pass_avg_ms, float(st.rays) / max(1, len(kernel_ms)), bpr, pass_avg_ms * len(kernel_ms)
        else:
            # dominant kernel: the BVH traversal.  Its algorithmic bytes are the traversal terms of SURVEY 8-d's formula
            # (nodes, triangles, terminators, index remap); the 44 B x spheres term belongs to k_wf_shade.
            k_total_ms, n_ext = ktimes["extend"]
            kname, k_avg_ms, rays_per_launch, k_bytes_per_ray = "k_wf_extend_ww", k_total_ms / n_ext, float(st.rays) / n_ext, bpr_trav
        achieved = k_bytes_per_ray * rays_per_launch / (k_avg_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": total_rays / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "scene": "Assets/cornell.obj + Assets/bunny_low.obj (5000 tris, reference SBVH) + 9 spheres",
                       "width": W, "height": H, "spp_per_gpu_per_step": SPP, "frame_spp": SPP * world, "depth": args.depth,
                       "pipeline": args.pipeline,
                       "collective": "none" if world == 1 else ("none on the data path: root reads the peers' float4 frames over NVLink inside the fused "
                                     "reduce+accumulate+resolve kernel (CUDA IPC); one 4-byte NCCL all_reduce per step as a stream-ordered barrier"
                                     if peer is not None else "1 NCCL reduce(float4 accum) per step"),
                       "l2": "256 MiB buffer written between steps (L2 flush); scene itself (0.5 MB) is cache-resident by design",
                       "rays_per_step": total_rays / args.steps, "ms_per_frame_16spp": ms / args.steps},
            "e2e": {"value": e2e_rays / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 296,
                    "d2h_bytes_per_step": W * H * 4, "ms_per_step": e2e_ms / args.steps,
                    "what": "gpt_render_frame_host: kernelInfo by value in, RGBA8 frame to pinned host memory out, sync per step"
                            if world == 1 else "render + NCCL reduce + resolve + RGBA8 D2H to pinned host on rank 0, sync per step"},
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm, "traffic": load_traffic(),
                         "kernel": kname, "kernel_ms_avg": k_avg_ms, "kernel_share_of_step": k_total_ms / ms,
                         "algorithmic_bytes_per_ray": k_bytes_per_ray, "rays_per_launch": rays_per_launch,
                         "algorithmic_bytes_per_launch": k_bytes_per_ray * rays_per_launch, "peak_source": peak_src,
                         "whole_pass": {"achieved": pass_achieved, "frac": pass_achieved / hbm, "bytes_per_ray": bpr, "ms_avg": pass_avg_ms,
                                        "what": "all kernels of one gpt_render pass (generate, extend, shade, collect, accumulate)"},
                         "kernel_class_ms_per_step": {k: v[0] / args.steps for k, v in ktimes.items() if v[1]},
                         "kernel_class_launches_per_step": {k: v[1] / args.steps for k, v in ktimes.items() if v[1]},
                         "note": "scene (0.4 MB packed) is L1/L2 resident: algorithmic bytes are served by cache, so the fraction can exceed 1 "
                                 "of HBM peak; the binding resource is issue slots x SIMT efficiency (profiles/)"},
            "clocks": sampler.summary(),
        }
        ip = load_issue_profile()
        if ip is not None and kname == "k_wf_extend_ww":
            # second view of the same kernel: the resource that actually binds it.  Peak = 4 warp instructions per SM per clock.
            sm_count = torch.cuda.get_device_properties(0).multi_processor_count
            mhz = (line["clocks"] or {}).get("sm_mhz") or 1965.0
            peak = sm_count * 4 * mhz * 1e6
            per_ray = ip["warp_instructions"] / 33177011.875
            ach = per_ray * rays_per_launch / (k_avg_ms * 1e-3)
            line["roofline"]["issue"] = {"bound": "issue slots", "warp_instructions_per_ray": per_ray, "lanes_per_instruction": ip["lanes_per_instruction"],
                                         "achieved": ach / 1e12, "peak": peak / 1e12, "unit": "T warp-inst/s", "frac": ach / peak,
                                         "ncu_issue_active_pct": ip["issue_active_pct"],
                                         "what": "warp instructions per ray from the committed ncu capture x rays per launch / live launch time, against "
                                                 "SMs x 4 schedulers x SM clock; the fraction the HBM figure cannot show"}
        # SURVEY 8-d "secondary figure": traversal steps/s = (inner nodes + triangle tests) per ray (oracle counters on the reference
        # arrays, profiles/r01_algorithmic_bytes.json) x rays/s, and rays/s per bounce index (coherent primaries vs incoherent
        # secondaries) from depth-limited renders of the same samples: bounce b = render(depth b+1) - render(depth b)
        try:
            with open(os.path.join(ROOT, "profiles", "r01_algorithmic_bytes.json")) as fh:
                ab = json.load(fh)[WORKLOAD]
            steps_per_ray = ab["inner_nodes_per_ray"] + ab["tri_tests_per_ray"]
            line["traversal_steps"] = {"per_ray": steps_per_ray, "value": steps_per_ray * line["value"] / 1e3, "unit": "G steps/s",
                                       "what": "(inner-node visits + triangle tests) of the reference algorithm per ray x rays/s"}
        except Exception as ex:
            line["traversal_steps"] = {"unavailable": repr(ex)}
        if world == 1:
            try:
                cum = []
                for dpt in range(1, args.depth + 1):
                    inf_d = bench_info(gpt, args)
                    inf_d.depth = dpt
                    ctx.render(inf_d, frame, args.seed, 0, SPP, stream.cuda_stream)      # warm-up of this depth's buffers
                    torch.cuda.synchronize(dev)
                    ctx.reset_stats()
                    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    reps = 3
                    t0.record(stream)
                    for r in range(reps):
                        ctx.render(inf_d, frame, args.seed, 0, SPP, stream.cuda_stream)
                    t1.record(stream)
                    torch.cuda.synchronize(dev)
                    cum.append((float(ctx.stats().rays) / reps, t0.elapsed_time(t1) / reps))
                pb, prev = [], (0.0, 0.0)
                for b, c in enumerate(cum):
                    dr, dt = c[0] - prev[0], c[1] - prev[1]
                    pb.append({"bounce": b, "rays": dr, "ms": dt, "Mrays_s": dr / (dt * 1e-3) / 1e6 if dt > 0 else None})
                    prev = c
                line["per_bounce"] = {"rows": pb, "what": "differences of depth-limited renders of the same 16 samples (bounce 0 also carries "
                                      "generate + collect + accumulate); no L2 flush between repetitions"}
            except Exception as ex:
                line["per_bounce"] = {"unavailable": repr(ex)}
        if world == 1 and args.pipeline != "mega":
            # secondary figure, NOT the headline: the same workload on the tree this library's own device builder makes from the
            # same mesh (gpt_build_bvh, PLOC) instead of the reference's host-built SBVH the headline traverses
            try:
                with gpt.Context(local, 0) as c2:
                    built = c2.build_bvh(sc["verts"], sc["faces"])
                    cold_ms = built.build_ms          # first build of the process: CUDA loads the builder's kernels lazily
                    c2.free_bvh(built)
                    built = c2.build_bvh(sc["verts"], sc["faces"])
                    c2.upload_built(built, sph)
                    for i in range(max(3, args.warmup)):
                        c2.render(info, frame, args.seed, i * SPP, SPP, stream.cuda_stream)
                    torch.cuda.synchronize(dev)
                    c2.reset_stats()
                    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    t0.record(stream)
                    for i in range(args.steps):
                        flush.zero_()
                        c2.render(info, frame, args.seed, (i + 8) * SPP, SPP, stream.cuda_stream)
                        c2.resolve(frame, W, H, rgba, None, stream.cuda_stream)
                    t1.record(stream)
                    torch.cuda.synchronize(dev)
                    ms2 = t0.elapsed_time(t1)
                    line["device_built_tree"] = {"value": float(c2.stats().rays) / (ms2 * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": ms2 / args.steps,
                                                 "builder": "gpt_build_bvh (PLOC clustering, SAH leaf collapse)", "build_ms": built.build_ms, "build_ms_first_in_process": cold_ms,
                                                 "inner_nodes": built.n_inner_nodes, "depth": built.depth,
                                                 "what": "same mesh, spheres, camera and samples; only the tree differs from the headline's"}
                    c2.free_bvh(built)
            except Exception as ex:
                line["device_built_tree"] = {"unavailable": repr(ex)}
        if world == 1 and not args.no_cpu_baseline:
            base, port = cpu_baseline(gpt, sc, args)
            line["cpu_baseline"] = base
            if port is not None:
                line["cpu_oracle_port"] = port
            try:
                rg = reference_gpu_kernel(gpt, torch, sc, sph, args, total_rays / args.steps)
            except Exception as ex:      # baseline leg must never take the product line down
                rg = {"unavailable": repr(ex)}
            if rg is not None:
                line["reference_gpu_kernel"] = rg
        emit_line(line)
    if peer is not None:
        peer.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--depth", type=int, default=8)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--pipeline", default="wavefront", choices=["mega", "wavefront"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--reduce", default="peer", choices=["peer", "nccl"], help="multi-GPU frame reduce: fused P2P kernel (default) or torch.distributed.reduce")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    # stdout carries exactly ONE JSON line: anything libraries print to fd 1 meanwhile (NCCL's version banner, the
    # reference CPU tracer's load messages) is routed to stderr, and the line is written to the real stdout at the end
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
