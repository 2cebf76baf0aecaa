#!/usr/bin/env python
"""bench.py -- Mrays/s (primary+secondary) and ms/frame of the progressive path-tracing hot path.

Workload (BASELINE.json configs[1]): Stanford bunny (Assets/bunny_low.obj, translated
(0,-15.2,-40)) inside Assets/cornell.obj, BVH built by the reference's SplitBVHBuilder/CudaBVH (committed
fixture tests/golden/scene_bunny_cornell.npz), the 8 default spheres of BasicScene.cpp:177-218 plus one METAL
sphere ("mixed diffuse/metal/specular"), 1920x1080, 16 spp per step per GPU, depth 8, default camera.

One step = one progressive frame: every rank traces its slice of the frame's GLOBAL sample indices, the root sums the ranks'
float4 frames (fused peer-memory reduce + accumulate + resolve kernel; --reduce nccl: one NCCL reduce) and resolves to RGBA8.
Weak scaling (default): 16 spp per GPU, the frame carries 16*N spp.  --scaling strong (or GPT_BENCH_SCALING=strong): the frame is
fixed at --frame-spp (default 16) samples whatever N.  `value` = rays traced by all ranks / device time (max over ranks).

  python bench.py [--gpus N --steps K --warmup W]        our CUDA path
  python bench.py --impl reference ...                   the reference's CPU tracer (oracle/_ref/libcpuref.so)

The roofline block: `achieved` / `frac` are SURVEY 8-d's ALGORITHMIC bytes of the dominant kernel over its live launch time
against the measured HBM copy peak (the contract's figure; the scene is cache-resident, so it can exceed 1), and beside it the
fractions that can bind -- DRAM bytes actually moved, L2 and L1 bytes, issue slots -- each = per-ray counter of a committed ncu
pass over one frame of this workload (profiles/r02_frame_counters.json) x live rays / live kernel time, against peaks MEASURED in
this run by micro-benchmarks (gpt_measure_peaks).  `configs` repeats value + fractions for BASELINE configs 2 and 4.
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


def load_algo_bytes(scene="bench"):
    """-> dict of the oracle's per-ray counters for a scene (profiles/r02_algorithmic_bytes.json, made by
    profiles/tools/algo_bytes.py; the round-1 full-frame count of the bench workload as the fallback), or None"""
    p = os.path.join(ROOT, "profiles", "r02_algorithmic_bytes.json")
    if os.path.exists(p):
        with open(p) as fh:
            a = json.load(fh).get(scene)
        if a:
            return a
    if scene == "bench":
        with open(os.path.join(ROOT, "profiles", "r01_algorithmic_bytes.json")) as fh:
            a = json.load(fh)[WORKLOAD]
        a["traversal_bytes_per_ray"] = (64 * a["inner_nodes_per_ray"] + 48 * a["tri_tests_per_ray"] + 16 * a["leaf_terminators_per_ray"]
                                        + 4 * a["tri_hits_per_ray"])
        return a
    return None


def load_frame_counters(scene="bench"):
    """per-kernel-class counters of one frame, per ray, from the committed ncu counter pass (profiles/tools/ncu_frame.py +
    ncu_frame_summary.py -> profiles/r02_frame_counters.json), or None"""
    p = os.path.join(ROOT, "profiles", "r02_frame_counters.json")
    if not os.path.exists(p):
        return None
    with open(p) as fh:
        return json.load(fh).get(scene)


def bound_fractions(cls, rays, ms, peaks, hbm):
    """The limits a kernel class can actually run into: committed per-ray counters x live rays / live time against peaks measured
    in this run.  cls = one class of load_frame_counters(); every fraction is <= 1 by construction of its peak."""
    if not cls or not rays or not ms:
        return None
    s = ms * 1e-3
    out = {"lanes_per_instruction": cls.get("lanes_per_instruction"), "ncu_issue_active_pct": cls.get("issue_active_pct"),
           "ncu_l1_hit_pct": cls.get("l1_hit_pct"), "ncu_l2_hit_pct": cls.get("l2_hit_pct")}

    def frac(key, peak, unit, scale):
        if cls.get(key) is None or not peak:
            return None
        ach = cls[key] * rays / s / scale
        return {"per_ray": cls[key], "achieved": ach, "peak": peak, "unit": unit, "frac": ach / peak}
    out["dram"] = frac("dram_bytes_per_ray", hbm, "GB/s", 1e9)
    # ceilings = coalesced streaming out of that level; the random-64-byte-gather rates (the traversal's access shape) ride along
    out["l2"] = frac("l2_bytes_per_ray", peaks.get("stream_l2_gbs"), "GB/s", 1e9)
    out["l1"] = frac("l1_bytes_per_ray", peaks.get("stream_l1_gbs"), "GB/s", 1e9)
    for k, g in (("l2", "gather_l2_gbs"), ("l1", "gather_l1_gbs")):
        if out.get(k) and peaks.get(g):
            out[k]["frac_of_random_gather_rate"] = out[k]["achieved"] / peaks[g]
    out["issue"] = frac("warp_inst_per_ray", peaks.get("warp_inst_per_s", 0) / 1e12, "T warp-inst/s", 1e12)
    fr = {k: out[k]["frac"] for k in ("dram", "l2", "l1", "issue") if out.get(k)}
    if fr:
        out["binding"] = max(fr, key=fr.get)
    return out


def workload_config(args, world):
    """the workload description both arms print (the driver compares the two dicts)"""
    strong = args.scaling == "strong"
    frame_spp = args.frame_spp if strong else args.spp * world
    return {"workload": WORKLOAD, "scene": "Assets/cornell.obj + Assets/bunny_low.obj (5000 tris, reference SBVH) + 9 spheres",
            "width": args.width, "height": args.height, "depth": args.depth, "frame_spp": frame_spp,
            "spp_per_gpu_per_step": frame_spp / world if strong else args.spp, "scaling": args.scaling,
            "l2": "256 MiB buffer written between steps (L2 flush); the scene itself (0.5 MB) is cache-resident by design"}


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
            "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f64" if kind == "reference" else "f32", "data": "synthetic",
            "config": workload_config(args, max(1, args.gpus)),
            "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit_line(line)
    return 0


# ---------------------------------------------------------------- other BASELINE configs, one record each
def config_records(gpt, torch, local, args, peaks, hbm, flush):
    """BASELINE configs 2 and 4 on this GPU at reduced sample counts (throughput per sample does not depend on the count), under the
    traversal plan the library picks by itself: value, ms, SURVEY 8-d algorithmic bytes per ray (committed oracle counters,
    profiles/r02_algorithmic_bytes.json), the fractions that can bind (committed ncu counters of the same scene x live time,
    against the peaks measured in this run) and the reference's own CUDA kernel on the very same arrays."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from support import config_check, procedural, refgpu, scenes
    refk = refgpu.RefGpuKernel(local) if (refgpu.available() and not args.no_cpu_baseline) else None
    sph = gpt.default_spheres()
    out = {}
    for name, spp, ref_spp in (("cfg2_dragon100k_sbvh", 16, 4), ("cfg2_dragon1p6M", 16, 2), ("cfg4_20M", 4, 1)):
        t_wall = time.time()
        verts, faces, ref_arrays, W, H, mat = config_check.make_scene(name, gpt, procedural, scenes)
        info = gpt.default_kernel_info(W, H)
        info.depth = args.depth
        info.triCurrentMat = mat
        rec = {"triangles": int(len(faces)), "width": W, "height": H, "spp_timed": spp, "depth": args.depth}
        with gpt.Context(local, gpt.FLAG_TIME_KERNELS) as ctx:
            built = None
            if ref_arrays is not None:
                ctx.upload_scene(ref_arrays["nodes"], ref_arrays["tris"], ref_arrays["indices"], sph)
                rec["tree"] = "reference SBVH (SplitBVHBuilder + CudaBVH, fixture)"
            else:
                built = ctx.build_bvh(verts, faces)
                ctx.upload_built(built, sph)
                rec["tree"] = "device-built (gpt_build_bvh): depth %d, %.1f ms" % (built.depth, built.build_ms)
            acc = torch.zeros(H, W, 4, dtype=torch.float32, device="cuda")
            for w in range(16):                               # buffers + the per-scene trial (<= 12 calls)
                ctx.render(info, acc, 0, 1_000_000 + w * spp, spp)
                if w >= 2 and not ctx.trial_info()[0]:
                    break
            torch.cuda.synchronize()
            ctx.reset_stats()
            reps = 3
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for r in range(reps):
                flush.zero_()
                ctx.render(info, acc, 0, r * spp, spp)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            st, kt = ctx.stats(), ctx.kernel_times()
            rays = float(st.rays) / reps
            rec.update({"value": rays / ms / 1e3, "unit": "Mrays/s", "ms_per_frame": ms, "rays_per_frame": rays,
                        "tree_in_use": ctx.tree_info()["in_use"], "trees": ctx.tree_info(),
                        "rebuilt_tree": ctx.rebuild_info(), "trial_ns_per_ray": ctx.trial_info()[1], "ray_reordering": ctx.reorder_info()[0],
                        "kernel_class_ms": {k: v[0] / reps for k, v in kt.items() if v[1]}})
            ab = load_algo_bytes(name)
            if ab:
                rec["algorithmic_bytes_per_ray"] = {"traversal": ab["traversal_bytes_per_ray"], "whole_bounce": ab["bytes_per_ray"],
                                                    "inner_nodes": ab["inner_nodes_per_ray"], "tri_tests": ab["tri_tests_per_ray"]}
                ext_ms = kt["extend"][0] / reps
                if ext_ms > 0:
                    ach = ab["traversal_bytes_per_ray"] * rays / (ext_ms * 1e-3) / 1e9
                    rec["roofline"] = {"bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm, "kernel": "extend"}
            fc = (load_frame_counters(name) or {}).get("classes", {})
            if fc:
                rec["bounds"] = {"extend": bound_fractions(fc.get("extend"), rays, kt["extend"][0] / reps, peaks, hbm),
                                 "shade": bound_fractions(fc.get("shade"), rays, kt["shade"][0] / reps, peaks, hbm)}
            if refk is not None:
                arr = ref_arrays if ref_arrays is not None else ctx.built_to_host(built)
                i2 = type(info).from_buffer_copy(info)
                rs = refgpu.RefScene(torch, i2, arr, sph, local)

                def frames(first, n):
                    for f in range(n):
                        i2.hash, i2.constantPdf = gpt.hash64(first + f), f + 1
                        refk.launch(i2)
                frames(0, 1)
                torch.cuda.synchronize()
                e0.record()
                frames(1, ref_spp)
                e1.record()
                torch.cuda.synchronize()
                rms = e0.elapsed_time(e1) / ref_spp * spp
                rec["reference_gpu_kernel"] = {"ms_per_frame": rms, "speedup": rms / ms,
                                               "what": "reference __global__ trace on the same arrays, %d one-sample launches timed, scaled to %d spp" % (ref_spp, spp)}
                del rs
            if built is not None:
                ctx.free_bvh(built)
        rec["wall_s"] = time.time() - t_wall
        out[name] = rec
    if refk is not None:
        refk.close()
    return out


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
    W, H = args.width, args.height
    strong = args.scaling == "strong"
    FRAME_SPP = args.frame_spp if strong else args.spp * world      # samples of one progressive frame (all ranks together)
    SPP = args.spp                                                   # single-GPU legs below (e2e, per-bounce, device tree)
    if strong and FRAME_SPP < world:
        raise SystemExit("bench.py: --frame-spp %d cannot be split over %d ranks" % (FRAME_SPP, world))
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
        # every rank renders its block of the frame's FRAME_SPP global sample indices (weak: SPP each; strong: FRAME_SPP / world)
        if peer is not None:
            # fused path: root P2P-reads the peers' frames inside ONE reduce + accumulate + resolve kernel, on a side stream
            peer.frame(render_ptr, i * FRAME_SPP, FRAME_SPP, total, rgba if rank == 0 else None, None, stream)
        else:
            mg.render_progressive_frame(render_fn, frame, total, i * FRAME_SPP, FRAME_SPP, dist=dist)
            if rank == 0:
                ctx.resolve(total, W, H, rgba, None, stream.cuda_stream)
        if rank == 0 and e2e:
            if peer is not None:
                peer.wait_reduced(stream)
            host_rgba.copy_(rgba, non_blocking=True)
            stream.synchronize()

    def reduce_check():
        """N > 1, untimed: one frame through the multi-GPU path (every rank its slice, root sums the ranks) against the SAME global
        sample range rendered by rank 0 alone.  Sample indices are global, so the two frames hold the same samples; they differ
        only in the association of the float sums.  Raises on mismatch."""
        n = FRAME_SPP
        chk = torch.zeros(H, W, 4, dtype=torch.float32, device=dev) if rank == 0 else None
        if peer is not None:
            peer.frame(render_ptr, 0, n, chk, None, None, stream)
            if rank == 0:
                peer.wait_reduced(stream)
        else:
            mg.render_progressive_frame(render_fn, frame, chk, 0, n, dist=dist)
        torch.cuda.synchronize(dev)
        res = None
        if rank == 0:
            alone = torch.zeros(H, W, 4, dtype=torch.float32, device=dev)
            ctx.render(info, alone, args.seed, 0, n, stream.cuda_stream)
            torch.cuda.synchronize(dev)
            diff = float((chk[..., :3] - alone[..., :3]).abs().max().item())
            count_ok = bool(torch.equal(chk[..., 3], alone[..., 3]) and float(alone[..., 3].min().item()) == float(n))
            tol = 2e-6 * n * max(1.0, float(alone[..., :3].max().item()) / n)
            res = {"frame_spp": n, "ranks": world, "max_abs_diff": diff, "tolerance": tol, "count_plane_exact": count_ok,
                   "mean_radiance_alone": float(alone[..., :3].sum().item()) / (3.0 * W * H * n),
                   "what": "peer-reduced float4 frame of all ranks vs the same global sample range rendered by rank 0 alone"}
            if not (count_ok and diff <= tol):
                raise SystemExit("bench.py: multi-GPU frame differs from the single-GPU render of the same samples: %r" % (res,))
        k_events.clear()
        return res

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
        if peer is not None and rank == 0:
            peer.wait_reduced(stream)              # the last frame's reduce runs on the root's side stream: it belongs to the timed region
        t1.record(stream)
        sync_all()
        ms = torch.tensor([t0.elapsed_time(t1)], device=dev)
        if dist is not None:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def finish_trial():
        """The per-scene trial (the tree as uploaded against the rebuilt one, gpt_rebuild_info) belongs in front of the timed region:
        its calls synchronise.  Every rank runs the same number of calls; if the ranks' timings disagree about the tree (they would
        then resolve t-ties differently and reduce_check would see it) all ranks fall back to the tree as uploaded."""
        n = 0
        while n < 16:
            pend = torch.tensor([1.0 if ctx.trial_info()[0] else 0.0], device=dev)
            if dist is not None:
                dist.all_reduce(pend, op=dist.ReduceOp.MAX)
            if float(pend.item()) == 0.0:
                break
            step(100000 + n)
            n += 1
        sync_all()
        use = torch.tensor([float(gpt.Context.TREE_NAMES.index(ctx.tree_info()["in_use"]))], device=dev)
        lo, hi = use.clone(), use.clone()
        if dist is not None:
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        agreed = float(lo.item()) == float(hi.item())
        if not agreed:
            ctx.set_flags(flags | gpt.FLAG_KEEP_TREE)
        return {"trial_calls": n, "ranks_agreed": agreed, "in_use": ctx.tree_info()["in_use"] if agreed else "as uploaded", "trees": ctx.tree_info(),
                "rebuilt": ctx.rebuild_info(), "trial_ns_per_ray_binary_wide_sorted": ctx.trial_info()[1]}

    tree_rec = finish_trial()
    check = reduce_check() if world > 1 else None
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
            ctx.render_frame_host(info, args.seed, i * FRAME_SPP, FRAME_SPP, host_rgba, None)     # gpt_render_frame_host: host buffers
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
        sampler.join(timeout=8)          # a query started inside a short timed region is allowed to finish: at least one sample

    if rank == 0:
        peaks = {}
        pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
        peak_src = "fallback"
        hbm = 6650.0
        if os.path.exists(pk):
            peaks = json.load(open(pk))
            hbm = float(peaks.get("hbm_gbs", hbm))
            peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)"
        ab = load_algo_bytes("bench")
        bpr, bpr_trav = ab["bytes_per_ray"], ab["traversal_bytes_per_ray"]
        fc = load_frame_counters("bench") or {}
        fcls = fc.get("classes", {})
        try:
            live_peaks = ctx.measure_peaks()
        except Exception as ex:
            live_peaks = {"unavailable": repr(ex)}
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
        ext_ms, shade_ms = ktimes["extend"][0], ktimes["shade"][0]
        dram_launch = fcls.get("extend", {}).get("dram_bytes_per_ray")
        line = {
            "metric": METRIC, "value": total_rays / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, world),
            "run": {"pipeline": args.pipeline, "extend_schedule": "ww (persistent warps, while-while, per-lane node and leaf loops)", "tree": tree_rec,
                    "collective": "none" if world == 1 else ("none on the data path: root reads the peers' float4 frames over NVLink inside the fused "
                                  "reduce+accumulate+resolve kernel (CUDA IPC) on a side stream; one 4-byte NCCL all_reduce per step as a stream-ordered barrier"
                                  if peer is not None else "1 NCCL reduce(float4 accum) per step"),
                    "rays_per_step": total_rays / args.steps, "ms_per_frame": ms / args.steps},
            "e2e": {"value": e2e_rays / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 296,
                    "d2h_bytes_per_step": W * H * 4, "ms_per_step": e2e_ms / args.steps,
                    "what": "gpt_render_frame_host: kernelInfo by value in, RGBA8 frame to pinned host memory out, sync per step"
                            if world == 1 else ("every rank renders its slice; " + ("fused peer-memory reduce + accumulate + resolve on the root" if peer is not None
                                                else "NCCL reduce + accumulate + resolve") + "; RGBA8 D2H to pinned host on rank 0, sync per step")},
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
                         "traffic": dram_launch * rays_per_launch if dram_launch else None,
                         "kernel": kname, "kernel_ms_avg": k_avg_ms, "kernel_share_of_step": k_total_ms / ms,
                         "algorithmic_bytes_per_ray": k_bytes_per_ray, "rays_per_launch": rays_per_launch,
                         "algorithmic_bytes_per_launch": k_bytes_per_ray * rays_per_launch, "peak_source": peak_src,
                         "note": "the contract's figure: SURVEY 8-d algorithmic bytes / live launch time against the measured HBM copy peak.  The packed scene "
                                 "(0.4 MB) is L1/L2 resident, those bytes never reach DRAM and the fraction may exceed 1: it bounds nothing here.  `traffic` = DRAM "
                                 "bytes one launch really moves (ray state).  The fractions that can bind are under `bounds`.",
                         "peaks_measured_this_run": live_peaks,
                         "bounds": {"extend": bound_fractions(fcls.get("extend"), float(st.rays), ext_ms, live_peaks, hbm),
                                    "shade": bound_fractions(fcls.get("shade"), float(st.rays), shade_ms, live_peaks, hbm),
                                    "what": "per kernel class: committed ncu per-ray counters (profiles/r02_frame_counters.json: one frame of this workload, "
                                            "every launch) x live rays / live class time, against peaks measured in this run by gpt_measure_peaks: DRAM bytes vs "
                                            "the HBM copy peak, L2 / L1 bytes (lts__t_bytes / l1tex__t_bytes) vs coalesced streaming out of an L2- / L1-resident working set, "
                                            "warp instructions vs the measured issue rate; `binding` = the largest fraction"},
                         "whole_pass": {"achieved": pass_achieved, "frac": pass_achieved / hbm, "bytes_per_ray": bpr, "ms_avg": pass_avg_ms,
                                        "what": "all kernels of one gpt_render pass (extend, shade, accumulate), algorithmic bytes incl. 44 B x spheres"},
                         "kernel_class_ms_per_step": {k: v[0] / args.steps for k, v in ktimes.items() if v[1]},
                         "kernel_class_launches_per_step": {k: v[1] / args.steps for k, v in ktimes.items() if v[1]}},
            "clocks": sampler.summary(),
        }
        if check is not None:
            line["reduce_check"] = check
        # SURVEY 8-d "secondary figure": traversal steps/s = (inner nodes + triangle tests) per ray (oracle counters on the reference
        # arrays, profiles/r01_algorithmic_bytes.json) x rays/s, and rays/s per bounce index (coherent primaries vs incoherent
        # secondaries) from depth-limited renders of the same samples: bounce b = render(depth b+1) - render(depth b)
        try:
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
                                      "the accumulate pass); no L2 flush between repetitions"}
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
        if world == 1 and args.pipeline != "mega":
            # the same steps on the BVH exactly as uploaded (GPT_FLAG_KEEP_TREE): what the rebuilt tree is worth, and the figure to
            # compare with earlier rounds' headline
            try:
                with gpt.Context(local, gpt.FLAG_KEEP_TREE) as c4:
                    c4.upload_scene(sc["nodes"], sc["tris"], sc["indices"], sph)
                    for i in range(max(3, args.warmup)):
                        c4.render(info, frame, args.seed, i * SPP, SPP, stream.cuda_stream)
                    torch.cuda.synchronize(dev)
                    c4.reset_stats()
                    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    t0.record(stream)
                    for i in range(args.steps):
                        flush.zero_()
                        c4.render(info, frame, args.seed, (i + 8) * SPP, SPP, stream.cuda_stream)
                        c4.resolve(frame, W, H, rgba, None, stream.cuda_stream)
                    t1.record(stream)
                    torch.cuda.synchronize(dev)
                    ms4 = t0.elapsed_time(t1)
                    line["tree_as_uploaded"] = {"value": float(c4.stats().rays) / (ms4 * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": ms4 / args.steps,
                                                "what": "GPT_FLAG_KEEP_TREE: the reference's SBVH traversed as uploaded (no rebuilt tree); same steps"}
            except Exception as ex:
                line["tree_as_uploaded"] = {"unavailable": repr(ex)}
        if world == 1 and args.pipeline != "mega":
            # opt-in GPT_FLAG_FAST (MUFU shading arithmetic; statistical parity, tests/test_gpu_refkernel.py "/FAST" cases): same steps
            try:
                with gpt.Context(local, gpt.FLAG_FAST) as c3:
                    c3.upload_scene(sc["nodes"], sc["tris"], sc["indices"], sph)
                    f2 = torch.zeros_like(frame)
                    for i in range(16):                                   # warm-up + the per-scene trial
                        c3.render(info, f2, args.seed, i * SPP, SPP, stream.cuda_stream)
                        if i + 1 >= max(3, args.warmup) and not c3.trial_info()[0]:
                            break
                    torch.cuda.synchronize(dev)
                    c3.reset_stats()
                    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    t0.record(stream)
                    for i in range(args.steps):
                        flush.zero_()
                        f2.zero_()
                        c3.render(info, f2, args.seed, (i + 8) * SPP, SPP, stream.cuda_stream)
                        c3.resolve(f2, W, H, rgba, None, stream.cuda_stream)
                    t1.record(stream)
                    torch.cuda.synchronize(dev)
                    ms3 = t0.elapsed_time(t1)
                    rays3 = float(c3.stats().rays)
                    # the same samples in both modes: what the changed arithmetic does to a 16-spp frame
                    frame.zero_(); f2.zero_()
                    ctx.render(info, frame, args.seed, 0, SPP, stream.cuda_stream)
                    c3.render(info, f2, args.seed, 0, SPP, stream.cuda_stream)
                    torch.cuda.synchronize(dev)
                    a, b = frame[..., :3] / SPP, f2[..., :3] / SPP
                    line["fast_mode"] = {"value": rays3 / (ms3 * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": ms3 / args.steps,
                                         "parity": {"contract": "statistical, against the reference's GPU kernel at 1024 spp (tests/test_gpu_refkernel.py, cases "
                                                                "'/FAST': mean z^2 <= 1.6, |mean z| <= 0.05, box-filtered RMSE <= 1e-3); primary hits identical",
                                                    "same_samples_vs_compat": {"mean_radiance_compat": float(a.mean().item()), "mean_radiance_fast": float(b.mean().item()),
                                                                               "rmse": float(((a - b) ** 2).mean().sqrt().item()),
                                                                               "pixels_differing_by_more_than_1e-3": float(((a - b).abs().amax(dim=-1) > 1e-3).float().mean().item())}},
                                         "what": "GPT_FLAG_FAST: rsqrt normalisation and hardware sin/cos/pow/reciprocal in the shade kernel; traversal unchanged; opt-in"}
            except Exception as ex:
                line["fast_mode"] = {"unavailable": repr(ex)}
        if world == 1 and not args.no_configs:
            try:
                line["configs"] = config_records(gpt, torch, local, args, live_peaks, hbm, flush)
            except Exception as ex:      # a secondary leg must never take the headline down
                line["configs"] = {"unavailable": repr(ex)}
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
    ap.add_argument("--no-configs", action="store_true", help="skip the per-config records (BASELINE configs 2 and 4)")
    ap.add_argument("--scaling", default=os.environ.get("GPT_BENCH_SCALING", "weak"), choices=["weak", "strong"],
                    help="multi-GPU: weak = --spp samples per GPU per step (default); strong = a fixed frame of --frame-spp samples split over the GPUs")
    ap.add_argument("--frame-spp", type=int, default=int(os.environ.get("GPT_BENCH_FRAME_SPP", "16")), help="samples of the fixed frame under --scaling strong")
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
