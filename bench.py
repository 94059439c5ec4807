#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 Path-integrator backend (contract: see DESIGN.md §Measurement).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload c3|c1|c2|c4|c5] [--spp S]

One "step" = one full frame of the workload rendered by the hot path. Default workload C3 (BASELINE.json configs[2],
the configuration the Msamples/s metric is quoted on): synthetic 999 700-triangle terrain, 1920x1080, 64 spp, max depth
8, 60 % Diffuse / 30 % GGX / 10 % Dielectric, LCG sampler. The one JSON line also carries compact sub-records
(`configs`) for the other BASELINE configs that fit the run: C1, C2 and C5 at N = 1, C4 (tile split) at N = 8.

N > 1 (torchrun, one process per GPU): the scene is replicated, every rank renders its share of the frame and ONE NCCL
sum-reduce of the film per step combines it (strong scaling: the frame is fixed). Default share = the sample split
(every GPU renders all pixels with spp/N samples); the bit-identical interleaved-tile split is timed beside it
(`alt_split`) and its reduced film is checked bit for bit against a 1-GPU render (`checks`). Rank 0 then also drives
all N GPUs from ONE process through the C ABI (akb_init(N) + akb_render(n_gpus = N); `single_process`).

`--impl reference` times the reference's own CPU implementation (oracle/_ref, built from /root/reference; falls back to
the oracle port where that library is absent) on a bounded sample of the same workload, on all host threads.
"""
import argparse
import ctypes as C
import hashlib
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

HBM_FALLBACK_GBS = 6650.0  # /opt/skills/guides/B200_PROFILING.md fallback
BYTES_CLOSEST, BYTES_ANY = 52, 36  # algorithmic bytes per ray
BVH_KIND = {"default": 0, "host": 1, "device": 2}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return HBM_FALLBACK_GBS, "fallback"


def make_scene(pkg, workload, spp, ref_subset=False):
    s = pkg.scenes
    if workload in ("c3", "c2"):
        sc = s.terrain_scene(n=708, resolution=(1920, 1080), spp=spp or 64, max_depth=8,
                             materials="c3ref" if ref_subset else "c3")
        name = "C3: synthetic 999700-triangle terrain, 1920x1080, %d spp, Path max_depth 8, 60%% Diffuse / 30%% GGX(0.3) / 10%% Dielectric(1.5), LCG sampler" % sc.integrator.spp
    elif workload == "c1":
        sc = s.cornell_box((512, 512), spp or 16, 5)
        name = "C1: Cornell box 512x512, %d spp, Path max_depth 5" % sc.integrator.spp
    elif workload == "c4":
        sc = s.instanced_scene(n=708, resolution=(3840, 2160), spp=spp or 256, max_depth=8)
        name = "C4: 10 placed copies of the 999700-triangle terrain (9997000 triangles, 10 meshes), 3840x2160, %d spp, Path max_depth 8, 70%% Diffuse / 30%% GGX(0.3)" % sc.integrator.spp
    elif workload == "c5":
        sc = s.many_light_scene(n=708, resolution=(1920, 1080), spp=spp or 128, max_depth=16, n_quads=5000,
                                rr_start_depth=0 if ref_subset else 3)
        name = "C5: 999698-triangle terrain + 10000 emissive triangles, 1920x1080, %d spp, depth 16, RR from depth 3" % sc.integrator.spp
    else:
        raise SystemExit("unknown workload " + workload)
    return sc, name


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def cpu_reference_run(pkg, workload, spp, rows, steps=1, warmup=0):
    """Times the reference's CPU path tracer (own BVH, all host threads) on `rows` image rows of the workload: bands of
    rows spread evenly over the frame, ALL bands handed over in one call, which schedules them the way the reference
    does (16x16-pixel tiles, one shared work list, kernel/integrators/cpu/integrator.cpp:115-141)."""
    from oracle_libs import Oracle, Reference, have_ref
    sc, name = make_scene(pkg, workload, spp, ref_subset=True)
    W, H = sc.camera.resolution
    rows = min(rows, H)
    t0 = time.time()
    if have_ref():
        impl, kind = Reference(sc), "reference"
    else:
        impl, kind = Oracle(sc), "port"
    build_s = time.time() - t0
    # bands spread evenly over the frame: an unbiased sample of sky, horizon and terrain pixels
    # (one contiguous band would be all terrain, all sky - or, for C4, the void between the placed copies)
    band = 8 if rows >= 16 else 1
    nb = max(1, rows // band)
    rows = nb * band
    ys = [min(H - band, int((i + 0.5) * H / nb)) for i in range(nb)]
    rects = [(0, y, W, y + band) for y in ys]
    p = sc.integrator.params()
    threads = os.cpu_count()
    times, busy = [], []
    for i in range(warmup + steps):
        t = time.perf_counter()
        impl.render_rects(p, rects, threads=threads)
        if i >= warmup:
            times.append(time.perf_counter() - t)
            busy.append(impl.busy[0] / max(impl.busy[1], 1e-9))
    n_samples = W * rows * sc.integrator.spp
    sec = float(np.mean(times))
    threads_busy = float(np.mean(busy))
    notes = {"c3": "; reference has no Dielectric BSDF: those 10% of triangles are shaded as Diffuse",
             "c5": "; reference has no Russian roulette: run without it",
             "c4": "; NOTE the reference's own TLAS loses meshes from 5 meshes on (tests/test_round2_cpu.py): its rays mostly miss, so this number flatters the reference"}
    sample = ("%s; CPU sample = %d bands of %d image rows spread evenly over the frame (first rows y = %d, %d, ... %d; %d x %d px x %d spp = %d samples per step), one call, 16x16-pixel tiles over %d worker threads, %.1f of them busy on average (thread CPU time / wall)%s; BVH build %.1f s not included"
              % (name, nb, band, ys[0], ys[1] if nb > 1 else ys[0], ys[-1], W, rows, sc.integrator.spp, n_samples, threads, threads_busy,
                 notes.get(workload, ""), build_s))
    impl.close()
    return {"value": n_samples / sec / 1e6, "unit": "Msamples/s", "cores": int(round(threads_busy)), "threads": threads,
            "threads_busy": threads_busy, "kind": kind, "sample": sample, "workload": name,
            "seconds_per_step": sec, "samples_per_step": n_samples}


CPU_KEYS = ("value", "unit", "cores", "threads", "threads_busy", "kind", "sample")


def cpu_baseline_or_error(pkg, workload, spp, rows):
    try:
        cpu = cpu_reference_run(pkg, workload, spp, rows)
        return {k: cpu[k] for k in CPU_KEYS}
    except Exception as ex:  # the checker is optional for the product; report why
        return {"value": None, "unit": "Msamples/s", "cores": 0, "threads": os.cpu_count(), "kind": "port", "sample": "failed: %r" % (ex,)}


def reference_gpu_run(pkg, workload, spp):
    """The reference's OWN CUDA wavefront path tracer (oracle/_ref/libakari_ref_gpu.so: its gpu/cuda/integrator.cpp
    patched in transit to build for sm_100a) on the same GPU: a comparison point, bounded in spp."""
    from oracle_libs import ReferenceGPU, have_ref_gpu
    if not have_ref_gpu():
        return None
    sc, name = make_scene(pkg, workload, spp, ref_subset=True)
    W, H = sc.camera.resolution
    p = sc.integrator.params()
    p.ray_clamp = 10.0  # its GPU integrator always clamps
    ref = ReferenceGPU(sc)
    out = {"workload": name, "note": "reference gpu::PathTracer (tile 256, managed memory, --use_fast_math), render() + device sync, best of 2 after a warm-up"}
    for label, wf in (("wavefront", True), ("megakernel", False)):
        ref.render_gpu(p, 256, wf)
        ms = min(ref.render_gpu(p, 256, wf)[1] for _ in range(2))
        out[label] = {"value": W * H * sc.integrator.spp / ms / 1e3, "unit": "Msamples/s", "ms_per_step": ms}
    ref.close()
    return out


def ncu_traffic():
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def path_workload_one_gpu(pkg, torch, workload, spp, steps, warmup, bvh, mode, local, cpu_rows, with_cpu=True, with_clocks=False):
    """One Path workload on ONE GPU: device-timed frames (`value`), the host-film C-ABI call (`e2e`), per-stage roofline of
    the traversal and shading kernels (a separate profiled frame), and the CPU baseline on a bounded sample."""
    sc, name = make_scene(pkg, workload, spp)
    t0 = time.time()
    be = pkg.Backend(sc, local, bvh_build=BVH_KIND[bvh])
    build_s = time.time() - t0
    info = be.info()
    W, H = sc.camera.resolution
    film = torch.zeros(H * W * 4, dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    p = sc.integrator.params()
    p.mode = pkg.abi.MODE_WAVEFRONT if mode == "wavefront" else pkg.abi.MODE_MEGAKERNEL
    for _ in range(warmup):
        be.render_device(film.data_ptr(), p, stream)
    torch.cuda.synchronize()
    clocks = ClockSampler(local) if with_clocks else None
    if clocks:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    acc = [0, 0, 0, 0]
    for _ in range(steps):
        st = be.render_device(film.data_ptr(), p, stream)
        acc[0] += st.n_samples
        acc[1] += st.n_closest_rays
        acc[2] += st.n_shadow_rays
        acc[3] += st.n_kernel_launches + 2  # + the two memsets of the render call
    e1.record()
    torch.cuda.synchronize()
    clk = clocks.stop() if clocks else None
    ms_total = e0.elapsed_time(e1)
    samples, n_c, n_s, launches = acc
    # ---- end to end through the host-buffer C-ABI call a reference maintainer would bind (akb_render): every step
    # passes the parameters from host memory and reads the film back to a host buffer inside the timed region.
    host_film_t = torch.zeros((H, W, 4), dtype=torch.float32, pin_memory=True)  # page-locked: the library copies straight into it
    host_film = host_film_t.numpy()
    be.render(p, out=host_film)
    e2e_steps = max(1, min(steps, 3))
    t = time.perf_counter()
    e2e_samples = 0
    for _ in range(e2e_steps):
        _, st = be.render(p, out=host_film)
        e2e_samples += st.n_samples
    e2e_s = time.perf_counter() - t
    peak, peak_src = measured_peaks()
    # ---- roofline of the dominant kernel (wf_trace): per-launch device time measured live with CUDA events on the
    # launch stream (akb_set_option profile_stages pass, outside the timed region), algorithmic bytes = 52 B per
    # closest-hit ray + 36 B per any-hit ray.
    roof = None
    if mode == "wavefront":
        pkg.set_option("profile_stages", 1)
        be.render_device(film.data_ptr(), p, stream)
        stp = be.render_device(film.data_ptr(), p, stream)
        pkg.set_option("profile_stages", 0)
        n_launch = max(1, (stp.n_kernel_launches - 2) // 2)
        bytes_total = BYTES_CLOSEST * stp.n_closest_rays + BYTES_ANY * stp.n_shadow_rays
        achieved = bytes_total / (stp.ms_extend * 1e-3) / 1e9 if stp.ms_extend > 0 else None
        tj = ncu_traffic() if workload == "c3" else {}
        traffic = tj.get("wf_trace_dram_bytes_per_full_frame_launch")
        counters = {k[len("wf_trace_"):]: v for k, v in tj.items() if k.startswith("wf_trace_") and k.endswith(("_pct", "_instruction", "_of_peak"))}
        roof = {"bound": "hbm", "kernel": "wf_trace_kernel (closest-hit + any-hit CWBVH traversal)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": (achieved / peak) if achieved else None,
                "traffic": traffic, "peak_source": peak_src, "launches": n_launch,
                "avg_launch_ms": stp.ms_extend / n_launch, "bytes_per_launch": bytes_total / n_launch,
                "kernel_share_of_step": stp.ms_extend / stp.ms_total, "shade_share_of_step": stp.ms_shade / stp.ms_total,
                "mrays_per_s_in_kernel": (stp.n_closest_rays + stp.n_shadow_rays) / (stp.ms_extend * 1e-3) / 1e6,
                "ncu_counters": counters, "ncu_source": tj.get("source") if traffic else None,
                "note": "BVH traversal is latency/issue/L2-bound, not HBM-streaming-bound; the HBM fraction is low by "
                        "construction (SURVEY.md §8d): ncu_counters carries the issue-slot / pipe / lane figures of the "
                        "same kernel"}
        # second kernel of the step: wf_shade gathers/scatters path state, 216 algorithmic bytes per shaded vertex
        # (ray 32 + hit 20 + state 40 read, state 40 + extension ray 32 written, + 52 of the trace)
        if stp.ms_shade > 0:
            sb = 164.0 * stp.n_closest_rays + 48.0 * stp.n_shadow_rays  # 216 - the 52 B already counted for the trace kernel
            roof["shade"] = {"bound": "hbm", "kernel": "wf_shade_kernel (path_vertex per hit, regeneration, film accumulate)",
                             "achieved": sb / (stp.ms_shade * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                             "frac": sb / (stp.ms_shade * 1e-3) / 1e9 / peak,
                             "ncu_counters": {k[len("wf_shade_"):]: v for k, v in tj.items() if k.startswith("wf_shade_")}}
    cpu = cpu_baseline_or_error(pkg, workload, spp, cpu_rows) if with_cpu else None
    # path state = 10 float4 + one u64 per slot (Workspace, csrc/kernels.cuh) + 12 u32 queue entries per slot
    state_mb = W * H * (168 + 48) / 1e6
    rec = {"metric": "Msamples/s", "value": samples / (ms_total * 1e-3) / 1e6, "unit": "Msamples/s", "steps": steps, "warmup": warmup,
           "ms_per_step": ms_total / steps,
           "config": {"workload": name, "mode": mode, "parallelism": "single", "tile": 32,
                      "l2": "inputs larger than L2: BVH+triangles %.0f MB, shading records %.0f MB, path state + queues %.0f MB"
                            % ((info["bvh_bytes"] + info["triangle_bytes"]) / 1e6, info["n_triangles"] * 96 / 1e6, state_mb),
                      "bvh_nodes": info["n_bvh_nodes"], "scene_build_s": round(build_s, 2),
                      "bvh_build": {1: "host SAH", 2: "device PLOC"}.get(info["bvh_build"]), "bvh_build_s": round(info["build_us"] / 1e6, 3)},
           "e2e": {"value": e2e_samples / e2e_s / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": C.sizeof(pkg.abi.RenderParams),
                   "d2h_bytes_per_step": W * H * 16, "api": "akb_render (page-locked host film buffer)"},
           "gpu_launches": int(launches),
           "rays": {"closest_per_sample": n_c / samples, "shadow_per_sample": n_s / samples,
                    "mrays_per_s": (n_c + n_s) / (ms_total * 1e-3) / 1e6},
           "roofline": roof, "cpu_baseline": cpu}
    if clk is not None:
        rec["clocks"] = clk
    return rec, be


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=["c3", "c1", "c5", "c4", "c2"])
    ap.add_argument("--split", default="auto", choices=["auto", "tile", "sample"],
                    help="N>1: how ranks share the frame (akb_split). auto = sample split (every GPU keeps a full-size "
                         "wavefront; statistically equivalent film), with the bit-identical tile split timed beside it")
    ap.add_argument("--bvh", default="default", choices=["default", "host", "device"],
                    help="akb_bvh_build: library default (host SAH below 4 Mi triangles, device PLOC above), or force one")
    ap.add_argument("--spp", type=int, default=0)
    ap.add_argument("--mode", default="wavefront", choices=["wavefront", "megakernel"])
    ap.add_argument("--cpu-rows", type=int, default=96)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--reference-gpu", action="store_true",
                    help="also time the reference's own CUDA path tracer (oracle/_ref/libakari_ref_gpu.so) on this GPU")
    ap.add_argument("--no-extras", action="store_true", help="skip the ray microbench and the other configs' sub-records")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    pkg = importlib.import_module("akarirender-1_b200")
    # Libraries (NCCL's version banner, the reference's logger) write to fd 1; the contract is ONE JSON line on stdout.
    # Everything but that line goes to stderr.
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(line) + "\n").encode())

    if a.impl == "reference":
        if rank != 0:
            return
        wl = "c3" if a.workload == "c2" else a.workload
        r = cpu_reference_run(pkg, wl, a.spp, a.cpu_rows, steps=max(1, a.steps), warmup=min(a.warmup, 1))
        line = {"impl": "reference", "metric": "Msamples/s", "value": r["value"], "unit": "Msamples/s", "n_gpus": a.gpus,
                "steps": a.steps, "warmup": a.warmup, "ms_per_step": r["seconds_per_step"] * 1e3, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": r["workload"], "mode": "reference CPU path (own BVH), bounded sample: see cpu_baseline.sample"},
                "cpu_baseline": {k: r[k] for k in CPU_KEYS},
                "e2e": {"value": r["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        emit(line)
        return

    import torch
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world == 1:
        if a.workload == "c2":
            emit(bench_c2(pkg, a, local, torch, steps=a.steps, warmup=a.warmup, headline=True))
            return
        rec, be = path_workload_one_gpu(pkg, torch, a.workload, a.spp, a.steps, a.warmup, a.bvh, a.mode, local, a.cpu_rows,
                                        with_cpu=not a.no_cpu_baseline, with_clocks=True)
        line = {"metric": rec["metric"], "value": rec["value"], "unit": rec["unit"], "n_gpus": 1, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": rec["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic"}
        line.update({k: rec[k] for k in ("config", "clocks", "e2e", "gpu_launches", "rays", "roofline", "cpu_baseline")})
        if not a.no_extras:
            line.update(ray_microbench(pkg, be, torch))
        be.close()
        if not a.no_extras and a.workload == "c3" and not a.spp:
            # the other BASELINE configs that fit one GPU, each with its own roofline / e2e / cpu_baseline (fewer steps)
            subs = {}
            for wl in ("c1", "c5"):
                try:
                    r, b2 = path_workload_one_gpu(pkg, torch, wl, 0, 3, 3, a.bvh, a.mode, local, a.cpu_rows if wl != "c1" else 512,
                                                  with_cpu=not a.no_cpu_baseline)
                    b2.close()
                    subs[wl] = r
                except Exception as ex:
                    subs[wl] = {"error": repr(ex)}
            try:
                subs["c2"] = bench_c2(pkg, a, local, torch, steps=2, warmup=1, headline=False)
            except Exception as ex:
                subs["c2"] = {"error": repr(ex)}
            line["configs"] = subs
        if a.reference_gpu:
            # opt-in: the reference's own CUDA path tracer on this GPU (C1 as named; C3 at 16 spp: it needs 2 s per 16 spp)
            try:
                rg = {"c1": reference_gpu_run(pkg, "c1", 0), "c3_16spp": reference_gpu_run(pkg, "c3", 16)}
                if rg["c1"]:
                    line["reference_gpu"] = rg
            except Exception as ex:
                line["reference_gpu"] = {"error": repr(ex)}
        emit(line)
        return
    bench_multi(pkg, a, emit, rank, world, local, torch)


def bench_multi(pkg, a, emit, rank, world, local, torch):
    """N > 1 under torchrun: one process per GPU, scene replicated, one NCCL sum-reduce of the film per frame."""
    import torch.distributed as dist
    dmod = importlib.import_module("akarirender-1_b200.distributed")
    dmod.init_from_env()
    host_group = dist.new_group(backend="gloo")  # host-side waits that must not park a kernel on the GPUs
    sc, name = make_scene(pkg, a.workload, a.spp)
    t0 = time.time()
    be = pkg.Backend(sc, local, bvh_build=BVH_KIND[a.bvh])
    build_s = time.time() - t0
    info = be.info()
    W, H = sc.camera.resolution
    film = torch.zeros(H * W * 4, dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    SPLITS = {"tile": pkg.abi.SPLIT_TILE, "sample": pkg.abi.SPLIT_SAMPLE}
    split = "sample" if a.split == "auto" else a.split

    def params_for(split_name, r=rank, w=world):
        q = sc.integrator.params(rank=r, world_size=w, shard_tile=32, split=SPLITS[split_name])
        q.mode = pkg.abi.MODE_WAVEFRONT if a.mode == "wavefront" else pkg.abi.MODE_MEGAKERNEL
        return q

    p = params_for(split)

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    def timed_steps(q, steps, warmup):
        """`steps` frames: render this rank's share + the one film collective; device time, max over ranks."""
        for _ in range(warmup):
            be.render_device(film.data_ptr(), q, stream)
            dmod.combine_film(film)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        acc = [0, 0, 0, 0]
        for _ in range(steps):
            st = be.render_device(film.data_ptr(), q, stream)
            dmod.combine_film(film)
            acc[0] += st.n_samples
            acc[1] += st.n_closest_rays
            acc[2] += st.n_shadow_rays
            acc[3] += st.n_kernel_launches + 2  # + the two memsets of akb_render_device
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        tot = torch.tensor(acc, dtype=torch.float64, device="cuda")
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        return float(ms.item()), [float(x) for x in tot.tolist()]

    clocks = ClockSampler(local)
    timed_steps(p, 0, a.warmup)
    if rank == 0:
        clocks.start()
    ms_total, (samples, n_c, n_s, launches) = timed_steps(p, a.steps, 0)
    clk = clocks.stop() if rank == 0 else None
    value = samples / (ms_total * 1e-3) / 1e6
    alt = None
    if a.split == "auto":  # the bit-identical interleaved-tile split, timed the same way
        ms_alt, tot_alt = timed_steps(params_for("tile"), a.steps, min(a.warmup, 3))
        alt = {"split": "tile", "value": tot_alt[0] / (ms_alt * 1e-3) / 1e6, "unit": "Msamples/s",
               "ms_per_step": ms_alt / a.steps,
               "note": "interleaved 32x32 tiles: the combined film is bit-identical to the 1-GPU film (checks.tile_split_film_bit_identical_to_1gpu)"}

    # ---- end to end: each rank renders its share on the device, ONE reduce, rank 0 copies the combined film to pinned
    # host memory, all inside the timed region.
    e2e_steps = max(1, min(a.steps, 3))
    host_film = torch.empty(H * W * 4, dtype=torch.float32, pin_memory=True)
    barrier()
    t = time.perf_counter()
    e2e_samples = 0
    for _ in range(e2e_steps):
        st = be.render_device(film.data_ptr(), p, stream)
        dmod.combine_film(film)
        if rank == 0:
            host_film.copy_(film, non_blocking=False)
        e2e_samples += st.n_samples
    barrier()
    e2e_s = torch.tensor([time.perf_counter() - t], dtype=torch.float64, device="cuda")
    e2e_n = torch.tensor([float(e2e_samples)], dtype=torch.float64, device="cuda")
    dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    dist.all_reduce(e2e_n, op=dist.ReduceOp.SUM)
    e2e_value = float(e2e_n.item()) / float(e2e_s.item()) / 1e6

    # ---- correctness on the NVLink hardware: the tile-split film after the reduce, bit for bit against a 1-GPU render
    checks = {}
    be.render_device(film.data_ptr(), params_for("tile"), stream)
    dmod.combine_film(film)
    torch.cuda.synchronize()
    if rank == 0:
        h_multi = hashlib.sha256(film.cpu().numpy().tobytes()).hexdigest()
        be.render_device(film.data_ptr(), params_for("tile", 0, 1), stream)
        torch.cuda.synchronize()
        h_single = hashlib.sha256(film.cpu().numpy().tobytes()).hexdigest()
        checks = {"tile_split_film_bit_identical_to_1gpu": h_multi == h_single, "film_sha256": h_single[:16]}
    dist.barrier(group=host_group)

    # ---- C4 (BASELINE: 10 M triangles, 4K, 256 spp, tile split over 8 GPUs) as a sub-record of the 8-GPU run
    subs = {}
    if world == 8 and a.workload == "c3" and not a.spp and not a.no_extras:
        be.close()
        be = None
        try:
            subs["c4"] = c4_sub_record(pkg, a, rank, world, local, torch, dist, dmod)
        except Exception as ex:
            subs["c4"] = {"error": repr(ex)}
        dist.barrier(group=host_group)

    # ---- the same frame from ONE process through the C ABI: rank 0 binds all N GPUs (akb_init(N)), the library runs
    # one host thread + stream per GPU and combines the film with one ncclReduce. The other ranks wait on the host
    # (gloo), their GPUs idle.
    single = None
    if be is not None:
        be.close()
    del film
    torch.cuda.synchronize()
    dist.barrier(group=host_group)
    if rank == 0 and not a.no_extras:
        try:
            single = single_process_multi_gpu(pkg, sc, world, a, checks.get("film_sha256"))
        except Exception as ex:
            single = {"error": repr(ex)}
    dist.barrier(group=host_group)

    if rank == 0:
        line = {"metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": a.steps,
                "warmup": a.warmup, "ms_per_step": ms_total / a.steps, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": name, "mode": a.mode,
                           "parallelism": "%s-split x%d, one process per GPU, scene replicated, one NCCL sum-reduce of the film per frame" % (split, world),
                           "tile": 32, "l2": "inputs larger than L2: BVH+triangles %.0f MB, shading records %.0f MB, path state + queues %.0f MB"
                           % ((info["bvh_bytes"] + info["triangle_bytes"]) / 1e6, info["n_triangles"] * 96 / 1e6, W * H * (168 + 48) / 1e6),
                           "bvh_nodes": info["n_bvh_nodes"], "scene_build_s": round(build_s, 2),
                           "bvh_build": {1: "host SAH", 2: "device PLOC"}.get(info["bvh_build"]), "bvh_build_s": round(info["build_us"] / 1e6, 3)},
                "clocks": clk,
                "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": C.sizeof(pkg.abi.RenderParams),
                        "d2h_bytes_per_step": W * H * 16, "api": "akb_render_device + film reduce + D2H on rank 0"},
                "gpu_launches": int(launches),
                "rays": {"closest_per_sample": n_c / samples, "shadow_per_sample": n_s / samples,
                         "mrays_per_s": (n_c + n_s) / (ms_total * 1e-3) / 1e6},
                "roofline": None, "cpu_baseline": None, "checks": checks}
        if alt:
            line["alt_split"] = alt
        if single:
            line["single_process"] = single
        if subs:
            line["configs"] = subs
        emit(line)
    dist.barrier(group=host_group)
    dist.destroy_process_group()


def c4_sub_record(pkg, a, rank, world, local, torch, dist, dmod):
    """C4 as BASELINE names it: 10 placed copies of the 1 M-triangle mesh, 3840x2160, 256 spp, interleaved 32x32 tiles
    over 8 GPUs (bit-identical to one GPU), device-built BVH; 1 warm-up + 2 timed frames."""
    sc, name = make_scene(pkg, "c4", 0)
    t0 = time.time()
    be = pkg.Backend(sc, local, bvh_build=BVH_KIND[a.bvh])
    build_s = time.time() - t0
    W, H = sc.camera.resolution
    film = torch.zeros(H * W * 4, dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    q = sc.integrator.params(rank=rank, world_size=world, shard_tile=32, split=pkg.abi.SPLIT_TILE)
    be.render_device(film.data_ptr(), q, stream)
    dmod.combine_film(film)
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    n = 0
    steps = 2
    for _ in range(steps):
        st = be.render_device(film.data_ptr(), q, stream)
        dmod.combine_film(film)
        n += st.n_samples
    e1.record()
    dist.barrier()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    tot = torch.tensor([float(n)], dtype=torch.float64, device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    info = be.info()
    be.close()
    return {"metric": "Msamples/s", "value": float(tot.item()) / (float(ms.item()) * 1e-3) / 1e6, "unit": "Msamples/s", "n_gpus": world,
            "steps": steps, "warmup": 1, "ms_per_step": float(ms.item()) / steps,
            "config": {"workload": name, "parallelism": "tile-split x%d (interleaved 32x32 tiles), one NCCL sum-reduce per frame" % world,
                       "bvh_nodes": info["n_bvh_nodes"], "scene_build_s": round(build_s, 2),
                       "bvh_build": {1: "host SAH", 2: "device PLOC"}.get(info["bvh_build"])},
            "cpu_baseline": None,
            "note": "no CPU baseline: from 5 meshes on the reference's own TLAS loses meshes (tests/test_round2_cpu.py), so its C4 frame is not this frame"}


def single_process_multi_gpu(pkg, sc, n, a, want_sha):
    """akb_init(n) + akb_render(n_gpus = n) from this one process: sample split timed like the torchrun path, then the
    tile split, whose film must be the 1-GPU film bit for bit."""
    W, H = sc.camera.resolution
    t0 = time.time()
    be = pkg.Backend(sc, bvh_build=BVH_KIND[a.bvh], n_devices=n)
    create_s = time.time() - t0
    info = be.info()
    out = {"api": "akb_init(%d) + akb_render(n_gpus = %d): one process, one host thread + stream per GPU in the library, one ncclReduce of the film" % (n, n),
           "scene_create_s": round(create_s, 2), "bvh_build_s": round(info["build_us"] / 1e6, 3),
           "replica_upload_s": round(info["upload_us"] / 1e6, 3)}
    import torch
    host_film_t = torch.zeros((H, W, 4), dtype=torch.float32, pin_memory=True)  # page-locked: device 0's film arrives by one DMA
    host_film = host_film_t.numpy()
    for split_name, split in (("sample", pkg.abi.SPLIT_SAMPLE), ("tile", pkg.abi.SPLIT_TILE)):
        p = sc.integrator.params(shard_tile=32, split=split)
        p.n_gpus = n
        for _ in range(2):
            be.render(p, out=host_film)
        steps = max(1, min(a.steps, 5))
        t = time.perf_counter()
        samples, dev_ms, comb_ms = 0, 0.0, 0.0
        for _ in range(steps):
            _, st = be.render(p, out=host_film)
            samples += st.n_samples
            dev_ms += st.ms_total
            comb_ms += st.ms_combine
        sec = time.perf_counter() - t
        out[split_name] = {"e2e_value": samples / sec / 1e6, "device_value": samples / (dev_ms * 1e-3) / 1e6, "unit": "Msamples/s",
                           "ms_per_step_wall": sec / steps * 1e3, "ms_device_slowest_gpu": dev_ms / steps, "ms_combine": comb_ms / steps}
        if split_name == "tile":
            sha = hashlib.sha256(host_film.tobytes()).hexdigest()[:16]
            out["tile_split_film_bit_identical_to_1gpu"] = (sha == want_sha) if want_sha else None
    be.close()
    return out


def bench_c2(pkg, a, local, torch, steps, warmup, headline):
    """C2 (BASELINE.json configs[1]): closest-hit / any-hit ray microbench against the 999 700-triangle
    mesh M1: AKB_C2_RAYS (default 100 M) coherent rays = pinhole camera (0,20,70) pitch -15 deg, fov 60, 12500 x 8000 raster
    in row-major order, and as many incoherent rays = cosine-hemisphere bounces off the primary hit points (misses:
    uniform origins in the mesh box, uniform directions). tmin = 1e-3, tmax = inf; any-hit variant tmax = 50.
    One step = the four launches (coherent/incoherent x closest/any). Rays are generated on the device with torch (bench
    input, not product code). value = closest-hit rays per second over the coherent + incoherent sets; e2e = akb_trace
    over HOST buffers of the full incoherent set."""
    import math
    n_rays = int(os.environ.get("AKB_C2_RAYS", "100000000"))
    RW = 12500
    RH = max(1, n_rays // RW)
    n = RW * RH
    sc = pkg.scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=1, max_depth=1, materials="c3")
    t0 = time.time()
    be = pkg.Backend(sc, local, bvh_build=BVH_KIND[a.bvh])
    build_s = time.time() - t0
    info = be.info()
    dev = "cuda"
    g = torch.Generator(device=dev)
    g.manual_seed(1234)
    cam = sc.camera
    c2w = torch.tensor(cam.c2w(), device=dev)
    s_ = math.atan(math.radians(cam.fov) / 2)
    ix = torch.arange(n, device=dev, dtype=torch.int64)
    px, py = (ix % RW).float() + 0.5, (ix // RW).float() + 0.5
    x = (px / RW * 2 - 1) * s_
    y = -(py / RH * 2 - 1) * s_ * RH / RW
    dl = torch.stack([x, y, -torch.ones_like(x)], 1)
    dl = dl / dl.norm(dim=1, keepdim=True)
    d = dl @ c2w[:3, :3].T
    org = torch.empty(n, 4, device=dev)
    org[:, :3] = c2w[:3, 3]
    org[:, 3] = 1e-3
    dr = torch.empty(n, 4, device=dev)
    dr[:, :3] = d
    dr[:, 3] = float("inf")
    del px, py, x, y, dl, d, ix
    hits = torch.empty(n * 5, dtype=torch.int32, device=dev)
    occ = torch.empty(n, dtype=torch.int32, device=dev)

    def trace(o, dd, any_hit):
        return be.trace_device(o.data_ptr(), dd.data_ptr(), n, any_hit, hits.data_ptr(), occ.data_ptr())

    trace(org, dr, 0)
    h = hits.view(n, 5)
    prim = h[:, 3].long()
    hit = prim >= 0
    t = h[:, 0].contiguous().view(torch.float32)
    mesh = sc.shapes[0]
    V = torch.from_numpy(mesh.vertices.reshape(-1, 3)).to(dev)
    I = torch.from_numpy(mesh.indices.reshape(-1, 3).astype(np.int64)).to(dev)
    pc = prim.clamp(min=0)
    v0, v1, v2 = V[I[pc, 0]], V[I[pc, 1]], V[I[pc, 2]]
    ng = torch.linalg.cross(v1 - v0, v2 - v0)
    ng = ng / ng.norm(dim=1, keepdim=True).clamp(min=1e-30)
    del v0, v1, v2, pc
    p = org[:, :3] + dr[:, :3] * torch.where(hit, t, torch.zeros_like(t))[:, None] + 1e-3 * ng
    u = torch.rand(n, 2, device=dev, generator=g)
    r, phi = u[:, 0].sqrt(), 2 * math.pi * u[:, 1]
    lx, lz, ly = r * phi.cos(), r * phi.sin(), (1 - u[:, 0]).clamp(min=0).sqrt()
    tang = torch.linalg.cross(ng, torch.tensor([1.0, 0.0, 0.0], device=dev).expand_as(ng))
    tang = tang / tang.norm(dim=1, keepdim=True).clamp(min=1e-30)
    bit = torch.linalg.cross(ng, tang)
    d2 = lx[:, None] * tang + ly[:, None] * ng + lz[:, None] * bit
    lo, hi = V.min(0).values, V.max(0).values
    rnd_o = lo + (hi - lo) * torch.rand(n, 3, device=dev, generator=g)
    rnd_d = torch.randn(n, 3, device=dev, generator=g)
    rnd_d = rnd_d / rnd_d.norm(dim=1, keepdim=True)
    org2 = torch.empty(n, 4, device=dev)
    org2[:, :3] = torch.where(hit[:, None], p, rnd_o)
    org2[:, 3] = 1e-3
    dr2 = torch.empty(n, 4, device=dev)
    dr2[:, :3] = torch.where(hit[:, None], d2, rnd_d)
    dr2[:, 3] = float("inf")
    del p, u, r, phi, lx, ly, lz, tang, bit, d2, rnd_o, rnd_d, ng, t, prim
    dr_any, dr2_any = dr.clone(), dr2.clone()
    dr_any[:, 3] = 50.0
    dr2_any[:, 3] = 50.0
    hit_frac = float(hit.float().mean().item())
    sets = (("coherent_closest", org, dr, 0), ("incoherent_closest", org2, dr2, 0), ("coherent_any", org, dr_any, 1),
            ("incoherent_any", org2, dr2_any, 1))
    for _ in range(max(1, warmup)):
        for _, o, dd, ah in sets:
            trace(o, dd, ah)
    clocks = ClockSampler(local)
    clocks.start()
    ms = {k: 0.0 for k, *_ in sets}
    for _ in range(steps):
        for k, o, dd, ah in sets:
            ms[k] += trace(o, dd, ah)  # device time of the traversal launch (CUDA events on the launch stream)
    clk = clocks.stop()
    mrays = {k: n * steps / (v * 1e-3) / 1e6 for k, v in ms.items()}
    closest_ms = ms["coherent_closest"] + ms["incoherent_closest"]
    value = 2 * n * steps / (closest_ms * 1e-3) / 1e6
    peak, peak_src = measured_peaks()
    achieved = BYTES_CLOSEST * 2 * n * steps / (closest_ms * 1e-3) / 1e9
    # ---- end to end: akb_trace with HOST ray and hit buffers (H2D + kernel + D2H inside the timed region), the FULL
    # incoherent set
    m = n
    # page-locked host buffers (the contract's "inputs in pinned host memory"): akb_trace then copies straight from / into
    # them, chunk by chunk on two streams; pageable buffers would go through the library's own pinned staging slots
    ho_t, hd_t = org2.cpu().pin_memory(), dr2.cpu().pin_memory()
    hh_t = torch.zeros(m * 5, dtype=torch.int32, pin_memory=True)
    ho, hd, hh = ho_t.numpy(), hd_t.numpy(), hh_t.numpy().view(pkg.HIT_DTYPE)
    be.trace(ho[:1 << 22], hd[:1 << 22], out=hh[:1 << 22])  # warm-up: the library sizes and pins its staging slots on the first call
    e2e_s = float("inf")
    for _ in range(2):
        t0 = time.perf_counter()
        be.trace(ho, hd, out=hh)
        e2e_s = min(e2e_s, time.perf_counter() - t0)
    # ---- CPU baseline: the reference's own BVH (Scene::intersect) over a bounded sample of the incoherent set
    cpu = None
    if not a.no_cpu_baseline:
        try:
            from oracle_libs import Oracle, Reference, have_ref
            sc_ref = pkg.scenes.terrain_scene(n=708, resolution=(1920, 1080), spp=1, max_depth=1, materials="c3ref")  # same mesh
            impl, kind = (Reference(sc_ref), "reference") if have_ref() else (Oracle(sc_ref), "port")
            k = min(m, 16 * 1024 * 1024)
            t0 = time.perf_counter()
            impl.trace(ho[:k], hd[:k])
            cpu_s = time.perf_counter() - t0
            cpu = {"value": k / cpu_s / 1e6, "unit": "Mrays/s", "cores": os.cpu_count(), "threads": os.cpu_count(), "kind": kind,
                   "sample": "first %d incoherent closest-hit rays of the same set (Scene::intersect, the reference's own BVH), 4096-ray chunks over all %d host threads" % (k, os.cpu_count())}
            impl.close()
        except Exception as ex:
            cpu = {"value": None, "unit": "Mrays/s", "cores": os.cpu_count(), "kind": "port", "sample": "failed: %r" % (ex,)}
    line = {"metric": "closest-hit Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": 1, "steps": steps, "warmup": warmup,
            "ms_per_step": sum(ms.values()) / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "C2: %d coherent + %d incoherent rays vs the 999700-triangle mesh M1 (closest hit; any hit with tmax 50 beside it)" % (n, n),
                       "l2": "inputs larger than L2: %.1f GB of rays + %.1f GB of hits per set" % (n * 32 / 1e9, n * 20 / 1e9),
                       "bvh_nodes": info["n_bvh_nodes"], "scene_build_s": round(build_s, 2),
                       "bvh_build": {1: "host SAH", 2: "device PLOC"}.get(info["bvh_build"]), "bvh_build_s": round(info["build_us"] / 1e6, 3),
                       "primary_hit_fraction": hit_frac},
            "clocks": clk,
            "e2e": {"value": m / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": m * 32, "d2h_bytes_per_step": m * 20,
                    "api": "akb_trace (page-locked host ray + hit buffers), all %d incoherent closest-hit rays" % m},
            "gpu_launches": 4 * steps, "mrays_per_s": mrays,
            "roofline": {"bound": "hbm", "kernel": "trace_batch_kernel (closest hit)", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                         "bytes_per_launch": BYTES_CLOSEST * n, "avg_launch_ms": closest_ms / (2 * steps),
                         "note": "52 B per closest-hit ray (32 B ray + 20 B hit); traversal is issue/latency-bound, see profiles/"},
            "cpu_baseline": cpu}
    be.close()
    if not headline:
        for k in ("n_gpus", "higher_is_better", "scaling", "vs_baseline", "dtype", "data"):
            line.pop(k, None)
    return line


def ray_microbench(pkg, be, torch):
    """C2-shaped closest-hit / any-hit microbench on the resident scene: coherent (camera) and incoherent (cosine bounce)
    ray sets generated on the fly, rays resident in HBM, device time of the traversal kernel."""
    W, H = be.width, be.height
    ys, xs = np.mgrid[0:H, 0:W]
    xy = np.stack([xs.ravel(), ys.ravel()], -1).astype(np.int32)
    rng = np.random.default_rng(1234)
    reps = 4
    orgs, dirs = [], []
    for _ in range(reps):
        o, d = be.generate_rays(xy, rng.random((xy.shape[0], 2), dtype=np.float32))
        orgs.append(o), dirs.append(d)
    org, dr = np.concatenate(orgs), np.concatenate(dirs)
    n = org.shape[0]
    d_o, d_d = torch.from_numpy(org).cuda(), torch.from_numpy(dr).cuda()
    d_h = torch.empty(n * 5, dtype=torch.int32, device="cuda")
    d_occ = torch.empty(n, dtype=torch.int32, device="cuda")
    out = {}

    def run(o, d, any_hit):
        best = None
        for _ in range(4):
            ms = be.trace_device(o.data_ptr(), d.data_ptr(), n, any_hit, d_h.data_ptr(), d_occ.data_ptr())
            best = ms if best is None else min(best, ms)
        return n / best / 1e3
    out["closest_hit_coherent_mrays_per_s"] = run(d_o, d_d, 0)
    hits = d_h.cpu().numpy().view(pkg.HIT_DTYPE)
    hit = hits["prim_id"] >= 0
    p = org[:, :3] + dr[:, :3] * np.where(hit, hits["t"], 10.0)[:, None].astype(np.float32)
    dd = rng.normal(size=(n, 3)).astype(np.float32)
    dd /= np.linalg.norm(dd, axis=1, keepdims=True)
    dd[:, 1] = np.abs(dd[:, 1])
    o2 = np.concatenate([p + np.array([0, 1e-2, 0], np.float32), np.full((n, 1), 1e-3, np.float32)], 1).astype(np.float32)
    d2 = np.concatenate([dd, np.full((n, 1), np.inf, np.float32)], 1).astype(np.float32)
    d_o2, d_d2 = torch.from_numpy(o2).cuda(), torch.from_numpy(d2).cuda()
    out["closest_hit_incoherent_mrays_per_s"] = run(d_o2, d_d2, 0)
    d2[:, 3] = 50.0
    d_d2 = torch.from_numpy(d2).cuda()
    out["any_hit_incoherent_mrays_per_s"] = run(d_o2, d_d2, 1)
    out["ray_microbench"] = "%d rays per set vs the workload's mesh, rays resident in HBM, best of 4 launches" % n
    return out


if __name__ == "__main__":
    main()
