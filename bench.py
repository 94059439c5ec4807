#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 Path-integrator backend (contract: see DESIGN.md §Measurement).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload c3|c1|c5] [--spp S]

One "step" = one full frame of the workload rendered by the hot path. Default workload C3 (BASELINE.json configs[2],
the configuration the Msamples/s metric is quoted on): synthetic 999 700-triangle terrain, 1920x1080, 64 spp, max depth
8, 60 % Diffuse / 30 % GGX / 10 % Dielectric, LCG sampler. N > 1 (torchrun): interleaved 32x32 tiles, scene replicated,
one NCCL sum-reduce of the film per step (strong scaling: the frame is fixed).

`--impl reference` times the reference's own CPU implementation (oracle/_ref, built from /root/reference; falls back to
the oracle port where that library is absent) on a bounded sample of the same workload, on all host threads.
"""
import argparse
import ctypes as C
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


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return HBM_FALLBACK_GBS, "fallback"


def make_scene(pkg, workload, spp, ref_subset=False):
    s = pkg.scenes
    if workload == "c3":
        sc = s.terrain_scene(n=708, resolution=(1920, 1080), spp=spp or 64, max_depth=8,
                             materials="c3ref" if ref_subset else "c3")
        name = "C3: synthetic 999700-triangle terrain, 1920x1080, %d spp, Path max_depth 8, 60%% Diffuse / 30%% GGX(0.3) / 10%% Dielectric(1.5), LCG sampler" % sc.integrator.spp
    elif workload == "c1":
        sc = s.cornell_box((512, 512), spp or 16, 5)
        name = "C1: Cornell box 512x512, %d spp, Path max_depth 5" % sc.integrator.spp
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
    """Times the reference's CPU path tracer (own BVH, all host threads) on `rows` image rows of the workload."""
    from oracle_libs import Oracle, Reference, have_ref
    sc, name = make_scene(pkg, workload, spp, ref_subset=True)
    W, H = sc.camera.resolution
    rows = min(rows, H)
    y0 = max(0, (H * 5) // 8 - rows // 2)  # a band through the terrain, not the sky
    rect = (0, y0, W, y0 + rows)
    t0 = time.time()
    if have_ref():
        impl, kind = Reference(sc), "reference"
    else:
        impl, kind = Oracle(sc), "port"
    build_s = time.time() - t0
    p = sc.integrator.params()
    cores = os.cpu_count()
    times = []
    for i in range(warmup + steps):
        t = time.perf_counter()
        impl.render(p, rect=rect, threads=cores)
        if i >= warmup:
            times.append(time.perf_counter() - t)
    n_samples = W * rows * sc.integrator.spp
    sec = float(np.mean(times))
    sample = ("%s; CPU sample = image rows %d..%d (%d x %d px x %d spp = %d samples per step)%s; BVH build %.1f s not included"
              % (name, y0, y0 + rows, W, rows, sc.integrator.spp, n_samples,
                 "; reference has no Dielectric BSDF: those 10% of triangles are shaded as Diffuse" if workload == "c3" else "",
                 build_s))
    return {"value": n_samples / sec / 1e6, "unit": "Msamples/s", "cores": cores, "kind": kind, "sample": sample, "workload": name,
            "seconds_per_step": sec, "samples_per_step": n_samples}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=["c3", "c1", "c5"])
    ap.add_argument("--spp", type=int, default=0)
    ap.add_argument("--mode", default="wavefront", choices=["wavefront", "megakernel"])
    ap.add_argument("--cpu-rows", type=int, default=24)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    pkg = importlib.import_module("akarirender-1_b200")

    if a.impl == "reference":
        if rank != 0:
            return
        r = cpu_reference_run(pkg, a.workload, a.spp, a.cpu_rows, steps=max(1, a.steps), warmup=min(a.warmup, 1))
        line = {"impl": "reference", "metric": "Msamples/s", "value": r["value"], "unit": "Msamples/s", "n_gpus": a.gpus,
                "steps": a.steps, "warmup": a.warmup, "ms_per_step": r["seconds_per_step"] * 1e3, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": r["workload"], "mode": "reference CPU path (own BVH), bounded sample: see cpu_baseline.sample"},
                "cpu_baseline": {"value": r["value"], "unit": "Msamples/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
                "e2e": {"value": r["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line), flush=True)
        return

    import torch
    import torch.distributed as dist
    dmod = importlib.import_module("akarirender-1_b200.distributed")
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        dmod.init_from_env()
    torch.cuda.set_device(local)
    sc, name = make_scene(pkg, a.workload, a.spp)
    t0 = time.time()
    be = pkg.Backend(sc, local)
    build_s = time.time() - t0
    info = be.info()
    W, H = sc.camera.resolution
    film = torch.zeros(H * W * 4, dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    p = sc.integrator.params(rank=rank, world_size=world, shard_tile=32)
    p.mode = pkg.abi.MODE_WAVEFRONT if a.mode == "wavefront" else pkg.abi.MODE_MEGAKERNEL

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        st = be.render_device(film.data_ptr(), p, stream)
        dmod.combine_film(film)
        return st

    for _ in range(a.warmup):
        step()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    launches = samples = n_c = n_s = 0
    for _ in range(a.steps):
        st = step()
        launches += st.n_kernel_launches + 2  # + the two memsets of akb_render_device
        samples += st.n_samples
        n_c += st.n_closest_rays
        n_s += st.n_shadow_rays
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    tot = torch.tensor([samples, n_c, n_s, launches], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    clk = clocks.stop() if rank == 0 else None
    ms_total = float(ms.item())
    samples, n_c, n_s, launches = (float(x) for x in tot.tolist())
    value = samples / (ms_total * 1e-3) / 1e6

    # ---- end to end through the host-buffer C-ABI call a reference maintainer would bind (akb_render): every step
    # passes the parameters from host memory and reads the film back to a host buffer inside the timed region.
    host_film = np.zeros((H, W, 4), np.float32)
    e2e_steps = max(1, min(a.steps, 3))
    be.render(p, out=host_film)
    barrier()
    t = time.perf_counter()
    e2e_samples = 0
    for _ in range(e2e_steps):
        _, st = be.render(p, out=host_film)
        e2e_samples += st.n_samples
    torch.cuda.synchronize()
    e2e_s = torch.tensor([time.perf_counter() - t], dtype=torch.float64, device="cuda")
    e2e_n = torch.tensor([float(e2e_samples)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
        dist.all_reduce(e2e_n, op=dist.ReduceOp.SUM)
    e2e_value = float(e2e_n.item()) / float(e2e_s.item()) / 1e6

    line = None
    if rank == 0:
        peak, peak_src = measured_peaks()
        # ---- roofline of the dominant kernel (wf_trace): per-launch device time measured live with CUDA events on the
        # launch stream (AKB_PROFILE_STAGES pass, outside the timed region), algorithmic bytes = 52 B per closest-hit
        # ray + 36 B per any-hit ray.
        roof = None
        if a.mode == "wavefront" and world == 1:
            os.environ["AKB_PROFILE_STAGES"] = "1"
            pkg._initialised = None
            pkg.init(local)
            stp = be.render_device(film.data_ptr(), p, stream)
            os.environ["AKB_PROFILE_STAGES"] = "0"
            pkg._initialised = None
            pkg.init(local)
            n_launch = max(1, (stp.n_kernel_launches - 2) // 2)
            bytes_total = BYTES_CLOSEST * stp.n_closest_rays + BYTES_ANY * stp.n_shadow_rays
            achieved = bytes_total / (stp.ms_extend * 1e-3) / 1e9 if stp.ms_extend > 0 else None
            traffic = None
            try:
                with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                    traffic = json.load(f).get("wf_trace_dram_bytes_per_launch")
            except Exception:
                pass
            roof = {"bound": "hbm", "kernel": "wf_trace_kernel (closest-hit + any-hit CWBVH traversal)",
                    "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": (achieved / peak) if achieved else None,
                    "traffic": traffic, "peak_source": peak_src, "launches": n_launch,
                    "avg_launch_ms": stp.ms_extend / n_launch, "bytes_per_launch": bytes_total / n_launch,
                    "kernel_share_of_step": stp.ms_extend / stp.ms_total, "shade_share_of_step": stp.ms_shade / stp.ms_total,
                    "mrays_per_s_in_kernel": (stp.n_closest_rays + stp.n_shadow_rays) / (stp.ms_extend * 1e-3) / 1e6,
                    "note": "BVH traversal is latency/issue/L2-bound, not HBM-streaming-bound; the HBM fraction is low by "
                            "construction (SURVEY.md §8d) - see profiles/ for issue-slot and L2 figures"}
        extras = {}
        if not a.no_extras and world == 1:
            extras = ray_microbench(pkg, be, torch)
        cpu = None
        if not a.no_cpu_baseline and world == 1:
            try:
                cpu = cpu_reference_run(pkg, a.workload, a.spp, a.cpu_rows)
                cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}
            except Exception as ex:  # the checker is optional for the product; report why
                cpu = {"value": None, "unit": "Msamples/s", "cores": os.cpu_count(), "kind": "port", "sample": "failed: %r" % (ex,)}
        line = {"metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": a.steps,
                "warmup": a.warmup, "ms_per_step": ms_total / a.steps, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": name, "mode": a.mode, "parallelism": "tile%d" % world if world > 1 else "single",
                           "tile": 32, "l2": "inputs larger than L2: BVH+triangles %.0f MB, shading records %.0f MB, path state %.0f MB"
                           % ((info["bvh_bytes"] + info["triangle_bytes"]) / 1e6, info["n_triangles"] * 96 / 1e6, W * H * 208 / 1e6),
                           "bvh_nodes": info["n_bvh_nodes"], "scene_build_s": round(build_s, 2)},
                "clocks": clk,
                "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": C.sizeof(pkg.abi.RenderParams),
                        "d2h_bytes_per_step": W * H * 16, "api": "akb_render (host film buffer)"},
                "gpu_launches": int(launches),
                "rays": {"closest_per_sample": n_c / samples, "shadow_per_sample": n_s / samples,
                         "mrays_per_s": (n_c + n_s) / (ms_total * 1e-3) / 1e6},
                "roofline": roof, "cpu_baseline": cpu}
        line.update(extras)
        print(json.dumps(line), flush=True)
    be.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


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
