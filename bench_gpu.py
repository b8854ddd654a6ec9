"""GPU arm of bench.py (kept separate so `bench.py --impl reference` never imports torch.cuda)."""
import ctypes as C
import json
import os
import sys

import numpy as np

from bench import (fluxmap_l2_vs_ref, TRAFFIC_BYTES_PER_RAY_NCU, ALGO_BYTES_PER_HIT_HIT_KERNEL, ALGO_BYTES_PER_RAY_HIT_KERNEL, FLUX_RES, METRIC, REPS, UNIT, ClockSampler, cpu_baseline,
                   nsttf_source_geometry)
from tests import scenes

ROOT = os.path.dirname(os.path.abspath(__file__))


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def run_gpu(args):
    import torch
    import torch.distributed as dist
    from tracer_b200 import _lib as L
    from tracer_b200.fluxmap import receiver_energy_bins
    from tracer_b200.tracer_engine_mp import TracerEngineMP

    # stdout carries exactly one JSON line: anything libraries print meanwhile (NCCL's version banner
    # under NCCL_DEBUG, warnings) goes to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = L.load()
    T = scenes.NS("tracer_b200")
    asm, _ = scenes.c3_nsttf(T, n_rays=10)
    surfaces = asm.get_surfaces()
    recv = surfaces[-1]
    eng = TracerEngineMP(asm, rng="philox", seed=1234 + rank, device_accel=args.accel, leaf_size=args.leaf_size)
    centre, direction, radius = nsttf_source_geometry()
    n = args.rays
    src = T.sources.buie_sunshape_device(n, centre, direction, radius, 0.01, flux=1000., pre_process_CSR=False,
                                         seed=99, first_ray=rank * n, total_rays=n * world, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(bundle, read_back=False):
        asm.reset_all_optics()
        eng.ray_tracer(bundle, reps=REPS, min_energy=1e-10, tree=True)
        bins = receiver_energy_bins(recv, FLUX_RES, device=dev)
        absorbed, hits, bins = eng.reduce_tallies(bins)
        if read_back:
            return bins.cpu().numpy(), eng.stats["ray_bounces"]
        return bins, eng.stats["ray_bounces"]

    sampler = ClockSampler(local)
    sampler.start()                  # nvidia-smi needs ~1 s to start; it samples both timed regions (value and e2e)

    def timed(bundle, steps, warmup, read_back=False, after_warmup=None):
        for _ in range(warmup):
            step(bundle, read_back)
        if after_warmup:
            after_warmup()
        barrier()
        l0 = lib.trc_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        bounces, last = 0, None
        for _ in range(steps):
            last, b = step(bundle, read_back)
            bounces += b
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        tot = torch.tensor([float(bounces)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        return float(ms.item()), float(tot.item()), lib.trc_launch_count() - l0, last

    kernel_events = []

    def start_events():
        eng.kernel_events = kernel_events
    ms, bounces, launches, bins = timed(src, args.steps, args.warmup, after_warmup=start_events)
    eng.kernel_events = None
    value = bounces / (ms * 1e-3)
    per_bounce = eng.stats["per_bounce"]
    tests_equiv = sum(nb[0] for nb in per_bounce) * len(surfaces) * world * args.steps / (ms * 1e-3)

    # dominant kernel: trc_trace_hits of bounce 1 (nearest hit over the whole source bundle +
    # emission of the hit records), CUDA events recorded around it inside the timed region
    ev = [(e[1], e[2], e[3]()) for e in kernel_events if e[0] == 0]
    k_ms = float(np.mean([t for _, _, t in ev]))
    k_hits = float(np.mean([h for _, h, _ in ev]))
    algo_bytes = ALGO_BYTES_PER_RAY_HIT_KERNEL * n + ALGO_BYTES_PER_HIT_HIT_KERNEL * k_hits
    peak, peak_src = _peaks()
    achieved = algo_bytes / (k_ms * 1e-3) / 1e9
    V, D = src._props["vertices"].pieces[0], src._props["directions"].pieces[0]
    e2e = None
    if not args.no_e2e:
        hV = torch.empty((3, n), dtype=torch.float64, pin_memory=True)
        hD = torch.empty((3, n), dtype=torch.float64, pin_memory=True)
        hE = torch.empty(n, dtype=torch.float64, pin_memory=True)
        hV.copy_(V); hD.copy_(D); hE.copy_(src._props["energy"].pieces[0])
        host = T.ray_bundle.RayBundle(vertices=hV.numpy(), directions=hD.numpy(), energy=hE.numpy())
        e_steps = max(1, min(args.steps, 3))
        e_ms, e_b, _, _ = timed(host, e_steps, 1, read_back=True)
        e2e = {"value": e_b / (e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 56 * n * world,
               "d2h_bytes_per_step": FLUX_RES * FLUX_RES * 8 * world, "ms_per_step": e_ms / e_steps}
        del hV, hD, hE, host

    clocks = sampler.stop()
    cpu, flux_cmp = None, None
    if rank == 0 and not args.no_cpu:
        procs = max(1, min(os.cpu_count() or 1, 64))
        v, b, wall = cpu_baseline(args.cpu_rays, procs)
        cpu = {"value": v, "unit": UNIT, "cores": procs, "kind": "port",
               "sample": "%d processes x %d rays of the same NSTTF scene, reps=%d, %.1f s wall" % (procs, args.cpu_rays, REPS, wall)}
        # the all-reduced GPU map is the map of ONE source sharded over the ranks (energy per ray = flux x area /
        # (rays x world)), the same total power as each CPU process traces
        gpu_map = bins.cpu().numpy() if hasattr(bins, "cpu") else np.asarray(bins)
        flux_cmp = fluxmap_l2_vs_ref(gpu_map, cpu_baseline.maps)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C3 NSTTF heliostat field (218 heliostats + receiver), Buie CSR 0.01, reps=3",
                       "rays_per_gpu_per_step": n, "surfaces": len(surfaces), "accel": args.accel,
                       "rng": "philox", "tree": True, "flux_map": "%dx%d on device" % (FLUX_RES, FLUX_RES),
                       "l2": "inputs (5.6 GB/GPU at 1e8 rays) larger than L2",
                       "per_bounce_rays_in_hits_survivors": per_bounce},
            "ray_surface_tests_per_s_brute_equivalent": tests_equiv,
            "receiver_power_W": float(bins.sum().item()) if hasattr(bins, "sum") else None,
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": TRAFFIC_BYTES_PER_RAY_NCU * n if args.accel == "bvh" else None,
                         "kernel": "nearest_hit_walk_kernel (trc_trace_hits, bounce 1)" if args.accel == "bvh"
                         else "nearest_hit_brute_kernel (trc_trace_hits, bounce 1)",
                         "kernel_ms": k_ms, "launches_timed": len(ev), "algorithmic_bytes": algo_bytes,
                         "bytes_per_ray": ALGO_BYTES_PER_RAY_HIT_KERNEL, "bytes_per_hit": ALGO_BYTES_PER_HIT_HIT_KERNEL,
                         "hits": k_hits, "peak_source": peak_src + ", burst (kernel timed by its own events)",
                         "traffic_source": "ncu dram__bytes_read+write per ray of profiles/ (2e7-ray capture) x rays"},
            "cpu_baseline": cpu, "flux_map_vs_ref": flux_cmp,
        }
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)
    if world > 1:
        dist.destroy_process_group()
