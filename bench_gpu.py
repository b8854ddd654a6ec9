"""GPU arm of bench.py (kept separate so `bench.py --impl reference` never imports torch.cuda)."""
import ctypes as C
import hashlib
import json
import os
import sys

import numpy as np

from bench import (fluxmap_l2_vs_ref, ALGO_BYTES_PER_HIT_HIT_KERNEL, ALGO_BYTES_PER_RAY_HIT_KERNEL, FLUX_RES, METRIC,
                   REPS, UNIT, WORKLOAD, ClockSampler, cpu_baseline, nsttf_source_geometry)
from tracer_b200.models import scenes

ROOT = os.path.dirname(os.path.abspath(__file__))
WALK_NCU_JSON = os.path.join(ROOT, "profiles", "walk_ncu.json")
WALK_SOURCES = ("tracer_b200/csrc/nearest.cu", "tracer_b200/csrc/geom.cuh", "tracer_b200/csrc/common.cuh",
                "tracer_b200/csrc/shadow_grid.cu")


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def walk_source_hash():
    h = hashlib.sha1()
    for rel in WALK_SOURCES:
        with open(os.path.join(ROOT, rel), "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def walk_ncu_record():
    """DRAM traffic and warp instructions per ray of the tree-walk kernel from the committed ncu capture
    (scripts/ncu_to_json.py).  Only trusted while the kernel's sources are the ones that were profiled."""
    if not os.path.exists(WALK_NCU_JSON):
        return None, "no profiles/walk_ncu.json"
    rec = json.load(open(WALK_NCU_JSON))
    if rec.get("source_sha1") != walk_source_hash():
        return None, "profiles/walk_ncu.json is stale (kernel sources changed since the capture)"
    return rec, "profiles/walk_ncu.json (%s)" % rec.get("capture", "?")


def _event_ms(fn, torch, reps=5, warm=2):
    """Best-of device time of fn() by CUDA events (inputs resident)."""
    best = None
    for it in range(warm + reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        if it >= warm:
            ms = e0.elapsed_time(e1)
            best = ms if best is None else min(best, ms)
    return best


def config_rates(T, torch, dev):
    """Short device-timed rates of the other BASELINE configurations at their named sizes (device-resident
    Philox sources, CUDA events around ray_tracer, best of 5 after 2 warm-up traces)."""
    from tracer_b200.tracer_engine import TracerEngine
    out = {}

    def rate(asm, src, reps, accel, min_energy=1e-10, tree=True, n_timed=5):
        eng = TracerEngine(asm, rng="philox", seed=1, device_accel=accel)

        def go():
            asm.reset_all_optics()
            eng.ray_tracer(src, reps=reps, min_energy=min_energy, tree=tree)
        ms = _event_ms(go, torch, reps=n_timed)
        st = eng.stats
        return {"ms": ms, "ray_bounces": st["ray_bounces"], "value": st["ray_bounces"] / (ms * 1e-3), "unit": UNIT,
                "surfaces": st["surfaces"], "per_bounce_rays_in_hits_survivors": st["per_bounce"]}

    down = np.r_[0., 0., -1.]
    asm, _ = scenes.c1_flat_mirror(T, n_rays=10)
    src = T.sources.disk_bundle_device(100000, np.c_[[0., 0., 4.]], down, 0.8, 4.65e-3, flux=1000., seed=1)
    out["C1 flat mirror + receiver, 1e5 pillbox rays, reps=3"] = rate(asm, src, 3, "auto")
    asm, _ = scenes.c2_dish(T, n_rays=10)
    src = T.sources.buie_sunshape_device(10000000, np.c_[[0., 0., 10.]], down, 2.5, 0.05, flux=1000., seed=2)
    out["C2 dish, 1e7 Buie rays, slope-error optics, reps=3"] = rate(asm, src, 3, "auto")
    del src
    # C3 dense: the source disc only covers the middle of the field, so most rays hit a heliostat
    # (examples/Sandia_NSTTF_field example.py:67-77 aims one disc per heliostat for the same purpose)
    asm, _ = scenes.c3_nsttf(T, n_rays=10)
    centre, direction, radius = nsttf_source_geometry()
    src = T.sources.buie_sunshape_device(20000000, centre, direction, 0.25 * radius, 0.01, flux=1000.,
                                         pre_process_CSR=False, seed=3)
    out["C3 dense (source disc over the field centre), 2e7 Buie rays, reps=3"] = rate(asm, src, 3, "bvh")
    del src
    torch.cuda.empty_cache()
    # C4: 1 001 000-face cavity mesh, Lambertian 0.7, reps=5, min_energy = 1e-3 of a ray's energy
    asm, _ = scenes.c4_cavity_mesh(T, n_theta=1000, n_z=500)
    n4 = 20000000
    src = T.sources.disk_bundle_device(n4, np.c_[[0., 0., -0.01]], np.r_[0., 0., 1.], 0.9, np.pi / 4., flux=1000., seed=5)
    e_ray = float(src._props["energy"].pieces[0][0].item())
    r4 = rate(asm, src, 5, "bvh", min_energy=1e-3 * e_ray, tree=False, n_timed=2)
    r4["note"] = "2e7-ray slices of the 1e8-ray config (five make the named size)"
    out["C4 triangle-mesh cavity, 1 001 000 faces, Lambertian, reps=5"] = r4
    del src, asm
    torch.cuda.empty_cache()
    # C5: one pass of the view-factor driver on the Holman cavity (5 elements x 1e5 rays, one bounce each)
    asm, elems, areas, emit = scenes.c5_holman(T)
    eng = TracerEngine(asm, rng="philox", seed=4)
    bundles = [emit(i, 100000, seed=10 + i) for i in range(len(areas))]
    dev_bundles = []
    for b in bundles:                       # emitters of the bands draw on the host: make them resident first
        if isinstance(b._props["vertices"], np.ndarray):
            cols = {k: T.ray_bundle._DeviceColumns([torch.from_numpy(np.ascontiguousarray(b._props[k])).to(dev)])
                    for k in ("vertices", "directions", "energy")}
            b = T.ray_bundle.RayBundle(**cols)
        dev_bundles.append(b)
    rows = []

    def passes():
        del rows[:]
        for b in dev_bundles:
            asm.reset_all_optics()
            eng.ray_tracer(b, reps=1, min_energy=1e-15, tree=False)
            rows.append(eng.absorbed_per_surface.clone())
    ms = _event_ms(passes, torch)
    F = torch.stack(rows).cpu().numpy()
    exact, _ = scenes.holman_exact()
    out["C5 Holman cavity view factors, 5 elements x 1e5 rays, one bounce"] = {
        "ms": ms, "ray_bounces": 5 * 100000, "value": 5e5 / (ms * 1e-3), "unit": UNIT, "surfaces": 5,
        "max_abs_error_vs_closed_form": float(np.abs(F - exact).max())}
    return out


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
    eng = TracerEngineMP(asm, rng="philox", seed=1234, device_accel=args.accel, leaf_size=args.leaf_size)
    centre, direction, radius = nsttf_source_geometry()
    n_total = args.rays
    n = -(-n_total // world)                 # this rank's shard of the one source (strong scaling)
    n = -(-n // 256) * 256 if world > 1 else n
    n_total = n * world

    def source(rays, first, total):
        return T.sources.buie_sunshape_device(rays, centre, direction, radius, 0.01, flux=1000., pre_process_CSR=False,
                                              seed=99, first_ray=first, total_rays=total, device=dev)
    src = source(n, rank * n, n_total)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(bundle, read_back=False):
        asm.reset_all_optics()
        eng.ray_tracer(bundle, reps=REPS, min_energy=1e-10, tree=True)
        bins = receiver_energy_bins(recv, FLUX_RES, device=dev)
        # the all-reduce of the step's tallies and flux map is enqueued on the process group's stream; this rank's
        # next trace does not wait for the slowest rank (every collective is waited for before the timed region ends)
        pend = eng.reduce_tallies(bins, async_op=True)
        if read_back:
            return pend.result()[2].cpu().numpy(), eng.stats["ray_bounces"]
        return pend, eng.stats["ray_bounces"]

    sampler = ClockSampler(local)
    sampler.start()                  # nvidia-smi needs ~1 s to start; it samples every timed region

    def timed(bundle, steps, warmup, read_back=False, after_warmup=None):
        for _ in range(warmup):
            w, _ = step(bundle, read_back)
            if not read_back:
                w.result()
        if after_warmup:
            after_warmup()
        barrier()
        l0 = lib.trc_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        bounces, last, pending = 0, None, []
        for _ in range(steps):
            last, b = step(bundle, read_back)
            bounces += b
            if not read_back:
                pending.append(last)
        for pend in pending:
            last = pend.result()[2]             # the stream waits for every step's collective: inside the timed region
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

    # dominant kernel: the nearest-hit pass of bounce 1 over this rank's shard (+ emission of the hit records),
    # CUDA events recorded around it inside the timed region
    ev = [(e[1], e[2], e[3]()) for e in kernel_events if e[0] == 0]
    k_ms = float(np.mean([t for _, _, t in ev]))
    k_hits = float(np.mean([h for _, h, _ in ev]))
    algo_bytes = ALGO_BYTES_PER_RAY_HIT_KERNEL * n + ALGO_BYTES_PER_HIT_HIT_KERNEL * k_hits
    peak, peak_src = _peaks()
    achieved = algo_bytes / (k_ms * 1e-3) / 1e9
    ncu, ncu_src = walk_ncu_record() if args.accel == "bvh" else (None, "brute-force kernel: no capture")
    V, D = src._props["vertices"].pieces[0], src._props["directions"].pieces[0]

    e2e = None
    if not args.no_e2e:
        hV = torch.empty((3, n), dtype=torch.float64, pin_memory=True)
        hD = torch.empty((3, n), dtype=torch.float64, pin_memory=True)
        hE = torch.empty(n, dtype=torch.float64, pin_memory=True)
        hV.copy_(V); hD.copy_(D); hE.copy_(src._props["energy"].pieces[0])
        host = T.ray_bundle.RayBundle(vertices=hV.numpy(), directions=hD.numpy(), energy=hE.numpy())
        host.__dict__["_ray_id_base"] = rank * n
        e_steps = max(1, min(args.steps, 3))
        e_ms, e_b, _, _ = timed(host, e_steps, 1, read_back=True)
        # ceiling of that number: the same bytes copied host -> device by all ranks at once, nothing else running
        dV, dD, dE = torch.empty_like(V), torch.empty_like(D), torch.empty(n, dtype=torch.float64, device=dev)
        best = None
        for it in range(3):
            barrier()
            c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            c0.record()
            dV.copy_(hV, non_blocking=True); dD.copy_(hD, non_blocking=True); dE.copy_(hE, non_blocking=True)
            c1.record()
            barrier()
            t = torch.tensor([c0.elapsed_time(c1)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if it > 0:
                best = float(t.item()) if best is None else min(best, float(t.item()))
        del dV, dD, dE
        h2d = 56 * n * world
        e2e = {"value": e_b / (e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": FLUX_RES * FLUX_RES * 8 * world, "ms_per_step": e_ms / e_steps,
               "h2d_ceiling_ms": best, "h2d_ceiling_gbs": h2d / (best * 1e-3) / 1e9,
               "h2d_ceiling_gbs_per_gpu": 56 * n / (best * 1e-3) / 1e9,
               "frac_of_h2d_ceiling": best / (e_ms / e_steps)}
        del hV, hD, hE, host

    V = D = None
    weak = None
    if world > 1 and not args.no_weak:
        del src
        torch.cuda.empty_cache()
        n_w = args.rays
        wsrc = source(n_w, rank * n_w, n_w * world)
        w_steps = max(1, min(args.steps, 5))
        w_ms, w_b, _, _ = timed(wsrc, w_steps, 2)
        weak = {"value": w_b / (w_ms * 1e-3), "unit": UNIT, "ms_per_step": w_ms / w_steps, "rays_per_gpu_per_step": n_w,
                "scaling": "weak"}
        del wsrc
        torch.cuda.empty_cache()

    clocks = sampler.stop()
    configs = pipes = cpu = flux_cmp = None
    if rank == 0 and world == 1:
        torch.cuda.empty_cache()
        f32, f64 = C.c_double(0.), C.c_double(0.)
        L.check(lib.trc_fma_peak(0, C.byref(f32)))
        L.check(lib.trc_fma_peak(1, C.byref(f64)))
        pipes = {"fp32_fma_tflops": f32.value, "fp64_fma_tflops": f64.value,
                 "how": "trc_fma_peak: 8 independent FMA chains per thread, 8 CTAs of 256 threads per SM, best of 3"}
        if not args.no_configs:
            try:
                del src
            except NameError:
                pass
            torch.cuda.empty_cache()
            configs = config_rates(T, torch, dev)
    if rank == 0 and world == 1 and not args.no_cpu:
        procs = max(1, min(os.cpu_count() or 1, 64))
        v, b, wall = cpu_baseline(args.cpu_rays, procs)
        cpu = {"value": v, "unit": UNIT, "cores": procs, "kind": "port",
               "sample": "%d processes x %d rays of the same NSTTF scene, reps=%d, %.1f s wall" % (procs, args.cpu_rays, REPS, wall)}
        gpu_map = bins.cpu().numpy() if hasattr(bins, "cpu") else np.asarray(bins)
        flux_cmp = fluxmap_l2_vs_ref(gpu_map, cpu_baseline.maps)

    if rank == 0:
        sm_hz = (clocks.get("sm_mhz") or 1965.0) * 1e6
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": ncu["dram_bytes_per_ray"] * n if ncu else None,
                    "kernel": (("nearest_hit_grid_kernel (bounce 1: shadow grid of the coherent source bundle; + the tree walk "
                                "over the chunks it leaves)" if eng.stats.get("shadow_grid") else "nearest_hit_walk_kernel (bounce 1)")
                               if args.accel == "bvh" else "nearest_hit_brute_kernel (bounce 1)"),
                    "kernel_ms": k_ms, "launches_timed": len(ev), "algorithmic_bytes": algo_bytes,
                    "bytes_per_ray": ALGO_BYTES_PER_RAY_HIT_KERNEL, "bytes_per_hit": ALGO_BYTES_PER_HIT_HIT_KERNEL,
                    "hits": k_hits, "rays": n, "peak_source": peak_src + ", burst (kernel timed by its own events)",
                    "traffic_source": ncu_src}
        if ncu:
            # the kernel is issue-bound, not bandwidth-bound: its second roofline is warp instructions per ray
            # against what the SMs' schedulers can issue (4 per SM per clock)
            issue_peak = 148 * 4 * sm_hz
            inst = ncu["warp_inst_per_ray"] * n
            roofline["issue"] = {"warp_inst_per_ray": ncu["warp_inst_per_ray"], "issued_per_s": inst / (k_ms * 1e-3),
                                 "peak_per_s": issue_peak, "frac": inst / (k_ms * 1e-3) / issue_peak,
                                 "floor_ms_at_full_issue": inst / issue_peak * 1e3,
                                 "hbm_floor_ms": algo_bytes / (peak * 1e9) * 1e3}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "rays_per_step": n_total, "rays_per_gpu_per_step": n,
                       "sharding": "one source of rays_per_step rays split over the GPUs (global Philox ray ids)",
                       "surfaces": len(surfaces), "accel": args.accel, "rng": "philox", "tree": True,
                       "flux_map": "%dx%d on device" % (FLUX_RES, FLUX_RES),
                       "l2": "inputs (56 B x rays per GPU: 5.6 GB at 1e8, 0.7 GB at 1.25e7) larger than L2",
                       "per_bounce_rays_in_hits_survivors": per_bounce},
            "ray_surface_tests_per_s_brute_equivalent": tests_equiv,
            "receiver_power_W": float(bins.sum().item()) if hasattr(bins, "sum") else None,
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
            "pipe_peaks": pipes, "weak": weak, "configs": configs,
            "cpu_baseline": cpu, "flux_map_vs_ref": flux_cmp,
        }
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)
    if world > 1:
        dist.destroy_process_group()
