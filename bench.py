"""bench.py -- ray-bounces/s of the NSTTF heliostat-field trace (BASELINE.json).

A step = one ``TracerEngine.ray_tracer`` pass (reps=3, tree=True) over one synthetic Buie-sunshape bundle aimed
at the 218-heliostat Sandia NSTTF field + tower receiver, followed by on-device
flux-map binning of the receiver hits and (N>1) the NCCL all-reduce of the tallies.

The metric's own configuration is ONE source of 1e8 rays sharded over the N GPUs (BASELINE.json configs[2],
the reference's ``.../procs`` scheme: tracer_engine_mp.py:19-35, sources.py:238), so the headline is
strong scaling: every rank traces 1e8 / N rays of the same source (global Philox ray ids).

  value    : device-resident shards, CUDA-event timed, max over ranks, aggregate ray-bounces/s
  e2e      : same step through the public API with HOST bundles in pinned memory (H2D inside the timed
             region) and the flux map read back; next to it the bare concurrent H2D rate of the same bytes
  roofline : dominant kernel (nearest-hit BVH walk over the source shard) timed by CUDA events around its
             launch inside the timed region; algorithmic bytes = 48 B/ray + 72 B/hit; DRAM traffic
             and instructions per ray from the committed ncu capture when it matches the kernel source
  weak     : (N>1) every rank traces its own 1e8 rays -- the round-1 line, kept for comparison
  configs  : (N=1) short device-timed rates of the other BASELINE configurations (C1, C2, C3-dense, C4, C5)
  cpu_baseline / --impl reference: the NumPy oracle port of the reference engine on the host cores,
             bounded sample of the same workload
"""
import os
os.environ.setdefault("OMP_NUM_THREADS", "1")       # CPU baseline: one BLAS thread per process
os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
import argparse
import json
import multiprocessing as mp
import subprocess
import sys
import threading
import time
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from tracer_b200.models import scenes  # noqa: E402  (scene builders; no oracle import here)

METRIC = "ray-bounces/sec, NSTTF field 1e8 rays"
UNIT = "ray-bounces/s"
REPS = 3
FLUX_RES = 50
FLUX_CMP_RES = 10                       # bins per side of the flux-map comparison with the CPU reference
ALGO_BYTES_PER_RAY_HIT_KERNEL = 48      # vertex + direction read once per ray
ALGO_BYTES_PER_HIT_HIT_KERNEL = 72      # per hit: 8 B energy read + 64 B hit record written
WORKLOAD = "C3 NSTTF heliostat field (218 heliostats + receiver), Buie CSR 0.01, reps=3"


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx = float(r[2])
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except (ValueError, IndexError):
                pass
        busy = sorted(sm)[len(sm) // 2:] if sm else []      # upper half = samples under load
        return {"sm_mhz": float(np.median(busy)) if busy else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


def nsttf_source_geometry():
    pos = scenes.nsttf_positions()
    sun = scenes.sun_vector(0., np.radians(35.05))
    lo, hi = pos.min(axis=0), pos.max(axis=0)
    centre = 0.5 * (lo + hi)
    radius = 1.1 * 0.5 * np.sqrt(np.sum((hi[:2] - lo[:2]) ** 2))
    return np.vstack(300. * sun + centre), -sun, radius


# ------------------------------------------------------------ CPU baseline
def _cpu_worker(args):
    n_rays, seed = args
    os.environ["OMP_NUM_THREADS"] = "1"
    from oracle.numpy_trace import OracleEngine          # checker timed as the CPU baseline
    T = scenes.NS('tracer_b200')
    np.random.seed(seed)
    asm, source = scenes.c3_nsttf(T, n_rays=n_rays)
    bundle = source()
    eng = OracleEngine(asm)
    t0 = time.perf_counter()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        eng.ray_tracer(bundle, reps=REPS, min_energy=1e-10, tree=True)
    wall = time.perf_counter() - t0
    # receiver flux map of this process (outside the timed region): absorbed energy per bin
    recv_idx = len(asm.get_surfaces()) - 1
    recv = asm.get_surfaces()[recv_idx]
    absorbed, hits = eng.get_all_hits(recv_idx)
    hx, hy = np.asarray(recv.get_geometry_manager()._half_dims, float).reshape(-1)
    local = np.dot(np.linalg.inv(recv._temp_frame), np.vstack([hits, np.ones(hits.shape[1])]))
    fmap, _, _ = np.histogram2d(local[0], local[1], bins=FLUX_CMP_RES, range=[[-hx, hx], [-hy, hy]], weights=absorbed)
    return eng.ray_bounces, wall, fmap


def cpu_baseline(rays_per_proc, procs, seed=0):
    """P independent processes, one source each (the multi_ray_sim scheme of
    tracer_engine_mp.py:28-35), wall time around the whole pool."""
    ctx = mp.get_context("fork")
    t0 = time.perf_counter()
    with ctx.Pool(procs) as pool:
        res = pool.map(_cpu_worker, [(rays_per_proc, seed + k) for k in range(procs)])
    wall = time.perf_counter() - t0
    bounces = sum(r[0] for r in res)
    trace_wall = max(r[1] for r in res)
    cpu_baseline.maps = np.array([r[2] for r in res])     # one independent flux map per process
    return bounces / trace_wall, bounces, wall


def fluxmap_l2_vs_ref(gpu_bins, cpu_maps):
    """BASELINE metric "flux-map L2 vs ref": relative L2 distance between the receiver map of
    the GPU trace (Philox) and the map of the CPU reference port (numpy.random), both as energy
    per bin on FLUX_CMP_RES^2 bins, next to the distance Monte-Carlo noise of the CPU sample
    alone explains (standard error over its independent processes)."""
    g = np.asarray(gpu_bins, float)
    f = g.shape[0] // FLUX_CMP_RES
    g = g.reshape(FLUX_CMP_RES, f, FLUX_CMP_RES, f).sum(axis=(1, 3))
    mean = cpu_maps.mean(axis=0)
    norm = np.sqrt(np.sum(mean ** 2))
    out = {"l2_rel": float(np.sqrt(np.sum((g - mean) ** 2)) / norm), "bins": "%dx%d" % (FLUX_CMP_RES, FLUX_CMP_RES),
           "ref": "CPU port, %d independent processes" % len(cpu_maps)}
    if len(cpu_maps) > 1:
        sem = cpu_maps.std(axis=0, ddof=1) / np.sqrt(len(cpu_maps))
        out["l2_rel_expected_from_mc_noise"] = float(np.sqrt(np.sum(sem ** 2)) / norm)
        live = sem > 0
        out["max_bin_z"] = float(np.max(np.abs(g - mean)[live] / sem[live]))
    return out


def run_reference(args, rank):
    """--impl reference: the oracle port of the reference engine on all host cores."""
    if rank != 0:
        return
    procs = max(1, min(os.cpu_count() or 1, 64))
    rays = args.cpu_rays
    vals = []
    for it in range(args.warmup + args.steps):
        v, bounces, wall = cpu_baseline(rays, procs, seed=100 * it)
        if it >= args.warmup:
            vals.append((v, bounces, wall))
    value = float(np.mean([v[0] for v in vals]))
    ms = float(np.mean([v[1] / v[0] for v in vals])) * 1e3
    sample = "%d processes x %d rays of the NSTTF scene (219 surfaces), reps=%d" % (procs, rays, REPS)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "impl": "reference",
        "config": {"workload": WORKLOAD, "rays_per_step": procs * rays, "surfaces": 219},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": procs, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rays", type=float, default=1e8, help="rays of the whole source per step (sharded over the GPUs)")
    ap.add_argument("--cpu-rays", type=float, default=2e5, help="rays per host process for the CPU baseline")
    ap.add_argument("--accel", default="bvh", choices=["bvh", "brute"])
    ap.add_argument("--leaf-size", type=int, default=1)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-configs", action="store_true")
    ap.add_argument("--no-weak", action="store_true")
    args = ap.parse_args()
    args.rays, args.cpu_rays = int(args.rays), int(args.cpu_rays)
    rank = int(os.environ.get("RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    from bench_gpu import run_gpu
    run_gpu(args)


if __name__ == "__main__":
    main()
