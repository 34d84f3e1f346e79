#!/usr/bin/env python
"""Benchmark of the B200-native MPAS-Atmosphere SRK3 dycore step (BASELINE.json metric: dycore timesteps/sec).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--level 7] [--levels 55] [--scalars 4]

One "step" = one full atm_srk3 dycore step (dynamics substeps + acoustic loop + monotonic scalar transport + every
halo exchange) of the Jablonowski-Williamson baroclinic wave on an in-harness icosahedral Voronoi mesh.
N = 1 workload: BASELINE.json configs[1] (x1.163842 = icosahedral level 7, 60 km, 55 levels, qv + 3 passive tracers);
the 15 km mesh the metric is quoted on needs ~160 GB of resident fields in double as built and does not fit one GPU
next to its host-side generation (DESIGN.md "Memory").  N > 1: the same mesh METIS-partitioned over N ranks (strong scaling),
one process per GPU, halo exchange over NCCL.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "tests"))

LEVEL_CELLS = {l: 10 * 4 ** l + 2 for l in range(3, 12)}
# algorithmic words (8-byte reals read+written once) per owned cell-level and call, SURVEY.md section 8(a)/(d);
# mixed entries are the per-call average over one dt with the Registry defaults
def words_per_call(ns):
    return {
        "atm_compute_dyn_tend": (79 + 39 + 39) / 3.0,
        "atm_advance_acoustic_step": (3 * 31 + 48) / 4.0,
        "atm_divergence_damping_3d": 9.0,
        "atm_recover_large_step_variables": (47 + 47 + 51) / 3.0,
        "atm_advance_scalars_mono": 25.0 * ns + 11,
        "atm_advance_scalars": 10.0 * ns + 7,
        "atm_compute_solve_diagnostics": (39 + 39 + 42) / 3.0,
        "atm_compute_vert_imp_coefs": 16.0,
        "small_step_prep": 18.0,
        "atm_rk_dynamics_substep_finish": 30.0,
        "atm_rk_integration_setup": 20.0 + 2 * ns,
        "atm_compute_moist_coefficients": 6.0 + ns,
        "mpas_reconstruct": 8.0,  # u over the 3 edges per cell + 5 outputs (mpas_vector_reconstruction.F:254-288)
    }


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(smax) if smax else None, "reasons": sorted(reasons), "samples": len(sm)}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--level", type=int, default=7, help="icosahedral level: 6=120 km, 7=60 km, 8=30 km, 9=15 km")
    ap.add_argument("--levels", type=int, default=55)
    ap.add_argument("--scalars", type=int, default=4)
    ap.add_argument("--lloyd", type=int, default=5)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--ktime", action="store_true", help="also print per-kernel CUDA-event times of one eager step to stderr")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=20.0, help="target CPU time of the bounded cpu_baseline sample")
    return ap.parse_args()


def workload_name(a):
    dx = {6: 120, 7: 60, 8: 30, 9: 15, 10: 7.5, 11: 3.75}.get(a.level, 0)
    return "JW baroclinic wave, icosahedral level %d (x1.%d, %s km), %d levels, %d scalars (qv + %d passive tracers), dycore only" % (
        a.level, LEVEL_CELLS[a.level], dx, a.levels, a.scalars, a.scalars - 1)


def set_tracers(block, ns):
    import numpy as np
    s = block.field("scalars@1")
    lat, lon = block.field("latCell"), block.field("lonCell")
    if ns > 1:
        r = np.arccos(np.clip(np.sin(0.6) * np.sin(lat) + np.cos(0.6) * np.cos(lat) * np.cos(lon - 1.0), -1, 1))
        s[:-1, :, 1] = np.where(r < 0.8, 0.5 * (1 + np.cos(np.pi * r / 0.8)), 0.0)[:-1, None]
    if ns > 2:
        s[:-1, :, 2] = 1.0
    if ns > 3:
        s[:-1, :, 3] = np.where(np.cos(lon - 1.0) * np.cos(lat) > 0.7, 1.0, 0.1)[:-1, None]


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def run_cpu_sample(a, level, nsteps_max, target_seconds):
    """Time the CPU restatement (oracle perf build: -O3 -march=native, OpenMP over all host cores) on `level`."""
    import common
    cores = os.cpu_count()
    case = common.host.HostCase(level, lloyd_iters=a.lloyd, sfc=True)
    block = case.build_block(0, nVertLevels=a.levels, init_case=2, num_scalars=a.scalars, full_fields=True)
    set_tracers(block, a.scalars)
    m = common.OracleModel([block], common.level_config(level), perf=True)
    m.init()
    t0 = time.perf_counter()
    n = 0
    while n < nsteps_max:
        m.step(1)
        n += 1
        if time.perf_counter() - t0 >= target_seconds:
            break
    dt = time.perf_counter() - t0
    return n, dt, cores, case.nCells


def run_reference(a):
    """--impl reference: the reference's CPU implementation of the path.  The Fortran reference cannot be built on this
    image (no gfortran / MPI / netCDF: SURVEY.md F1), so this arm times the C++/OpenMP line-by-line restatement of the
    same loops (oracle/, perf build) on all host cores."""
    rank, local_rank, world = dist_env()
    if rank != 0:
        return
    total = a.steps + a.warmup
    # bounded sample: pick the largest mesh whose (steps + warmup) fit ~150 s, probing from the small side
    level = min(a.level, 5)
    n, dt, cores, ncells = run_cpu_sample(a, level, 1, 0.0)
    per_cell_step = dt / ncells
    while level < a.level and per_cell_step * LEVEL_CELLS[level + 1] * total < 150.0:
        level += 1
    import common
    case = common.host.HostCase(level, lloyd_iters=a.lloyd, sfc=True)
    block = case.build_block(0, nVertLevels=a.levels, init_case=2, num_scalars=a.scalars, full_fields=True)
    set_tracers(block, a.scalars)
    m = common.OracleModel([block], common.level_config(level), perf=True)
    m.init()
    m.step(a.warmup)
    t0 = time.perf_counter()
    m.step(a.steps)
    dt = time.perf_counter() - t0
    scale = case.nCells / float(LEVEL_CELLS[a.level])
    value = a.steps / dt * scale
    sample = "%d steps of the same case on icosahedral level %d (%d of %d cells), steps/s scaled by the cell ratio %.4f" % (
        a.steps, level, case.nCells, LEVEL_CELLS[a.level], scale)
    line = {
        "impl": "reference", "metric": "dycore timesteps/sec", "value": value, "unit": "steps/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": 1000.0 / value, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(a), "note": "reference Fortran (gfortran+MPI) unavailable on this image; C++/OpenMP restatement of the same loops"},
        "cpu_baseline": {"value": value, "unit": "steps/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


E2E_FIELDS = ["u@1", "w@1", "theta_m@1", "rho_zz@1", "scalars@1", "ru", "rw", "rho_p", "rtheta_p", "exner", "pressure_p", "rho_edge", "pv_edge", "v", "ke",
              "divergence", "vorticity"]  # the restart set the step reads at a step boundary (Registry.xml restart stream)


def run_b200(a):
    import numpy as np
    import torch
    import common
    host = common.host
    rank, local_rank, world = dist_env()
    N = a.gpus
    if N != world:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch with torch.distributed.run --nproc-per-node %d" % (N, world, N))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the dycore has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if N > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    t_setup = time.perf_counter()
    case = host.HostCase(a.level, lloyd_iters=a.lloyd, sfc=True)
    if N > 1:
        case.partition(N, "metis")
    # "lean": the work / tendency arrays (all zero at start, 83 % of a block's words) get no host storage, their device mirrors
    # are created zero-filled from the metadata -- what lets the 15 km mesh be set up by 2..8 ranks within the host's memory
    block = case.build_block(rank, nVertLevels=a.levels, init_case=2, num_scalars=a.scalars, full_fields="lean")
    set_tracers(block, a.scalars)
    cfg = common.level_config(a.level)
    dy = host.Dycore([block], cfg, rank=rank, nranks=N, device=local_rank)
    if N > 1:
        obj = [host.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        dy.comm_init(obj[0])
    dy.init()
    use_graph = not a.no_graph
    dy.set_option("graph", 1 if use_graph else 0)
    t_setup = time.perf_counter() - t_setup
    bcfg = block.cfg()
    nCS, L = int(bcfg["nCellsSolve"]), a.levels

    def barrier():
        dy.sync()
        if N > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if N == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    L_ = host.lib()
    dy.step(a.warmup)
    barrier()
    clocks = ClockSampler(local_rank)
    clocks.start()
    l0 = dy.launch_count()
    x0 = dy.exchange_count()
    L_.mpas_b200_timer_start()
    dy.step(a.steps)
    ms = L_.mpas_b200_timer_stop()
    barrier()
    clk = clocks.stop()
    launches = dy.launch_count() - l0
    exchanges = dy.exchange_count() - x0
    ms = max_over_ranks(ms)
    value = a.steps / (ms / 1000.0)
    wmax, umax = dy.max_abs("w@1"), dy.max_abs("u@1")
    if not (np.isfinite(wmax) and np.isfinite(umax) and umax < 300.0):
        raise SystemExit("integration blew up: max|u| = %r max|w| = %r" % (umax, wmax))

    # ---- per-routine timing pass (CUDA events around every routine call, eager launches) -> roofline
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_gbs = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if "hbm_gbs" in peaks else "B200_PROFILING.md fallback 6650 GB/s (of fallback)"
    dy.set_option("timing", 1)
    dy.reset_timers()
    nt = 2
    dy.step(nt)
    timers = dy.timers()
    dy.set_option("timing", 0)
    if a.ktime and rank == 0:
        import ctypes as C
        L_.mpas_b200_ktime_enable(1)
        dy.set_option("graph", 0)
        dy.step(2)
        buf = C.create_string_buffer(1 << 16)
        L_.mpas_b200_ktime_report(buf, 1 << 16)
        L_.mpas_b200_ktime_enable(0)
        dy.set_option("graph", 1 if use_graph else 0)
        rows = [ln.split() for ln in buf.value.decode().splitlines()]
        tot = sum(float(r[1]) for r in rows)
        for r in sorted(rows, key=lambda r: -float(r[1])):
            print("KTIME %-28s %8.3f ms/step %5.1f calls/step %8.1f us/call %5.1f%%" % (r[0], float(r[1]) / 2, int(r[2]) / 2, 1e3 * float(r[1]) / int(r[2]), 100 * float(r[1]) / tot), file=sys.stderr)
    words = words_per_call(a.scalars)
    nCS_total = nCS
    phases = {}
    for name, (tms, calls) in timers.items():
        per_call = tms / max(calls, 1)
        e = {"ms_per_step": tms / nt, "calls_per_step": calls / nt, "ms_per_call": per_call}
        if name in words and per_call > 0:
            gb = words[name] * 8.0 * nCS_total * L / 1e9
            e["algorithmic_GB_per_call"] = gb
            e["achieved_GBs"] = gb / (per_call / 1000.0)
            e["frac"] = e["achieved_GBs"] / peak_gbs
        phases[name] = e
    dom = max((n for n in phases if n in words), key=lambda n: phases[n]["ms_per_step"])
    roofline = {"bound": "hbm", "kernel": dom, "achieved": phases[dom]["achieved_GBs"], "peak": peak_gbs, "unit": "GB/s", "frac": phases[dom]["frac"],
                "traffic": None, "peak_source": peak_src,
                "note": "achieved = SURVEY 8(d) words x 8 B x owned cell-levels per call / CUDA-event time per call (eager pass after the timed region)"}
    # traffic: DRAM bytes (read + write) per call of the dominant routine from the committed ncu capture of this same command
    # (profiles/r01_d_dram_traffic.json, tools/dram_traffic.py); only valid for the default workload on one GPU
    try:
        if N == 1 and a.level == 7 and a.levels == 55 and a.scalars == 4:
            tr = json.load(open(os.path.join(ROOT, "profiles", "r01_d_dram_traffic.json")))
            roofline["traffic"] = tr[dom]["dram_GB_per_step"] / phases[dom]["calls_per_step"]
            roofline["traffic_unit"] = "GB per call (ncu dram__bytes_read.sum + dram__bytes_write.sum, profiles/r01_d_dram_traffic.csv)"
            roofline["algorithmic_GB_per_call"] = phases[dom]["algorithmic_GB_per_call"]
            roofline["whole_step_traffic_GB"] = tr["_whole_step"]["dram_GB_per_step"]
    except Exception:
        pass
    step_gb = sum(phases[n]["algorithmic_GB_per_call"] * phases[n]["calls_per_step"] for n in phases if "algorithmic_GB_per_call" in phases[n])
    roofline["whole_step"] = {"algorithmic_GB": step_gb, "achieved_GBs": step_gb * value * (1 if N == 1 else 1), "frac": step_gb * value / peak_gbs}

    # ---- e2e: the same step through the public API with HOST buffers (pinned), copies inside the timed region
    e2e = None
    if not a.no_e2e:
        bufs = {}
        h2d = 0
        for n in E2E_FIELDS:
            kind, loc, dims = block.info(n)
            cnt = int(np.prod(dims))
            t = torch.empty(cnt, dtype=torch.float64).pin_memory()
            arr = t.numpy()
            dy.field(0, n, out=arr.reshape(tuple(reversed(dims))))
            bufs[n] = (t, arr)
            h2d += arr.nbytes
        ke = max(1, min(a.steps, 5))
        barrier()
        L_.mpas_b200_timer_start()
        for _ in range(ke):
            for n, (t, arr) in bufs.items():
                rc = L_.mpas_b200_dycore_set_field(dy.h, 0, n.encode(), arr.ctypes.data, arr.nbytes)
                assert rc == 0
            dy.step(1)
            for n, (t, arr) in bufs.items():
                rc = L_.mpas_b200_dycore_get_field(dy.h, 0, n.encode(), arr.ctypes.data, arr.nbytes)
                assert rc == 0
        ems = max_over_ranks(L_.mpas_b200_timer_stop())
        e2e = {"value": ke / (ems / 1000.0), "unit": "steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": h2d, "steps": ke,
               "what": "per step: upload the %d-field restart set from pinned host memory, one dycore step, download the same set" % len(E2E_FIELDS)}

    # ---- CPU baseline (rank 0, N = 1): bounded sample of the same workload on the host cores
    cpu = None
    if rank == 0 and N == 1 and not a.no_cpu_baseline:
        lvl = a.level
        n, dt_, cores, ncells = run_cpu_sample(a, min(lvl, 5), 1, 0.0)
        while lvl > 5 and dt_ / ncells * LEVEL_CELLS[lvl] > a.cpu_seconds:
            lvl -= 1
        n, dt_, cores, ncells = run_cpu_sample(a, lvl, 3, a.cpu_seconds)
        scale = ncells / float(LEVEL_CELLS[a.level])
        cpu = {"value": n / dt_ * scale, "unit": "steps/s", "cores": cores, "kind": "port",
               "sample": "%d step(s) of the same case on icosahedral level %d (%d of %d cells) in %.1f s, scaled by the cell ratio; C++/OpenMP restatement, -O3 -march=native"
                         % (n, lvl, ncells, LEVEL_CELLS[a.level], dt_)}

    if rank == 0:
        dt_model = cfg["config_dt"]
        line = {
            "metric": "dycore timesteps/sec", "value": value, "unit": "steps/s", "n_gpus": N, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms / a.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(a), "nCells": LEVEL_CELLS[a.level], "nVertLevels": a.levels, "num_scalars": a.scalars, "config_dt_s": dt_model,
                       "simulated_days_per_day": value * dt_model, "partition": "METIS k-way, %d blocks" % N if N > 1 else "single block",
                       "l2_policy": "no flush needed: one step streams ~%.0f GB through a 126 MB L2" % step_gb, "cuda_graph": use_graph,
                       "halo_exchanges_per_step": exchanges / a.steps, "halo_transport": dy.halo_transport(), "setup_seconds": t_setup, "max_abs_u": umax, "max_abs_w": wmax},
            "roofline": roofline, "phases": phases, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clk,
        }
        print(json.dumps(line), flush=True)
    dy.close()
    if N > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
