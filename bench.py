#!/usr/bin/env python
"""Benchmark of the B200-native MPAS-Atmosphere SRK3 dycore step (BASELINE.json metric: dycore timesteps/sec).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--level 9] [--levels 55] [--scalars 1]

One "step" = one full atm_srk3 dycore step (dynamics substeps + acoustic loop + monotonic scalar transport + every
halo exchange) of the Jablonowski-Williamson baroclinic wave on an in-harness icosahedral Voronoi mesh.
Default workload at EVERY N: the mesh the metric is quoted on, BASELINE.json configs[3] -- x1.2621442 = icosahedral
level 9, 15 km, 55 levels, qv only (configs[3] names no tracers) -- one block at N = 1, METIS-partitioned over N ranks for
N > 1 (strong scaling), one process per GPU, halo exchange over NVLink peer stores / NCCL.  If the device cannot hold it
(a caught out-of-memory) the run falls back one icosahedral level at a time and says so in config.workload.
`--level 7 --scalars 4` is BASELINE.json configs[1] (x1.163842, 60 km, qv + 3 passive tracers).
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "tests"))

LEVEL_CELLS = {l: 10 * 4 ** l + 2 for l in range(3, 12)}
LEVEL_KM = {5: 240, 6: 120, 7: 60, 8: 30, 9: 15, 10: 7.5, 11: 3.75}
DEFAULT_LEVEL = 9


def words_per_call(ns, L, minimal=False):
    """algorithmic words (reals read + written once) per owned cell-level and call, generated from the per-routine array table
    roofline/arrays.yaml (tools/roofline_words.py; it also cross-checks them against ncu dram bytes).  Default = the arrays the
    REFERENCE's loops touch, the figure of SURVEY.md section 8(a)/(d) (dyn_tend 74 / 41 against the survey's hand count 79 / 39);
    minimal = without the physics-coupling / output-only arrays the library does not even allocate (memory diet)."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import roofline_words
    return roofline_words.words_per_call(ns, L, with_optional=not minimal)


def kernel_source_sha():
    """identifies the kernel sources a profile capture belongs to (a capture of other sources is refused)"""
    h = hashlib.sha256()
    d = os.path.join(ROOT, "mpas-model_b200", "csrc", "cuda")
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh", ".h")):
            h.update(f.encode())
            h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


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
    ap.add_argument("--level", type=int, default=None, help="icosahedral level: 6=120 km, 7=60 km, 8=30 km, 9=15 km (default 9, the metric's mesh)")
    ap.add_argument("--levels", type=int, default=55)
    ap.add_argument("--scalars", type=int, default=None, help="num_scalars (default: 1 = qv at level >= 9 as configs[3], 4 = qv + 3 tracers below as configs[1])")
    ap.add_argument("--lloyd", type=int, default=5)
    ap.add_argument("--case", default="jw", choices=["jw", "supercell"],
                    help="supercell = BASELINE.json configs[2]: doubly periodic planar hex mesh, 2 km, --nx x --nx cells, qv qc qr transported (an extra line, not the metric)")
    ap.add_argument("--nx", type=int, default=200, help="supercell: cells per side")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--ktime", action="store_true", help="also print per-kernel CUDA-event times of one eager step to stderr")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=30.0, help="target CPU time of the bounded cpu_baseline sample")
    ap.add_argument("--precision", default=None, choices=["double", "single"], help="single = the PRECISION=single build (dtype f32), an extra line")
    a = ap.parse_args()
    if a.case == "supercell":  # "level" 7 = the requested --nx, every level below halves the side (the bounded CPU sample)
        a.auto_level = False
        a.level = 7
        a.auto_scalars = False
        a.scalars = a.scalars or 3
        if "--levels" not in sys.argv:
            a.levels = 40
        return a
    a.auto_level = a.level is None
    if a.level is None:
        a.level = DEFAULT_LEVEL
    a.auto_scalars = a.scalars is None
    if a.scalars is None:
        a.scalars = 1 if a.level >= 9 else 4
    return a


def side(a, level):
    return max(8, (a.nx >> (7 - level)) & ~1)


def n_cells(a, level):
    return side(a, level) ** 2 if a.case == "supercell" else LEVEL_CELLS[level]


def host_case(a, level):
    import common
    if a.case == "supercell":
        return common.host.HostCase(0, planar=(side(a, level), side(a, level), 2000.0), sfc=True)
    return common.host.HostCase(level, lloyd_iters=a.lloyd, sfc=True)


def model_config(a, level):
    import common
    if a.case == "supercell":
        return dict(config_dt=12.0, config_len_disp=2000.0)  # 2 km -> 12 s
    return common.level_config(level)


def init_case(a):
    return 5 if a.case == "supercell" else 2


def workload_name(level, levels, ns, note="", a=None):
    if a is not None and a.case == "supercell":
        return "idealised supercell, doubly periodic planar hex mesh %d x %d cells, 2 km, %d levels, %d moist scalars (qv, qc, qr) transported, physics off%s" % (
            side(a, level), side(a, level), levels, ns, note)
    tr = "qv only" if ns == 1 else "qv + %d passive tracers" % (ns - 1)
    return "JW baroclinic wave, icosahedral level %d (x1.%d, %s km), %d levels, %d scalars (%s), dycore only%s" % (
        level, LEVEL_CELLS[level], LEVEL_KM.get(level, 0), levels, ns, tr, note)


def config_dict(a, level, ns, note=""):
    """identical in the b200 and the reference arm for the same command line"""
    import common
    return {"workload": workload_name(level, a.levels, ns, note, a), "nCells": n_cells(a, level), "nVertLevels": a.levels, "num_scalars": ns,
            "config_dt_s": model_config(a, level)["config_dt"],
            "partition": "METIS k-way, %d blocks" % a.gpus if a.gpus > 1 else "single block", "l2_policy": "no flush needed: one step streams far more than the 126 MB L2"}


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


def use_all_cores(share=1):
    """torch.distributed.run exports OMP_NUM_THREADS=1 to its workers: set the OpenMP thread count of the host harness and of
    the CPU oracle explicitly (all host cores / `share` processes)"""
    import common
    n = max(1, (os.cpu_count() or 1) // share)
    common.host.hostlib().mpas_host_set_num_threads(n)
    return n


def cpu_model(a, level, ns, threads):
    """the CPU restatement (oracle perf build: -O3 -march=native, OpenMP) of the same case on `level`, initialised"""
    import common
    common.oracle(perf=True).oracle_set_num_threads(threads)
    case = host_case(a, level)
    block = case.build_block(0, nVertLevels=a.levels, init_case=init_case(a), num_scalars=ns, full_fields=True)
    if a.case == "jw":
        set_tracers(block, ns)
    m = common.OracleModel([block], model_config(a, level), perf=True)
    m.init()
    return case, block, m


def cpu_sample_level(a, ns, threads, nsteps, budget_s, max_level):
    """largest icosahedral level <= max_level whose `nsteps` steps fit `budget_s` of CPU time, from a one-step probe on level 5"""
    level = min(5, max_level)
    case, block, m = cpu_model(a, level, ns, threads)
    m.step(1)
    t0 = time.perf_counter()
    m.step(1)
    per_cell = (time.perf_counter() - t0) / case.nCells
    while level < max_level and per_cell * n_cells(a, level + 1) * nsteps <= budget_s:
        level += 1
    return level


def phase_ms(m, nsteps):
    return {k: {"ms_per_step": v[0] / nsteps, "calls_per_step": v[1] / nsteps} for k, v in m.timers().items()}


def run_reference(a):
    """--impl reference: the reference's CPU implementation of the path.  The Fortran reference cannot be built on this
    image (no gfortran / MPI / netCDF: SURVEY.md F1), so this arm times the C++/OpenMP line-by-line restatement of the
    same loops (oracle/, perf build) on ALL host cores; only libmpas_host.so (mesh, JW state: no CUDA) and the oracle are
    loaded.  Each step is a bounded sample of the workload: the same case on a smaller icosahedral level, steps/s scaled by
    the cell ratio (the loops are linear in the number of cells)."""
    rank, local_rank, world = dist_env()
    if rank != 0:
        return
    threads = use_all_cores(1)
    total = a.steps + a.warmup
    level = cpu_sample_level(a, a.scalars, threads, total, 150.0, min(a.level, 7))
    case, block, m = cpu_model(a, level, a.scalars, threads)
    m.step(a.warmup)
    m.reset_timers()
    t0 = time.perf_counter()
    m.step(a.steps)
    dt = time.perf_counter() - t0
    scale = case.nCells / float(n_cells(a, a.level))
    value = a.steps / dt * scale
    sample = "%d steps of the same case on icosahedral level %d (%d of %d cells) in %.1f s on %d OpenMP threads, steps/s scaled by the cell ratio %.5f" % (
        a.steps, level, case.nCells, n_cells(a, a.level), dt, threads, scale)
    line = {
        "impl": "reference", "metric": "dycore timesteps/sec", "value": value, "unit": "steps/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": 1000.0 / value, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_dict(a, a.level, a.scalars),
        "note": "reference Fortran (gfortran+MPI) unavailable on this image; C++/OpenMP restatement of the same loops, all host cores",
        "cpu_baseline": {"value": value, "unit": "steps/s", "cores": threads, "kind": "port", "sample": sample,
                         "phases_ms_on_sample_mesh": phase_ms(m, a.steps)},
        "e2e": {"value": value, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


E2E_FIELDS = ["u@1", "w@1", "theta_m@1", "rho_zz@1", "scalars@1", "ru", "rw", "rho_p", "rtheta_p", "exner", "pressure_p", "rho_edge", "pv_edge", "v", "ke",
              "divergence", "vorticity"]  # the restart set the step reads at a step boundary (Registry.xml restart stream)
CHECKSUM_FIELDS = ["u@1", "w@1", "theta_m@1", "rho_zz@1", "scalars@1"]
NS_DEPENDENT = ("atm_advance_scalars", "atm_advance_scalars_mono", "atm_rk_integration_setup", "atm_compute_moist_coefficients")


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
    dist = None
    if N > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    threads = use_all_cores(N)

    def all_ok(ok):
        if N == 1:
            return ok
        t = torch.tensor([1 if ok else 0], dtype=torch.int32, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return bool(t.item())

    # ---- set-up, with the out-of-memory fallback (one icosahedral level at a time, decided collectively)
    level, ns, note = a.level, a.scalars, ""
    while True:
        t_setup = time.perf_counter()
        case = host_case(a, level)
        if N > 1:
            case.partition(N, "metis")
        # "lean": the work / tendency arrays (all zero at start, 83 % of a block's words) get no host storage, their device mirrors
        # are created zero-filled from the metadata
        block = case.build_block(rank, nVertLevels=a.levels, init_case=init_case(a), num_scalars=ns, full_fields="lean")
        if a.case == "jw":
            set_tracers(block, ns)
        cfg = model_config(a, level)
        dy, err = None, ""
        try:
            dy = host.Dycore([block], cfg, rank=rank, nranks=N, device=local_rank)
        except RuntimeError as e:
            err = str(e)
            if "out of memory" not in err:
                raise
        if all_ok(dy is not None):
            break
        if dy is not None:
            dy.close()
        dy = None
        if not a.auto_level or level <= 7:
            raise SystemExit("level %d does not fit the device memory: %s" % (level, err))
        del block, case
        note = " [fallback: level %d did not fit the device memory of %d GPU(s)]" % (a.level, N)
        level -= 1
        if a.auto_scalars:
            ns = 1 if level >= 9 else 4
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

    def sum_over_ranks_u64(x):
        if N == 1:
            return x
        t = torch.tensor([x - (1 << 64) if x >= (1 << 63) else x], dtype=torch.int64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)  # two's complement wrap-around = addition mod 2^64
        return int(t.item()) & ((1 << 64) - 1)

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
    # owned prognostic state after warmup + steps steps as one integer per field: equal for every N (decomposition test)
    checksum = {n.split("@")[0]: "0x%016x" % sum_over_ranks_u64(dy.checksum(n)) for n in CHECKSUM_FIELDS}
    transport = dy.halo_transport()
    device_gb = max_over_ranks(dy.device_bytes() / 1e9)

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
    words = words_per_call(ns, L)
    words_min = words_per_call(ns, L, minimal=True)
    phases = {}
    for name, (tms, calls) in timers.items():
        per_call = tms / max(calls, 1)
        e = {"ms_per_step": tms / nt, "calls_per_step": calls / nt, "ms_per_call": per_call}
        if name in words and per_call > 0:
            gb = words[name] * (4.0 if host.SINGLE else 8.0) * nCS * L / 1e9
            e["algorithmic_GB_per_call"] = gb
            e["achieved_GBs"] = gb / (per_call / 1000.0)
            e["frac"] = e["achieved_GBs"] / peak_gbs
            e["words_per_cell_level"] = words[name]
            e["frac_minimal_words"] = e["frac"] * words_min[name] / words[name]
        phases[name] = e
    # The device-resident driver runs every acoustic edge pass inside a divergence-damping launch (DESIGN.md section 3 rule 12), so the two
    # routines' timers are only meaningful together: their combined algorithmic bytes over their combined time.
    ph_ac, ph_dd = phases.get("atm_advance_acoustic_step"), phases.get("atm_divergence_damping_3d")
    if ph_ac and ph_dd and "algorithmic_GB_per_call" in ph_ac and "algorithmic_GB_per_call" in ph_dd:
        gb_pair = ph_ac["algorithmic_GB_per_call"] * ph_ac["calls_per_step"] + ph_dd["algorithmic_GB_per_call"] * ph_dd["calls_per_step"]
        ms_pair = ph_ac["ms_per_step"] + ph_dd["ms_per_step"]
        ph_ac["note"] = ph_dd["note"] = "edge pass folded into the damping launch: read with the other routine, see acoustic_plus_divergence_damping"
        phases["acoustic_plus_divergence_damping"] = {"ms_per_step": ms_pair, "algorithmic_GB_per_step": gb_pair, "achieved_GBs": gb_pair / (ms_pair / 1000.0),
                                                      "frac": gb_pair / (ms_pair / 1000.0) / peak_gbs}
    dom = max((n for n in phases if n in words), key=lambda n: phases[n]["ms_per_step"])
    roofline = {"bound": "hbm", "kernel": dom, "achieved": phases[dom]["achieved_GBs"], "peak": peak_gbs, "unit": "GB/s", "frac": phases[dom]["frac"],
                "traffic": None, "peak_source": peak_src,
                "note": "achieved = roofline/arrays.yaml words x sizeof(real) x owned cell-levels per call / CUDA-event time per call (eager pass after the timed region)"}
    # traffic: DRAM bytes (read + write) per call of the dominant routine from the committed ncu capture of these kernel sources
    # (profiles/r02_dram_traffic.json, tools/dram_traffic.py).  The capture is of configs[1] (level 7, 4 scalars); a capture of other
    # kernel sources is refused, other workloads scale it by the cell count (traffic per cell-level does not depend on the mesh size)
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "r02_dram_traffic.json")))
        if tr.get("_kernel_source_sha") != kernel_source_sha():
            roofline["traffic_note"] = "committed ncu capture is of other kernel sources (%s, now %s): not used" % (tr.get("_kernel_source_sha"), kernel_source_sha())
        elif N == 1 and a.case == "jw" and a.levels == int(tr["_nVertLevels"]) and not host.SINGLE and (
                ns == int(tr["_num_scalars"]) or dom not in NS_DEPENDENT):  # the dynamics routines move the same bytes for any num_scalars
            scale = n_cells(a, level) / float(tr["_nCells"])
            roofline["traffic"] = tr[dom]["dram_GB_per_step"] / phases[dom]["calls_per_step"] * scale
            roofline["traffic_unit"] = "GB per call (ncu dram__bytes_read.sum + dram__bytes_write.sum, profiles/r02_dram_traffic.csv%s)" % (
                "" if scale == 1.0 else ", captured on %d cells and scaled by the cell ratio" % int(tr["_nCells"]))
            roofline["algorithmic_GB_per_call"] = phases[dom]["algorithmic_GB_per_call"]
            if ns == int(tr["_num_scalars"]):
                roofline["whole_step_traffic_GB"] = tr["_whole_step"]["dram_GB_per_step"] * scale
        else:
            roofline["traffic_note"] = "the committed ncu capture is 1 GPU, %s levels, %s scalars, f64" % (tr["_nVertLevels"], tr["_num_scalars"])
    except Exception as e:
        roofline["traffic_note"] = "no usable capture: %s" % (e,)
    step_gb = sum(phases[n]["algorithmic_GB_per_call"] * phases[n]["calls_per_step"] for n in phases if "algorithmic_GB_per_call" in phases[n])
    # whole job: `value` steps/s, every rank streams its own step_gb per step (max-over-ranks time), roofline per GPU
    roofline["whole_step"] = {"algorithmic_GB": step_gb, "achieved_GBs": step_gb * value, "frac": step_gb * value / peak_gbs,
                              "note": "per GPU: this rank's owned cell-levels; time = the max-over-ranks step"}

    # ---- e2e: the same step through the public API with HOST buffers (pinned), copies inside the timed region
    e2e = None
    if not a.no_e2e:
        bufs = {}
        h2d = 0
        real_np = np.float32 if host.SINGLE else np.float64
        for n in E2E_FIELDS:
            kind, loc, dims = block.info(n)
            cnt = int(np.prod(dims))
            t = torch.empty(cnt, dtype=torch.float32 if host.SINGLE else torch.float64).pin_memory()
            arr = t.numpy()
            rc = L_.mpas_b200_dycore_get_field(dy.h, 0, n.encode(), arr.ctypes.data, arr.nbytes)
            assert rc == 0, host.last_error()
            bufs[n] = (t, arr)
            h2d += arr.nbytes
        ke = max(1, min(a.steps, 5 if level < 9 else 3))
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
               "what": "per step: upload the %d-field restart set from pinned host memory, one dycore step, download the same set (bytes: this rank)" % len(E2E_FIELDS)}
        del bufs

    # ---- CPU baseline (rank 0, N = 1): bounded sample of the same workload on the host cores, >= 10 steps, per-phase times
    cpu = None
    if rank == 0 and N == 1 and not a.no_cpu_baseline:
        nst = 10
        lvl = cpu_sample_level(a, ns, threads, nst + 1, a.cpu_seconds, min(level, 7))
        ccase, cblock, m = cpu_model(a, lvl, ns, threads)
        m.step(1)
        m.reset_timers()
        t0 = time.perf_counter()
        m.step(nst)
        dt_ = time.perf_counter() - t0
        scale = ccase.nCells / float(n_cells(a, level))
        cpu = {"value": nst / dt_ * scale, "unit": "steps/s", "cores": threads, "kind": "port",
               "sample": "%d steps (after 1 warm-up step) of the same case on icosahedral level %d (%d of %d cells) in %.1f s, scaled by the cell ratio; "
                         "C++/OpenMP restatement, -O3 -march=native; the Fortran reference cannot be built on this box" % (nst, lvl, ccase.nCells, n_cells(a, level), dt_),
               "phases_ms_on_sample_mesh": phase_ms(m, nst)}

    if rank == 0:
        dt_model = cfg["config_dt"]
        line = {
            "metric": "dycore timesteps/sec", "value": value, "unit": "steps/s", "n_gpus": N, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms / a.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32" if host.SINGLE else "f64", "data": "synthetic",
            "config": config_dict(a, level, ns, note),
            "run": {"simulated_days_per_day": value * dt_model, "cuda_graph": use_graph, "halo_exchanges_per_step": exchanges / a.steps, "halo_transport": transport,
                    "setup_seconds": t_setup, "host_threads": threads, "device_GB_per_rank": device_gb, "max_abs_u": umax, "max_abs_w": wmax,
                    "state_checksum": checksum, "state_checksum_after_steps": a.warmup + a.steps,
                    "state_checksum_what": "sum mod 2^64 of the bit patterns of the owned u, w, theta_m, rho_zz, scalars over all ranks: equal for every N"},
            "roofline": roofline, "phases": phases, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clk,
        }
        print(json.dumps(line), flush=True)
    dy.close()
    if N > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.precision == "single":
        os.environ["MPAS_B200_PRECISION"] = "single"
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
