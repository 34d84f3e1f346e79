"""Per-kernel and per-routine DRAM traffic from an ncu CSV with dram__bytes_read.sum, dram__bytes_write.sum and
gpu__time_duration.sum of every launch of `bench.py --steps 2 --warmup 3 --no-graph` (7 steps after init).
usage: python tools/dram_traffic.py gpurun_out/dram.csv profiles/r02_dram_traffic.json [level levels scalars] > profiles/r02_dram_traffic.csv
The json carries the hash of the kernel sources it was captured from (bench.py refuses a capture of other sources) and the
workload it was captured on (bench.py scales it by the cell ratio for other meshes)."""
import csv
import json
import re
import sys
from collections import OrderedDict

ROUTINE = [("dt_", "atm_compute_dyn_tend"), ("acoustic_", "atm_advance_acoustic_step"), ("divdamp", "atm_divergence_damping_3d"),
           ("recover_", "atm_recover_large_step_variables"), ("set_garbage", "atm_recover_large_step_variables"), ("smlstep", "small_step_prep"),
           ("sd_", "atm_compute_solve_diagnostics"), ("vert_imp", "atm_compute_vert_imp_coefs"), ("cofrz", "atm_compute_vert_imp_coefs"),
           ("as_", "atm_advance_scalars"), ("mono_", "atm_advance_scalars_mono"), ("monoN_", "atm_advance_scalars_mono"), ("ax_", "atm_rk_dynamics_substep_finish"),
           ("moist_", "atm_compute_moist_coefficients"), ("substep_finish", "atm_rk_dynamics_substep_finish"), ("halo_", "halo_exchange"),
           ("rk_setup", "atm_rk_integration_setup"), ("reconstruct", "mpas_reconstruct"), ("zero_cols", "scratch_garbage_columns"), ("clamp_nonneg", "atm_advance_scalars_mono")]
NSTEPS = 7  # 3 warm-up + 2 timed + 2 steps of the eager per-routine timing pass


def main(path, json_out):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    iname, imet, ival, iunit, iid = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit"), hdr.index("ID")
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "nsecond": 1e-3}
    launches = OrderedDict()
    for r in rows[1:]:
        name = re.sub(r"\(.*", "", r[iname]).replace("<unnamed>::", "").replace("void ", "")
        d = launches.setdefault(r[iid], {"name": name})
        d[r[imet]] = float(r[ival].replace(",", "")) * scale.get(r[iunit], 1.0)
    per = OrderedDict()
    for d in launches.values():
        if not any(d["name"].startswith(p) or ("::" + p) in d["name"] for p, _ in ROUTINE):
            continue
        a = per.setdefault(d["name"], [0, 0.0, 0.0, 0.0])
        a[0] += 1
        a[1] += d.get("dram__bytes_read.sum", 0.0)
        a[2] += d.get("dram__bytes_write.sum", 0.0)
        a[3] += d.get("gpu__time_duration.sum", 0.0)
    w = csv.writer(sys.stdout)
    w.writerow(["kernel", "launches", "dram_read_MB_per_launch", "dram_write_MB_per_launch", "mean_us", "dram_GBs"])
    rout = OrderedDict()
    for name, (n, rd, wr, t) in sorted(per.items(), key=lambda kv: -kv[1][3]):
        w.writerow([name, n, "%.1f" % (rd / n / 1e6), "%.1f" % (wr / n / 1e6), "%.1f" % (t / n), "%.0f" % ((rd + wr) / t * 1e-3 if t else 0)])
        for p, rname in ROUTINE:
            if name.startswith(p):
                a = rout.setdefault(rname, [0.0, 0.0])
                a[0] += rd + wr
                a[1] += t
                break
    out = {k: {"dram_GB_per_step": v[0] / NSTEPS / 1e9, "ncu_ms_per_step": v[1] / NSTEPS / 1e3} for k, v in rout.items()}
    out["_whole_step"] = {"dram_GB_per_step": sum(v[0] for v in rout.values()) / NSTEPS / 1e9, "ncu_ms_per_step": sum(v[1] for v in rout.values()) / NSTEPS / 1e3}
    import os
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
    import bench
    level, levels, scalars = (int(x) for x in (sys.argv[3:6] if len(sys.argv) >= 6 else (7, 55, 4)))
    out["_kernel_source_sha"] = bench.kernel_source_sha()
    out["_nCells"] = bench.LEVEL_CELLS[level]
    out["_nVertLevels"] = levels
    out["_num_scalars"] = scalars
    out["_command"] = "ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none python bench.py --level %d --levels %d --scalars %d --steps 2 --warmup 3 --no-graph --no-cpu-baseline --no-e2e" % (level, levels, scalars)
    json.dump(out, open(json_out, "w"), indent=1)


main(sys.argv[1], sys.argv[2])
