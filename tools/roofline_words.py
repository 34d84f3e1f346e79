#!/usr/bin/env python
"""Algorithmic words per owned cell-level and call of every routine, generated from roofline/arrays.yaml.

  python tools/roofline_words.py [--ns 4] [--levels 55] [--with-optional] [--dram profiles/r02_dram_traffic.json --cells 163842]

bench.py imports words_per_call(); run as a script it prints the table next to the hand counts of SURVEY.md section 8(d)
and, given an ncu DRAM capture (tools/dram_traffic.py), the measured DRAM bytes over the algorithmic bytes per routine.
"""
import argparse
import json
import os

import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TABLE = os.path.join(ROOT, "roofline", "arrays.yaml")
MAX_EDGES = 6  # icosahedral Voronoi mesh

# hand counts of SURVEY.md section 8(a)/(d) (they include the physics-coupling arrays), kept for comparison
SURVEY_WORDS = {
    "atm_compute_dyn_tend": {"rk1": 79, "rk23": 39}, "atm_advance_acoustic_step": {"first_substep": 31, "later_substep": 48},
    "atm_divergence_damping_3d": {"all": 9}, "atm_recover_large_step_variables": {"rk12": 47, "rk3": 51}, "small_step_prep": {"all": 18},
    "atm_compute_vert_imp_coefs": {"all": 16}, "atm_compute_solve_diagnostics": {"rk12": 39, "rk3": 42},
    "atm_rk_dynamics_substep_finish": {"substep1": 30, "substep2": 30, "last": 30},
}


def _mult(loc, L):
    return {"cell": 1.0, "cell1": (L + 1.0) / L, "edge": 3.0, "vertex": 2.0, "cell_mE": MAX_EDGES * (L + 1.0) / L}[loc]


def _x(e, ns, n_moist):
    x = e.get("x", 1)
    if x == "ns":
        return ns
    if x == "2ns":
        return 2 * ns
    if x == "n_moist":
        return n_moist
    return float(x)


def _resolve(variants, name):
    v = dict(variants[name])
    if "like" in v:
        base = _resolve(variants, v["like"])
        read = [e for e in base["read"] if e["a"] not in v.get("read_remove", [])] + v.get("read_extra", [])
        write = [e for e in base["write"] if e["a"] not in v.get("write_remove", [])] + v.get("write_extra", [])
        v["read"], v["write"] = read, write
    return v


def load():
    return yaml.safe_load(open(TABLE))


def variant_words(ns=4, L=55, with_optional=False, n_moist=1):
    """{routine: {variant: (words, calls_per_dt)}}"""
    out = {}
    for routine, spec in load().items():
        out[routine] = {}
        for vname in spec["variants"]:
            v = _resolve(spec["variants"], vname)
            w = 0.0
            for e in v["read"] + v["write"]:
                if e.get("optional") and not with_optional:
                    continue
                w += _mult(e["loc"], L) * _x(e, ns, n_moist)
            out[routine][vname] = (w, v["calls_per_dt"])
    return out


def words_per_call(ns=4, L=55, with_optional=False):
    """{routine: words per owned cell-level, averaged over the calls of one dt} -- what bench.py multiplies by
    sizeof(real) x owned cell-levels and divides by the measured time per call"""
    out = {}
    for routine, vs in variant_words(ns, L, with_optional).items():
        calls = sum(c for _, c in vs.values())
        out[routine] = sum(w * c for w, c in vs.values()) / calls
    return out


def words_per_dt(ns=4, L=55, with_optional=False):
    return sum(sum(w * c for w, c in vs.values()) for vs in variant_words(ns, L, with_optional).values())


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--ns", type=int, default=4)
    ap.add_argument("--levels", type=int, default=55)
    ap.add_argument("--with-optional", action="store_true")
    ap.add_argument("--dram", default=None, help="json of tools/dram_traffic.py: prints measured DRAM bytes / algorithmic bytes per routine")
    ap.add_argument("--cells", type=int, default=163842)
    a = ap.parse_args()
    vw = variant_words(a.ns, a.levels, a.with_optional)
    full = variant_words(a.ns, a.levels, True)
    print("%-36s %-14s %9s %9s %8s %6s" % ("routine", "variant", "words", "+optional", "SURVEY", "calls"))
    for r, vs in vw.items():
        for v, (w, c) in vs.items():
            print("%-36s %-14s %9.2f %9.2f %8s %6d" % (r, v, w, full[r][v][0], SURVEY_WORDS.get(r, {}).get(v, "-"), c))
    print("per dt: %.0f words (%.0f with the optional physics / output arrays) per owned cell-level, ns = %d" % (
        words_per_dt(a.ns, a.levels), words_per_dt(a.ns, a.levels, True), a.ns))
    if a.dram:
        tr = json.load(open(a.dram))
        wpc = words_per_call(a.ns, a.levels, a.with_optional)
        print("\n%-36s %12s %12s %8s" % ("routine", "algo GB/dt", "DRAM GB/dt", "ratio"))
        for r, vs in vw.items():
            if r not in tr:
                continue
            calls = sum(c for _, c in vs.values())
            gb = wpc[r] * calls * 8.0 * a.cells * a.levels / 1e9
            print("%-36s %12.2f %12.2f %8.2f" % (r, gb, tr[r]["dram_GB_per_step"], tr[r]["dram_GB_per_step"] / gb))
