"""PRECISION=single check (north_star: 1e-5 relative L2 for a PRECISION=single build).

Run as its own process (one precision per process, mpas-model_b200/host.py):

    MPAS_B200_PRECISION=single python tools/sp_check.py [--level 4] [--levels 26] [--scalars 2] [--blocks 1]

The float32 device dycore (libmpas_b200_sp.so, the same kernel sources with -DSINGLE_PRECISION) and the double CPU
oracle start from the same JW state (the oracle's in double, the device's rounded to float on upload) and take 1 and
10 steps; one JSON line with the relative L2 difference of every prognostic field after each is printed, plus the
difference to the FLOAT build of the oracle (oracle/_build/liboracle_sp.so) under "vs_float_oracle".
The reference's single build rounds the same way: RKIND reals everywhere (src/framework/mpas_kind_types.F:25-29).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--level", type=int, default=4)
    ap.add_argument("--levels", type=int, default=26)
    ap.add_argument("--scalars", type=int, default=2)
    ap.add_argument("--blocks", type=int, default=1)
    a = ap.parse_args()
    if os.environ.get("MPAS_B200_PRECISION", "").lower() != "single":
        raise SystemExit("set MPAS_B200_PRECISION=single (the precision is chosen when the library is loaded)")
    import numpy as np
    import common

    assert common.host.SINGLE and common.host.lib().mpas_b200_real_size() == 4
    case, blocks = common.make_case(a.level, a.blocks, nVertLevels=a.levels, num_scalars=a.scalars)
    for b in blocks:  # a smooth tracer in the second scalar slot
        if a.scalars > 1:
            s = b.field("scalars@1")
            lat = b.field("latCell")
            s[:-1, :, 1] = (np.cos(lat) ** 2)[:-1, None]
    cfg = common.level_config(a.level)
    gpu = common.GpuModel(blocks, cfg)
    ora32 = common.OracleModel(blocks, cfg, single=True)  # float build of the oracle: its own float copies of the initial pools
    ora = common.OracleModel(blocks, cfg)                 # double oracle, in place on the host pools
    gpu.init()
    ora32.init()
    ora.init()
    names = ["u@1", "w@1", "theta_m@1", "rho_zz@1", "scalars@1"]
    out = {"real_size": 4, "level": a.level, "nVertLevels": a.levels, "num_scalars": a.scalars, "blocks": a.blocks, "after": {}}
    done = 0
    for target in (1, 10):
        gpu.step(target - done)
        ora.step(target - done)
        ora32.step(target - done)
        done = target
        res = {}
        # float kernels against the FLOAT oracle (same rounding of the inputs; the oracle evaluates expressions with double
        # literals in double, the kernels in float, so this is close but not bitwise)
        res["vs_float_oracle"] = {n: max(float(common.rel_l2(gpu.field(ib, n)[:int(b.cfg()[{1: "nCellsSolve", 2: "nEdgesSolve", 3: "nVerticesSolve"}[b.info(n)[1]]])],
                                                             ora32.field(ib, n)[:int(b.cfg()[{1: "nCellsSolve", 2: "nEdgesSolve", 3: "nVerticesSolve"}[b.info(n)[1]]])]))
                                         for ib, b in enumerate(blocks)) for n in names if n != "w@1"}
        for n in names:
            worst = 0.0
            for ib, b in enumerate(blocks):
                kind, loc, dims = b.info(n)
                nown = int(b.cfg()[{1: "nCellsSolve", 2: "nEdgesSolve", 3: "nVerticesSolve"}[loc]])
                # the time levels swap every step: "@1" of the model objects is always the current state
                g, o = gpu.field(ib, n)[:nown], ora.field(ib, n)[:nown]
                if not np.isfinite(g).all():
                    worst = float("nan")
                    break
                worst = max(worst, float(common.rel_l2(g, o)))
            res[n] = worst
        # w is a small residual (|w| ~ 1e-3 m/s): also report its error against the velocity scale max|u|
        res["w_abs_over_max_u"] = float(max(np.abs(gpu.field(ib, "w@1") - ora.field(ib, "w@1")).max() for ib in range(len(blocks))) /
                                        max(np.abs(ora.field(ib, "u@1")).max() for ib in range(len(blocks))))
        out["after"][str(target)] = res
    out["launches"] = int(gpu.launch_count())
    print(json.dumps(out))


if __name__ == "__main__":
    main()
