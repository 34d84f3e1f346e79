"""Print GPU-vs-oracle relative L2 errors of every prognostic / diagnostic field after N steps (run on a GPU box)."""
import sys, os, json
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import numpy as np
import common
from test_gpu_model import _tracers, STATE, DIAG

def main():
    level = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    L = int(sys.argv[2]) if len(sys.argv) > 2 else 26
    ns = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    nb = int(sys.argv[4]) if len(sys.argv) > 4 else 1
    case, blocks = common.make_case(level, nb, nVertLevels=L, num_scalars=ns)
    _tracers(blocks, ns)
    cfg = common.level_config(level)
    gpu = common.GpuModel(blocks, cfg); ora = common.OracleModel(blocks, cfg)
    gpu.init(); ora.init()
    done = 0
    out = {}
    for target in (1, 2, 5, 10):
        gpu.step(target - done); ora.step(target - done); done = target
        errs = {}
        for n in STATE + DIAG:
            e = 0.0
            for ib, b in enumerate(blocks):
                kind, loc, dims = b.info(n)
                nown = int(b.cfg()[{1: "nCellsSolve", 2: "nEdgesSolve", 3: "nVerticesSolve"}[loc]])
                e = max(e, common.rel_l2(gpu.field(ib, n)[:nown], ora.field(ib, n)[:nown]))
            errs[n] = e
        out[target] = errs
        print("after %2d steps: " % target + "  ".join("%s=%.1e" % (k, v) for k, v in errs.items() if k in STATE), flush=True)
        print("     worst diag: ", sorted(((v, k) for k, v in errs.items() if k in DIAG), reverse=True)[:4], flush=True)
    print(json.dumps({"level": level, "L": L, "ns": ns, "blocks": nb, "errors": out}))

main()
