"""Generate the golden fixture tests/golden/jw_l3_L8.npz from the CPU oracle (regression pin; the reference itself ships no
golden vectors for this path -- SURVEY.md F4 -- and cannot be run here, so this pins the oracle against drift and gives the
GPU tests a fixture that does not need the oracle at run time).   python tests/golden/make_golden.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import common  # noqa: E402

LEVEL, L, NS, STEPS = 3, 8, 2, 2
FIELDS = ["u@1", "w@1", "theta_m@1", "rho_zz@1", "scalars@1", "ru", "rw", "exner", "pressure_p", "pv_edge", "rho_edge"]


def make_blocks():
    case, blocks = common.make_case(LEVEL, 1, nVertLevels=L, num_scalars=NS)
    b = blocks[0]
    b.field("scalars@1")[:-1, :, 1] = (np.sin(2 * b.field("latCell")) ** 2)[:-1, None]
    return case, blocks


if __name__ == "__main__":
    case, blocks = make_blocks()
    m = common.OracleModel(blocks, common.level_config(LEVEL))
    m.init()
    out = {"init:" + n: m.field(0, n).copy() for n in FIELDS}
    m.step(STEPS)
    out.update({"step%d:" % STEPS + n: m.field(0, n).copy() for n in FIELDS})
    np.savez_compressed(os.path.join(HERE, "jw_l3_L8.npz"), **out)
    print("wrote", os.path.join(HERE, "jw_l3_L8.npz"), sum(v.nbytes for v in out.values()), "bytes raw")
