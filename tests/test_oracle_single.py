"""Cross pin of SURVEY.md section 8(c): the oracle built with mpas_real = float (the reference's PRECISION=single,
src/framework/mpas_kind_types.F:25-29) against the double oracle, same JW case, 1 and 10 steps: 1e-5 relative L2
(north_star's tolerance for a single build).  w is a small residual of large cancelling terms:
it is held against the velocity scale, like in tests/test_gpu_single.py."""
import numpy as np

import common

TOL_SINGLE = 1e-5


def test_float_oracle_within_1e5_of_the_double_oracle():
    lvl, L, ns = 3, 12, 2
    cfg = common.level_config(lvl)
    case_d, blocks_d = common.make_case(lvl, 1, nVertLevels=L, num_scalars=ns)
    case_s, blocks_s = common.make_case(lvl, 1, nVertLevels=L, num_scalars=ns)
    for blocks in (blocks_d, blocks_s):
        s = blocks[0].field("scalars@1")
        s[:-1, :, 1] = (np.cos(blocks[0].field("latCell")) ** 2)[:-1, None]
    dbl = common.OracleModel(blocks_d, cfg)
    sgl = common.OracleModel(blocks_s, cfg, single=True)
    dbl.init()
    sgl.init()
    done = 0
    for target in (1, 10):
        dbl.step(target - done)
        sgl.step(target - done)
        done = target
        for n in ("u@1", "theta_m@1", "rho_zz@1", "scalars@1"):
            a, b = sgl.field(0, n), dbl.field(0, n)
            assert a.dtype == np.float32 and np.isfinite(a).all()
            e = common.rel_l2(a, b)
            print(target, n, e)
            assert e <= TOL_SINGLE, (target, n, e)
        werr = np.abs(sgl.field(0, "w@1").astype(np.float64) - dbl.field(0, "w@1")).max() / np.abs(dbl.field(0, "u@1")).max()
        print(target, "w_abs_over_max_u", werr)
        assert werr <= TOL_SINGLE, (target, werr)
