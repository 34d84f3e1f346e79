"""Lean host blocks (pool.h Field::zero): the dycore work arrays that are all zero once a block has been built get no host
storage; everything else -- names, shapes, contents once looked at, exchange lists -- is identical to a full block.  This is
what lets 2..8 ranks set up the 15 km mesh inside the host's memory."""
import numpy as np

import common


def test_lean_block_equals_full_block_and_only_zero_fields_lack_storage():
    host = common.host
    case = host.HostCase(3, lloyd_iters=3, sfc=True)
    case.partition(2, "sfc")
    case2 = host.HostCase(3, lloyd_iters=3, sfc=True)
    case2.partition(2, "sfc")
    full = case.build_block(1, nVertLevels=8, num_scalars=3, full_fields=True)
    lean = case2.build_block(1, nVertLevels=8, num_scalars=3, full_fields="lean")
    assert full.names() == lean.names()
    lean_names = [n for n in lean.names() if lean.is_lean(n)]
    assert not any(full.is_lean(n) for n in full.names())
    words = lambda b, names: sum(int(np.prod(b.info(n)[2])) for n in names if b.info(n)[0] == 0)
    frac = words(lean, lean_names) / words(lean, lean.names())
    print("fields without host storage: %d of %d, %.0f %% of the real words" % (len(lean_names), len(lean.names()), 100 * frac))
    assert frac > 0.6
    for n in ("tend_u", "ru_p", "rtheta_pp", "scalars@2", "flux_arr", "scale_arr"):
        assert n in lean_names
    for n in ("u@1", "rho", "theta", "zgrid", "zz", "adv_coefs", "scalars@1"):
        assert n not in lean_names
    for n in full.names():
        assert full.info(n) == lean.info(n), n
        a = full.field(n)
        if n in lean_names:
            assert not a.any(), n  # exactly the all-zero arrays
        b = lean.field(n)  # gives a lean field its storage
        assert np.array_equal(a, b), n
        assert not lean.is_lean(n)
    assert full.cfg() == lean.cfg()
