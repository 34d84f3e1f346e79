"""PRECISION=single build on the GPU: libmpas_b200_sp.so (float32 kernels, the same sources with -DSINGLE_PRECISION)
against the DOUBLE oracle after 1 and 10 steps.  north_star tolerance for a single build: 1e-5 relative L2.

The precision is chosen when the library is loaded, so the check runs in its own process (tools/sp_check.py).
w is a small residual of large cancelling terms (|w| ~ 1e-3 m/s against a hydrostatic balance of ~10 m/s^2, DESIGN.md
section 4): its float32 round-off is asserted against the velocity scale, the other prognostic fields at 1e-5."""
import json
import os
import subprocess
import sys

import pytest

import common

pytestmark = pytest.mark.gpu

TOL_SINGLE = 1e-5


def _run(*args):
    env = dict(os.environ, MPAS_B200_PRECISION="single")
    r = subprocess.run([sys.executable, os.path.join(common.ROOT, "tools", "sp_check.py"), *args], env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return json.loads(r.stdout.strip().splitlines()[-1])


@pytest.mark.parametrize("blocks", [1, 2])
def test_single_precision_build_within_1e5_of_the_double_oracle(blocks):
    out = _run("--level", "4", "--levels", "26", "--scalars", "2", "--blocks", str(blocks))
    print(json.dumps(out))
    assert out["real_size"] == 4 and out["launches"] > 0
    for nsteps in ("1", "10"):
        res = out["after"][nsteps]
        for name in ("u@1", "theta_m@1", "rho_zz@1", "scalars@1"):
            assert res[name] <= TOL_SINGLE, (nsteps, name, res[name])
        assert res["w_abs_over_max_u"] <= TOL_SINGLE, (nsteps, res)
        for name, e in res["vs_float_oracle"].items():  # float kernels against the float oracle
            assert e <= TOL_SINGLE, (nsteps, name, e)
