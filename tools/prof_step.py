"""Two eager dycore steps of configs[1] (level 7, 55 levels, 4 scalars) for ncu:  ncu ... python tools/prof_step.py [level] [scalars]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import common  # noqa: E402

level = int(sys.argv[1]) if len(sys.argv) > 1 else 7
ns = int(sys.argv[2]) if len(sys.argv) > 2 else 4
case = common.host.HostCase(level, lloyd_iters=5, sfc=True)
block = case.build_block(0, nVertLevels=55, init_case=2, num_scalars=ns, full_fields="lean")
dy = common.host.Dycore([block], common.level_config(level))
dy.init()
dy.step(2)
dy.sync()
print("ok", dy.max_abs("u@1"))
