"""Multi-rank correctness on real GPUs (skipped with fewer than two): one process per GPU under torch.distributed.run runs
tools/multigpu_check.py -- (a) the decomposition test (N ranks == 1 block, bitwise, on every owned element of the prognostic
state; COMPASS decomp_test idea), (b) the reference's halo test idiom (mpas_test_core_halo_exch.F:631, 1010) across
processes -- over NCCL grouped send/recv, over the peer-store transport (IPC windows, sequence flags), eager and with the
step captured in CUDA graphs."""
import os
import socket
import subprocess
import sys

import pytest

import common

pytestmark = pytest.mark.gpu
ROOT = common.ROOT


def _ngpus():
    try:
        return common.host.lib().mpas_b200_device_count()
    except Exception:
        return 0


def _port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(nranks, level, steps, graph, env_extra):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nranks), "--master-addr", "127.0.0.1",
           "--master-port", str(_port()), os.path.join(ROOT, "tools", "multigpu_check.py"), str(level), str(steps), str(graph)]
    env = dict(os.environ, OMP_NUM_THREADS="4", **env_extra)
    return subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)


@pytest.mark.parametrize("transport,graph", [("nccl", 0), ("p2p", 0), ("p2p", 1), ("nccl", 1)])
def test_ranks_reproduce_one_block_bitwise_and_pass_the_halo_idiom(transport, graph):
    n = _ngpus()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    nranks = 4 if n >= 4 else 2
    r = _run(nranks, 5, 5, graph, {"MPAS_B200_TUNE_p2p": "1" if transport == "p2p" else "0"})
    out = r.stdout[-3000:] + r.stderr[-3000:]
    assert r.returncode == 0, out
    assert "BITWISE EQUAL" in r.stdout and "HALO IDIOM OK" in r.stdout, out
    assert ("transport=%s" % transport) in r.stdout, out
