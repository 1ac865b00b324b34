"""N>1 path on CPU: hovl.dist's exchange layer (all-to-all-v, all-gather-v, gather, stat
reduction) over gloo with world_size 2 and 3, the oracle standing in for the device stages."""
import os
import socket
import subprocess
import sys

import pytest

from conftest import ROOT


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_path_over_gloo(oracle, world):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "dist", "cpu_worker.py")]
    env = dict(os.environ, OMP_NUM_THREADS="2")
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert "DIST_CPU_OK world=%d" % world in p.stdout


def test_read_slices_cover_everything():
    from hovl import dist as hdist
    for n in (0, 1, 7, 200000):
        for w in (1, 2, 3, 8):
            cuts = [hdist.read_slice(n, r, w) for r in range(w)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
