"""Sharded path on real GPUs: world_size 1 always, world_size 2 when the box has two GPUs
(gpurun --gpus 2).  Rank 0 compares with the single-GPU path record for record."""
import os
import socket
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [1, 2, 4])
def test_sharded_equals_single_gpu(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "dist", "gpu_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert "DIST_GPU_OK world=%d" % world in p.stdout


@pytest.mark.parametrize("world", [2, 4, 8])
def test_one_handle_over_several_devices_equals_single_gpu(world, oracle):
    """hovl_create(devices, n > 1) + hovl_run: the sharded path inside ONE process (csrc/multi.cu, peer memory, one host
    thread per device) returns exactly the one-device records; run twice (buffers are reused), ragged reads included."""
    import numpy as np
    import torch
    import hovl
    from conftest import synth_reads
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    cases = [(600, 100, 6000, 12, 0, {}), (3000, 150, 40000, 12, 40, {}),
             (2500, 1000, 90000, 16, 0, {"max_ignore": 1000, "max_coll": 2**31 - 1})]
    with hovl.Context(list(range(world))) as multi, hovl.Context(0) as single:
        for n, rl, glen, k, ragged, kw in cases:
            bases, offs = synth_reads(n + rl, n, rl, glen, sub=0.01, ragged=ragged)
            s = hovl.default_settings(k=k, order=1, id_mode=1, **kw)
            ref = single.run(s, bases, offs)
            for _ in range(2):
                res = multi.run(s, bases, offs)
                assert len(res.records) == len(ref.records) > 0
                assert np.ascontiguousarray(res.records).tobytes() == np.ascontiguousarray(ref.records).tobytes()
                for key in ("n_kmers", "n_pair_instances", "n_distinct_pairs", "n_dispatched", "n_leads", "n_duds",
                            "n_align_valid", "n_ovl_valid", "dp_cells", "n_records"):
                    assert res.stats[key] == ref.stats[key], key
        with pytest.raises(hovl.HovlError) as e:               # reference order needs one table
            multi.run(hovl.default_settings(), bases, offs)
        assert e.value.code == hovl.E_UNSUPPORTED


def test_cli_over_several_devices_writes_the_same_file(tmp_path):
    """hovl_cli --hovl-devices 0,1: the C++ host over one handle on several GPUs (no Python, no launcher) writes the same
    .ovl as the one-GPU run with --hovl-wide-ids (canonical record order)."""
    import subprocess
    import torch
    import hovl
    from conftest import synth_reads
    from hovl import io as hio
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    bases, offs = synth_reads(9, 3000, 150, 40000, sub=0.01, ragged=40)
    fa = str(tmp_path / "reads.fa")
    hio.write_fasta(fa, bases, offs)
    one, two = str(tmp_path / "one.ovl"), str(tmp_path / "two.ovl")
    for out, extra in ((one, ["--hovl-wide-ids"]), (two, ["--hovl-devices", "0,1"])):
        p = subprocess.run([hovl.CLI_PATH, "-i", fa, "-o", out] + extra, capture_output=True, timeout=300)
        assert p.returncode == 0, p.stderr
    a, b = open(one, "rb").read(), open(two, "rb").read()
    assert a == b and a.count(b"{OVL") > 1000
