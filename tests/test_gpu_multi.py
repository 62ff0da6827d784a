"""Multi-GPU sharding on real devices; skipped on a one-GPU box.  The N = 2 gloo twin that runs on
CPU is tests/test_cpu_dist.py."""
import json
import os
import socket
import subprocess
import sys

import pytest

from conftest import ROOT


@pytest.mark.gpu
@pytest.mark.timeout(600)
def test_sample_and_tile_sharding_over_nccl(tmp_path):
    from sail_b200 import capi
    n = min(capi.device_count(), 4)
    if n < 2:
        pytest.skip("needs >= 2 GPUs (found %d)" % n)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "multi.json")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tools", "multigpu_check.py"), "--out", out]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=570)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    res = json.load(open(out))
    assert res["world"] == n and res["ok"], res
