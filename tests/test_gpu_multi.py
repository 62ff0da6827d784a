"""Multi-GPU sharding on real devices, driven through the C ABI: sail_create_info.ndev (one context over
several GPUs of this process; peer-memory and NCCL combine) and sail_group_join (one GPU per process).  Skipped on a one-GPU
box.  The host-side split logic has a CPU twin in tests/test_cpu_dist.py."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT


def _gpus():
    from sail_b200 import capi
    return min(capi.device_count(), 4)


@pytest.mark.gpu
@pytest.mark.timeout(600)
@pytest.mark.parametrize("mode", ["one_process", "processes"])
def test_sample_and_tile_sharding_through_the_c_abi(tmp_path, mode):
    n = _gpus()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (found %d)" % n)
    out = str(tmp_path / "multi.json")
    cmd = [sys.executable, os.path.join(ROOT, "tools", "multigpu_check.py"), "--gpus", str(n), "--out", out]
    if mode == "processes":
        cmd.append("--processes")
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=570)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-4000:]
    res = json.load(open(out))
    assert res["world"] == n and res["ok"], res


@pytest.mark.gpu
def test_sharded_context_argument_checks():
    from sail_b200 import capi
    with pytest.raises(capi.SailError) as ei:
        capi.Context(64, 64, devices=[0, 0], shard=capi.SHARD_SAMPLE)
    assert ei.value.code == capi.ERR_INVALID and "twice" in str(ei.value)
    with pytest.raises(capi.SailError):
        capi.Context(64, 64, devices=[0, 99], shard=capi.SHARD_TILE)
    with capi.Context(64, 64) as ctx:
        assert ctx.group_info()[:2] == (1, 0)
        with pytest.raises(capi.SailError):
            ctx.last_combine_ms()
        with pytest.raises(capi.SailError):
            ctx.group_join(1, 0, bytes(128))
