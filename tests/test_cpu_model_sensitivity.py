"""How far the oracle's arithmetic model (a / b := a * rcp(b), normalize(v) := v * rcp(length v); oracle/glsl.h) is from the
other defensible reading of GLSL's '/' (IEEE division; liboracle_ieee.so, -DSAIL_ORACLE_IEEE_DIV): the bar BASELINE.json's
north_star states -- PSNR >= 40 dB between the images, hit ids equal -- on a small, fast sample.  The full measurement
(512x512, 1024 spp, configs 1-3: PSNR 80-106 dB, <= 10 of 1e6 paths on another object by bounce 5) is
baseline/model_sensitivity.json, made by tools/model_sensitivity.py."""
import json
import os
import sys

import pytest

from conftest import ROOT

sys.path.insert(0, os.path.join(ROOT, "tools"))


@pytest.mark.parametrize("cfg", [1, 2, 3])
def test_the_two_division_models_agree_far_inside_the_tolerance(cfg):
    import model_sensitivity
    r = model_sensitivity.compare(cfg, 96, 96, 24, dump_frames=2)
    assert r["psnr_clamped_db"] >= 40.0 and r["psnr_unclamped_db"] >= 40.0, r
    assert r["hit_id_agreement_per_bounce"][0] >= 0.9999, r          # primary rays: the same object
    assert min(r["hit_id_agreement_per_bounce"]) >= 0.999, r         # later bounces: a handful of paths at silhouettes
    assert r["mean_abs"] <= 1e-3, r


def test_committed_full_measurement_meets_the_bar():
    with open(os.path.join(ROOT, "baseline", "model_sensitivity.json")) as f:
        full = json.load(f)
    assert sorted(r["config"] for r in full["results"]) == [1, 2, 3]
    for r in full["results"]:
        assert r["size"] == [512, 512] and r["spp"] == 1024
        assert r["psnr_clamped_db"] >= 40.0 and r["psnr_unclamped_db"] >= 40.0
        assert min(r["hit_id_agreement_per_bounce"]) >= 0.9999
