#!/usr/bin/env python
"""Renders the scene script of the reference's demo page (/root/reference/ui/ui.js:10-44: area-light rectangle, Cornell box,
mirror sphere, glass sphere, `tonemapping` filter) at 512x512 through sail_b200.Renderer and saves the displayed image as PNG --
the headless stand-in for the canvas, for a QUALITATIVE comparison with the reference's README screenshot
(img/ex_cornellbox.png, 516x515, unknown sample count, captured from an RGB8-accumulating WebGL context: SURVEY.md Q1).

    python tools/render_ui_scene.py --spp 1024 --out profiles/r02_ui_scene_512.png [--reference /root/reference/img/ex_cornellbox.png]

With --reference (only present in the build container) it also prints a coarse comparison: per-channel means and the correlation
of 16x16 block means after cropping / resampling the screenshot to 512x512.
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import sail_b200 as Sail  # noqa: E402
from sail_b200 import image  # noqa: E402
from sail_b200.renderer import Renderer  # noqa: E402


def ui_scene():
    scene = Sail.Scene()
    camera = Sail.Camera([2.78, 2.73, -6], [2.78, 2.73, 2.79])
    matte = Sail.Matte(0.7)
    scene.add(Sail.AreaLight(Sail.Rectangle([2.13, 5.48, 2.27], [3.43, 5.48, 3.32], matte, Sail.Color.BLACK), [1, 1, 1]))
    scene.add(Sail.Cornellbox())
    scene.add(Sail.Sphere([1.5, 1.25, 2.70], 1.2, Sail.Mirror(1.0), Sail.Color.WHITE))
    scene.add(Sail.Sphere([3.9, 1.25, 1.70], 1.2, Sail.Glass(1, 1, 1.5), Sail.Color.WHITE))
    scene.add(camera)
    scene.filter = 'tonemapping'
    scene.trace = 'path'
    return scene


def block_means(img, n=16):
    H, W = img.shape[:2]
    return img[: H // n * n, : W // n * n, :3].reshape(n, H // n, n, W // n, 3).mean(axis=(1, 3))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--spp", type=int, default=1024)
    ap.add_argument("--out", default="profiles/r02_ui_scene_512.png")
    ap.add_argument("--reference", default="")
    ap.add_argument("--compare", default="", help="skip rendering: compare this PNG (rendered earlier on the GPU box) with --reference")
    args = ap.parse_args()
    if args.compare:
        ours = image.read_png(args.compare)[..., :3].astype(np.float64)
        ref = image.read_png(args.reference)[..., :3].astype(np.float64)[:512, :512]
        a, b = block_means(ours), block_means(ref)
        print(json.dumps({"ours": args.compare, "reference": args.reference, "mean_rgb": [float(v) for v in ours.mean(axis=(0, 1))],
                          "reference_mean_rgb": [float(v) for v in ref.mean(axis=(0, 1))],
                          "block16_correlation": float(np.corrcoef(a.reshape(-1), b.reshape(-1))[0, 1]),
                          "block16_mean_abs_diff_8bit": float(np.abs(a - b).mean())}))
        return
    scene = ui_scene()
    r = Renderer((512, 512))
    r.update(scene)
    img = None
    for i in range(args.spp // 32):
        img = np.array(r.render(scene, frames=32))
        if i == 0:
            r.ctx.specialise_wait()
    r.save(args.out, img)
    res = {"out": args.out, "spp": scene.sampleCount, "kernel": r.ctx.kernel_info(), "mean_rgb": [float(v) for v in img[..., :3].mean(axis=(0, 1))]}
    r.close()
    if args.reference and os.path.exists(args.reference):
        ref = image.read_png(args.reference)[..., :3].astype(np.float64)
        ref = ref[:512, :512] if ref.shape[0] >= 512 and ref.shape[1] >= 512 else ref
        ours = img[::-1, :, :3].astype(np.float64)  # PNG rows run top to bottom
        a, b = block_means(ours), block_means(ref)
        res["reference_mean_rgb"] = [float(v) for v in ref.mean(axis=(0, 1))]
        res["block16_correlation"] = float(np.corrcoef(a.reshape(-1), b.reshape(-1))[0, 1])
    print(json.dumps(res))


if __name__ == "__main__":
    main()
