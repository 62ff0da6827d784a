"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/sail_b200.h declares, its structs have the documented layout, validation works without a
device, and the product path fails loudly (no CPU fallback) when no GPU is present."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT, scene_tables

HEADER = os.path.join(ROOT, "include", "sail_b200.h")


@pytest.fixture(scope="module")
def capi():
    from sail_b200 import capi as c
    if not os.path.exists(c.LIB_PATH):
        c.build()
    return c


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sail_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_and_library_exports(capi):
    names = declared_functions()
    assert len(names) >= 25
    lib = ctypes.CDLL(capi.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, "declared in include/sail_b200.h but not exported: %s" % missing
    # the Python binding covers exactly the declared surface
    assert sorted(capi.EXPORTED_SYMBOLS) == names


def test_struct_layouts(capi):
    assert ctypes.sizeof(capi.CreateInfo) == 5 * 4 + 3 * 4 + 8 * 4 + 4 + 3 * 4 and capi.CreateInfo.device_ids.offset == 32
    assert ctypes.sizeof(capi.Config) == 4 * 4 + 5 * 4 + 4 + 8 + 64 + 16
    assert ctypes.sizeof(capi.Frame) == 84
    assert capi.Frame.eye.offset == 64 and capi.Frame.textureWeight.offset == 76
    assert capi.Config.filter_weights.offset == 48


def test_version_and_defaults(capi):
    lib = capi.load()
    assert lib.sail_abi_version() == capi.ABI_VERSION == 2
    assert b"sm_100a" in lib.sail_version()
    cfg = capi.Config()
    lib.sail_default_config(ctypes.byref(cfg))
    assert cfg.max_bounces == 5 and cfg.trace_id == 0 and cfg.accum_mode == capi.ACCUM_MEAN
    assert cfg.filter_id == 0 and cfg.kernel == capi.KERNEL_AUTO


def test_no_cpu_fallback(capi):
    """Without a B200 the library must refuse to create a context (and never render on the host)."""
    if capi.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(capi.SailError) as ei:
        capi.Context(64, 64)
    assert ei.value.code == capi.ERR_CUDA
    assert "no CPU path" in str(ei.value) or "CUDA" in str(ei.value)


def test_product_does_not_touch_oracle():
    """Nothing under sail_b200/ (the product) may import, link or execute oracle/."""
    bad = []
    for dirpath, _, files in os.walk(os.path.join(ROOT, "sail_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) or f == "Makefile":
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                if re.search(r"^\s*(import|from)\s+oracle\b", txt, flags=re.M) or "liboracle" in txt or '"../../oracle' in txt:
                    bad.append(f)
    assert not bad, bad


@pytest.mark.parametrize("cfg,expect", [(1, 0), (2, 2 | 8), (3, 15), (5, 2 | 8)])
def test_scene_features_select_kernel_variant(capi, cfg, expect):
    _, o, t, l, _ = scene_tables(cfg, 64, 64)
    assert capi.scene_features(o, t, l) == expect


def test_upload_validation_without_device(capi):
    _, o, t, l, _ = scene_tables(2, 64, 64)
    bad = o.copy()
    bad[1, 0] = 17  # unknown shape category
    with pytest.raises(capi.SailError) as ei:
        capi.scene_features(bad, t, l)
    assert ei.value.code == capi.ERR_INVALID and "object 1" in str(ei.value)
    many = np.tile(o[:1], (65, 1))
    with pytest.raises(capi.SailError):
        capi.scene_features(many, t, l)
    # empty scene and single-row tables (the n-1 = 0 divisions of SURVEY.md Q5) are accepted
    z18, z16 = np.zeros((0, 18), np.float32), np.zeros((0, 16), np.float32)
    assert capi.scene_features(z18, z16, z18) == 0
    assert capi.scene_features(o[:1], t[:2], l[:1]) >= 0


def test_napi_addon_source_compiles_against_the_header():
    """napi/sail_napi.cc cannot be built here (no Node.js); check at least that it is valid C++ against
    include/sail_b200.h and the N-API signatures it uses (tests/napi_stub/node_api.h)."""
    import subprocess
    r = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-I", os.path.join(ROOT, "tests", "napi_stub"), "-I",
                        os.path.join(ROOT, "include"), os.path.join(ROOT, "napi", "sail_napi.cc")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_js_host_is_complete_on_paper():
    """The JS host cannot run here (no Node.js).  What can be checked: every `native.<name>` the files under js/ use is a
    function napi/sail_napi.cc registers; the addon is loaded from exactly one place (js/native.js); every relative import
    resolves to a file under js/ or -- for `../x/y` -- to a module of the reference tree the host is dropped into
    (js/ -> <sail>/src/native/); js/index.js exports the thirty names of the reference's namespace
    (/root/reference/index.js:15-45) without `window`."""
    js_dir = os.path.join(ROOT, "js")
    napi = open(os.path.join(ROOT, "napi", "sail_napi.cc")).read()
    registered = set(re.findall(r'\{"(\w+)", nullptr, \w+, nullptr', napi))
    assert len(registered) >= 15
    used, loaders = set(), []
    ref_modules = {"../scene/scene", "../scene/geometry", "../scene/light", "../scene/camera", "../scene/material", "../scene/texture",
                   "../utils/matrix", "../core/color", "../core/pickup", "../shader/filter/box", "../shader/filter/gaussian",
                   "../shader/filter/mitchell", "../shader/filter/sinc", "../shader/filter/triangle"}
    for f in sorted(os.listdir(js_dir)):
        txt = open(os.path.join(js_dir, f)).read()
        used |= set(re.findall(r"\bnative\.(\w+)\(", txt))
        if "sail_napi.node" in txt:
            loaders.append(f)
        assert not re.search(r"\bwindow\b", re.sub(r"//.*", "", txt)), f
        for mod in re.findall(r"from '([^']+)'", re.sub(r"//.*", "", txt)):
            if mod.startswith("./"):
                assert os.path.exists(os.path.join(js_dir, mod[2:] + ".js")), "%s imports %s" % (f, mod)
            elif mod.startswith("../"):
                assert mod in ref_modules, "%s imports %s, not a module of the reference tree" % (f, mod)
            else:
                assert mod == "module", (f, mod)
    assert loaders == ["native.js"], loaders
    assert used and used <= registered, "used but not registered in napi/sail_napi.cc: %s" % sorted(used - registered)
    index = open(os.path.join(js_dir, "index.js")).read()
    names = ["Renderer", "Scene", "Cube", "Sphere", "Rectangle", "Cone", "Cylinder", "Disk", "Hyperboloid", "Paraboloid", "AreaLight",
             "PointLight", "SpotLight", "Cornellbox", "Camera", "Control", "Matte", "Mirror", "Metal", "Glass", "UniformColor",
             "Checkerboard", "Checkerboard2", "Bilerp", "Mix", "Scale", "UV", "Color", "Matrix", "Vector"]
    body = index[index.index("const Sail = {"):index.index("};")]
    for n in names:
        assert re.search(r"\b%s\b" % n, body), n
    if os.path.isdir("/root/reference/src"):  # in the build container: the reference modules really exist
        for mod in ref_modules:
            assert os.path.exists(os.path.join("/root/reference/src", mod[3:] + ".js")), mod


@pytest.mark.parametrize("cfg", [1, 2, 3, 5])
def test_scene_specialised_kernels_compile_without_a_gpu(capi, cfg):
    """The kernels that NVRTC builds per scene structure at run time must compile for every
    BASELINE config -- checked here with sail_jit_compile_check, which needs NVRTC but no device."""
    _, o, t, l, _ = scene_tables(cfg, 64, 64)
    for kernel in (capi.KERNEL_MEGA, capi.KERNEL_WAVEFRONT):
        try:
            info = capi.jit_compile_check(o, t, l, kernel)
        except capi.SailError as e:
            if "NVRTC not found" in str(e):
                pytest.skip("no NVRTC on this machine: the library would run its ahead-of-time variants")
            raise
        assert info.startswith("jit ") and "n=%d" % o.shape[0] in info and "cubin" in info, info


def test_kernel_build_knobs_compile(capi, monkeypatch):
    """The knobs sail_jit.cpp reads from the environment when it builds a kernel: extra #defines (the literal last bounce
    that bench.py reports next to the default), the wavefront's paths per CTA."""
    _, o, t, l, _ = scene_tables(3, 64, 64)
    try:
        monkeypatch.setenv("SAIL_B200_JIT_DEFINES", "SAIL_LITERAL_LAST_BOUNCE")
        assert "-DSAIL_LITERAL_LAST_BOUNCE" in capi.jit_compile_check(o, t, l, capi.KERNEL_MEGA)
        monkeypatch.delenv("SAIL_B200_JIT_DEFINES")
        assert "512 paths" in capi.jit_compile_check(o, t, l, capi.KERNEL_WAVEFRONT)  # quadric scene: the measured default
        monkeypatch.setenv("SAIL_B200_WF_THREADS", "256")
        assert "256 paths" in capi.jit_compile_check(o, t, l, capi.KERNEL_WAVEFRONT)
        monkeypatch.setenv("SAIL_B200_WF_THREADS", "100")
        with pytest.raises(capi.SailError):
            capi.jit_compile_check(o, t, l, capi.KERNEL_WAVEFRONT)
    except capi.SailError as e:
        if "NVRTC not found" in str(e):
            pytest.skip("no NVRTC on this machine")
        raise


def test_scene_specialisation_limits(capi):
    """A run of equal shapes stays one loop (64 spheres compile); more than 16 runs of distinct shapes are refused, which
    makes the library use its ahead-of-time variant."""
    import sail_b200 as Sail
    m, w = Sail.Matte(0.7), Sail.Color.WHITE
    from sail_b200.wire import serialize_scene

    def tables(objs):
        sc = Sail.Scene()
        for ob in objs:
            sc.add(ob)
        sc.add(Sail.Camera([0, 1, -5], [0, 1, 0]))
        return serialize_scene(sc)
    try:
        info = capi.jit_compile_check(*tables([Sail.Sphere([i * 0.1, 1, 2], 0.04, m, w) for i in range(64)]))
    except capi.SailError as e:
        if "NVRTC not found" in str(e):
            pytest.skip("no NVRTC on this machine")
        raise
    assert "shapes=2x64" in info, info
    alternating = []
    for i in range(17):
        alternating.append(Sail.Sphere([i * 0.2, 1, 2], 0.05, m, w) if i % 2 else Sail.Cube([i * 0.2, 0, 0], [i * 0.2 + 0.1, 0.1, 0.1], m, w))
    with pytest.raises(capi.SailError) as ei:
        capi.jit_compile_check(*tables(alternating))
    assert "16 runs" in str(ei.value)


def test_host_may_exit_while_a_kernel_is_compiling(capi):
    """SAIL_SPECIALISE_AUTO compiles on a worker thread; a host that exits before NVRTC is done must terminate promptly and
    cleanly (round 2: a static destructor joining that thread deadlocked against libnvrtc's own teardown)."""
    import subprocess
    import sys
    try:
        _, o, t, l, _ = scene_tables(3, 64, 64)
        capi.jit_compile_check(o, t, l)
    except capi.SailError as e:
        if "NVRTC not found" in str(e):
            pytest.skip("no NVRTC on this machine")
        raise
    code = ("import os, sys; sys.path.insert(0, %r); sys.path.insert(0, %r); os.environ['SAIL_B200_JIT_CHECK_NOWAIT'] = '1'\n"
            "from sail_b200 import capi; from conftest import scene_tables\n"
            "_, o, t, l, _ = scene_tables(3, 64, 64)\n"
            "print(capi.jit_compile_check(o, t, l, capi.KERNEL_WAVEFRONT))\n") % (ROOT, os.path.join(ROOT, "tests"))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and "compiling" in r.stdout, (r.returncode, r.stdout, r.stderr[-500:])
