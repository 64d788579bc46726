"""The C++ drop-in classes of include/vr_b200.hpp (VolumeReader, VolumeKdtree, Renderer) driven by
tests/cpp/dropin_main.cpp, a miniature of the reference's main.cpp pipeline (main.cpp:141-290,404).
CPU part (no GPU): brick files -> LoadBricksToTexture placement -> build -> save -> open; checked against
numpy placement, the Python mirror's stream and (when present) the unmodified reference's stream.
GPU part (-m gpu): + levelCut bit-exact vs the oracle, the reference's crash case rejected, frames."""
import os
import subprocess

import numpy as np
import pytest

import oracle_api as oa
import volumerenderer_b200 as vr
from volumerenderer_b200 import render, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def fnv_fast(a):
    a = np.frombuffer(np.ascontiguousarray(a).tobytes(), np.uint8)
    h = 2166136261
    for x in a.tolist():
        h = ((h ^ x) * 16777619) & 0xFFFFFFFF
    return h


@pytest.fixture(scope="module")
def exe(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("cpp") / "dropin_main")
    lib = os.path.join(ROOT, "volumerenderer_b200")
    r = subprocess.run(["/usr/bin/g++", "-std=c++14", "-O2", "-I" + os.path.join(ROOT, "include"),
                        os.path.join(ROOT, "tests", "cpp", "dropin_main.cpp"), "-o", out, "-L" + lib, "-lvr_b200",
                        "-Wl,-rpath," + lib], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return out


@pytest.fixture(scope="module")
def scene(tmp_path_factory):
    d = tmp_path_factory.mktemp("bricks")
    full = synth.rm_region(700, 1100, 940, 64, 64, 32)  # (Z, Y, X) = 2x2x2 bricks of 32x32x16
    for b in range(8):
        i, j, k = b % 2, (b // 2) % 2, b // 4
        full[k * 16:(k + 1) * 16, j * 32:(j + 1) * 32, i * 32:(i + 1) * 32].tofile(str(d / f"brick_{b}_t7.raw"))
    return str(d), full


def run(exe, mode, d):
    r = subprocess.run([exe, mode, d], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    kv = {}
    for tok in r.stdout.split():
        if "=" in tok:
            k, v = tok.split("=", 1)
            kv[k] = v
    return kv


def check_cpu(kv, full):
    assert kv["dims"] == "64,64,32"
    assert int(kv["assembled"]) == fnv_fast(full)          # VolumeReader.h:184-205 placement
    assert kv["missing_brick_returns_false"] == "1" and kv["input_intact"] == "1"
    s, st = vr.Stream.compress(full, 1, 2)
    hd = s.header
    assert int(kv["nodes"]) == hd.numActiveNodes and kv["depth"] == f"{hd.origTreeDepth},{hd.maxTreeDepth}"
    assert int(kv["maxerr"]) == st.maxError <= 1
    assert int(kv["stream"]) == fnv_fast(s.bits) and int(kv["dmap"]) == fnv_fast(s.distance_map)
    if oa.REF is not None:  # the unmodified reference produces the same stream
        ref = oa.ref_build(full, 1, 2)
        assert int(kv["stream"]) == fnv_fast(ref["bits"][: (ref["n"] + 3) // 4]) and int(kv["nodes"]) == ref["n"]
    assert kv["reopened_equal"] == "1"
    assert kv["open_missing"] == "error2"                  # VR_ERR_IO instead of exit(-1) (VolumeKdTree_recover.cpp:560-565)
    return s


def test_cpp_pipeline_cpu(exe, scene):
    d, full = scene
    check_cpu(run(exe, "cpu", d), full)


@pytest.mark.gpu
def test_cpp_pipeline_gpu(exe, scene):
    d, full = scene
    kv = run(exe, "gpu", d)
    s = check_cpu(kv, full)
    hd = s.header
    rc, want = oa.oracle_levelcut(s.bits, hd.numActiveNodes, s.distance_map, hd.origTreeDepth, hd.maxTreeDepth, 64, 64, 32)
    assert rc == 0 and int(kv["decoded"]) == fnv_fast(want)
    assert int(kv["maxerr_measured"]) == int(np.abs(want.astype(int) - full.astype(int)).max())
    assert kv["levelcut_ctx_equal"] == "1"                 # two-argument levelCut (shared context) == explicit context
    assert kv["cut10"] == "error5"                         # VR_ERR_UNSUPPORTED; the reference segfaults
    assert kv["gpu_build_equal"] == "1"                    # VolumeKdtree::build(Context&): GPU producer, same stream
    # LoadBricksToTexture(toGPU=true): pipelined loader, same placement on the host copy and in the device slab
    assert kv["streamed_ok"] == "1" and int(kv["streamed_host"]) == fnv_fast(full) and int(kv["streamed_device"]) == fnv_fast(full)
    assert kv["streamed_dealloc"] == f"1,0,{fnv_fast(full)}" and kv["streamed_missing_returns_false"] == "1"
    p = render.default_params(96, 72, render.COMPOSITE, step=(1 / 64, 1 / 64, 1 / 32))
    img, S, _ = oa.oracle_render(want, render.oracle_param_blob(p), 4)
    assert int(kv["composite_samples"]) == S
    p.mode = render.ISOSURFACE; p.isoValue = 120 / 255
    _, S2, _ = oa.oracle_render(want, render.oracle_param_blob(p), 4)
    assert int(kv["iso_samples"]) == S2
