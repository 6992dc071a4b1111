"""The C++ host side over the C ABI: er::RenderingManager (include/er_manager.hpp, the reference's
RenderingManager interface, src/Managers.h:41-67) driven by tests/cpp/manager_main.cpp.

CPU: the header compiles against include/er_cuda.h, links against liber_cuda.so alone, and without a B200
start_rendering fails loudly (exit code 3) instead of falling back.
GPU: the C++ call sequence (start_rendering -> poll get_render_info -> get_pass) produces the passes of the
Python mirror of the same interface bit for bit, and get_render_info has the reference's N+1 semantics."""
import os
import subprocess

import numpy as np
import pytest

import common
from elevenrender_b200 import capi, host

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "elevenrender_b200")


@pytest.fixture(scope="module")
def exe(tmp_path_factory):
    capi.lib()  # builds nothing, but fails early when the library is missing
    out = str(tmp_path_factory.mktemp("cpp") / "manager_main")
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "manager_main.cpp"), "-o", out, "-L", LIBDIR, "-l:liber_cuda.so",
           f"-Wl,-rpath,{LIBDIR}", "-lpthread"]
    subprocess.check_call(cmd)
    return out


def write_scene(path, spec, spp, rng_order):
    hd = spec["hdri"]
    assert len(hd) == 5 and not hd[4], "the driver takes an unfiltered environment texture"
    mats = [host.material_words(**kw) for kw in spec["materials"]] or [host.material_words()]
    tris = np.ascontiguousarray(spec["tris"], dtype=np.float32)
    with open(path, "wb") as f:
        np.array([spec["x_res"], spec["y_res"], spp, rng_order, tris.shape[0], len(mats), hd[0], hd[1], hd[2]],
                 dtype=np.int32).tofile(f)
        tris.tofile(f)
        np.stack(mats).astype(np.float32).tofile(f)
        host.camera_words(**spec["camera"]).astype(np.float32).tofile(f)
        np.ascontiguousarray(hd[3], dtype=np.float32).tofile(f)


def test_cpp_manager_fails_loudly_without_device(exe, tmp_path):
    try:
        have = capi.device_count() > 0
    except capi.ErError:
        have = False
    if have:
        pytest.skip("a CUDA device is present: covered by the gpu test")
    spec = common.make_spec("grid")
    write_scene(tmp_path / "scene.bin", spec, 2, host.ER_RNG_LTR)
    r = subprocess.run([exe, str(tmp_path / "scene.bin"), str(tmp_path / "out.bin")], capture_output=True, text=True)
    assert r.returncode == 3, r.stderr
    assert "error:" in r.stderr and not os.path.exists(tmp_path / "out.bin")


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["grid", "spheres"])
def test_cpp_manager_matches_python_mirror(gpu, exe, tmp_path, name):
    spec = common.make_spec(name)
    spp = 5
    write_scene(tmp_path / "scene.bin", spec, spp, host.ER_RNG_LTR)
    r = subprocess.run([exe, str(tmp_path / "scene.bin"), str(tmp_path / "out.bin")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    raw = np.fromfile(tmp_path / "out.bin", dtype=np.uint32)
    assert raw[0] == spp + 1  # dev_samples starts at 1
    got = raw[1:].view(np.float32).reshape(4, spec["y_res"], spec["x_res"], 4)

    rm = host.RenderingManager(host.RenderParameters(spec["x_res"], spec["y_res"], spp, rng_order=host.ER_RNG_LTR))
    rm.upload(common.to_ours(spec))
    rm.render_samples(spp)
    want = np.stack([rm.get_pass(p) for p in ("beauty", "normal", "tangent", "bitangent")])
    rm.close()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
