"""The C++ drop-in boundary (include/rtp/*.h over the C ABI): a caller written against the reference's
camera / hittable_list / sphere / material API compiles unchanged and renders through librt_b200.so.

CPU part: the example builds, exports exactly the scene the library's generator makes (485 spheres,
byte for byte), fails loudly without a GPU, and -- where /root/reference exists -- the reference's OWN
src/main.cpp compiles against include/rtp with only the camera type swapped (nothing is copied: the
file is patched into a temp dir by sed at test time).
GPU part: the example's stdout is byte-identical to the P3 text of the reference harness (fixture)."""
import os
import subprocess

import numpy as np
import pytest

from conftest import golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "examples", "final_scene")
INC = os.path.join(ROOT, "include", "rtp")
LIBDIR = os.path.join(ROOT, "raytracingproject_b200")
REF_MAIN = "/root/reference/src/main.cpp"


@pytest.fixture(scope="module")
def exe():
    if not os.path.exists(EXE):
        import __graft_entry__ as g
        g.build()
    return EXE


def test_example_exports_the_library_scene(exe, rt):
    raw = subprocess.run([exe, "--dump-scene"], capture_output=True, check=True).stdout
    n_s, n_m = np.frombuffer(raw[:16], "u8")
    s, m = rt.scene_rtiow()
    assert (n_s, n_m) == (485, 485)
    assert raw[16:16 + 64 * n_s] == s.tobytes()
    assert raw[16 + 64 * n_s:] == m.tobytes()


@pytest.mark.skipif(not os.path.exists(REF_MAIN), reason="reference tree not present on this box")
def test_reference_main_compiles_with_only_the_camera_type_swapped(tmp_path, rt):
    src = open(REF_MAIN).read()
    assert src.count("CPUImpl::Camera") == 1 and src.count('"camera_cpu.h"') == 1
    patched = tmp_path / "main_b200.cpp"
    patched.write_text(src.replace("CPUImpl::Camera", "B200Impl::Camera").replace('"camera_cpu.h"', '"camera_b200.h"'))
    out = tmp_path / "main_b200"
    r = subprocess.run(["/usr/bin/g++", "-std=c++17", "-O2", "-I" + INC, str(patched), "-L" + LIBDIR, "-lrt_b200",
                        "-Wl,-rpath," + LIBDIR, "-o", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]


def test_no_gpu_is_a_loud_error_in_cpp_too(exe):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = subprocess.run([exe, "64", "1", "5"], capture_output=True, text=True)
    assert r.returncode != 0 and "no CPU fallback" in r.stderr and r.stdout == ""


@pytest.mark.gpu
def test_cpp_render_is_byte_identical_to_the_reference_harness(exe, O):
    """BASELINE config 1 through the C++ camera in parity mode: stdout == the P3 text of the reference
    harness (keyed Philox, seed 0) for the same scene."""
    r = subprocess.run([exe, "400", "10", "50", "f64"], capture_output=True, check=True)
    g = golden("ref_c1_philox_seed0.npz")
    assert r.stdout == O.ppm_p3(g["rgb8"])


@pytest.mark.gpu
def test_cpp_render_fast_mode_within_tolerance(exe):
    r = subprocess.run([exe, "400", "10", "50"], capture_output=True, check=True)
    lines = r.stdout.decode().split("\n")
    assert lines[0] == "P3" and lines[1] == "400 225" and lines[2] == "255"
    img = np.array([[int(v) for v in ln.split()] for ln in lines[3:3 + 400 * 225]], dtype=np.uint8).reshape(225, 400, 3)
    g = golden("ref_c1_philox_seed0.npz")["rgb8"]
    d = np.abs(img.astype(int) - g.astype(int)).max(-1)
    assert (d <= 1).mean() >= 0.999


@pytest.mark.gpu
def test_cpp_binary_formats_and_progressive_passes(exe, O):
    """8f-2 through the C++ camera: P6 carries the bytes of the P3 text; four progressive passes end on the same
    image (sum order differs in the last bit only); PFM holds the mean radiance whose write_color is that image."""
    g = golden("ref_c1_philox_seed0.npz")["rgb8"]
    p6 = subprocess.run([exe, "400", "10", "50", "f64", "p6"], capture_output=True, check=True).stdout
    assert p6 == O.ppm_p6(g)
    prog = subprocess.run([exe, "400", "10", "50", "f64", "p6", "4"], capture_output=True, check=True)
    assert b"Samples so far: 10" in prog.stderr
    hl = len(b"P6\n400 225\n255\n")
    a = np.frombuffer(prog.stdout[hl:], np.uint8).astype(int); b = g.reshape(-1).astype(int)
    assert prog.stdout[:hl] == p6[:hl] and np.abs(a - b).max() <= 1 and (a != b).mean() < 1e-3
    pfm = subprocess.run([exe, "400", "10", "50", "f64", "pfm"], capture_output=True, check=True).stdout
    hl = len(b"PF\n400 225\n-1.0\n")
    assert pfm[:hl] == b"PF\n400 225\n-1.0\n" and len(pfm) == hl + 400 * 225 * 12
    mean = np.frombuffer(pfm[hl:], "<f4").reshape(225, 400, 3)[::-1].astype(np.float64)
    byte = (256 * np.clip(np.sqrt(mean), 0.0, 0.999)).astype(int)
    assert (np.abs(byte - g.astype(int)) <= 1).all() and (byte != g).mean() < 1e-3      # float rounding can cross a byte boundary
