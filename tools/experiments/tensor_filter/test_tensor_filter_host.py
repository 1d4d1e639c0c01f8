"""Groundwork for the tensor-core pre-filter: EXPERIMENTAL code that the shipped library does
not contain (-DSPT_TENSOR_FILTER=1 builds only).  What can be checked without a GPU is checked here:

* tools/tensor_filter_check.cpp -- the sphere-matrix builder the experimental build uses, the ray rows restated operation for
  operation, the UMMA canonical layout and two accumulator models that bracket the tensor core: no pair the reference's
  separately rounded sequence accepts may be filtered out, and the error must stay inside the slack;
* the experimental variant still compiles for sm_100a, and the DEFAULT build's device code does not change with it
  (the macro really is off).
"""
import os
import shutil
import struct
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def checker(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("tfc") / "tensor_filter_check")
    subprocess.run(["g++", "-O2", "-std=c++17", "-frounding-math", "-ffp-contract=off", "-o", out,
                    os.path.join(ROOT, "tools", "tensor_filter_check.cpp")], check=True)
    return out


def write_scene(path, centers, radii):
    g = np.concatenate([np.asarray(centers, np.float32).reshape(-1, 3), np.asarray(radii, np.float32).reshape(-1, 1)], axis=1)
    with open(path, "wb") as f:
        f.write(struct.pack("<I", len(g)))
        f.write(np.ascontiguousarray(g, np.float32).tobytes())


def run(checker, scene, rays, *extra):
    r = subprocess.run([checker, scene, str(rays), *map(str, extra)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.strip().endswith("OK"), r.stdout + r.stderr
    return r.stdout


def test_config3_scene_no_true_hit_filtered_out(checker, tmp_path):
    z = np.load(os.path.join(ROOT, "tests", "golden", "scene_random_seed1.npz"))
    scene = str(tmp_path / "scene.bin")
    write_scene(scene, z["centers"].reshape(-1, 4)[:, :3], z["radii"])
    out = run(checker, scene, 4096)
    assert "128 on the tensor core, 128 matrix rows" in out  # 152 slots: the last 24 stay with the FFMA2 loop


@pytest.mark.parametrize("eye", [None, (300.5, -2.0, -1800.0)])
def test_camera_form_no_true_hit_filtered_out(checker, tmp_path, eye):
    # primary_hits_tc: six products of the direction + the constant (K = 32), matrix specialised for the eye; the default camera
    # of the reference, and an eye far from the scene (slack scales with |E|^2)
    z = np.load(os.path.join(ROOT, "tests", "golden", "scene_random_seed1.npz"))
    cam = np.load(os.path.join(ROOT, "tests", "golden", "camera_default.npz"))
    scene = str(tmp_path / "scene.bin")
    write_scene(scene, z["centers"].reshape(-1, 4)[:, :3], z["radii"])
    e = [float(v) for v in cam["eye"][:3]] if eye is None else list(eye)
    out = run(checker, scene, 4096, "cam", *e)
    assert "camera form" in out and "K = 32" in out


@pytest.mark.parametrize("name", ["scene_random_seed2", "scene_init_seed1"])
def test_other_golden_scenes(checker, tmp_path, name):
    z = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    scene = str(tmp_path / "scene.bin")
    write_scene(scene, z["centers"].reshape(-1, 4)[:, :3], z["radii"])
    run(checker, scene, 2048)


def test_centres_and_radii_near_the_filter_limits(checker, tmp_path):
    # |C| up to ~4000 (kBigCentre 4096) and radii up to 60 (kBigRadius 64): the error scales with A^2 + B^2 + r^2 and so does the slack
    rng = np.random.default_rng(9)
    scene = str(tmp_path / "far.bin")
    write_scene(scene, rng.uniform(-2300, 2300, size=(120, 3)), rng.uniform(0.5, 60, size=120))
    out = run(checker, scene, 2048)
    assert "120 slots (120 on the tensor core, 128 matrix rows)" in out


def test_short_list_is_padded_to_sixteen_rows(checker, tmp_path):
    # 21 spheres -> 24 slots -> 32 matrix rows, the padding never a candidate; includes a negative radius (the hollow glass
    # sphere of the reference's scenes) and centres far from the origin
    rng = np.random.default_rng(3)
    cen = rng.uniform(-300, 300, size=(21, 3))
    rad = rng.uniform(0.2, 6.0, size=21)
    rad[5] = -rad[5]
    scene = str(tmp_path / "small.bin")
    write_scene(scene, cen, rad)
    out = run(checker, scene, 2048)
    assert "24 slots (24 on the tensor core, 32 matrix rows)" in out


@pytest.mark.skipif(shutil.which("nvcc") is None or shutil.which("cuobjdump") is None, reason="needs the CUDA toolchain")
def test_experimental_variant_compiles_and_default_device_code_is_untouched(tmp_path):
    import simplepathtracer_b200 as spt

    shipped = spt.build_library()
    variant = str(tmp_path / "tensor.so")
    r = subprocess.run([os.path.join(ROOT, "tools", "build_variant.sh"), variant, "-DSPT_TENSOR_FILTER=1"], capture_output=True, text=True)
    assert r.returncode == 0 and os.path.exists(variant), r.stderr[-2000:]

    def kernels(path):
        out = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True, check=True).stdout
        return {line.split()[-1] for line in out.splitlines() if line.strip().startswith("Function :")}

    base, exp = kernels(shipped), kernels(variant)
    tc = {k for k in exp if "trace_secondary_tc" in k or "primary_hits_tc" in k}
    assert len(tc) == 2 and not any(k in base for k in tc)
    assert exp - tc == base  # the variant only ADDS the experimental kernels
    sass = subprocess.run(["cuobjdump", "-sass", variant], capture_output=True, text=True).stdout
    assert "UTCMMA" in sass or "UTCHMMA" in sass or "TCMMA" in sass, "no tcgen05.mma in the experimental kernel"


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="needs the CUDA toolchain")
def test_time_shared_variant_fits_sixty_four_registers(tmp_path):
    # four ray tiles taking turns on two accumulators: 2 CTAs x 16 warps per SM only if ptxas stays within 64 registers
    variant = str(tmp_path / "tensor42.so")
    r = subprocess.run([os.path.join(ROOT, "tools", "build_variant.sh"), variant, "-DSPT_TENSOR_FILTER=1", "-DSPT_TENSOR_TILES=4",
                        "-DSPT_TENSOR_ACC=2", "-Xptxas", "-v"], capture_output=True, text=True)
    assert r.returncode == 0 and os.path.exists(variant), r.stderr[-2000:]
    lines = r.stderr.splitlines()
    at = next(i for i, l in enumerate(lines) if "Compiling entry function" in l and "trace_secondary_tc" in l)
    info = " ".join(lines[at:at + 4])
    import re

    assert int(re.search(r"Used (\d+) registers", info).group(1)) <= 64, info
    assert int(re.search(r"(\d+) bytes spill stores", info).group(1)) <= 16, info


def test_gpu_harness_and_scene_exporter_build(tmp_path):
    """tools/tensor_filter_ab.cpp (the C A/B harness next round's first GPU call runs) compiles against include/spt_cuda.h, and
    tools/export_scene.py writes the layout it reads: n | SptCamera (96 B) | centers4 | radii | colors4 | fuzz | materials."""
    import sys

    exe = str(tmp_path / "tensor_filter_ab")
    subprocess.run(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "include"), "-o", exe,
                    os.path.join(ROOT, "tools", "tensor_filter_ab.cpp"), "-ldl"], check=True)
    scene = str(tmp_path / "c3.bin")
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "export_scene.py"), scene], check=True, capture_output=True)
    raw = open(scene, "rb").read()
    (n,) = struct.unpack_from("<I", raw)
    z = np.load(os.path.join(ROOT, "tests", "golden", "scene_random_seed1.npz"))
    assert n == len(z["radii"]) and len(raw) == 4 + 96 + n * (16 + 4 + 16 + 4 + 1)
    assert np.array_equal(np.frombuffer(raw, np.float32, n * 4, 4 + 96), z["centers"].astype(np.float32).ravel())
    assert np.array_equal(np.frombuffer(raw, np.uint8, n, 4 + 96 + n * 40), z["materials"])
    # without a library argument the harness refuses (exit 2) instead of touching a GPU
    assert subprocess.run([exe]).returncode == 2
