"""The product's host-side mirror (include/spt_host/*.hpp through include/spt_host.h) against golden vectors
produced by the UNMODIFIED reference: scene generators, camera basis, BMP bytes.  No GPU needed."""
import os

import numpy as np
import pytest

import simplepathtracer_b200 as spt
from conftest import golden_path


@pytest.fixture(scope="module", autouse=True)
def lib():
    spt.build_library()
    return spt.load_library()


@pytest.mark.parametrize("seed", [1, 2])
def test_generators_match_reference(seed):
    for name, gen in (("init", spt.init_spheres), ("random", spt.generate_spheres)):
        g = np.load(golden_path(f"scene_{name}_seed{seed}.npz"))
        s = gen(seed)
        for k in ("centers", "radii", "materials", "colors", "fuzz"):
            assert np.array_equal(getattr(s, k), g[k]), (name, seed, k)


def test_camera_basis_matches_reference():
    g = np.load(golden_path("camera_default.npz"))
    c = spt.make_camera()
    assert np.array_equal(c.view, g["view"]) and np.array_equal(c.eye, g["eye"]) and np.array_equal(c.sky, g["sky"])
    g = np.load(golden_path("camera_offaxis.npz"))  # goes through the reference's Cross() z-lane bug
    c = spt.make_camera(tuple(g["eye_in"]), tuple(g["look_in"]))
    assert np.array_equal(c.view, g["view"])


def test_bmp_bytes_match_stb(tmp_path):
    rgb = np.load(golden_path("tiny_7x5_rgb.npy"))
    out = str(tmp_path / "t.bmp")
    spt.save_bmp(out, rgb)
    assert open(out, "rb").read() == open(golden_path("tiny_7x5.bmp"), "rb").read()


def test_synthetic_scenes():
    mg = spt.generate_metal_glass(4, 500)
    assert mg.n == 501 and set(np.unique(mg.materials[1:])) == {spt.REFLECTIVE, spt.REFRACTIVE}
    assert 0.4 < (mg.materials[1:] == spt.REFLECTIVE).mean() < 0.6
    lat = spt.generate_lattice(5, 128)
    assert lat.n == 128 * 128 and lat.radii[0] == 1e6
    assert (lat.radii[1:] >= 0.3).all() and (lat.radii[1:] <= 0.5).all()
    assert np.array_equal(spt.generate_lattice(5, 128).centers, lat.centers)  # deterministic
    assert (lat.centers[:, 3] == 0).all() and (mg.centers[:, 3] == 0).all()


def test_exports_declared_in_header(lib):
    import re

    from simplepathtracer_b200 import binding

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = re.sub(r"/\*.*?\*/", "", open(os.path.join(root, "include", "spt_host.h")).read(), flags=re.S)
    declared = sorted(set(re.findall(r"\b(spt_host_\w+)\s*\(", text)))
    assert declared == sorted(binding.HOST_EXPORTS)
    for name in declared:
        assert hasattr(lib, name)


def test_pfm_and_accumulator_checkpoint_files(tmp_path):
    rng = np.random.default_rng(3)
    acc = np.zeros((5, 7, 4), np.float32)
    acc[..., :3] = rng.uniform(0, 255 * 9, size=(5, 7, 3))
    acc[..., 3] = 9
    pfm = str(tmp_path / "a.pfm")
    spt.save_pfm(pfm, acc)
    raw = open(pfm, "rb").read()
    assert raw.startswith(b"PF\n7 5\n-1.0\n")
    body = np.frombuffer(raw[len(b"PF\n7 5\n-1.0\n"):], "<f4").reshape(5, 7, 3)
    want = (acc[..., :3] * (np.float32(1) / acc[..., 3:4])) / np.float32(255)
    assert np.array_equal(body, want)
    ck = str(tmp_path / "a.sptacc")
    spt.save_accum(ck, acc, samples_done=9, depth=4, seed=77)
    assert os.path.getsize(ck) == 48 + acc.nbytes  # "SPTACC2" header: ... seed, identity, reserved
    got = spt.load_accum(ck, 7, 5)
    assert got is not None and np.array_equal(got[0], acc) and got[1:] == (9, 4, 77)
    assert spt.load_accum(ck, 8, 5) is None  # another image size
    assert spt.load_accum(str(tmp_path / "missing"), 7, 5) is None


def test_hdr_file_decodes_to_the_linear_image(tmp_path):
    """Radiance RGBE next to the PFM (SURVEY 8f-3 names .hdr / stbi_write_hdr): header, row order, and every pixel within
    the format's own resolution (8-bit mantissa under the largest channel's exponent, truncated)."""
    rng = np.random.default_rng(5)
    acc = np.zeros((6, 9, 4), np.float32)
    acc[..., :3] = rng.uniform(0, 255 * 4, size=(6, 9, 3)) * rng.choice([1e-6, 1e-2, 1.0, 40.0], size=(6, 9, 1))
    acc[..., 3] = 4
    acc[0, 0, :3] = 0  # black pixel: (0, 0, 0, 0)
    path = str(tmp_path / "a.hdr")
    spt.save_hdr(path, acc)
    raw = open(path, "rb").read()
    head = b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\nEXPOSURE=1.0\n\n-Y 6 +X 9\n"
    assert raw.startswith(head) and len(raw) == len(head) + 6 * 9 * 4
    px = np.frombuffer(raw[len(head):], np.uint8).reshape(6, 9, 4)[::-1]  # file rows run top to bottom
    want = (acc[..., :3] * (np.float32(1) / acc[..., 3:4])) / np.float32(255)
    assert tuple(px[0, 0]) == (0, 0, 0, 0)
    scale = np.ldexp(1.0, px[..., 3].astype(np.int32) - 136)[..., None]  # mantissa * 2^(e - 128 - 8)
    got = px[..., :3].astype(np.float64) * scale
    top = want.max(axis=-1, keepdims=True).astype(np.float64)
    lit = top[..., 0] > 0
    assert np.all(got[lit] <= want[lit] * (1 + 1e-6)) and np.all(want[lit] - got[lit] <= top[lit] / 128 + 1e-30)
    assert np.all(px[..., :3].max(axis=-1)[lit] >= 128)  # normalised: the largest channel uses the full mantissa


def test_scene_file_round_trip(tmp_path):
    for sc in (spt.generate_spheres(2), spt.generate_metal_glass(4, 500), spt.generate_lattice(5, 16)):
        path = str(tmp_path / "scene.sptscene")
        spt.save_scene_file(path, sc, eye=(1, 2, -4), look=(0, 1, 0), sky=(10, 20, 30))
        size = 80 + sum((b + 15) // 16 * 16 for b in (sc.n * 16, sc.n * 4, sc.n, sc.n * 16, sc.n * 4))
        assert os.path.getsize(path) == size and open(path, "rb").read(8) == b"SPTSCN1\0"
        got, cam = spt.load_scene_file(path)
        for name in ("centers", "radii", "materials", "colors", "fuzz"):
            assert np.array_equal(getattr(got, name), getattr(sc, name)), name
        assert tuple(cam["eye"]) == (1, 2, -4) and tuple(cam["sky"]) == (10, 20, 30)
    bad = str(tmp_path / "bad.sptscene")
    open(bad, "wb").write(b"SPTSCN1\0" + b"\x05\0\0\0" * 10)
    with pytest.raises(spt.SptError):
        spt.load_scene_file(bad)
