"""The C++ host side on a GPU: the headless program over include/CudaPathTracer.hpp + the spt_host mirror, and
the drop-in compiled INSIDE the reference's own translation unit (oracle/_ref/ref_integration, built in the
container where /root/reference exists; only the binary travels)."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def read_bmp(path):
    raw = open(path, "rb").read()
    assert raw[:2] == b"BM"
    off = int.from_bytes(raw[10:14], "little")
    w = int.from_bytes(raw[18:22], "little")
    h = int.from_bytes(raw[22:26], "little")
    stride = (w * 3 + 3) // 4 * 4
    rows = np.frombuffer(raw, np.uint8, offset=off).reshape(h, stride)[:, : w * 3].reshape(h, w, 3)
    return rows[::-1, :, ::-1].copy()  # top-down RGB


@pytest.mark.parametrize("tracer,pass_cap", [("recursive", 0), ("task", 10)])
def test_headless_program_matches_c_abi(tmp_path, tracer, pass_cap, oracle):
    """spt_headless == RenderSegmentCuda over the host mirror's globals: its BMP equals the oracle's bytes."""
    exe = os.path.join(ROOT, "simplepathtracer_b200", "host", "spt_headless")
    if not os.path.exists(exe):
        subprocess.run(["make", "-s", "-C", os.path.dirname(exe)], check=True)
    out = str(tmp_path / "img.bmp")
    w, h, spp, depth, seed = 96, 64, 8, 6, 3
    r = subprocess.run([exe, "--scene", "random", "--seed", str(seed), "--width", str(w), "--height", str(h), "--spp", str(spp),
                        "--depth", str(depth), "--tracer", tracer, "--tiles", "2", "--out", out], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Duration:" in r.stdout and "4/4" in r.stdout
    scene = oracle.generate_spheres(seed)
    acc, _ = oracle.render_accum(scene, oracle.camera(), w, h, spp, depth, pass_cap=pass_cap, seed=seed)
    assert np.array_equal(read_bmp(out), oracle.resolve(acc))


@pytest.mark.parametrize("scene", [0, 1])
def test_drop_in_inside_the_reference_translation_unit(tmp_path, scene):
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_integration")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/ref_integration not built (needs /root/reference at build time)")
    prefix = str(tmp_path / "ri")
    r = subprocess.run([exe, str(scene), "160", "90", "512", "8", prefix], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    psnr = float(r.stdout.split("psnr8bit")[1].split()[0])
    cpu, gpu = read_bmp(prefix + "_cpu.bmp"), read_bmp(prefix + "_cuda.bmp")
    assert cpu.shape == gpu.shape == (90, 160, 3)
    # two independent 512-spp estimates of the same image, compared after 8-bit gamma encoding
    assert psnr > 30.0, r.stdout
    assert abs(float(cpu.mean()) - float(gpu.mean())) < 1.0


def read_pfm(path):
    raw = open(path, "rb").read()
    head, dims, scale, body = raw.split(b"\n", 3)
    assert head == b"PF" and float(scale) < 0  # little-endian
    w, h = (int(v) for v in dims.split())
    return np.frombuffer(body, "<f4").reshape(h, w, 3)  # bottom row first == the tracers' y


def test_headless_progressive_checkpoint_resume(tmp_path, oracle):
    """--progressive passes leave the bits of a one-shot render; a render stopped at spp/2 with --checkpoint and
    resumed to spp gives the same BMP, PFM and accumulators (the running sums continue across launches)."""
    exe = os.path.join(ROOT, "simplepathtracer_b200", "host", "spt_headless")
    if not os.path.exists(exe):
        subprocess.run(["make", "-s", "-C", os.path.dirname(exe)], check=True)
    w, h, spp, depth, seed = 80, 48, 12, 6, 5
    base = [exe, "--scene", "reference", "--seed", str(seed), "--width", str(w), "--height", str(h), "--depth", str(depth), "--tracer", "task"]
    one, prog, half = str(tmp_path / "one.bmp"), str(tmp_path / "prog.bmp"), str(tmp_path / "half.bmp")
    pfm, ck, hdr = str(tmp_path / "img.pfm"), str(tmp_path / "render.sptacc"), str(tmp_path / "img.hdr")
    assert subprocess.run(base + ["--spp", str(spp), "--out", one]).returncode == 0
    r = subprocess.run(base + ["--spp", str(spp), "--out", prog, "--progressive", "4", "--pfm", pfm, "--hdr", hdr], capture_output=True, text=True)
    assert r.returncode == 0 and "4/4" in r.stdout and f"({spp} spp)" in r.stdout, r.stdout + r.stderr
    assert open(one, "rb").read() == open(prog, "rb").read()
    # stop half way, then resume from the checkpoint
    assert subprocess.run(base + ["--spp", str(spp // 2), "--out", half, "--checkpoint", ck]).returncode == 0
    r = subprocess.run(base + ["--spp", str(spp), "--out", half, "--checkpoint", ck, "--progressive", "2"], capture_output=True, text=True)
    assert r.returncode == 0 and f"Resuming at sample {spp // 2}" in r.stdout, r.stdout + r.stderr
    assert open(one, "rb").read() == open(half, "rb").read()
    # the oracle agrees with all of it, bytes and floats
    import simplepathtracer_b200 as spt

    scene = oracle.init_spheres(seed)
    acc, _ = oracle.render_accum(scene, oracle.camera(), w, h, spp, depth, pass_cap=10, seed=seed)
    assert np.array_equal(read_bmp(one), oracle.resolve(acc))
    got = spt.load_accum(ck, w, h)
    assert got is not None and got[1:] == (spp, depth, seed)
    assert np.array_equal(got[0], acc)
    want = (acc[..., :3] * (np.float32(1) / acc[..., 3:4])) / np.float32(255)
    assert np.array_equal(read_pfm(pfm), want.astype(np.float32))
    # the Radiance file holds the same picture at the format's resolution (top row first, 8-bit mantissa, truncated)
    raw = open(hdr, "rb").read()
    body = raw[raw.index(b"+X %d\n" % w) + len(b"+X %d\n" % w):]
    px = np.frombuffer(body, np.uint8).reshape(h, w, 4)[::-1]
    dec = px[..., :3].astype(np.float64) * np.ldexp(1.0, px[..., 3].astype(np.int32) - 136)[..., None]
    top = want.max(axis=-1, keepdims=True).astype(np.float64)
    assert np.all(dec <= want * (1 + 1e-6) + 1e-30) and np.all(want - dec <= top / 128 + 1e-30)


def test_headless_scene_file(tmp_path):
    """--dump-scene writes the generated scene; rendering from the file gives the same image."""
    exe = os.path.join(ROOT, "simplepathtracer_b200", "host", "spt_headless")
    common = ["--seed", "4", "--width", "64", "--height", "40", "--spp", "4", "--depth", "6", "--tracer", "task"]
    a, b, scn = str(tmp_path / "a.bmp"), str(tmp_path / "b.bmp"), str(tmp_path / "mg.sptscene")
    assert subprocess.run([exe, "--scene", "metalglass", "--dump-scene", scn, "--out", a] + common).returncode == 0
    assert subprocess.run([exe, "--scene", scn, "--out", b] + common).returncode == 0
    assert open(a, "rb").read() == open(b, "rb").read()
    import simplepathtracer_b200 as spt

    sc, _ = spt.load_scene_file(scn)
    assert sc.n == 501


def test_headless_grid_flag_same_image(tmp_path):
    """spt_headless --grid 1 (SPT_FLAG_GRID through spt_cuda::SetFlags): the 16 384-sphere lattice and the mirror / glass scene
    render to the same bytes as through the linear loop."""
    exe = os.path.join(ROOT, "simplepathtracer_b200", "host", "spt_headless")
    for scene, depth in (("lattice", "16"), ("metalglass", "50")):
        common = ["--scene", scene, "--seed", "5", "--width", "96", "--height", "56", "--spp", "4", "--depth", depth, "--tracer", "task"]
        a, b = str(tmp_path / "lin.bmp"), str(tmp_path / "grid.bmp")
        assert subprocess.run([exe, "--out", a] + common).returncode == 0
        assert subprocess.run([exe, "--out", b, "--grid", "1"] + common).returncode == 0
        assert open(a, "rb").read() == open(b, "rb").read()


def test_headless_multi_device_and_checkpoint_identity(tmp_path, oracle):
    """spt_headless --devices a,b: ONE render spread over two device contexts by the library (RenderImageParallelMain's role,
    Renderer.hpp:257-302).  8 samples over 2 devices = sample ranges: the image is the in-order sum of two partial renders,
    which the oracle reproduces exactly; --tiles with ragged, partly empty tiles works (both tile ends are clamped); and a
    checkpoint written for another scene is refused."""
    import torch

    exe = os.path.join(ROOT, "simplepathtracer_b200", "host", "spt_headless")
    subprocess.run(["make", "-s", "-C", os.path.dirname(exe)], check=True)
    devs = "0,1" if torch.cuda.device_count() >= 2 else "0,0"
    w, h, spp, depth, seed = 90, 52, 8, 6, 3
    out = str(tmp_path / "multi.bmp")
    base = ["--seed", str(seed), "--width", str(w), "--height", str(h), "--depth", str(depth), "--tracer", "task"]
    r = subprocess.run([exe, "--scene", "random", "--spp", str(spp), "--devices", devs, "--out", out] + base, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    scene = oracle.generate_spheres(seed)
    cam = oracle.camera()
    a, _ = oracle.render_accum(scene, cam, w, h, spp // 2, depth, pass_cap=10, seed=seed)
    b, _ = oracle.render_accum(scene, cam, w, h, spp - spp // 2, depth, pass_cap=10, seed=seed, sample_begin=spp // 2)
    assert np.array_equal(read_bmp(out), oracle.resolve(a + b))
    # width 10 in 8 tiles: ceil-sized tiles run past the image, the last ones are empty
    small = str(tmp_path / "small.bmp")
    r = subprocess.run([exe, "--scene", "reference", "--seed", "2", "--width", "10", "--height", "9", "--spp", "2", "--depth", "4",
                        "--tracer", "task", "--tiles", "8", "--out", small], capture_output=True, text=True)
    assert r.returncode == 0 and "64/64" in r.stdout, r.stdout + r.stderr
    acc, _ = oracle.render_accum(oracle.init_spheres(2), cam, 10, 9, 2, 4, pass_cap=10, seed=2)
    assert np.array_equal(read_bmp(small), oracle.resolve(acc))
    # checkpoint identity: same size / depth / seed, another scene -> not resumed
    ck = str(tmp_path / "ck.sptacc")
    assert subprocess.run([exe, "--scene", "reference", "--spp", "4", "--checkpoint", ck, "--out", out] + base).returncode == 0
    r = subprocess.run([exe, "--scene", "random", "--spp", "8", "--checkpoint", ck, "--out", out] + base, capture_output=True, text=True)
    assert r.returncode == 0 and "Resuming" not in r.stdout, r.stdout
    r = subprocess.run([exe, "--scene", "random", "--spp", "8", "--checkpoint", ck, "--out", out] + base, capture_output=True, text=True)
    assert r.returncode == 0 and "Resuming at sample 8" in r.stdout, r.stdout  # the file now belongs to the random scene, complete
