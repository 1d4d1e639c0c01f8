#!/usr/bin/env python
"""Generate the golden fixtures in tests/golden/ from the UNMODIFIED reference.

Run in the build container only (needs /root/reference to build oracle/_ref/libspt_ref.so):

    python tests/golden/make_golden.py                # everything but the two slow renders below
    python tests/golden/make_golden.py --converged2   # converged_task2.npz: configs 2 and 4 (minutes of CPU)

Every array below is an output of the reference's own code (compiled by oracle/Makefile from
/root/reference/include with shim headers, no source edits), with its clock-seeded RNG handed a
chosen seed.  The fixtures travel with the repo; /root/reference and this script's inputs do not.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import Reference, build_oracle  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def scene_dict(sc):
    return dict(centers=sc.centers, radii=sc.radii, materials=sc.materials, colors=sc.colors, fuzz=sc.fuzz)


def config2_scene():
    """BASELINE config 2: one diffuse sphere over a ground sphere (the scene of test_first_hit_config2_full_resolution)."""
    from oracle import Scene

    return Scene(np.array([[0, -1000.5, 0, 0], [0, 1, 0, 0]], np.float32), np.array([1000.0, 0.5], np.float32),
                 np.array([3, 3], np.uint8), np.array([[30, 144, 255, 0], [255, 100, 230, 0]], np.float32), np.zeros(2, np.float32))


def converged_configs_2_and_4():
    """Converged renders of the UNMODIFIED reference for BASELINE config 2 (its "convergence check") and for a config-4
    style metal/glass scene (501 spheres, through the wide-index build: Collision.hpp:87-92 holds a uint8_t index):
    RenderSegmentTask over square tiles, all host threads.  Written to converged_task2.npz (minutes of CPU)."""
    import simplepathtracer_b200 as spt  # host-side generators only (no GPU needed)

    threads = os.cpu_count() or 8
    cv = {}
    jobs = (
        ("c2", False, config2_scene(), (480, 270, 8192, 4, 30)),
        ("c4", True, spt.generate_metal_glass(4, 500), (160, 90, 8192, 50, 10)),
    )
    for name, wide, scene, (w, h, spp, depth, tile) in jobs:
        R = Reference(wide=wide)
        R.set_camera()
        R.set_scene(scene)
        R.set_config(w, h, spp, depth)
        mean, rgb, sec = R.render_mt(1, tile, threads, 4321)
        print(f"converged {name}: {w}x{h}x{spp} depth {depth} in {sec:.1f}s = {w*h*spp/sec/1e6:.2f} Msamples/s on {threads} threads")
        cv[f"{name}_mean"] = mean[..., :3].astype(np.float32)
        cv[f"{name}_config"] = np.array([w, h, spp, depth], np.int64)
        for k, v in scene_dict(scene).items():
            cv[f"{name}_scene_{k}"] = v
    np.savez_compressed(os.path.join(OUT, "converged_task2.npz"), **cv)


def main():
    build_oracle(ref=True)
    if "--converged2" in sys.argv:
        converged_configs_2_and_4()
        return
    R = Reference()
    cam = R.set_camera()  # Globals.hpp:22-24 defaults through Renderer.hpp:321
    np.savez_compressed(os.path.join(OUT, "camera_default.npz"), view=cam.view, eye=cam.eye, sky=cam.sky)
    cam2 = R.set_camera((1, 2, -5), (0.5, 1, 0.3))
    np.savez_compressed(os.path.join(OUT, "camera_offaxis.npz"), view=cam2.view, eye=cam2.eye, sky=cam2.sky,
                        eye_in=np.array([1, 2, -5], np.float32), look_in=np.array([0.5, 1, 0.3], np.float32))
    R.set_camera()

    # RNG front end (Random.hpp:86-93, :115-127)
    np.savez_compressed(
        os.path.join(OUT, "rng.npz"),
        uniform_m1_1_seed77=R.uniform_seq(77, -1, 1, 4096),
        uniform_half_seed78=R.uniform_seq(78, -0.5, 0.5, 4096),
        uniform_0_255_seed79=R.uniform_seq(79, 0, 255, 1024),
        shell_seed5=R.shell_seq(5, 2048),
    )

    # scenes (SceneGenerators.hpp) -- seed 1 is the scene every benchmark config uses
    for seed in (1, 2):
        np.savez_compressed(os.path.join(OUT, f"scene_init_seed{seed}.npz"), **scene_dict(R.init_spheres(seed)))
        np.savez_compressed(os.path.join(OUT, f"scene_random_seed{seed}.npz"), **scene_dict(R.generate_spheres(seed)))

    # first hit on jitter-free pinhole rays (Collision.hpp:87-109 + helpers)
    fh = {}
    for name, gen in (("init", R.init_spheres), ("random", R.generate_spheres)):
        gen(1)
        for (w, h) in ((160, 90), (96, 96)):
            R.set_config(w, h, 1, 8)
            idx, t, nrm, d = R.first_hit()
            fh[f"{name}_{w}x{h}_idx"] = idx.astype(np.uint16)
            fh[f"{name}_{w}x{h}_t"] = t
            fh[f"{name}_{w}x{h}_normal"] = nrm
            fh[f"{name}_{w}x{h}_dir"] = d
    np.savez_compressed(os.path.join(OUT, "first_hit.npz"), **fh)

    # whole renders with the reference's RNG sequence seeded with 1234, one thread, both tracers
    rd = {}
    for name, gen in (("init", R.init_spheres), ("random", R.generate_spheres)):
        gen(1)
        w, h, spp, depth = 48, 48, 4, 8
        R.set_config(w, h, spp, depth)
        for tracer in (0, 1):
            mean, rgb = R.render(tracer, 1234)
            rd[f"{name}_tracer{tracer}_mean"] = mean
            rd[f"{name}_tracer{tracer}_rgb"] = rgb
        # a non-square segment through the recursive tracer
        mean, rgb = R.render(0, 99, seg=(5, 30, 7, 47))
        rd[f"{name}_seg_mean"] = mean
        rd[f"{name}_seg_rgb"] = rgb
    rd["config"] = np.array([48, 48, 4, 8, 1234], np.int64)
    np.savez_compressed(os.path.join(OUT, "renders_refseq.npz"), **rd)

    # converged reference renders for the statistical gate (PSNR >= 45 dB / RMSE <= 1e-3):
    # RenderSegmentTask over square tiles, all host threads, high spp.
    cv = {}
    threads = os.cpu_count() or 8
    for name, gen, (w, h, spp, depth, tile) in (
        ("init", R.init_spheres, (128, 128, 16384, 10, 32)),
        ("random", R.generate_spheres, (160, 90, 8192, 8, 10)),
    ):
        gen(1)
        R.set_config(w, h, spp, depth)
        mean, rgb, sec = R.render_mt(1, tile, threads, 4321)
        print(f"converged {name}: {w}x{h}x{spp} in {sec:.1f}s = {w*h*spp/sec/1e6:.2f} Msamples/s on {threads} threads")
        cv[f"{name}_mean"] = mean[..., :3].astype(np.float32)
        cv[f"{name}_config"] = np.array([w, h, spp, depth], np.int64)
    np.savez_compressed(os.path.join(OUT, "converged_task.npz"), **cv)

    # io::WritePixel known answers (IOHelpers.hpp:17-22)
    vals = np.concatenate([np.linspace(0, 255, 511, dtype=np.float32), np.array([0.2, 0.5, 1.0, 63.74, 63.76, 254.99], np.float32)])
    out = np.array([R.write_pixel((v, v * 0.5, 255 - v)) for v in vals], np.uint8)
    np.savez_compressed(os.path.join(OUT, "write_pixel.npz"), values=vals, bytes=out)

    # the BMP serialiser the reference calls (stbi_write_bmp)
    rgb = (np.arange(7 * 5 * 3) % 251).astype(np.uint8).reshape(5, 7, 3)
    path = os.path.join(OUT, "tiny_7x5.bmp")
    R.save_bmp(path, rgb)
    np.save(os.path.join(OUT, "tiny_7x5_rgb.npy"), rgb)
    print("golden fixtures written to", OUT)


if __name__ == "__main__":
    main()
