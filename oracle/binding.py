"""ctypes bindings for oracle/libspt_oracle.so and oracle/_ref/libspt_ref*.so.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = "/root/reference"

f32p = C.POINTER(C.c_float)
u32p = C.POINTER(C.c_uint32)
u8p = C.POINTER(C.c_uint8)


def _p(a, typ):
    return None if a is None else a.ctypes.data_as(typ)


@dataclass
class Scene:
    """SoA scene exactly as the reference globals (include/Globals.hpp:31-37)."""

    centers: np.ndarray  # (n,4) f32  g_spheres
    radii: np.ndarray  # (n,)  f32  g_radii
    materials: np.ndarray  # (n,)  u8   g_materials (Definitions.hpp:7-13)
    colors: np.ndarray  # (n,4) f32  g_colors
    fuzz: np.ndarray  # (n,)  f32  g_diffuses

    @property
    def n(self) -> int:
        return int(self.radii.shape[0])

    @staticmethod
    def of(obj) -> "Scene":
        return Scene(
            np.ascontiguousarray(obj.centers, np.float32).reshape(-1, 4),
            np.ascontiguousarray(obj.radii, np.float32).reshape(-1),
            np.ascontiguousarray(obj.materials, np.uint8).reshape(-1),
            np.ascontiguousarray(obj.colors, np.float32).reshape(-1, 4),
            np.ascontiguousarray(obj.fuzz, np.float32).reshape(-1),
        )


@dataclass
class Camera:
    view: np.ndarray = field(default_factory=lambda: np.diag([1, 1, 1, 0]).astype(np.float32))
    eye: np.ndarray = field(default_factory=lambda: np.array([0, 1, -3, 0], np.float32))
    sky: np.ndarray = field(default_factory=lambda: np.array([137, 207, 240, 0], np.float32))

    @staticmethod
    def of(obj) -> "Camera":
        return Camera(
            np.ascontiguousarray(obj.view, np.float32).reshape(4, 4),
            np.ascontiguousarray(obj.eye, np.float32).reshape(4),
            np.ascontiguousarray(obj.sky, np.float32).reshape(4),
        )


class _CScene(C.Structure):
    _fields_ = [("n", C.c_uint32), ("centers4", f32p), ("radii", f32p), ("materials", u8p), ("colors4", f32p), ("fuzz", f32p)]


class _CCamera(C.Structure):
    _fields_ = [("view", C.c_float * 16), ("eye", C.c_float * 4), ("sky", C.c_float * 4)]


class _CParams(C.Structure):
    _fields_ = [
        ("width", C.c_uint32),
        ("height", C.c_uint32),
        ("depth", C.c_uint32),
        ("pass_cap", C.c_uint32),
        ("sample_begin", C.c_uint32),
        ("sample_end", C.c_uint32),
        ("seed", C.c_uint64),
        ("pow_mode", C.c_int),
    ]


class Counters(C.Structure):
    _fields_ = [
        ("segments", C.c_uint64),
        ("tests", C.c_uint64),
        ("passed", C.c_uint64),
        ("samples", C.c_uint64),
        ("dropped", C.c_uint64),
        ("rng_draws", C.c_uint64),
    ]

    def asdict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


def build_oracle(ref: bool = True) -> None:
    """Compile the C restatement and, when /root/reference is present, the reference itself."""
    targets = ["all"]
    if ref and os.path.isdir(os.path.join(REF_ROOT, "include")):
        targets.append("ref")
        if os.path.exists(os.path.join(os.path.dirname(HERE), "simplepathtracer_b200", "libspt_cuda.so")):
            targets.append("integration")  # the drop-in shown inside the reference's own translation unit
    subprocess.run(["make", "-s", "-C", HERE] + targets, check=True)


def ref_available(wide: bool = False) -> bool:
    return os.path.exists(os.path.join(HERE, "_ref", "libspt_ref_wide.so" if wide else "libspt_ref.so"))


def _cscene(sc: Scene) -> _CScene:
    return _CScene(sc.n, _p(sc.centers, f32p), _p(sc.radii, f32p), _p(sc.materials, u8p), _p(sc.colors, f32p), _p(sc.fuzz, f32p))


def _ccam(cam: Camera) -> _CCamera:
    c = _CCamera()
    c.view[:] = [float(v) for v in cam.view.reshape(-1)]
    c.eye[:] = [float(v) for v in cam.eye.reshape(-1)]
    c.sky[:] = [float(v) for v in cam.sky.reshape(-1)]
    return c


class Oracle:
    """The plain-C restatement."""

    POW_EXACT, POW_LIBM = 0, 1
    RECURSIVE, TASK = 0, 1

    def __init__(self):
        path = os.path.join(HERE, "libspt_oracle.so")
        if not os.path.exists(path):
            build_oracle(ref=False)
        L = self.L = C.CDLL(path)
        L.spo_splitmix_next.restype = C.c_uint32
        L.spo_splitmix_next.argtypes = [C.POINTER(C.c_uint64)]
        L.spo_splitmix_seed.restype = C.c_uint64
        L.spo_splitmix_seed.argtypes = [C.c_uint32]
        L.spo_u32_to_unit.restype = C.c_float
        L.spo_u32_to_unit.argtypes = [C.c_uint32]
        L.spo_stream_key.restype = C.c_uint64
        L.spo_stream_key.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32]
        L.spo_event_state.restype = C.c_uint64
        L.spo_event_state.argtypes = [C.c_uint64, C.c_uint32]
        L.spo_pow5.restype = C.c_float
        L.spo_pow5.argtypes = [C.c_float, C.c_int]
        L.spo_closest_hit.restype = C.c_uint32
        L.spo_init_spheres.restype = C.c_uint32
        L.spo_generate_spheres.restype = C.c_uint32
        L.spo_sample_radiance.restype = C.c_int

    # -- RNG ---------------------------------------------------------------
    def uniform_seq(self, seed32, lo, hi, count):
        out = np.empty(count, np.float32)
        self.L.spo_rng_uniform_seq(C.c_uint32(seed32), C.c_float(lo), C.c_float(hi), C.c_uint32(count), _p(out, f32p))
        return out

    def shell_seq(self, seed32, count):
        out = np.empty((count, 3), np.float32)
        self.L.spo_rng_shell_seq(C.c_uint32(seed32), C.c_uint32(count), _p(out, f32p))
        return out

    def splitmix_seq(self, state, count):
        s = C.c_uint64(state)
        return np.array([self.L.spo_splitmix_next(C.byref(s)) for _ in range(count)], np.uint32)

    # -- camera ------------------------------------------------------------
    def camera(self, eye=(0, 1, -3), look=(0, 1, 0), up=(0, 1, 0), sky=(137, 207, 240)) -> Camera:
        view = np.zeros(16, np.float32)
        self.L.spo_camera_basis((C.c_float * 3)(*eye), (C.c_float * 3)(*look), (C.c_float * 3)(*up), _p(view, f32p))
        return Camera(view.reshape(4, 4), np.array([*eye, 0], np.float32), np.array([*sky, 0], np.float32))

    # -- scenes ------------------------------------------------------------
    def _gen(self, fn, seed32) -> Scene:
        cap = 1024
        sc = Scene(np.zeros((cap, 4), np.float32), np.zeros(cap, np.float32), np.zeros(cap, np.uint8),
                   np.zeros((cap, 4), np.float32), np.zeros(cap, np.float32))
        n = fn(C.c_uint32(seed32), C.c_uint32(cap), _p(sc.centers, f32p), _p(sc.radii, f32p), _p(sc.materials, u8p),
               _p(sc.colors, f32p), _p(sc.fuzz, f32p))
        return Scene(sc.centers[:n].copy(), sc.radii[:n].copy(), sc.materials[:n].copy(), sc.colors[:n].copy(), sc.fuzz[:n].copy())

    def init_spheres(self, seed32=1) -> Scene:
        return self._gen(self.L.spo_init_spheres, seed32)

    def generate_spheres(self, seed32=1) -> Scene:
        return self._gen(self.L.spo_generate_spheres, seed32)

    # -- collision ---------------------------------------------------------
    def closest_hit_many(self, scene, dirs, orgs):
        sc = Scene.of(scene)
        dirs = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        orgs = np.ascontiguousarray(orgs, np.float32).reshape(-1, 3)
        idx = np.empty(dirs.shape[0], np.uint32)
        cs = _cscene(sc)
        self.L.spo_closest_hit_many(C.byref(cs), C.c_uint32(dirs.shape[0]), _p(dirs, f32p), _p(orgs, f32p), _p(idx, u32p))
        return idx

    def first_hit(self, scene, cam, width, height, seg=None):
        sc, cm = Scene.of(scene), Camera.of(cam)
        y0, y1, x0, x1 = seg if seg else (0, height, 0, width)
        npx = (y1 - y0) * (x1 - x0)
        idx = np.empty(npx, np.uint32)
        t = np.empty(npx, np.float32)
        nrm = np.empty((npx, 3), np.float32)
        d = np.empty((npx, 3), np.float32)
        cs, cc = _cscene(sc), _ccam(cm)
        self.L.spo_first_hit(C.byref(cs), C.byref(cc), C.c_uint32(width), C.c_uint32(height), C.c_uint32(x0), C.c_uint32(x1),
                             C.c_uint32(y0), C.c_uint32(y1), _p(idx, u32p), _p(t, f32p), _p(nrm, f32p), _p(d, f32p))
        return idx, t, nrm, d

    # -- renders -----------------------------------------------------------
    def render_accum(self, scene, cam, width, height, spp, depth, pass_cap=0, seed=1, seg=None, sample_begin=0,
                     pow_mode=0, accum=None):
        """Counter-RNG render; returns (accum (H,W,4) f32 unflipped, counters dict)."""
        sc, cm = Scene.of(scene), Camera.of(cam)
        y0, y1, x0, x1 = seg if seg else (0, height, 0, width)
        if accum is None:
            accum = np.zeros((height, width, 4), np.float32)
        p = _CParams(width, height, depth, pass_cap, sample_begin, sample_begin + spp, seed, pow_mode)
        ctr = Counters()
        cs, cc = _cscene(sc), _ccam(cm)
        self.L.spo_render_accum(C.byref(cs), C.byref(cc), C.byref(p), C.c_uint32(y0), C.c_uint32(y1), C.c_uint32(x0),
                                C.c_uint32(x1), _p(accum, f32p), C.byref(ctr))
        return accum, ctr.asdict()

    def sample_radiance(self, scene, cam, width, height, depth, x, y, sample, pass_cap=0, seed=1, pow_mode=0):
        sc, cm = Scene.of(scene), Camera.of(cam)
        p = _CParams(width, height, depth, pass_cap, 0, 0, seed, pow_mode)
        rgb = np.zeros(4, np.float32)
        cs, cc = _cscene(sc), _ccam(cm)
        ok = self.L.spo_sample_radiance(C.byref(cs), C.byref(cc), C.byref(p), C.c_uint32(x), C.c_uint32(y), C.c_uint32(sample),
                                        _p(rgb, f32p), None)
        return bool(ok), rgb

    def render_refseq(self, scene, cam, width, height, spp, depth, tracer, seed32, seg=None, pow_mode=1):
        """Replay of RenderSegment / RenderSegmentTask with the reference RNG sequence.
        Returns (mean (H,W,4) f32 at flipped slots, rgb8 (H,W,3), counters)."""
        sc, cm = Scene.of(scene), Camera.of(cam)
        y0, y1, x0, x1 = seg if seg else (0, height, 0, width)
        mean = np.zeros((height, width, 4), np.float32)
        rgb = np.zeros((height, width, 3), np.uint8)
        ctr = Counters()
        cs, cc = _cscene(sc), _ccam(cm)
        self.L.spo_render_refseq(C.byref(cs), C.byref(cc), C.c_uint32(width), C.c_uint32(height), C.c_uint32(spp),
                                 C.c_uint32(depth), C.c_int(tracer), C.c_int(pow_mode), C.c_uint32(y0), C.c_uint32(y1),
                                 C.c_uint32(x0), C.c_uint32(x1), C.c_uint32(seed32), _p(mean, f32p), _p(rgb, u8p), C.byref(ctr))
        return mean, rgb, ctr.asdict()

    def resolve(self, accum, seg=None):
        accum = np.ascontiguousarray(accum, np.float32)
        height, width = accum.shape[:2]
        y0, y1, x0, x1 = seg if seg else (0, height, 0, width)
        rgb = np.zeros((height, width, 3), np.uint8)
        self.L.spo_resolve(_p(accum, f32p), C.c_uint32(width), C.c_uint32(height), C.c_uint32(y0), C.c_uint32(y1),
                           C.c_uint32(x0), C.c_uint32(x1), _p(rgb, u8p))
        return rgb

    def write_pixel(self, rgb):
        a = np.ascontiguousarray(rgb, np.float32).reshape(3)
        out = np.zeros(3, np.uint8)
        self.L.spo_write_pixel(_p(a, f32p), _p(out, u8p))
        return out

    def pow5(self, x, mode=0):
        return float(self.L.spo_pow5(C.c_float(x), C.c_int(mode)))


class Reference:
    """The unmodified reference, compiled from /root/reference by oracle/Makefile (target ref)."""

    RECURSIVE, TASK = 0, 1

    def __init__(self, wide: bool = False):
        path = os.path.join(HERE, "_ref", "libspt_ref_wide.so" if wide else "libspt_ref.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path + " (build with: make -C oracle ref; needs /root/reference)")
        L = self.L = C.CDLL(path)
        L.sptref_render_mt.restype = C.c_double
        L.sptref_scene_size.restype = C.c_uint32
        L.sptref_set_scene.restype = C.c_int
        self.wide = bool(L.sptref_wide_index())

    def set_config(self, width, height, spp, depth):
        self.width, self.height = width, height
        self.L.sptref_set_config(C.c_uint32(width), C.c_uint32(height), C.c_uint32(spp), C.c_uint32(depth))

    def set_camera(self, eye=(0, 1, -3), look=(0, 1, 0)) -> Camera:
        self.L.sptref_set_camera((C.c_float * 3)(*eye), (C.c_float * 3)(*look))
        view = np.zeros(16, np.float32)
        e = np.zeros(4, np.float32)
        sky = np.zeros(4, np.float32)
        self.L.sptref_get_camera(_p(view, f32p), _p(e, f32p))
        self.L.sptref_get_sky(_p(sky, f32p))
        return Camera(view.reshape(4, 4), e, sky)

    def get_scene(self) -> Scene:
        n = self.L.sptref_scene_size()
        sc = Scene(np.zeros((n, 4), np.float32), np.zeros(n, np.float32), np.zeros(n, np.uint8), np.zeros((n, 4), np.float32),
                   np.zeros(n, np.float32))
        self.L.sptref_get_scene(_p(sc.centers, f32p), _p(sc.radii, f32p), _p(sc.materials, u8p), _p(sc.colors, f32p), _p(sc.fuzz, f32p))
        return sc

    def init_spheres(self, seed32=1) -> Scene:
        self.L.sptref_init_spheres(C.c_uint32(seed32))
        return self.get_scene()

    def generate_spheres(self, seed32=1) -> Scene:
        self.L.sptref_generate_spheres(C.c_uint32(seed32))
        return self.get_scene()

    def set_scene(self, scene):
        sc = Scene.of(scene)
        self._keep = sc
        rc = self.L.sptref_set_scene(C.c_uint32(sc.n), _p(sc.centers, f32p), _p(sc.radii, f32p), _p(sc.materials, u8p),
                                     _p(sc.colors, f32p), _p(sc.fuzz, f32p))
        if rc != 0:
            raise ValueError("scene has > 255 spheres: needs the wide-index build (Collision.hpp:87-92)")

    def first_hit(self, seg=None):
        y0, y1, x0, x1 = seg if seg else (0, self.height, 0, self.width)
        npx = (y1 - y0) * (x1 - x0)
        idx = np.empty(npx, np.uint32)
        t = np.empty(npx, np.float32)
        nrm = np.empty((npx, 3), np.float32)
        d = np.empty((npx, 3), np.float32)
        self.L.sptref_first_hit(C.c_uint32(x0), C.c_uint32(x1), C.c_uint32(y0), C.c_uint32(y1), _p(idx, u32p), _p(t, f32p),
                                _p(nrm, f32p), _p(d, f32p))
        return idx, t, nrm, d

    def closest_hit_many(self, dirs, orgs):
        dirs = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        orgs = np.ascontiguousarray(orgs, np.float32).reshape(-1, 3)
        idx = np.empty(dirs.shape[0], np.uint32)
        self.L.sptref_closest_hit(C.c_uint32(dirs.shape[0]), _p(dirs, f32p), _p(orgs, f32p), _p(idx, u32p))
        return idx

    def render(self, tracer, seed32, seg=None):
        y0, y1, x0, x1 = seg if seg else (0, self.height, 0, self.width)
        mean = np.zeros((self.height, self.width, 4), np.float32)
        rgb = np.zeros((self.height, self.width, 3), np.uint8)
        self.L.sptref_render(C.c_int(tracer), C.c_uint32(y0), C.c_uint32(y1), C.c_uint32(x0), C.c_uint32(x1), C.c_uint32(seed32),
                             _p(mean, f32p), _p(rgb, u8p))
        return mean, rgb

    def render_mt(self, tracer, tile, threads, seed32):
        mean = np.zeros((self.height, self.width, 4), np.float32)
        rgb = np.zeros((self.height, self.width, 3), np.uint8)
        sec = self.L.sptref_render_mt(C.c_int(tracer), C.c_uint32(tile), C.c_uint32(threads), C.c_uint32(seed32), _p(mean, f32p),
                                      _p(rgb, u8p))
        return mean, rgb, float(sec)

    def uniform_seq(self, seed32, lo, hi, count):
        out = np.empty(count, np.float32)
        self.L.sptref_rng_uniform(C.c_uint32(seed32), C.c_float(lo), C.c_float(hi), C.c_uint32(count), _p(out, f32p))
        return out

    def shell_seq(self, seed32, count):
        out = np.empty((count, 3), np.float32)
        self.L.sptref_rng_shell(C.c_uint32(seed32), C.c_uint32(count), _p(out, f32p))
        return out

    def write_pixel(self, rgb):
        out = np.zeros(3, np.uint8)
        self.L.sptref_write_pixel(C.c_float(rgb[0]), C.c_float(rgb[1]), C.c_float(rgb[2]), _p(out, u8p))
        return out

    def save_bmp(self, path, rgb8):
        rgb8 = np.ascontiguousarray(rgb8, np.uint8)
        h, w = rgb8.shape[:2]
        return self.L.sptref_save_bmp(path.encode(), C.c_uint32(w), C.c_uint32(h), _p(rgb8, u8p))
