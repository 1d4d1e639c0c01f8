"""ctypes binding of include/spt_cuda.h (plumbing; see package docstring)."""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess
from dataclasses import dataclass, field
from typing import Optional, Sequence

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB_NAME = "libspt_cuda.so"

ABI_VERSION = 3  # SPT_ABI_VERSION of include/spt_cuda.h
PARTITION_AUTO, PARTITION_SAMPLES, PARTITION_BANDS = 0, 1, 2
DEFAULT_STAGING_BYTES = 8 << 30  # SPT_DEFAULT_STAGING_BYTES
STATIC_SCRATCH_RAYS = 65536  # SPT_STATIC_SCRATCH_RAYS
MEGAKERNEL, WAVEFRONT = 0, 1
FLAG_STRICT_LOOP, FLAG_COUNTERS, FLAG_CROSS_CHECK, FLAG_STREAMED, FLAG_RESIDENT, FLAG_SPLIT, FLAG_SINGLE_KERNEL = 1, 2, 4, 8, 16, 32, 64
FLAG_SERIAL_CHUNKS = 128
FLAG_OWN_PIXELS = 256
FLAG_GRID = 512
SKYBOX, REFLECTIVE, REFRACTIVE, DIFFUSE = 0, 1, 2, 3  # include/Definitions.hpp:7-13

f32p = C.POINTER(C.c_float)
u32p = C.POINTER(C.c_uint32)
u8p = C.POINTER(C.c_uint8)


class SptError(RuntimeError):
    pass


def library_path() -> str:
    return os.environ.get("SPT_CUDA_LIB") or os.path.join(HERE, LIB_NAME)


NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "--fmad=false",  # belt and braces: every rounded op is spelled out with __f*_rn as well
    "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-Xcompiler", "-fvisibility=hidden", "-shared",
]


def build_library(verbose: bool = False, force: bool = False) -> str:
    """Compile the CUDA library for sm_100a in-tree (nvcc cross-compiles without a GPU)."""
    out = library_path()
    srcs = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cuh", ".cpp", ".hpp", ".h"))]
    srcs += [os.path.join(ROOT, "include", f) for f in sorted(os.listdir(os.path.join(ROOT, "include"))) if f.endswith((".h", ".hpp"))]
    hostdir = os.path.join(ROOT, "include", "spt_host")
    if os.path.isdir(hostdir):
        srcs += [os.path.join(hostdir, f) for f in sorted(os.listdir(hostdir))]
    if not force and os.path.exists(out):
        # an explicitly named library (A/B experiments, tools/ab_bench.sh) is never rebuilt behind the caller's back
        if os.environ.get("SPT_CUDA_LIB") or all(os.path.getmtime(out) >= os.path.getmtime(s) for s in srcs):
            return out
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    units = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cpp"))]
    extra = os.environ.get("SPT_NVCC_EXTRA", "").split()  # experiments only, e.g. -DSPT_FILTER_TC=1
    cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-I", os.path.join(ROOT, "include")] + units + ["-o", out]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise SptError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return out


@dataclass
class Scene:
    """SoA scene laid out as the reference globals (include/Globals.hpp:31-37)."""

    centers: np.ndarray  # (n,4) f32  g_spheres
    radii: np.ndarray  # (n,)  f32  g_radii
    materials: np.ndarray  # (n,)  u8   g_materials
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

    def save(self, path: str) -> None:
        np.savez_compressed(path, centers=self.centers, radii=self.radii, materials=self.materials, colors=self.colors, fuzz=self.fuzz)

    @staticmethod
    def load(path: str) -> "Scene":
        z = np.load(path)
        return Scene.of(type("S", (), {k: z[k] for k in ("centers", "radii", "materials", "colors", "fuzz")}))


@dataclass
class Camera:
    """viewMatrix / eyePos / initColor (include/Globals.hpp:21-29); defaults = the reference's camera."""

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


class Segment(C.Structure):
    """== RenderSegmentData (include/Definitions.hpp:15-21)"""

    _fields_ = [("yBegin", C.c_uint32), ("yEnd", C.c_uint32), ("xBegin", C.c_uint32), ("xEnd", C.c_uint32)]


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
        ("variant", C.c_uint32),
        ("flags", C.c_uint32),
        ("band_count", C.c_uint32),
        ("band_index", C.c_uint32),
        ("split_chunk", C.c_uint32),
        ("strip_rows", C.c_uint32),
    ]


class Counters(C.Structure):
    _fields_ = [("segments", C.c_uint64), ("tests", C.c_uint64), ("samples", C.c_uint64), ("dropped", C.c_uint64), ("exact", C.c_uint64),
                ("warp_iterations", C.c_uint64)]

    def asdict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


class Timing(C.Structure):
    _fields_ = [("trace_ms", C.c_float), ("resolve_ms", C.c_float), ("launches", C.c_uint32), ("reserved", C.c_uint32),
                ("primary_ms", C.c_float), ("secondary_ms", C.c_float)]


class DeviceInfo(C.Structure):
    _fields_ = [
        ("name", C.c_char * 64),
        ("sm_count", C.c_int),
        ("cc_major", C.c_int),
        ("cc_minor", C.c_int),
        ("clock_khz", C.c_int),
        ("smem_per_block_optin", C.c_size_t),
        ("global_mem", C.c_size_t),
    ]


EXPORTS = [
    "spt_cuda_create", "spt_cuda_create_multi", "spt_cuda_device_count", "spt_cuda_set_partition", "spt_cuda_set_staging_limit",
    "spt_cuda_destroy", "spt_cuda_set_stream", "spt_cuda_device_info", "spt_cuda_set_scene",
    "spt_cuda_render", "spt_cuda_accumulate", "spt_cuda_resolve", "spt_cuda_synchronize", "spt_cuda_first_hit",
    "spt_cuda_progressive_begin", "spt_cuda_progressive_add", "spt_cuda_progressive_image", "spt_cuda_progressive_load",
    "spt_cuda_closest_hit", "spt_cuda_closest_hit_static", "spt_cuda_get_counters", "spt_cuda_reset_counters", "spt_cuda_get_timing",
    "spt_cuda_measure_fp32_peak", "spt_cuda_self_check_division", "spt_cuda_last_error", "spt_cuda_version", "spt_cuda_abi_version",
]
HOST_EXPORTS = [
    "spt_host_init_spheres", "spt_host_generate_spheres", "spt_host_generate_metal_glass", "spt_host_generate_lattice",
    "spt_host_save_scene", "spt_host_load_scene",
    "spt_host_camera_basis", "spt_host_save_bmp", "spt_host_save_pfm", "spt_host_save_hdr", "spt_host_save_accum", "spt_host_load_accum",
]

_LIB = None


def load_library():
    """dlopen the in-tree CUDA library.  Raises (never falls back) when it is missing."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = library_path()
    if not os.path.exists(path):
        raise SptError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc). "
                       "There is no CPU fallback.")
    L = C.CDLL(path)
    # struct layouts below are revision ABI_VERSION of include/spt_cuda.h: a library built from another revision is refused
    if not hasattr(L, "spt_cuda_abi_version") or (setattr(L.spt_cuda_abi_version, "restype", C.c_uint32) or L.spt_cuda_abi_version()) != ABI_VERSION:
        raise SptError(f"{path} was built from another revision of include/spt_cuda.h (want ABI {ABI_VERSION}): rebuild it")
    L.spt_cuda_last_error.restype = C.c_char_p
    L.spt_cuda_version.restype = C.c_char_p
    L.spt_cuda_create.argtypes = [C.POINTER(C.c_void_p), C.c_int]
    L.spt_cuda_create_multi.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.c_int]
    L.spt_cuda_device_count.argtypes = [C.c_void_p]
    L.spt_cuda_set_partition.argtypes = [C.c_void_p, C.c_int]
    L.spt_cuda_set_staging_limit.argtypes = [C.c_void_p, C.c_uint64]
    L.spt_cuda_destroy.argtypes = [C.c_void_p]
    L.spt_cuda_destroy.restype = None
    L.spt_cuda_set_stream.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    L.spt_cuda_device_info.argtypes = [C.c_void_p, C.POINTER(DeviceInfo)]
    L.spt_cuda_set_scene.argtypes = [C.c_void_p, C.POINTER(_CScene)]
    L.spt_cuda_render.argtypes = [C.c_void_p, C.POINTER(_CCamera), C.POINTER(_CParams), Segment, C.c_void_p, C.c_void_p]
    L.spt_cuda_accumulate.argtypes = [C.c_void_p, C.POINTER(_CCamera), C.POINTER(_CParams), Segment, C.c_void_p]
    L.spt_cuda_resolve.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, Segment, C.c_void_p]
    L.spt_cuda_synchronize.argtypes = [C.c_void_p]
    L.spt_cuda_progressive_begin.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
    L.spt_cuda_progressive_add.argtypes = [C.c_void_p, C.POINTER(_CCamera), C.POINTER(_CParams), Segment]
    L.spt_cuda_progressive_image.argtypes = [C.c_void_p, Segment, C.c_void_p, C.c_void_p]
    L.spt_cuda_progressive_load.argtypes = [C.c_void_p, C.c_void_p]
    L.spt_cuda_first_hit.argtypes = [C.c_void_p, C.POINTER(_CCamera), C.c_uint32, C.c_uint32, Segment, C.c_uint32, C.c_void_p,
                                     C.c_void_p, C.c_void_p]
    L.spt_cuda_closest_hit.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
    if hasattr(L, "spt_cuda_closest_hit_static"):  # absent from libraries built before it existed (A/B runs via SPT_CUDA_LIB)
        L.spt_cuda_closest_hit_static.argtypes = L.spt_cuda_closest_hit.argtypes
    L.spt_cuda_get_counters.argtypes = [C.c_void_p, C.POINTER(Counters)]
    L.spt_cuda_reset_counters.argtypes = [C.c_void_p]
    L.spt_cuda_get_timing.argtypes = [C.c_void_p, C.POINTER(Timing)]
    L.spt_cuda_measure_fp32_peak.argtypes = [C.c_void_p, f32p, f32p]
    if hasattr(L, "spt_cuda_self_check_division"):  # absent from libraries built before it existed (A/B runs via SPT_CUDA_LIB)
        L.spt_cuda_self_check_division.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    gen = [C.c_uint32, C.c_uint32, f32p, f32p, u8p, f32p, f32p]
    L.spt_host_init_spheres.argtypes = gen
    L.spt_host_generate_spheres.argtypes = gen
    L.spt_host_generate_metal_glass.argtypes = [C.c_uint32] + gen
    L.spt_host_generate_lattice.argtypes = [C.c_uint32] + gen
    for f in (L.spt_host_init_spheres, L.spt_host_generate_spheres, L.spt_host_generate_metal_glass, L.spt_host_generate_lattice):
        f.restype = C.c_uint32
    L.spt_host_camera_basis.argtypes = [f32p, f32p, f32p, f32p]
    L.spt_host_camera_basis.restype = None
    L.spt_host_save_bmp.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, u8p]
    L.spt_host_save_scene.argtypes = [C.c_char_p, C.c_uint32, f32p, f32p, u8p, f32p, f32p, f32p]
    L.spt_host_load_scene.argtypes = [C.c_char_p, C.c_uint32, f32p, f32p, u8p, f32p, f32p, f32p]
    L.spt_host_load_scene.restype = C.c_uint32
    L.spt_host_save_pfm.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, f32p]
    L.spt_host_save_hdr.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, f32p]
    L.spt_host_save_accum.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64, f32p]
    L.spt_host_load_accum.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, u32p, u32p, C.POINTER(C.c_uint64), f32p]
    _LIB = L
    return L


def _ccam(cam: Camera) -> _CCamera:
    c = _CCamera()
    c.view[:] = [float(v) for v in cam.view.reshape(-1)]
    c.eye[:] = [float(v) for v in cam.eye.reshape(-1)]
    c.sky[:] = [float(v) for v in cam.sky.reshape(-1)]
    return c


def _seg(seg, width, height) -> Segment:
    if seg is None:
        return Segment(0, height, 0, width)  # RenderImage's {0, g_height, 0, g_width} (Renderer.hpp:306)
    if isinstance(seg, Segment):
        return seg
    return Segment(*[int(v) for v in seg])


def _devptr(x) -> int:
    """Device pointer of a torch tensor / anything with data_ptr(), or a raw int."""
    return int(x.data_ptr()) if hasattr(x, "data_ptr") else int(x)


class PathTracer:
    """One context: RenderSegment / RenderSegmentTask as CUDA (include/spt_cuda.h).  `device` is one CUDA ordinal, or a
    sequence of them for a multi-device context (spt_cuda_create_multi: one host process, one render spread over the
    GPUs, accumulators combined over NVLink by the library's fused reduce-and-resolve kernel)."""

    def __init__(self, device=0):
        self.L = load_library()
        h = C.c_void_p()
        if isinstance(device, (list, tuple)):
            devs = (C.c_int * len(device))(*[int(d) for d in device])
            self._check(self.L.spt_cuda_create_multi(C.byref(h), devs, len(device)))
            self.devices = [int(d) for d in device]
        else:
            self._check(self.L.spt_cuda_create(C.byref(h), device))
            self.devices = [int(device)]
        self.h = h
        self.device = self.devices[0]
        self.scene: Optional[Scene] = None
        self.split_chunk = 0  # SptParams.split_chunk of every call made through this object (0 = as the staging limit allows)
        self.strip_rows = 0  # SptParams.strip_rows (0 = chosen by the library)

    def device_count(self) -> int:
        return int(self.L.spt_cuda_device_count(self.h))

    def set_partition(self, partition: int) -> None:
        self._check(self.L.spt_cuda_set_partition(self.h, int(partition)))

    def set_staging_limit(self, nbytes: int) -> None:
        """Bound of the first-hit staging buffers of split rendering, per device (0 = SPT_DEFAULT_STAGING_BYTES)."""
        self._check(self.L.spt_cuda_set_staging_limit(self.h, int(nbytes)))

    def _check(self, rc: int):
        if rc != 0:
            raise SptError(f"spt_cuda error {rc}: {self.L.spt_cuda_last_error().decode()}")

    def close(self):
        if getattr(self, "h", None):
            self.L.spt_cuda_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ------------------------------------------------------------------
    def device_info(self) -> dict:
        d = DeviceInfo()
        self._check(self.L.spt_cuda_device_info(self.h, C.byref(d)))
        return {"name": d.name.decode(), "sm_count": d.sm_count, "cc": (d.cc_major, d.cc_minor), "clock_khz": d.clock_khz,
                "smem_per_block_optin": d.smem_per_block_optin, "global_mem": d.global_mem}

    def set_stream(self, cuda_stream: Optional[int]):
        """Launch on the given cudaStream_t (0 = legacy default stream); None restores the context's own stream."""
        if cuda_stream is None:
            self._check(self.L.spt_cuda_set_stream(self.h, None, 1))
        else:
            self._check(self.L.spt_cuda_set_stream(self.h, C.c_void_p(cuda_stream), 0))

    def set_scene(self, scene) -> None:
        sc = Scene.of(scene)
        cs = _CScene(sc.n, sc.centers.ctypes.data_as(f32p), sc.radii.ctypes.data_as(f32p), sc.materials.ctypes.data_as(u8p),
                     sc.colors.ctypes.data_as(f32p), sc.fuzz.ctypes.data_as(f32p))
        self._check(self.L.spt_cuda_set_scene(self.h, C.byref(cs)))
        self.scene = sc

    def _params(self, width, height, depth, pass_cap, sample_begin, sample_end, seed, variant, flags, bands=None) -> _CParams:
        count, index = bands if bands else (0, 0)  # (band_count, band_index): interleaved 8-row bands, include/spt_cuda.h
        return _CParams(width, height, depth, pass_cap, sample_begin, sample_end, seed, variant, flags, count, index, int(self.split_chunk), int(self.strip_rows))

    def render(self, cam, width, height, spp, depth, pass_cap=0, seed=1, seg=None, sample_begin=0, variant=MEGAKERNEL,
               flags=0, rgb8: Optional[np.ndarray] = None, want_accum: bool = False):
        """Whole hot path with host buffers (spt_cuda_render).  Returns (rgb8 (H,W,3) u8, accum (H,W,4) f32 or None)."""
        cm = _ccam(Camera.of(cam))
        if rgb8 is None:
            rgb8 = np.zeros((height, width, 3), np.uint8)
        accum = np.zeros((height, width, 4), np.float32) if want_accum else None
        p = self._params(width, height, depth, pass_cap, sample_begin, sample_begin + spp, seed, variant, flags)
        self._check(self.L.spt_cuda_render(self.h, C.byref(cm), C.byref(p), _seg(seg, width, height), rgb8.ctypes.data,
                                           accum.ctypes.data if accum is not None else None))
        return rgb8, accum

    def accumulate(self, cam, width, height, sample_begin, sample_end, depth, d_accum, pass_cap=0, seed=1, seg=None,
                   variant=MEGAKERNEL, flags=0, bands=None):
        """Adds samples [sample_begin, sample_end) into a DEVICE float4 accumulator (async).  bands = (count, index)
        restricts the call to the 8-row bands b of the segment with b % count == index."""
        cm = _ccam(Camera.of(cam))
        p = self._params(width, height, depth, pass_cap, sample_begin, sample_end, seed, variant, flags, bands)
        self._check(self.L.spt_cuda_accumulate(self.h, C.byref(cm), C.byref(p), _seg(seg, width, height), C.c_void_p(_devptr(d_accum))))

    def resolve(self, d_accum, width, height, d_rgb8, seg=None):
        self._check(self.L.spt_cuda_resolve(self.h, C.c_void_p(_devptr(d_accum)), width, height, _seg(seg, width, height),
                                            C.c_void_p(_devptr(d_rgb8))))

    # progressive rendering on the context-owned accumulator (spt_cuda_progressive_*)
    def progressive_begin(self, width, height):
        self._check(self.L.spt_cuda_progressive_begin(self.h, width, height))
        self._prog = (width, height)

    def progressive_add(self, cam, sample_begin, sample_end, depth, pass_cap=0, seed=1, seg=None, variant=MEGAKERNEL, flags=0,
                        bands=None):
        width, height = self._prog
        cm = _ccam(Camera.of(cam))
        p = self._params(width, height, depth, pass_cap, sample_begin, sample_end, seed, variant, flags, bands)
        self._check(self.L.spt_cuda_progressive_add(self.h, C.byref(cm), C.byref(p), _seg(seg, width, height)))

    def progressive_image(self, seg=None, want_accum=False, rgb8: Optional[np.ndarray] = None):
        width, height = self._prog
        if rgb8 is None:
            rgb8 = np.zeros((height, width, 3), np.uint8)
        accum = np.zeros((height, width, 4), np.float32) if want_accum else None
        self._check(self.L.spt_cuda_progressive_image(self.h, _seg(seg, width, height), rgb8.ctypes.data,
                                                      accum.ctypes.data if want_accum else None))
        return rgb8, accum

    def progressive_load(self, accum: np.ndarray):
        width, height = self._prog
        accum = np.ascontiguousarray(accum, np.float32)
        assert accum.shape == (height, width, 4)
        self._check(self.L.spt_cuda_progressive_load(self.h, accum.ctypes.data))

    def synchronize(self):
        self._check(self.L.spt_cuda_synchronize(self.h))

    def first_hit(self, cam, width, height, seg=None, flags=0):
        cm = _ccam(Camera.of(cam))
        s = _seg(seg, width, height)
        npx = (s.yEnd - s.yBegin) * (s.xEnd - s.xBegin)
        idx = np.empty(npx, np.uint32)
        t = np.empty(npx, np.float32)
        nrm = np.empty((npx, 3), np.float32)
        self._check(self.L.spt_cuda_first_hit(self.h, C.byref(cm), width, height, s, flags, idx.ctypes.data, t.ctypes.data, nrm.ctypes.data))
        return idx, t, nrm

    def closest_hit(self, dirs: Sequence, orgs: Sequence, flags=0, want_evals=False):
        dirs = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        orgs = np.ascontiguousarray(orgs, np.float32).reshape(-1, 3)
        idx = np.empty(dirs.shape[0], np.uint32)
        ev = np.empty(dirs.shape[0], np.uint32) if want_evals else None
        self._check(self.L.spt_cuda_closest_hit(self.h, dirs.shape[0], dirs.ctypes.data, orgs.ctypes.data, flags, idx.ctypes.data,
                                                ev.ctypes.data if want_evals else None))
        return (idx, ev) if want_evals else idx

    def closest_hit_static(self, dirs: Sequence, orgs: Sequence, flags=0, want_evals=False):
        """spt_cuda_closest_hit_static: the same through the library's static device scratch -- no device allocation."""
        dirs = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        orgs = np.ascontiguousarray(orgs, np.float32).reshape(-1, 3)
        idx = np.empty(dirs.shape[0], np.uint32)
        ev = np.empty(dirs.shape[0], np.uint32) if want_evals else None
        self._check(self.L.spt_cuda_closest_hit_static(self.h, dirs.shape[0], dirs.ctypes.data, orgs.ctypes.data, flags, idx.ctypes.data,
                                                       ev.ctypes.data if want_evals else None))
        return (idx, ev) if want_evals else idx

    def counters(self) -> dict:
        c = Counters()
        self._check(self.L.spt_cuda_get_counters(self.h, C.byref(c)))
        return c.asdict()

    def reset_counters(self):
        self._check(self.L.spt_cuda_reset_counters(self.h))

    def timing(self) -> dict:
        t = Timing()
        self._check(self.L.spt_cuda_get_timing(self.h, C.byref(t)))
        return {"trace_ms": float(t.trace_ms), "resolve_ms": float(t.resolve_ms), "launches": int(t.launches),
                "primary_ms": float(t.primary_ms), "secondary_ms": float(t.secondary_ms)}

    def self_check_division(self, sets: int, seed: int = 1):
        """(operand sets tested, mismatches against __fdiv_rn) of the shared-reciprocal division (spt_cuda_self_check_division)."""
        t, m = C.c_uint64(), C.c_uint64()
        self._check(self.L.spt_cuda_self_check_division(self.h, seed, sets, C.byref(t), C.byref(m)))
        return int(t.value), int(m.value)

    def measure_fp32_peak(self):
        a, b = C.c_float(), C.c_float()
        self._check(self.L.spt_cuda_measure_fp32_peak(self.h, C.byref(a), C.byref(b)))
        return float(a.value), float(b.value)


# ---------------------------------------------------------------------------------------------------------
# host mirror (include/spt_host.h): the reference's scene generators / camera / BMP output, no GPU needed
def _generate(fn, *head) -> Scene:
    n = fn(*head, 0, None, None, None, None, None)
    sc = Scene(np.zeros((n, 4), np.float32), np.zeros(n, np.float32), np.zeros(n, np.uint8), np.zeros((n, 4), np.float32),
               np.zeros(n, np.float32))
    got = fn(*head, n, sc.centers.ctypes.data_as(f32p), sc.radii.ctypes.data_as(f32p), sc.materials.ctypes.data_as(u8p),
             sc.colors.ctypes.data_as(f32p), sc.fuzz.ctypes.data_as(f32p))
    assert got == n
    return sc


def init_spheres(seed: int = 1) -> Scene:
    """InitSpheres() (SceneGenerators.hpp:68-133) with the RNG seeded by `seed`."""
    return _generate(load_library().spt_host_init_spheres, seed)


def generate_spheres(seed: int = 1) -> Scene:
    """GenerateSpheres() (SceneGenerators.hpp:6-66) with the RNG seeded by `seed`."""
    return _generate(load_library().spt_host_generate_spheres, seed)


def generate_metal_glass(seed: int = 4, count: int = 500) -> Scene:
    """BASELINE config 4: ground + `count` mirror / glass spheres."""
    return _generate(load_library().spt_host_generate_metal_glass, seed, count)


def generate_lattice(seed: int = 5, side: int = 128) -> Scene:
    """BASELINE config 5: ground + side*side - 1 spheres on a jittered lattice."""
    return _generate(load_library().spt_host_generate_lattice, seed, side)


def make_camera(eye=(0, 1, -3), look=(0, 1, 0), up=(0, 1, 0), sky=(137, 207, 240)) -> Camera:
    """viewMatrix = Transpose(CreateCameraBasisMatrix(eyePos, lookAt, upDir)) (Renderer.hpp:321)."""
    view = np.zeros(16, np.float32)
    load_library().spt_host_camera_basis((C.c_float * 3)(*eye), (C.c_float * 3)(*look), (C.c_float * 3)(*up), view.ctypes.data_as(f32p))
    return Camera(view.reshape(4, 4), np.array([*eye, 0], np.float32), np.array([*sky, 0], np.float32))


def save_bmp(path: str, rgb8: np.ndarray) -> None:
    """io::SaveImage's serialiser (stbi_write_bmp-compatible)."""
    rgb8 = np.ascontiguousarray(rgb8, np.uint8)
    h, w = rgb8.shape[:2]
    if not load_library().spt_host_save_bmp(path.encode(), w, h, rgb8.ctypes.data_as(u8p)):
        raise SptError("cannot write " + path)


def save_pfm(path: str, accum: np.ndarray) -> None:
    """Linear float image (mean radiance / 255) as a Portable Float Map; accum = (H, W, 4) sums and counts."""
    accum = np.ascontiguousarray(accum, np.float32)
    h, w = accum.shape[:2]
    if not load_library().spt_host_save_pfm(path.encode(), w, h, accum.ctypes.data_as(f32p)):
        raise SptError("cannot write " + path)


def save_hdr(path: str, accum: np.ndarray) -> None:
    """The same linear image as Radiance RGBE (.hdr, flat scanlines, top row first)."""
    accum = np.ascontiguousarray(accum, np.float32)
    h, w = accum.shape[:2]
    if not load_library().spt_host_save_hdr(path.encode(), w, h, accum.ctypes.data_as(f32p)):
        raise SptError("cannot write " + path)


def save_accum(path: str, accum: np.ndarray, samples_done: int, depth: int, seed: int) -> None:
    accum = np.ascontiguousarray(accum, np.float32)
    h, w = accum.shape[:2]
    if not load_library().spt_host_save_accum(path.encode(), w, h, samples_done, depth, seed, accum.ctypes.data_as(f32p)):
        raise SptError("cannot write " + path)


def load_accum(path: str, width: int, height: int):
    """Returns (accum (H, W, 4), samples_done, depth, seed) or None when the file does not match."""
    accum = np.zeros((height, width, 4), np.float32)
    sd, dp, seed = C.c_uint32(), C.c_uint32(), C.c_uint64()
    if not load_library().spt_host_load_accum(path.encode(), width, height, C.byref(sd), C.byref(dp), C.byref(seed), accum.ctypes.data_as(f32p)):
        return None
    return accum, int(sd.value), int(dp.value), int(seed.value)


def save_scene_file(path: str, scene, eye=(0, 1, -3), look=(0, 1, 0), up=(0, 1, 0), sky=(137, 207, 240)) -> None:
    """The scene + camera globals as a .sptscene file (include/spt_host/SceneFile.hpp)."""
    sc = Scene.of(scene)
    cam = np.array([*eye, 0, *look, 0, *up, 0, *sky, 0], np.float32)
    ok = load_library().spt_host_save_scene(path.encode(), sc.n, sc.centers.ctypes.data_as(f32p), sc.radii.ctypes.data_as(f32p),
                                            sc.materials.ctypes.data_as(u8p), sc.colors.ctypes.data_as(f32p), sc.fuzz.ctypes.data_as(f32p),
                                            cam.ctypes.data_as(f32p))
    if not ok:
        raise SptError("cannot write " + path)


def load_scene_file(path: str):
    """Returns (Scene, camera dict with eye / look / up / sky) or raises on a malformed file."""
    L = load_library()
    cam = np.zeros(16, np.float32)
    n = L.spt_host_load_scene(path.encode(), 0, None, None, None, None, None, cam.ctypes.data_as(f32p))
    if n == 0:
        raise SptError("not a scene file: " + path)
    sc = Scene(np.zeros((n, 4), np.float32), np.zeros(n, np.float32), np.zeros(n, np.uint8), np.zeros((n, 4), np.float32),
               np.zeros(n, np.float32))
    got = L.spt_host_load_scene(path.encode(), n, sc.centers.ctypes.data_as(f32p), sc.radii.ctypes.data_as(f32p),
                                sc.materials.ctypes.data_as(u8p), sc.colors.ctypes.data_as(f32p), sc.fuzz.ctypes.data_as(f32p),
                                cam.ctypes.data_as(f32p))
    assert got == n
    return sc, {"eye": cam[0:3].copy(), "look": cam[4:7].copy(), "up": cam[8:11].copy(), "sky": cam[12:15].copy()}
