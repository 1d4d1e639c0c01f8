"""simplepathtracer_b200 -- B200-native (sm_100a) path-tracing hot path of SimplePathTracer.

Python here is plumbing only: a ctypes binding of the C ABI declared in ``include/spt_cuda.h``
(built in-tree as ``simplepathtracer_b200/libspt_cuda.so`` by ``__graft_entry__.build()``), used
by the parity tests and by ``bench.py``.  The product is the CUDA library and the C++ host mirror
``include/CudaPathTracer.hpp``.

There is no CPU fallback: importing works anywhere, but creating a ``PathTracer`` raises when the
library or a CUDA device is missing.
"""
from .binding import (  # noqa: F401
    Camera,
    PathTracer,
    Scene,
    Segment,
    SptError,
    build_library,
    library_path,
    load_library,
    MEGAKERNEL,
    WAVEFRONT,
    FLAG_STRICT_LOOP,
    FLAG_COUNTERS,
    FLAG_CROSS_CHECK,
    FLAG_STREAMED,
    FLAG_RESIDENT,
    FLAG_SPLIT,
    FLAG_SINGLE_KERNEL,
    FLAG_SERIAL_CHUNKS,
    FLAG_OWN_PIXELS,
    FLAG_GRID,
    DIFFUSE,
    REFLECTIVE,
    REFRACTIVE,
    SKYBOX,
    PARTITION_AUTO,
    PARTITION_SAMPLES,
    PARTITION_BANDS,
    DEFAULT_STAGING_BYTES,
    STATIC_SCRATCH_RAYS,
    init_spheres,
    generate_spheres,
    generate_metal_glass,
    generate_lattice,
    make_camera,
    save_bmp,
    save_scene_file,
    load_scene_file,
    save_pfm,
    save_hdr,
    save_accum,
    load_accum,
)

__version__ = "0.2.0"
