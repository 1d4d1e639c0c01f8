#!/usr/bin/env python3
"""Write a golden scene + camera as the flat binary tools/tensor_filter_ab.cpp reads (no oracle, no GPU):
    python tools/export_scene.py build_ab/c3_scene.bin [scene_random_seed1] [camera_default]
Layout: uint32 n | SptCamera (view[16], eye[4], sky[4]) | centers4 | radii | colors4 | fuzz | materials (uint8)."""
import os
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = sys.argv[1]
scene = np.load(os.path.join(ROOT, "tests", "golden", (sys.argv[2] if len(sys.argv) > 2 else "scene_random_seed1") + ".npz"))
cam = np.load(os.path.join(ROOT, "tests", "golden", (sys.argv[3] if len(sys.argv) > 3 else "camera_default") + ".npz"))
n = len(scene["radii"])
with open(out, "wb") as f:
    f.write(struct.pack("<I", n))
    for k in ("view", "eye", "sky"):
        f.write(np.ascontiguousarray(cam[k], np.float32).tobytes())
    for k, dt in (("centers", np.float32), ("radii", np.float32), ("colors", np.float32), ("fuzz", np.float32), ("materials", np.uint8)):
        f.write(np.ascontiguousarray(scene[k], dt).tobytes())
print(out, n, "spheres")
