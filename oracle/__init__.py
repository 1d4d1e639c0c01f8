"""TEST INFRASTRUCTURE ONLY -- ctypes bindings for the CPU oracle.

* ``Oracle``  : the plain-C restatement (oracle/spt_oracle.c -> libspt_oracle.so)
* ``Reference``: the UNMODIFIED reference headers compiled with shims
  (oracle/ref_shim -> oracle/_ref/libspt_ref*.so), when it has been built.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference``
legs may import this package.  The product (simplepathtracer_b200) never does.
"""
from .binding import Oracle, Reference, Scene, Camera, build_oracle, ref_available  # noqa: F401
