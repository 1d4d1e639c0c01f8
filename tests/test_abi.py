"""The C-ABI library loads and exports every symbol include/spt_cuda.h declares (no GPU needed)."""
import ctypes
import os
import re

import pytest

import simplepathtracer_b200 as spt
from simplepathtracer_b200 import binding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    spt.build_library()
    return spt.load_library()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "spt_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(spt_cuda_\w+)\s*\(", text)))


def test_header_declares_what_binding_uses():
    assert set(binding.EXPORTS) == set(declared_symbols())


def test_every_declared_symbol_is_exported(lib):
    for name in declared_symbols():
        assert hasattr(lib, name), name


def test_struct_sizes_match_header():
    # natural C layout of the structs in include/spt_cuda.h on LP64
    assert ctypes.sizeof(binding.Segment) == 16
    assert ctypes.sizeof(binding._CCamera) == 96
    assert ctypes.sizeof(binding._CParams) == 56  # revision 2: band_count, band_index; revision 3: split_chunk, reserved
    assert ctypes.sizeof(binding._CScene) == 48
    assert ctypes.sizeof(binding.Counters) == 48
    assert ctypes.sizeof(binding.Timing) == 24


def test_struct_layouts_match_the_c_compiler(tmp_path):
    """sizeof / offsetof as gcc sees include/spt_cuda.h against the ctypes mirror, field by field."""
    import shutil
    import subprocess

    if not shutil.which("gcc"):
        pytest.skip("no gcc")
    structs = {"SptSegment": binding.Segment, "SptCamera": binding._CCamera, "SptParams": binding._CParams, "SptScene": binding._CScene,
               "SptCounters": binding.Counters, "SptTiming": binding.Timing, "SptDeviceInfo": binding.DeviceInfo}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "spt_cuda.h"', "int main(void) {",
             'printf("abi %u\\n", (unsigned)SPT_ABI_VERSION);']
    for cname, cls in structs.items():
        lines.append(f'printf("{cname} %zu\\n", sizeof({cname}));')
        for fname, _ in cls._fields_:
            lines.append(f'printf("{cname}.{fname} %zu\\n", offsetof({cname}, {fname}));')
    lines.append("return 0; }")
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = dict(line.split() for line in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    assert int(got["abi"]) == binding.ABI_VERSION
    for cname, cls in structs.items():
        assert int(got[cname]) == ctypes.sizeof(cls), cname
        for fname, _ in cls._fields_:
            assert int(got[f"{cname}.{fname}"]) == getattr(cls, fname).offset, f"{cname}.{fname}"


def test_version_string(lib):
    assert b"sm_100a" in lib.spt_cuda_version()
    lib.spt_cuda_abi_version.restype = ctypes.c_uint32
    assert lib.spt_cuda_abi_version() == binding.ABI_VERSION


def test_no_cpu_fallback_without_device(lib):
    """Without a CUDA device the product path must fail loudly, never fall back."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(spt.SptError):
        spt.PathTracer(0)


def test_product_does_not_touch_the_oracle():
    """Nothing under simplepathtracer_b200/ or include/ may reference oracle/."""
    bad = []
    for base in ("simplepathtracer_b200", "include"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                    text = open(os.path.join(dirpath, f), errors="ignore").read()
                    if re.search(r"(from|import)\s+oracle|oracle/|spt_oracle|libspt_ref", text):
                        bad.append(os.path.join(dirpath, f))
    assert not bad, bad
