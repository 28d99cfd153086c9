"""CPU: libvoxelb200.so loads and exports exactly the entry points include/voxelb200.h declares,
and the Python binding covers all of them. No compute call is made (no GPU here)."""
import ctypes
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "voxelb200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return set(re.findall(r"\b(vxb_[a-z0-9_]+)\s*\(", src))


def test_library_exports_every_declared_symbol():
    from voxel_game_b200 import _lib, build

    build.build()
    out = subprocess.check_output(["nm", "-D", "--defined-only", _lib.LIB_PATH], text=True)
    exported = {line.split()[-1] for line in out.splitlines() if " T " in line and line.split()[-1].startswith("vxb_")}
    declared = declared_symbols()
    assert declared, "no declarations parsed"
    assert declared - exported == set(), "declared but not exported: %s" % sorted(declared - exported)
    assert exported - declared == set(), "exported but not declared: %s" % sorted(exported - declared)


def test_python_binding_covers_the_header():
    from voxel_game_b200 import _lib

    assert set(_lib.SIGNATURES) == declared_symbols()
    lib = _lib.load()  # resolves every symbol or raises
    for name in _lib.SIGNATURES:
        assert getattr(lib, name) is not None


def test_struct_sizes_match_the_reference_layouts():
    from voxel_game_b200 import _lib

    assert _lib.PROJECTILE_DTYPE.itemsize == 112  # common.glsl:8-25
    assert _lib.PLAYER_DTYPE.itemsize == 128      # common.glsl:27-38
    assert _lib.COLLISION_DTYPE.itemsize == 16    # common.glsl:40-45
    assert ctypes.sizeof(_lib.Settings) == 13 * 4
    assert ctypes.sizeof(_lib.RayCaps) == 5 * 4


def test_no_cpu_fallback_without_a_device():
    """On a box without a GPU the product path refuses to run instead of computing on the CPU."""
    import pytest
    import voxel_game_b200 as vg

    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("a CUDA device is present")
    with pytest.raises(vg.VxbError, match="no CUDA device|cudaSetDevice|CUDA"):
        vg.VoxelComputePipeline(vg.GameSettings())


def test_product_does_not_reference_the_oracle():
    pkg = os.path.join(ROOT, "voxel_game_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_lib" not in text and "liboracle" not in text and "oracle/" not in text, os.path.join(dirpath, f)
