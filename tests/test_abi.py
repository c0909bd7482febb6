"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/rtxnerf.h declares (no compute calls without a GPU), and the Params mirror has the
reference's layout (rtx/include/params.h:14-42 as seen in the reference's own PTX)."""
import ctypes as C
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "rtxnerf.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rtxnerf_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from rtx_nerf_b200 import _lib, build
    build.build()
    lib = C.CDLL(_lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), "librtxnerf_b200.so lacks " + n
    assert set(names) == set(_lib.SIGNATURES), "ctypes table and header disagree"


def test_abi_version_and_error_strings():
    from rtx_nerf_b200 import _lib
    lib = _lib.load()
    assert lib.rtxnerf_abi_version() == 1
    assert lib.rtxnerf_error_string(0) == b"ok"
    assert b"power of two" in lib.rtxnerf_error_string(-2)


def test_params_layout_matches_reference():
    """Offsets read from the reference's PTX (params+88 focal/aspect, +116 min.z, +128 max.z,
    +132 width, +136 height) — see DESIGN.md."""
    from rtx_nerf_b200._lib import Params
    assert Params.look_at.offset == 8
    assert Params.num_hits.offset == 72
    assert Params.viewing_direction.offset == 80
    assert Params.focal_length.offset == 88
    assert Params.aspect_ratio.offset == 92
    assert Params.min_point.offset + 8 == 116
    assert Params.max_point.offset + 8 == 128
    assert Params.width.offset == 132
    assert Params.height.offset == 136
    assert C.sizeof(Params) == 144


def test_argument_errors_need_no_gpu():
    """Argument validation happens before any CUDA call."""
    from rtx_nerf_b200 import _lib
    lib = _lib.load()
    h = C.c_void_p()
    assert lib.rtxnerf_grid_create(None, 8, C.byref(h)) == -1
    bits = (C.c_uint32 * 32)()
    assert lib.rtxnerf_grid_create(bits, 10, C.byref(h)) == -2
    assert lib.rtxnerf_launch_sampler(None, None, None, None, None, 4, 8, None, None, 0, 32, None) == -1
    assert lib.rtxnerf_launch_volrender(None, None, None, None, None, 4, 32, None, None) == -1
    # a sub-buffer that breaks the vector alignment of its layout is refused on the host (never a device fault):
    # addresses below are never dereferenced, the check precedes every CUDA call
    ok, odd = 0x10000, 0x10004
    assert lib.rtxnerf_launch_volrender(ok, odd, ok, ok, ok, 4, 32, ok, None) == -1          # [N][4] fp32 at +4
    assert lib.rtxnerf_launch_sampler(ok, ok, ok, ok, odd, 4, 8, ok, ok, 0, 32, None) == -1   # sample rows at +4
    assert lib.rtxnerf_launch_volrender_backward(ok, ok, ok, ok, ok, ok, 4, 32, odd, None) == -1


def test_product_never_imports_oracle():
    """The product package must not reference oracle/ (a CPU fallback would void parity)."""
    pkg = os.path.join(ROOT, "rtx_nerf_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "liboracle" not in text and "from oracle" not in text and "import oracle" not in text, f
