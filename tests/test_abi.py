"""CPU-side checks of the drop-in boundary: libb200splat.so loads, exports every symbol include/b200splat.h
declares, the ctypes struct layouts match, and the argument validation returns the documented codes without
touching a device (no compute calls)."""
import ctypes as C
import re
from pathlib import Path

import pytest

from neuralgaussiansplatting_b200 import _lib

ROOT = Path(__file__).resolve().parent.parent
HEADER = (ROOT / "include" / "b200splat.h").read_text()


@pytest.fixture(scope="module")
def lib():
    return _lib.load()


def declared_symbols():
    # every function prototype in the header: `<type> b200s_xxx(`
    names = set(re.findall(r"\b(b200s_[a-z0-9_]+)\s*\(", HEADER))
    return sorted(names)


def test_header_declares_what_binding_binds():
    bound = {n for n, _, _ in _lib.SYMBOLS}
    assert set(declared_symbols()) == bound


def test_every_declared_symbol_is_exported(lib):
    for name in declared_symbols():
        assert hasattr(lib, name), name


def test_abi_version_and_build_info(lib):
    assert lib.b200s_abi_version() == int(re.search(r"#define B200S_ABI_VERSION (\d+)", HEADER).group(1))
    assert b"sm_100a" in lib.b200s_build_info()


def test_struct_layouts_match(lib):
    for which, st in enumerate((_lib.Header, _lib.ForwardArgs, _lib.BackwardArgs, _lib.IdxmapArgs,
                                _lib.PeerReduceArgs)):
        assert lib.b200s_sizeof(which) == C.sizeof(st)
    assert C.sizeof(_lib.Header) == 64
    assert lib.b200s_sizeof(99) == 0


def test_header_cites_reference_for_every_entry_point():
    for tag in ("rasterizer_impl.cu:198-336", "rasterizer_impl.cu:340-434", "rasterizer_impl.cu:141-153",
                "rasterizer_impl.cu:851-1008", "rasterize_points.cu:35-115", "rasterize_points.cu:117-196"):
        assert tag in HEADER, tag


def test_workspace_sizes_monotone_and_aligned(lib):
    prev = 0
    for P in (0, 1, 255, 256, 257, 4096, 50_000, 3_000_000):
        n = lib.b200s_geometry_workspace_size(P)
        assert n % 256 == 0 and n >= prev
        prev = n
    # 48-B record + 9 B/Gaussian + 20 B/Gaussian for the depth sort (depth bits, 2 x (key, id)) + look-back
    assert lib.b200s_geometry_workspace_size(3_000_000) < 3_000_000 * 80
    assert lib.b200s_image_workspace_size(1920, 1080) >= 1920 * 1080 * 8
    assert lib.b200s_image_workspace_size(17, 5) % 256 == 0


def test_binning_capacity_is_inverse_of_size(lib):
    for cap in (1, 2, 4095, 4096, 4097, 1_000_000, 17_225_097):
        nbytes = lib.b200s_binning_workspace_size(cap, 800, 800)
        got = lib.b200s_binning_capacity_for_bytes(nbytes)
        assert got >= cap
        assert lib.b200s_binning_workspace_size(got, 800, 800) == nbytes
    assert lib.b200s_binning_capacity_for_bytes(0) == -1


def test_backward_workspace_size(lib):
    a = lib.b200s_backward_workspace_size(1000, 0)
    b = lib.b200s_backward_workspace_size(1000, 50_000)
    assert a >= 1000 * 9 * 4 and b > a


def test_argument_errors_without_a_device(lib):
    # NULL argument block / inconsistent arguments are rejected before any CUDA call
    assert lib.b200s_forward(None, None) == -1
    assert lib.b200s_backward(None, None) == -1
    assert lib.b200s_idxmap_forward(None, None) == -1
    assert lib.b200s_idxmap_forward_static(None, None) == -1
    a = _lib.ForwardArgs()
    a.P, a.W, a.H = 10, 0, 16
    assert lib.b200s_forward_preprocess(C.byref(a), None) == -1          # W <= 0
    a.W = 16
    assert lib.b200s_forward_preprocess(C.byref(a), None) == -1          # NULL means3D
    assert lib.b200s_mark_visible(-1, None, None, None, None, None) == -1
    assert lib.b200s_mark_visible(5, None, None, None, None, None) == -1
    assert lib.b200s_mark_visible(0, None, None, None, None, None) == 0
    assert lib.b200s_idxmap_backward(-1, 4, 4, None, None, None, None) == -1
    assert lib.b200s_read_header(None, None, None) == -1


def test_peer_allreduce_argument_checks(lib):
    assert lib.b200s_peer_allreduce_flag_bytes(32, 8) == 2 * 32 * 8 * 4
    assert lib.b200s_peer_allreduce_flag_bytes(0, 8) == 0
    assert lib.b200s_peer_allreduce_sum_f32(None, None) == -1
    a = _lib.PeerReduceArgs()
    a.rank, a.world, a.ctas, a.epoch, a.n_floats = 0, 2, 16, 1, 1024
    assert lib.b200s_peer_allreduce_sum_f32(C.byref(a), None) == -1      # NULL peer buffers
    a.peer_bufs[0] = a.peer_bufs[1] = 4096
    a.peer_flags[0] = a.peer_flags[1] = 8192
    a.n_floats = 1022
    assert lib.b200s_peer_allreduce_sum_f32(C.byref(a), None) == -1      # not a multiple of 4 floats
    a.n_floats, a.epoch = 1024, 0
    assert lib.b200s_peer_allreduce_sum_f32(C.byref(a), None) == -1      # epochs start at 1
    a.epoch, a.rank = 1, 2
    assert lib.b200s_peer_allreduce_sum_f32(C.byref(a), None) == -1      # rank outside the world
    a.rank, a.use_multimem = 0, 1
    assert lib.b200s_peer_allreduce_sum_f32(C.byref(a), None) == -1      # multimem without a multicast mapping
    a.use_multimem, a.world = 0, 1
    assert lib.b200s_peer_allreduce_sum_f32(C.byref(a), None) == 0       # world of one: nothing to do, no launch


def test_check_turns_status_into_runtime_error():
    _lib.check(0, "ok")
    with pytest.raises(RuntimeError, match="B200S_E_WORKSPACE"):
        _lib.check(-2, "x")
    with pytest.raises(RuntimeError, match="CUDA error 700"):
        _lib.check(700, "x")


def test_header_is_plain_c99(tmp_path):
    """The drop-in boundary must be bindable from C (cgo / JNI / ctypes / a torch extension): no C++ in the header."""
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    src = tmp_path / "h.c"
    src.write_text('#include "b200splat.h"\n'
                   'int main(void) { B200sForwardArgs f; B200sBackwardArgs b; B200sIdxmapArgs i; B200sPeerReduceArgs p;\n'
                   '  B200sAdamArgs a; B200sLossArgs l; (void)f; (void)b; (void)i; (void)p; (void)a; (void)l;\n'
                   '  return (int)sizeof(B200sHeader) - 64; }\n')
    inc = Path(__file__).resolve().parent.parent / "include"
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", str(inc), "-c", str(src), "-o",
                        str(tmp_path / "h.o")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
