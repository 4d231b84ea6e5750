"""The C-ABI library loads without a GPU and exports every symbol include/das3r_raster.h declares; the pure-host entry
points behave (no compute calls here)."""
import ctypes
import os
import re

import _harness as Hn

ROOT = Hn.ROOT


def declared_symbols():
    names = set()
    for h in sorted(os.listdir(os.path.join(ROOT, "include"))):
        if not h.endswith(".h"):
            continue
        text = open(os.path.join(ROOT, "include", h)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names |= set(re.findall(r"\b(das3r_[a-z0-9_]+)\s*\(", text))
    return sorted(names)


def test_header_symbols_are_exported():
    lib = Hn.cabi()
    names = declared_symbols()
    assert "das3r_raster_forward" in names and "das3r_raster_backward" in names and "das3r_knn_dist2" in names
    assert "das3r_mark_visible" in names
    assert "das3r_frontend_forward" in names and "das3r_frontend_backward" in names
    assert "das3r_loss_forward" in names and "das3r_loss_backward" in names
    assert "das3r_adam_step" in names
    assert "das3r_raster_backward_blend" in names and "das3r_raster_backward_finish" in names
    for n in names:
        assert hasattr(lib, n), f"{n} is declared in include/das3r_raster.h but not exported by libdas3r_raster.so"


def test_every_entry_point_cites_the_reference():
    text = open(os.path.join(ROOT, "include", "das3r_raster.h")).read()
    for n in ("das3r_raster_forward", "das3r_raster_backward", "das3r_mark_visible", "das3r_knn_dist2"):
        i = text.index(f"int {n}(")
        comment = text[text.rfind("/*", 0, i):i]
        assert re.search(r"\.(cu|h|py):\d+", comment), f"{n} lacks a reference file:line citation"


def test_version_and_error_string():
    lib = Hn.cabi()
    assert lib.das3r_abi_version() == 3
    assert lib.das3r_last_error() in (b"",) or isinstance(lib.das3r_last_error(), bytes)


def test_layout_query_is_consistent():
    P, W, H, R = 1000, 70, 50, 4321
    L = Hn.layout(P, W, H, R)
    T = ((W + 15) // 16) * ((H + 15) // 16)
    assert L.g_recB == L.g_recA + 16 and L.g_recC == L.g_recA + 32
    assert L.g_tiles_touched >= L.g_recA + 48 * P
    assert L.g_clamped >= L.g_tiles_touched + 4 * P
    assert L.geom_bytes >= L.g_clamped + P
    assert L.i_n_contrib >= L.i_final_T + 4 * W * H
    assert L.i_ranges >= L.i_n_contrib + 4 * W * H
    assert L.i_tile_counts >= L.i_ranges + 8 * T
    assert L.b_point_list >= L.b_unsorted + 8 * R
    assert L.b_rect_off >= L.b_point_list + 4 * R
    assert L.b_rect_gid >= L.b_rect_list + 64 * R
    assert L.binning_bytes >= L.b_keys_sorted + 8 * R
    for f, _ in L._fields_:
        assert getattr(L, f) % 16 == 0, f


def test_bad_arguments_are_rejected_without_touching_the_gpu():
    lib = Hn.cabi()
    rc = lib.das3r_raster_forward(None, None, None, None, None, None, None, None, None, None, None)
    assert rc == -1
    assert b"forward" in lib.das3r_last_error()
    rc = lib.das3r_raster_backward(None, None, None, None, None, None, None)
    assert rc == -1
    assert lib.das3r_mark_visible(ctypes.c_int32(-1), None, None, None, None, None) == -1
    assert lib.das3r_knn_dist2(ctypes.c_int32(-5), None, None, None, ctypes.c_size_t(0), None) == -1
    # zero-sized problems are no-ops
    assert lib.das3r_mark_visible(ctypes.c_int32(0), None, None, None, None, None) == 0
    assert lib.das3r_knn_dist2(ctypes.c_int32(0), None, None, None, ctypes.c_size_t(0), None) == 0
    assert lib.das3r_raster_backward_scratch_bytes(ctypes.c_int32(1000)) >= 48 * 1000


def test_size_queries_and_argument_errors_need_no_gpu():
    """Pure-host behaviour of the newer entry points: scratch-size queries, and argument validation that refuses a call
    before any CUDA work is attempted (negative DAS3R_E* code + message)."""
    import ctypes

    lib = Hn.cabi()
    for name in ("das3r_pointmap_scratch_bytes", "das3r_frontend_backward_scratch_bytes", "das3r_loss_state_bytes"):
        getattr(lib, name).restype = ctypes.c_size_t
    n = 50 * 288 * 512
    assert lib.das3r_pointmap_scratch_bytes(ctypes.c_int64(n)) >= 4 * ((n + 1023) // 1024)
    assert lib.das3r_pointmap_scratch_bytes(ctypes.c_int64(0)) > 0
    assert lib.das3r_frontend_backward_scratch_bytes(ctypes.c_int32(1_000_000), ctypes.c_int32(8)) >= 3907 * 8 * 16 * 4
    assert lib.das3r_raster_backward_scratch_bytes(ctypes.c_int32(1000)) >= 48 * 1000
    assert lib.das3r_knn_scratch_bytes(ctypes.c_int32(1000)) > 0
    assert lib.das3r_loss_state_bytes(ctypes.c_int32(512), ctypes.c_int32(288)) >= 15 * 512 * 288 * 4
    EINVAL = -1
    # in-switch all-reduce: null multicast pointer, bad rank, misaligned length
    lib.das3r_nvls_allreduce_sum.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_int32,
                                             ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p]
    assert lib.das3r_nvls_allreduce_sum(None, 0, 16, 0, 2, 0, None) == EINVAL
    assert b"nvls_allreduce" in lib.das3r_last_error()
    assert lib.das3r_nvls_allreduce_sum(ctypes.c_void_p(4096), 0, 16, 2, 2, 0, None) == EINVAL
    assert lib.das3r_nvls_allreduce_sum(ctypes.c_void_p(4096), 0, 18, 0, 2, 0, None) == EINVAL
    assert lib.das3r_nvls_allreduce_sum(ctypes.c_void_p(4096), 0, 0, 0, 2, 0, None) == 0  # nothing to do
    # multi-view entry points: no views
    assert lib.das3r_raster_backward_finish_ex(0, None, None, 0, 0, -1, 0, None) == EINVAL
    assert lib.das3r_raster_backward_blend_batch(0, None, None) == EINVAL
    assert lib.das3r_raster_forward_render_batch(0, None, None, None) == EINVAL
    assert b"forward_render_batch" in lib.das3r_last_error()
    # pointmap initialisation: null description
    lib.das3r_pointmap_select.restype = ctypes.c_int64
    assert lib.das3r_pointmap_select(None, None, None, None, ctypes.c_size_t(0), None) == EINVAL
    assert lib.das3r_pointmap_init(None, None, ctypes.c_int64(0), None, None, ctypes.c_size_t(0), None, ctypes.c_size_t(0),
                                   None) == EINVAL
    assert lib.das3r_frontend_forward_ex(None, 1, None, None, None, None) == EINVAL
