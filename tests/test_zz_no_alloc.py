"""das3r_raster_backward_static - the backward for hosts that have no way to allocate device memory (a plugin or C
host without an allocator of its own, a process whose device memory belongs to another framework's pool, a loop that
must not risk the device-wide stall of cudaMalloc/cudaFree). Its gradient-accumulator scratch is a fixed array inside
the library instead of a caller buffer.

The test drives both entry points through the bare C ABI (ctypes, raw device pointers): the forward once, then the
regular backward with caller scratch as the yardstick, then the allocation-free backward twice in a row on that state
and once more on a different, smaller scene - every result must agree with the regular entry point within the
gradient tolerance of the parity tests (1e-3 relative; the atomics make summation order differ run to run). The
back-to-back calls prove that the shared array is cleared at the start of each call rather than assumed clean.

Named test_zz_* so that pytest collects it after every other file; it needs nothing from them and passes in a fresh
process as well.
"""
import ctypes

import pytest
import torch

import _harness as Hn
from cases import make_case

pytestmark = pytest.mark.gpu

STATIC_MAX_GAUSSIANS = 131072  # DAS3R_STATIC_MAX_GAUSSIANS in include/das3r_raster.h
DAS3R_ELIMIT = -4


class Params(ctypes.Structure):
    _fields_ = [("P", ctypes.c_int32), ("D", ctypes.c_int32), ("M", ctypes.c_int32), ("width", ctypes.c_int32),
                ("height", ctypes.c_int32), ("tan_fovx", ctypes.c_float), ("tan_fovy", ctypes.c_float),
                ("scale_modifier", ctypes.c_float), ("prefiltered", ctypes.c_int32), ("debug", ctypes.c_int32),
                ("background", ctypes.c_void_p), ("viewmatrix", ctypes.c_void_p), ("projmatrix", ctypes.c_void_p),
                ("campos", ctypes.c_void_p)]


class Inputs(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations",
                                               "cov3D_precomp")]


class Outputs(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in ("out_color", "radii", "out_depth", "out_alpha")]


class State(ctypes.Structure):
    _fields_ = [("geom", ctypes.c_void_p), ("geom_bytes", ctypes.c_size_t), ("binning", ctypes.c_void_p),
                ("binning_bytes", ctypes.c_size_t), ("image", ctypes.c_void_p), ("image_bytes", ctypes.c_size_t),
                ("num_rendered", ctypes.c_int32)]


GRAD_NAMES = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
              "dL_drotations")


class Grads(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in GRAD_NAMES + ("scratch",)] + [("accumulate", ctypes.c_int32)]


class Scene:
    """Forward pass of one seeded case through the C ABI; keeps every buffer alive for the backward calls."""

    def __init__(self, lib, name, dev):
        vi, dL = make_case(name)
        self.vd = {k: (v.contiguous() if isinstance(v, torch.Tensor) else v) for k, v in Hn.to_dev(vi, dev).items()}
        vd = self.vd
        self.P, self.W, self.H = vd["means3D"].shape[0], vd["image_width"], vd["image_height"]
        self.M = vd["shs"].shape[1]
        self.dL = dL.to(dev).contiguous()
        self.blobs = []
        self.dev = dev

        def alloc(nbytes, _ctx):
            t = torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=dev)
            self.blobs.append(t)
            return t.data_ptr()

        self.cb = ctypes.CFUNCTYPE(ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p)(alloc)
        self.params = Params(self.P, int(vd["sh_degree"]), self.M, self.W, self.H, vd["tanfovx"], vd["tanfovy"],
                             float(vd.get("scale_modifier", 1.0)), 0, 0, vd["bg"].data_ptr(),
                             vd["viewmatrix"].data_ptr(), vd["projmatrix"].data_ptr(), vd["campos"].data_ptr())
        self.inputs = Inputs(vd["means3D"].data_ptr(), vd["shs"].data_ptr(), None, vd["opacities"].data_ptr(),
                             vd["scales"].data_ptr(), vd["rotations"].data_ptr(), None)
        self.color = torch.empty(3, self.H, self.W, device=dev)
        self.radii = torch.empty(self.P, dtype=torch.int32, device=dev)
        out = Outputs(self.color.data_ptr(), self.radii.data_ptr(), None, None)
        self.state = State()
        self.stream = torch.cuda.current_stream().cuda_stream
        R = lib.das3r_raster_forward(ctypes.byref(self.params), ctypes.byref(self.inputs), ctypes.byref(out), self.cb,
                                     None, self.cb, None, self.cb, None, ctypes.byref(self.state),
                                     ctypes.c_void_p(self.stream))
        assert R > 0, lib.das3r_last_error()

    def backward(self, lib, static: bool):
        P, M, dev = self.P, self.M, self.dev
        shapes = ((P, 3), (P, 3), (P, 1), (P, 3), (P, 6), (P, M, 3), (P, 3), (P, 4))
        # poison the outputs: every element has to be written by the call
        g = {n: torch.full(s, float("nan"), device=dev) for n, s in zip(GRAD_NAMES, shapes)}
        if static:
            scratch_ptr = None  # the entry point under test must not need it
            fn = lib.das3r_raster_backward_static
        else:
            self._scratch = torch.empty(int(lib.das3r_raster_backward_scratch_bytes(ctypes.c_int32(P))),
                                        dtype=torch.uint8, device=dev)
            scratch_ptr = self._scratch.data_ptr()
            fn = lib.das3r_raster_backward
        grads = Grads(*[g[n].data_ptr() for n in GRAD_NAMES], scratch_ptr, 0)
        rc = fn(ctypes.byref(self.params), ctypes.byref(self.inputs), ctypes.c_void_p(self.radii.data_ptr()),
                ctypes.byref(self.state), ctypes.c_void_p(self.dL.data_ptr()), ctypes.byref(grads),
                ctypes.c_void_p(self.stream))
        assert rc == 0, lib.das3r_last_error()
        torch.cuda.synchronize()
        return g


def assert_same(got, want, what):
    for n in GRAD_NAMES:
        assert bool(torch.isfinite(got[n]).all()), f"{what}: {n} has elements the call did not write"
        err = Hn.rel_err(got[n], want[n])
        assert err <= Hn.GRAD_RTOL, f"{what}: {n} differs from das3r_raster_backward by {err}"


def test_allocation_free_backward_matches_the_regular_one():
    assert torch.cuda.is_available()
    dev = torch.device("cuda:0")
    lib = Hn.cabi()
    assert hasattr(lib, "das3r_raster_backward_static")

    big = Scene(lib, "sh3_small", dev)
    want_big = big.backward(lib, static=False)
    first = big.backward(lib, static=True)       # also performs the one-time address look-up on this device
    second = big.backward(lib, static=True)      # immediately again: the array still holds the first call's sums
    assert_same(first, want_big, "first call")
    assert_same(second, want_big, "second call in a row")

    # a different scene with fewer Gaussians next: stale accumulators beyond or inside its range must not leak in
    small = Scene(lib, "behind_and_big", dev)
    want_small = small.backward(lib, static=False)
    assert_same(small.backward(lib, static=True), want_small, "after a larger scene")
    assert_same(big.backward(lib, static=True), want_big, "back to the first scene")


def test_too_many_gaussians_are_refused_up_front():
    lib = Hn.cabi()
    p = Params()
    p.P = STATIC_MAX_GAUSSIANS + 1
    p.width, p.height = 16, 16
    g = Grads()
    # nothing else is filled in: the size check has to come before any pointer is looked at
    rc = lib.das3r_raster_backward_static(ctypes.byref(p), None, None, None, None, ctypes.byref(g), None)
    assert rc == DAS3R_ELIMIT
    msg = lib.das3r_last_error().decode()
    assert str(STATIC_MAX_GAUSSIANS) in msg and "exceed" in msg
