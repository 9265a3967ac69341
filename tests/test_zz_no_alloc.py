"""ngsx_integrate_static -- the allocation-free variant of ngsx_integrate (scratch in a static __device__ array of the library;
no counterpart in the reference, include/ngsx_abi.h) -- against ngsx_integrate through the C ABI.  Collected last on purpose.
Passes alone in a fresh process and after the other tests: it needs nothing but the library."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

RTOL = 1e-12
NEG, POS, IF = 0, 1, 2
STATIC_MAX_ELEMENTS = 4194304


def _problem(dim, n, radius):
    from ngsxfem_b200 import meshes, ngs, xfem
    m = meshes.MakeStructured2DMesh(n) if dim == 2 else meshes.MakeStructured3DMesh(n)
    mesh = ngs.Mesh(m)
    gf = ngs.GridFunction(ngs.H1(mesh, order=1))
    xfem.InterpolateToP1(ngs.SphereLevelset((0.5,) * dim, radius), gf)
    return mesh, gf


def _both(lib, L, mesh, gf, dt, order, terms, st):
    poly = lib.make_poly(terms)
    out = torch.full((2,), float("nan"), dtype=torch.float64, device="cuda")
    lib.check(L.ngsx_integrate(ctypes.byref(mesh.c_struct()), gf.vec.data_ptr(), dt, order, ctypes.byref(poly), out.data_ptr(), st))
    lib.check(L.ngsx_integrate_static(ctypes.byref(mesh.c_struct()), gf.vec.data_ptr(), dt, order, ctypes.byref(poly),
                                      out.data_ptr() + 8, st))
    torch.cuda.synchronize()
    return out.cpu().numpy()


@pytest.mark.parametrize("dim", [2, 3])
def test_integrate_static_equals_integrate_two_calls_in_a_row(dim):
    from ngsxfem_b200 import _lib as lib
    assert torch.cuda.is_available()
    L = lib.load()
    st = torch.cuda.current_stream().cuda_stream
    big = _problem(dim, 96 if dim == 2 else 24, 0.33)
    small = _problem(dim, 17 if dim == 2 else 7, 0.41)
    # two calls in a row on the shared scratch: a large mesh first, then a small one with another domain, order and integrand --
    # whatever the first call left behind must not reach the second result
    calls = [(big, NEG, 2, [(1.0, 0, 0, 0)]),
             (small, IF, 4, [(0.5, 2, 0, 0), (1.5, 0, 1, 0), (2.0, 0, 0, 0)]),
             (big, POS, 3, [(1.0, 1, 1, 0)]),
             (small, NEG, 1, None)]
    for (mesh, gf), dt, order, terms in calls:
        ref, got = _both(lib, L, mesh, gf, dt, order, terms, st)
        assert np.isfinite(ref) and ref != 0.0
        assert abs(got - ref) <= RTOL * abs(ref)
        assert got == ref                                                    # same kernels, same reduction tree
    # known answer on top: area / volume of the NEG part converges to the disc / ball
    (mesh, gf) = big
    _, vol = _both(lib, L, mesh, gf, NEG, 1, None, st)
    exact = np.pi * 0.33 ** 2 if dim == 2 else 4.0 / 3.0 * np.pi * 0.33 ** 3
    assert abs(vol - exact) < 0.02 * exact


def test_integrate_static_allocates_nothing_and_rejects_large_meshes():
    from ngsxfem_b200 import _lib as lib
    L = lib.load()
    st = torch.cuda.current_stream().cuda_stream
    mesh, gf = _problem(3, 16, 0.3)
    poly = lib.make_poly(None)
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    cs = mesh.c_struct()
    lib.check(L.ngsx_integrate_static(ctypes.byref(cs), gf.vec.data_ptr(), NEG, 2, ctypes.byref(poly), out.data_ptr(), st))   # set-up
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        lib.check(L.ngsx_integrate_static(ctypes.byref(cs), gf.vec.data_ptr(), NEG, 2, ctypes.byref(poly), out.data_ptr(), st))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
    # beyond the static scratch: a clear error before anything is touched
    huge = lib.Mesh()
    ctypes.memmove(ctypes.byref(huge), ctypes.byref(cs), ctypes.sizeof(lib.Mesh))
    huge.ne = STATIC_MAX_ELEMENTS + 1
    rc = L.ngsx_integrate_static(ctypes.byref(huge), gf.vec.data_ptr(), NEG, 2, ctypes.byref(poly), out.data_ptr(), st)
    assert rc == -5
    msg = L.ngsx_last_error().decode()
    assert "NGSX_STATIC_MAX_ELEMENTS" in msg and str(STATIC_MAX_ELEMENTS) in msg
