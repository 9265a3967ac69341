"""TEST INFRASTRUCTURE -- loader for oracle/_ref/libngsxref.so (the reference's own
sources compiled in place by oracle/build_ref.sh).  Not part of the product path."""
import ctypes
import glob
import os
import site

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def ref_path():
    return os.path.join(_HERE, "_ref", "libngsxref.so")


def available():
    return os.path.exists(ref_path())


def load():
    """dlopen the compiled reference; the bundled openblas needs its libgfortran/libquadmath
    preloaded (they sit next to it without an rpath)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    for sp in site.getsitepackages():
        d = os.path.join(sp, "opencv_python_headless.libs")
        for pat in ("libquadmath-*.so*", "libgfortran-*.so*", "libopenblasp-*.so*"):
            for f in sorted(glob.glob(os.path.join(d, pat))):
                try:
                    ctypes.CDLL(f, mode=ctypes.RTLD_GLOBAL)
                except OSError:
                    pass
    L = ctypes.CDLL(ref_path(), mode=ctypes.RTLD_GLOBAL)
    dp = ctypes.POINTER(ctypes.c_double)
    ip = ctypes.POINTER(ctypes.c_int)
    lp = ctypes.POINTER(ctypes.c_long)
    cp = ctypes.POINTER(ctypes.c_byte)
    c_int, c_long, c_double = ctypes.c_int, ctypes.c_long, ctypes.c_double
    L.ref_select_rule.argtypes = [c_int, c_int, dp, dp, c_int]
    L.ref_cut_rule_untrafo.argtypes = [c_int, dp, c_int, c_int, dp, dp, c_int]
    L.ref_cut_rule.argtypes = [c_int, dp, dp, c_int, c_int, dp, dp, c_int]
    L.ref_cutinfo_element.argtypes = [c_int, dp, dp, dp]
    L.ref_st_cut_rule.argtypes = [c_int, dp, dp, c_int, c_int, c_int, c_int, dp, dp, dp, c_int]
    if hasattr(L, "ref_st_elmat_h1"):
        L.ref_st_elmat_h1.argtypes = [c_int, c_int, c_int, dp, dp, c_int, c_int, c_int, c_int, c_double, c_double, c_double, dp]
    L.ref_h1_ndof.argtypes = [c_int, c_int]
    L.ref_h1_shape.argtypes = [c_int, c_int, ip, dp, dp, dp]
    L.ref_elmat_h1.argtypes = [c_int, c_int, ip, dp, dp, c_int, c_int, c_double, c_double, dp]
    L.ref_elvec_h1.argtypes = [c_int, c_int, ip, dp, dp, c_int, c_int, c_int, dp, dp]
    L.ref_cut_rule_iso.argtypes = [c_int, c_int, dp, dp, c_int, c_int, dp, dp, dp, dp, c_int]
    L.ref_elmat_h1_iso.argtypes = [c_int, c_int, c_int, dp, dp, c_int, c_int, c_double, c_double, dp]
    L.ref_elvec_h1_iso.argtypes = [c_int, c_int, c_int, dp, dp, c_int, c_int, c_int, dp, dp]
    L.ref_cutinfo_mesh.argtypes = [c_int, c_long, ip, dp, dp, cp, dp, c_int]
    L.ref_cutinfo_mesh.restype = c_long
    L.ref_integrate_mesh.argtypes = [c_int, c_long, ip, dp, dp, c_int, c_int, c_int, dp, c_int, lp]
    L.ref_integrate_mesh.restype = c_double
    L.ref_elmat_mesh.argtypes = [c_int, c_long, lp, ip, dp, dp, c_int, c_int, c_int, c_double, c_double, dp, ip, c_int]
    L.ref_elmat_mesh.restype = c_long
    L.ref_assemble_csr.argtypes = [c_int, ip, dp, dp, c_int, ip, c_int, c_int, c_double, c_double,
                                   c_int, lp, ip, lp, ip, dp, c_int]
    L.ref_assemble_csr.restype = c_long
    sp = ctypes.POINTER(ctypes.c_byte)
    L.ref_elmat_h1_form.argtypes = [c_int, c_int, c_int, ip, dp, dp, c_int, c_int, dp, dp]
    L.ref_facetpatch_elmat_h1.argtypes = [c_int, c_int, c_int, ip, dp, ip, dp, c_double, c_int, dp, dp]
    L.ref_nxfem_elmat.argtypes = [c_int, c_int, c_int, ip, dp, dp, sp, c_double, c_double, c_double, c_double, dp]
    L.ref_nxfem_elvec.argtypes = [c_int, c_int, c_int, ip, dp, dp, sp, c_int, c_int, dp, dp]
    _LIB = L
    return L
