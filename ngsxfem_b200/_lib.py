"""ctypes binding of libngsx_b200.so (the C ABI of include/ngsx_abi.h).

The CUDA library is the product: there is no CPU fallback.  ``load()`` raises if the shared object is missing
(build it with ``python -c 'import __graft_entry__ as g; g.build()'``) and every call that reaches a kernel
raises if no CUDA device is present."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libngsx_b200.so")

c_int, c_i32, c_i64, c_double, c_void_p = ctypes.c_int, ctypes.c_int32, ctypes.c_int64, ctypes.c_double, ctypes.c_void_p

NEG, POS, IF = 0, 1, 2
CDOM_NO, CDOM_NEG, CDOM_POS, CDOM_UNCUT, CDOM_IF, CDOM_HASNEG, CDOM_HASPOS, CDOM_ANY = range(8)
LSET_AFFINE, LSET_SPHERE, LSET_RING, LSET_CYLZ = range(4)
POLY_MAXTERMS = 8
INT64_MAX = 2 ** 63 - 1


class NgsxError(RuntimeError):
    """Raised for every non-zero status of the C ABI (message = ngsx_last_error())."""

    def __init__(self, status, msg):
        super().__init__(f"ngsx status {status}: {msg}")
        self.status = status


class Mesh(ctypes.Structure):
    _fields_ = [("dim", c_i32), ("_pad", c_i32), ("ne", c_i64), ("nv", c_i64), ("vert", c_void_p), ("el2v", c_void_p)]


class FormH1(ctypes.Structure):
    _fields_ = [("c_mass", c_double), ("c_stiff", c_double), ("c_nitsche", c_double), ("c_dndn", c_double),
                ("hpow_mass", c_i32), ("hpow_stiff", c_i32), ("hpow_nitsche", c_i32), ("hpow_dndn", c_i32)]


class XTerm(ctypes.Structure):
    _fields_ = [("coef", c_double), ("hpow", c_i32), ("kappa", c_i32), ("pair", c_i32), ("blk_u", c_i32), ("blk_v", c_i32), ("_pad", c_i32)]


class Deformation(ctypes.Structure):
    _fields_ = [("order", c_i32), ("nd", c_i32), ("ndof", c_i64), ("el2dof", c_void_p), ("coef", c_void_p)]


class Poly(ctypes.Structure):
    _fields_ = [("nterms", c_i32), ("_pad", c_i32), ("coef", c_double * POLY_MAXTERMS), ("ex", (c_i32 * 3) * POLY_MAXTERMS)]


# name -> (restype, argtypes); every symbol include/ngsx_abi.h declares
SIGNATURES = {
    "ngsx_abi_version": (c_int, []),
    "ngsx_last_error": (ctypes.c_char_p, []),
    "ngsx_launch_count": (c_i64, []),
    "ngsx_interpolate_p1_clamp": (c_int, [c_i64, c_void_p, c_double, c_void_p, c_void_p]),
    "ngsx_interpolate_p1_cf": (c_int, [ctypes.POINTER(Mesh), c_int, ctypes.POINTER(c_double), c_double, c_void_p, c_void_p]),
    "ngsx_classify": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ngsx_node_marks": (c_int, [c_i64, c_i32, c_void_p, c_void_p, c_void_p, c_i64, c_int, c_void_p, c_void_p, c_void_p]),
    "ngsx_rule_count": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "ngsx_exclusive_scan": (c_int, [c_i64, c_void_p, c_void_p, c_void_p]),
    "ngsx_rule_fill": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ngsx_elmat_h1": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_int, ctypes.POINTER(FormH1), c_i64, c_void_p,
                              c_void_p, c_void_p, c_void_p]),
    "ngsx_elvec_h1": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_int, ctypes.POINTER(Poly), c_i64, c_void_p,
                              c_void_p, c_void_p, c_void_p]),
    "ngsx_elmat_h1_cf": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_int, ctypes.POINTER(FormH1), c_int, c_void_p, c_void_p, c_i64,
                                 c_void_p, c_void_p, c_void_p, c_void_p]),
    "ngsx_elvec_h1_cf": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_int, ctypes.POINTER(Poly), c_int, c_void_p, c_void_p, c_i64,
                                 c_void_p, c_void_p, c_void_p, c_void_p]),
    "ngsx_integrate_cf": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, ctypes.POINTER(Poly), c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ngsx_gather_vec": (c_int, [c_i64, c_void_p, c_void_p, c_void_p, ctypes.c_double, c_int, c_void_p, c_void_p]),
    "ngsx_csr_graph_create": (c_int, [c_i64, ctypes.POINTER(c_void_p)]),
    "ngsx_csr_graph_set_lower": (c_int, [c_void_p, c_int]),
    "ngsx_csr_graph_add_table": (c_int, [c_void_p, c_i64, c_i32, c_void_p, c_void_p]),
    "ngsx_csr_graph_finalize": (c_int, [c_void_p, ctypes.POINTER(c_i64), c_void_p]),
    "ngsx_csr_graph_export": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "ngsx_csr_graph_destroy": (c_int, [c_void_p]),
    "ngsx_slot_map": (c_int, [c_i64, c_i32, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ngsx_color_groups": (c_int, [c_i64, c_i32, c_void_p, c_i64, c_i32, c_void_p, c_void_p, c_i32, ctypes.POINTER(c_i32)]),
    "ngsx_assemble_csr_h1": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_int, ctypes.POINTER(FormH1), c_i64, c_void_p,
                                     c_void_p, c_i32, c_void_p, c_i32, c_void_p, c_void_p]),
    "ngsx_assemble_vec_h1": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_int, ctypes.POINTER(Poly), c_i64, c_void_p,
                                     c_void_p, c_i32, c_void_p, c_i32, c_void_p, c_void_p]),
    "ngsx_rowplan_create": (c_int, [c_i64, c_i32, c_void_p, c_i64, c_void_p, c_void_p, c_i64, c_i64, ctypes.POINTER(c_void_p), c_void_p]),
    "ngsx_rowplan_create_lower": (c_int, [c_i64, c_i32, c_void_p, c_i64, c_void_p, c_void_p, c_i64, c_i64, ctypes.POINTER(c_void_p), c_void_p]),
    "ngsx_rowplan_create_owned": (c_int, [c_i64, c_i32, c_void_p, c_i64, c_void_p, c_void_p, c_void_p, c_i32, ctypes.POINTER(c_void_p), c_void_p]),
    "ngsx_rowplan_info": (c_int, [c_void_p, ctypes.POINTER(c_i64)]),
    "ngsx_rowplan_destroy": (c_int, [c_void_p]),
    "ngsx_rowplan_cut_usage": (c_int, [c_void_p, ctypes.POINTER(c_i64), ctypes.POINTER(c_i64), c_void_p]),
    "ngsx_rowplan_reserve_cut": (c_int, [c_void_p, c_i64]),
    "ngsx_rowplan_rowslots": (c_int, [c_void_p, ctypes.POINTER(c_void_p), ctypes.POINTER(c_i64), c_void_p]),
    "ngsx_assemble_csr_rows_packed": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "ngsx_set_packed_output": (None, [c_void_p]),
    "ngsx_classify_integrate": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_i32, ctypes.POINTER(c_i32),
                                        ctypes.POINTER(c_i32), ctypes.POINTER(Poly), c_void_p, c_void_p]),
    "ngsx_xfem_elmat_h1_deformed": (c_int, [ctypes.POINTER(Mesh), ctypes.POINTER(Deformation), c_void_p, c_int, c_int, c_int, c_i32, c_void_p,
                                            c_i64, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "ngsx_facetpatch_cmat_h1_deformed": (c_int, [ctypes.POINTER(Mesh), ctypes.POINTER(Deformation), c_int, ctypes.c_double, c_int, c_i64, c_void_p, c_void_p, c_i32, c_void_p, c_void_p]),
    "ngsx_classify_bnd": (c_int, [c_int, c_i64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ngsx_facets_neighbor_types_periodic": (c_int, [c_i64, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "ngsx_q_classify": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ngsx_q_rule_count": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "ngsx_q_rule_fill": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ngsx_q_integrate": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, ctypes.POINTER(Poly), c_void_p, c_void_p, c_void_p]),
    "ngsx_assemble_csr_rows_h1": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_int, ctypes.POINTER(FormH1), c_i64, c_void_p,
                                          c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "ngsx_set_gather_wait_event": (None, [c_void_p]),
    "ngsx_integrate": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, ctypes.POINTER(Poly), c_void_p, c_void_p]),
    "ngsx_integrate_static": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, ctypes.POINTER(Poly), c_void_p, c_void_p]),
    "ngsx_elmat_h1_deformed": (c_int, [ctypes.POINTER(Mesh), ctypes.POINTER(Deformation), c_void_p, c_int, c_int, c_int, ctypes.POINTER(FormH1),
                                       c_i64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ngsx_elvec_h1_deformed": (c_int, [ctypes.POINTER(Mesh), ctypes.POINTER(Deformation), c_void_p, c_int, c_int, c_int, ctypes.POINTER(Poly),
                                       c_i64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ngsx_integrate_deformed": (c_int, [ctypes.POINTER(Mesh), ctypes.POINTER(Deformation), c_void_p, c_int, c_int, ctypes.POINTER(Poly),
                                        c_void_p, c_void_p]),
    "ngsx_facets_neighbor_types": (c_int, [c_i64, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "ngsx_elements_with_neighbor_facets": (c_int, [c_i64, c_void_p, c_void_p, c_i64, c_void_p, c_void_p]),
    "ngsx_dofs_of_elements": (c_int, [c_i64, c_i32, c_void_p, c_void_p, c_i64, c_void_p, c_void_p]),
    "ngsx_facetpatch_elmat_h1": (c_int, [ctypes.POINTER(Mesh), c_int, c_double, c_int, c_i64, c_void_p, c_void_p, c_void_p]),
    "ngsx_xfem_elmat_h1": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_int, c_i32, ctypes.POINTER(XTerm), c_i64, c_void_p, c_void_p,
                                   c_void_p, c_int, c_void_p, c_void_p]),
    "ngsx_st_rule_count": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "ngsx_st_rule_fill": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ngsx_peer_mailbox_bytes": (c_int, [c_int]),
    "ngsx_peer_alloc": (c_int, [c_int, ctypes.POINTER(c_void_p), c_void_p]),
    "ngsx_peer_open": (c_int, [c_void_p, ctypes.POINTER(c_void_p)]),
    "ngsx_peer_close": (c_int, [c_void_p]),
    "ngsx_peer_free": (c_int, [c_void_p]),
    "ngsx_peer_exchange": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ngsx_peer_collect": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "ngsx_st_elmat_h1": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_double, c_double, c_double, c_i64,
                                 c_void_p, c_void_p, c_void_p]),
    "ngsx_st_classify": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "ngsx_st_integrate": (c_int, [ctypes.POINTER(Mesh), c_void_p, c_int, c_int, c_int, c_int, ctypes.POINTER(Poly), ctypes.POINTER(c_i32),
                                  c_void_p, c_void_p]),
    "ngsx_assemble_csr_rows_mats": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "ngsx_facetpatch_table_h1": (c_int, [c_i64, c_void_p, c_i32, c_void_p, c_i32, c_void_p, c_void_p, c_void_p]),
    "ngsx_facetpatch_cmat_h1": (c_int, [ctypes.POINTER(Mesh), c_int, c_double, c_int, c_i64, c_void_p, c_void_p, c_i32, c_void_p, c_void_p]),
}

_LIB = None


def load():
    """dlopen the CUDA library; fails loudly when it has not been built."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: the CUDA extension is the product and there is no CPU fallback. "
                          "Build it with  python -c 'import __graft_entry__ as g; g.build()'")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    lib.ngsx_debug_set_fast_path.argtypes = [c_int]
    lib.ngsx_debug_set_fast_path.restype = None
    lib.ngsx_debug_set_order3_fma.argtypes = [c_int]
    lib.ngsx_debug_set_order3_fma.restype = None
    lib.ngsx_debug_set_elvec_warp.argtypes = [c_int]
    lib.ngsx_debug_set_elvec_warp.restype = None
    lib.ngsx_debug_set_deformed_warp.argtypes = [c_int]
    lib.ngsx_debug_set_deformed_warp.restype = None
    lib.ngsx_debug_set_timing.argtypes = [c_int]
    lib.ngsx_debug_set_timing.restype = None
    lib.ngsx_debug_get_timing.argtypes = [ctypes.POINTER(c_double)]
    lib.ngsx_debug_get_timing.restype = c_int
    _LIB = lib
    return lib


def check(status):
    if status != 0:
        raise NgsxError(status, load().ngsx_last_error().decode("utf-8", "replace"))


def make_poly(terms):
    """terms: iterable of (coef, ex, ey, ez); None -> constant 1"""
    if terms is None:
        terms = [(1.0, 0, 0, 0)]
    terms = list(terms)
    if len(terms) > POLY_MAXTERMS:
        raise NgsxError(-2, f"polynomial coefficient with more than {POLY_MAXTERMS} terms is outside the CUDA path")
    p = Poly()
    p.nterms = len(terms)
    for m, (c, a, b, d) in enumerate(terms):
        p.coef[m] = float(c)
        for k, v in enumerate((a, b, d)):
            if int(v) != v or v < 0:
                raise NgsxError(-2, "only non-negative integer exponents are supported")
            p.ex[m][k] = int(v)
    return p
