"""ctypes bindings of the two in-tree shared libraries.

libfeelgpu.so  -- the CUDA engine behind the C ABI of include/feelgpu.h (the product).
libfeelhost.so -- the C++ host-side mirror (mesh generator, DofTable, graph, reference tables).

There is no fallback: if libfeelgpu.so is missing the import of the engine fails loudly, and every engine call
fails with FEELGPU_ERR_CUDA when no CUDA device is present.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBDIR = os.path.join(_HERE, "lib")

c_i32p = C.POINTER(C.c_int32)
c_i64p = C.POINTER(C.c_int64)
c_dblp = C.POINTER(C.c_double)


class FeelGpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("feelgpu error %d: %s" % (code, msg))
        self.code = code


class Coeff(C.Structure):
    _fields_ = [("kind", C.c_int), ("value", C.c_double * 3), ("field_space", C.c_void_p),
                ("field_values", C.c_void_p), ("expr_len", C.c_int), ("expr_ops", c_i32p),
                ("expr_consts", c_dblp), ("lambda_", C.c_double), ("mu", C.c_double)]


class Tables(C.Structure):
    _fields_ = [("nq", C.c_int), ("w", c_dblp), ("xi", c_dblp), ("n", C.c_int), ("phi", c_dblp),
                ("dphi", c_dblp), ("nf", C.c_int), ("phif", c_dblp)]


class Form2Desc(C.Structure):
    _fields_ = [("kind", C.c_int), ("coeff", Coeff), ("tables", Tables), ("algorithm", C.c_int)]


class Form1Desc(C.Structure):
    _fields_ = [("kind", C.c_int), ("coeff", Coeff), ("tables", Tables)]


# symbols declared in include/feelgpu.h (tests/test_cabi_symbols.py checks the header against this list)
GPU_SYMBOLS = {
    "feelgpu_context_create": [C.c_int, C.POINTER(C.c_void_p)],
    "feelgpu_context_destroy": [C.c_void_p],
    "feelgpu_sync": [C.c_void_p],
    "feelgpu_mesh_create": [C.c_void_p, C.c_int64, c_dblp, C.c_int64, c_i32p, C.POINTER(C.c_void_p)],
    "feelgpu_mesh_update_points": [C.c_void_p, c_dblp, C.c_void_p],
    "feelgpu_mesh_destroy": [C.c_void_p],
    "feelgpu_space_create": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, c_i32p, C.c_int64, C.c_int64, c_i64p,
                             C.POINTER(C.c_void_p)],
    "feelgpu_space_destroy": [C.c_void_p],
    "feelgpu_pattern_create": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, c_i64p, c_i64p, C.c_int,
                               C.POINTER(C.c_void_p)],
    "feelgpu_pattern_nnz": [C.c_void_p, c_i64p],
    "feelgpu_pattern_plans_ready": [C.c_void_p, C.c_int, C.POINTER(C.c_int), c_dblp],
    "feelgpu_pattern_destroy": [C.c_void_p],
    "feelgpu_pattern_plan_stats": [C.c_void_p, c_i64p],
    "feelgpu_pattern_eplan_stats": [C.c_void_p, c_i64p],
    "feelgpu_pattern_rplan_stats": [C.c_void_p, c_i64p],
    "feelgpu_eplan_host_build": [C.c_int, C.c_int, C.c_int64, c_i32p, C.c_int64, c_dblp, c_i32p, C.c_int64, c_i64p,
                                 C.c_void_p, C.POINTER(C.c_void_p), c_i64p],
    "feelgpu_eplan_host_copy": [C.c_void_p, C.c_void_p, c_i32p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, c_dblp],
    "feelgpu_eplan_host_free": [C.c_void_p],
    "feelgpu_eplan_host_coefficients": [C.c_int, C.POINTER(Tables), C.c_int, c_dblp, c_dblp],
    "feelgpu_rplan_host_build": [C.c_int, C.c_int, C.c_int64, c_i32p, C.c_int64, c_dblp, c_i32p, C.c_int64, c_i64p, C.c_int,
                                 C.c_int, C.c_int, C.POINTER(C.c_void_p), c_i64p],
    "feelgpu_rplan_host_copy": [C.c_void_p, C.c_void_p, c_i32p, C.c_void_p, C.c_void_p],
    "feelgpu_rplan_host_free": [C.c_void_p],
    "feelgpu_form2_assemble": [C.c_void_p, C.POINTER(Form2Desc), C.c_int, C.c_void_p, C.c_void_p],
    "feelgpu_form1_assemble": [C.c_void_p, C.POINTER(Form1Desc), C.c_int, C.c_void_p, C.c_void_p],
    "feelgpu_form2_assemble_host": [C.c_void_p, C.POINTER(Form2Desc), C.c_int, c_dblp],
    "feelgpu_form2_assemble_noalloc": [C.c_void_p, C.POINTER(Form2Desc), C.c_int, C.c_void_p, C.c_void_p],
    "feelgpu_form1_assemble_host": [C.c_void_p, C.POINTER(Form1Desc), C.c_int, c_dblp],
    "feelgpu_form2_zero_rows": [C.c_void_p, C.c_int64, c_i32p, c_dblp, C.c_int, C.c_double, C.c_void_p, C.c_void_p,
                                C.c_void_p],
    "feelgpu_malloc": [C.c_void_p, C.c_int64, C.POINTER(C.c_void_p)],
    "feelgpu_free": [C.c_void_p, C.c_void_p],
    "feelgpu_memcpy_h2d": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64],
    "feelgpu_memset_zero": [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p],
    "feelgpu_memcpy_d2h": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64],
    "feelgpu_host_alloc_pinned": [C.c_void_p, C.c_int64, C.POINTER(C.c_void_p)],
    "feelgpu_host_free_pinned": [C.c_void_p, C.c_void_p],
    "feelgpu_scatter_add": [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p],
    "feelgpu_graph_create": [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)],
    "feelgpu_graph_get": [C.c_void_p, c_i64p, c_i64p],
    "feelgpu_graph_destroy": [C.c_void_p],
    "feelgpu_pattern_create_from_graph": [C.c_void_p, C.c_int, C.POINTER(C.c_void_p)],
    "feelgpu_matvec": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p],
    "feelgpu_faces_create": [C.c_void_p, C.c_int64, c_i32p, c_i32p, C.POINTER(C.c_void_p)],
    "feelgpu_faces_destroy": [C.c_void_p],
    "feelgpu_form1_assemble_faces": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p],
    "feelgpu_form2_assemble_faces": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p],
    "feelgpu_comm_unique_id": [C.c_char_p],
    "feelgpu_comm_init": [C.c_void_p, C.c_char_p, C.c_int, C.c_int],
    "feelgpu_comm_destroy": [C.c_void_p],
    "feelgpu_exchange_create": [C.c_void_p, c_i32p, C.c_int64, C.c_int, C.POINTER(C.c_void_p)],
    "feelgpu_exchange_destroy": [C.c_void_p],
    "feelgpu_exchange_info": [C.c_void_p, c_i64p],
    "feelgpu_form2_close": [C.c_void_p, C.c_void_p, C.c_void_p],
    "feelgpu_form1_close": [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p],
    "feelgpu_launch_count": [C.c_void_p, c_i64p],
    "feelgpu_last_timing": [C.c_void_p, c_dblp],
    "feelgpu_set_timing": [C.c_void_p, C.c_int],
}

_gpu = None
_host = None


def gpu_lib_path():
    return os.path.join(_LIBDIR, "libfeelgpu.so")


def host_lib_path():
    return os.path.join(_LIBDIR, "libfeelhost.so")


def gpu():
    """libfeelgpu.so; raises if the CUDA extension has not been built (no fallback exists)."""
    global _gpu
    if _gpu is None:
        path = gpu_lib_path()
        if not os.path.exists(path):
            raise ImportError("feelpp_b200: %s is missing -- run `make` (or __graft_entry__.build()); "
                              "there is no CPU fallback" % path)
        lib = C.CDLL(path)
        for name, args in GPU_SYMBOLS.items():
            fn = getattr(lib, name)
            fn.argtypes = args
            fn.restype = C.c_int
        lib.feelgpu_last_error.argtypes = [C.c_void_p]
        lib.feelgpu_last_error.restype = C.c_char_p
        lib.feelgpu_eplan_host_error.argtypes = [C.c_void_p]
        lib.feelgpu_eplan_host_error.restype = C.c_char_p
        lib.feelgpu_rplan_host_error.argtypes = [C.c_void_p]
        lib.feelgpu_rplan_host_error.restype = C.c_char_p
        _gpu = lib
    return _gpu


def host():
    global _host
    if _host is None:
        path = host_lib_path()
        if not os.path.exists(path):
            raise ImportError("feelpp_b200: %s is missing -- run `make`" % path)
        lib = C.CDLL(path)
        lib.feelhost_last_error.restype = C.c_char_p
        lib.feelhost_partition_last_error.restype = C.c_char_p
        lib.feelhost_unit_cube_sizes.argtypes = [C.c_int, c_i64p, c_i64p]
        lib.feelhost_unit_cube.argtypes = [C.c_int, C.c_double, C.c_int, c_dblp, c_i32p]
        lib.feelhost_boundary_faces.argtypes = [C.c_int64, C.c_int64, c_i32p, c_i32p, c_i32p, c_i64p]
        lib.feelhost_gmsh_open.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_void_p), c_i64p, c_i64p, c_i64p, c_i64p,
                                           c_dblp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        lib.feelhost_gmsh_get.argtypes = [C.c_void_p, c_dblp, c_i64p, c_i32p, c_i32p, c_i32p, c_i32p, C.c_char_p,
                                          c_i32p]
        lib.feelhost_gmsh_free.argtypes = [C.c_void_p]
        lib.feelhost_doftable.argtypes = [C.c_int64, c_dblp, C.c_int64, c_i32p, C.c_int, C.c_int, c_i32p, c_i64p]
        lib.feelhost_graph_build.argtypes = [C.c_int64, C.c_int, C.c_int, c_i32p, C.c_int64, C.c_int,
                                             C.POINTER(C.c_void_p), c_i64p]
        lib.feelhost_graph_get.argtypes = [C.c_void_p, c_i64p, c_i64p]
        lib.feelhost_graph_free.argtypes = [C.c_void_p]
        lib.feelhost_quadrature.argtypes = [C.c_int, C.POINTER(C.c_int), c_dblp, c_dblp]
        lib.feelhost_face_quadrature.argtypes = [C.c_int, C.POINTER(C.c_int), c_dblp, c_dblp]
        lib.feelhost_nlocaldof.argtypes = [C.c_int]
        lib.feelhost_lagrange_points.argtypes = [C.c_int, c_dblp]
        lib.feelhost_tabulate.argtypes = [C.c_int, C.c_int, c_dblp, c_dblp, c_dblp]
        lib.feelhost_parse_expr.argtypes = [C.c_char_p, C.c_int, c_i32p, c_dblp, C.POINTER(C.c_int),
                                            C.POINTER(C.c_int)]
        lib.feelhost_csr_select_remap_sort.argtypes = [C.c_int64, C.c_int64, c_i64p, c_i64p, c_i64p, c_i64p, c_i64p,
                                                       c_i64p, c_i64p, c_i64p]
        lib.feelhost_dual_graph.argtypes = [C.c_int64, C.c_int64, c_i32p, c_i32p, c_i32p]
        if hasattr(lib, "feelhost_partition_metis"):      # absent when the host mirror was built without METIS
            lib.feelhost_partition_metis.argtypes = [C.c_int64, C.c_int64, c_i32p, C.c_int, C.c_int, c_i32p]
        _host = lib
    return _host


def check(rc, ctx=None):
    if rc != 0:
        msg = gpu().feelgpu_last_error(ctx)
        raise FeelGpuError(rc, msg.decode() if msg else "")


def hcheck(rc):
    if rc != 0:
        msg = host().feelhost_last_error().decode() or host().feelhost_partition_last_error().decode()
        raise RuntimeError("feelhost: " + msg)


def ptr(a, typ):
    return a.ctypes.data_as(typ)
