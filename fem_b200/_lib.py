"""ctypes loader for fem_b200/libfemx.so (the C ABI of include/femx.h).

There is no fallback: if the shared library is missing this raises, and if it loads but
no CUDA device exists femx_create fails with FEMX_ERR_CUDA.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FEMX_LIB") or os.path.join(_HERE, "libfemx.so")   # FEMX_LIB: a development build of the same library
_lib = None

c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)
c_i64_p = C.POINTER(C.c_int64)
c_i32_p = C.POINTER(C.c_int32)
c_u16_p = C.POINTER(C.c_uint16)


class FemxOpts(C.Structure):
    _fields_ = [("quirks", C.c_int), ("assembly_variant", C.c_int), ("reserved", C.c_int * 6)]


# every symbol include/femx.h declares: (restype, argtypes)
V = C.c_void_p
SIGNATURES = {
    "femx_create": (C.c_int, [C.POINTER(V), C.c_int, C.POINTER(FemxOpts)]),
    "femx_destroy": (C.c_int, [V]),
    "femx_last_error": (C.c_char_p, [V]),
    "femx_host_alloc": (C.c_int, [C.POINTER(V), C.c_int64]),
    "femx_host_free": (C.c_int, [V]),
    "femx_device_alloc": (C.c_int, [V, C.POINTER(V), C.c_int64]),
    "femx_device_free": (C.c_int, [V, V]),
    "femx_memcpy": (C.c_int, [V, V, V, C.c_int64, C.c_int, C.c_int]),
    "femx_synchronize": (C.c_int, [V]),
    "femx_set_mesh": (C.c_int, [V, C.c_int, C.c_int, c_int_p, c_int_p, c_int_p, c_int_p, c_int_p, C.c_int,
                                c_double_p, C.c_int, c_int_p, C.c_int]),
    "femx_set_materials": (C.c_int, [V, C.c_int, c_double_p, C.c_int, c_double_p]),
    "femx_set_bcs": (C.c_int, [V, C.c_int, c_int_p, c_double_p]),
    "femx_build_plan": (C.c_int, [V]),
    "femx_pattern_size": (C.c_int, [V, c_i64_p, c_i64_p]),
    "femx_get_pattern": (C.c_int, [V, V, V, C.c_int]),
    "femx_get_coo": (C.c_int, [V, V, V, C.c_int]),
    "femx_mpiaij_sizes": (C.c_int, [V, C.c_int64, C.c_int64, V, V, V]),
    "femx_get_pattern_mpiaij": (C.c_int, [V, C.c_int64, C.c_int64, V, V, V, V, V]),
    "femx_set_stream": (C.c_int, [V, V]),
    "femx_set_async": (C.c_int, [V, C.c_int]),
    "femx_get_slot_map": (C.c_int, [V, c_i64_p, c_u16_p]),
    "femx_element_ke": (C.c_int, [V, C.c_int, C.c_int, c_double_p, c_double_p, c_double_p]),
    "femx_element_stress": (C.c_int, [V, C.c_int, C.c_int, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p]),
    "femx_element_cohesive": (C.c_int, [V, C.c_int, c_double_p, c_int_p, c_double_p, C.c_int, c_double_p, C.c_double,
                                        c_double_p, c_double_p]),
    "femx_assemble_K": (C.c_int, [V, V, C.c_int]),
    "femx_assemble_K_fint": (C.c_int, [V, V, V, V, C.c_int]),
    "femx_apply_kbc": (C.c_int, [V, V, C.c_int]),
    "femx_jacobian": (C.c_int, [V, V, V, C.c_double, V, C.c_int]),
    "femx_jacobian_update": (C.c_int, [V, V, V, C.c_double, V, C.c_int]),
    "femx_residual": (C.c_int, [V, V, V, V, V, C.c_double, C.c_double, V, C.c_int]),
    "femx_residual_forces": (C.c_int, [V, V, V, C.c_double, C.c_double, V, C.c_int]),
    "femx_stress": (C.c_int, [V, V, V, C.c_int, C.c_int]),
    "femx_stress_noalloc": (C.c_int, [V, V, V, C.c_int]),
    "femx_stress_noalloc_max_doubles": (C.c_int64, []),
    "femx_set_ownership": (C.c_int, [V, C.c_int, C.c_int, C.c_int64, C.c_int64]),
    "femx_gather_sum": (C.c_int, [V, V, C.c_int64, V, V, V, C.c_int64, V]),
    "femx_stress_partial": (C.c_int, [V, V, V, V, C.c_int]),
    "femx_set_loads": (C.c_int, [V, C.c_int, c_int_p, c_double_p, C.c_int, c_int_p, c_double_p]),
    "femx_form_rhs": (C.c_int, [V, C.c_double, C.c_double, C.c_double, V, C.c_int]),
    "femx_peer_arena_create": (C.c_int, [V, C.c_int64, C.c_int64, V]),
    "femx_peer_connect": (C.c_int, [V, C.c_int, C.c_int, V, V, V, V, V, V, V, V]),
    "femx_peer_connect_local": (C.c_int, [V, C.c_int, C.c_int, V, V, V, V, V, V, V, V]),
    "femx_peer_set_merge": (C.c_int, [V, C.c_int64, V, V, V, C.c_int64, V, V, V]),
    "femx_peer_merged": (C.c_int, [V, V, V]),
    "femx_peer_complete": (C.c_int, [V]),
    "femx_peer_set_deferred": (C.c_int, [V, C.c_int]),
    "femx_peer_last_error": (C.c_int, [V]),
    "femx_peer_set_scatter": (C.c_int, [V, V, V, V, V]),
    "femx_scatter_u": (C.c_int, [V, V, V, C.c_int]),
    "femx_peer_recv": (C.c_int, [V, V, V]),
    "femx_peer_disconnect": (C.c_int, [V]),
    "femx_launch_count": (C.c_int64, [V]),
    "femx_plan_seconds": (C.c_double, [V]),
    "femx_last_assembly_kernel_ms": (C.c_double, [V]),
    "femx_last_kernel_ms": (C.c_double, [V]),
    "femx_version": (C.c_int, []),
}


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `make -C fem_b200/csrc` "
                               "(python -c 'import __graft_entry__ as g; g.build()'); femx has no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the ABI lost a symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib
