"""Loader for libkgcuda.so (the C-ABI library, include/kgcuda.h). There is no CPU fallback: if the library is
missing or no CUDA device is present, calls fail loudly."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("KGCUDA_LIB", os.path.join(_HERE, "libkgcuda.so"))     # override: debug builds of the same library
KG_MAX_TERMS = 6
dp = C.POINTER(C.c_double)
ip = C.POINTER(C.c_int32)

# every symbol include/kgcuda.h declares
EXPORTS = [
    "kg_model_create", "kg_model_destroy", "kg_last_error", "kg_model_npairs", "kg_chi_points", "kg_omega_batch",
    "kg_omega_plan_create", "kg_omega_plan_run", "kg_omega_plan_sync", "kg_omega_plan_fetch", "kg_omega_plan_omega_dev",
    "kg_omega_plan_destroy", "kg_omega_plan_counters", "kg_measure_fp64_peak", "kg_model_set_stream", "kg_omega_plan_export_dev", "kg_transfer_lengths_batch",
    "kg_model_create_ex", "kg_potential_points", "kg_genpot_describe", "kg_omega_plan_quadrature_stats", "kg_set_exact_quadrature",
    "kg_omega_plan_create_masked", "kg_omega_table_create", "kg_omega_table_eval", "kg_omega_table_eval_dev", "kg_omega_table_save",
    "kg_omega_table_load", "kg_omega_table_lr", "kg_omega_table_info", "kg_omega_table_destroy", "kg_cross_sections", "kg_second_virial", "kg_model_set_temperature_dependent", "kg_build_info", "kg_host_alloc", "kg_host_free", "kg_tcache_create", "kg_tcache_destroy", "kg_tcache_lr", "kg_tcache_fetch", "kg_transport_batch", "kg_last_timings",
    "kg_chi_points_noalloc", "kg_chi_points_noalloc_max",
    "kg_eos_saftvrmie_create", "kg_eos_destroy", "kg_eos_last_error", "kg_eos_tv_batch", "kg_eos_specific_volume_batch", "kg_eos_last_kernel_ms", "kg_eos_tv_batch_dev", "kg_quantum_phase_shifts", "kg_quantum_cross_sections", "kg_quantum_omega", "kg_model_set_quantum", "kg_model_set_jkwb_limits",
]


class PotentialDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("nterms", C.c_int32), ("sigma", C.c_double), ("eps", C.c_double),
                ("C", C.c_double * KG_MAX_TERMS), ("lam", C.c_double * KG_MAX_TERMS)]


KG_MAX_POT_PARAMS = 24


class PotentialDescEx(C.Structure):
    _fields_ = [("kind", C.c_int32), ("fh_order", C.c_int32), ("sigma", C.c_double), ("eps", C.c_double), ("fh_D", C.c_double),
                ("nparams", C.c_int32), ("reserved", C.c_int32), ("params", C.c_double * KG_MAX_POT_PARAMS)]


_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(the CUDA path has no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.kg_model_create.argtypes = [C.POINTER(PotentialDesc), C.c_int, dp, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.kg_model_create_ex.argtypes = [C.POINTER(PotentialDescEx), C.c_int, dp, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.kg_potential_points.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_int, dp, dp, dp]
        L.kg_genpot_describe.argtypes = [C.POINTER(PotentialDescEx), C.c_double, dp, C.c_int]
        L.kg_omega_plan_quadrature_stats.argtypes = [C.c_void_p, dp]
        L.kg_set_exact_quadrature.argtypes = [C.c_int]
        L.kg_model_destroy.argtypes = [C.c_void_p]
        L.kg_last_error.restype = C.c_char_p
        L.kg_eos_saftvrmie_create.argtypes = [C.c_int, C.c_int, dp, dp, dp, dp, dp, ip, dp, dp, C.POINTER(C.c_void_p)]
        L.kg_eos_destroy.argtypes = [C.c_void_p]
        L.kg_eos_destroy.restype = None
        L.kg_eos_last_error.argtypes = [C.c_void_p]
        L.kg_eos_last_error.restype = C.c_char_p
        L.kg_eos_tv_batch.argtypes = [C.c_void_p, C.c_int, dp, dp, dp, dp, dp, dp, dp]
        L.kg_quantum_phase_shifts.argtypes = [C.c_void_p, C.c_int, C.c_int, ip, dp, C.c_int, dp, ip]
        L.kg_quantum_cross_sections.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_int, ip, C.c_int, dp, dp, ip]
        L.kg_quantum_omega.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_int, dp, C.c_int, ip, ip, dp, ip]
        L.kg_model_set_jkwb_limits.argtypes = [C.c_void_p, C.c_double, C.c_int]
        L.kg_model_set_quantum.argtypes = [C.c_void_p, C.c_int, dp, dp]
        L.kg_eos_tv_batch_dev.argtypes = [C.c_void_p, C.c_int, dp, dp, dp, C.c_uint, C.POINTER(C.c_void_p)]
        L.kg_eos_specific_volume_batch.argtypes = [C.c_void_p, C.c_int, dp, dp, dp, dp, ip]
        L.kg_eos_last_kernel_ms.argtypes = [C.c_void_p]
        L.kg_eos_last_kernel_ms.restype = C.c_double
        L.kg_build_info.restype = C.c_char_p
        L.kg_build_info.argtypes = []
        L.kg_host_alloc.restype = C.c_void_p
        L.kg_host_alloc.argtypes = [C.c_size_t]
        L.kg_host_free.restype = None
        L.kg_host_free.argtypes = [C.c_void_p]
        L.kg_last_error.argtypes = [C.c_void_p]
        L.kg_model_npairs.argtypes = [C.c_void_p]
        L.kg_omega_table_create.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_int, C.c_int, ip, ip, C.POINTER(C.c_void_p)]
        L.kg_omega_table_eval.argtypes = [C.c_void_p, C.c_int, dp, dp]
        L.kg_omega_table_eval_dev.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.kg_omega_table_save.argtypes = [C.c_void_p, C.c_char_p]
        L.kg_omega_table_load.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_void_p)]
        L.kg_omega_table_lr.argtypes = [C.c_void_p, ip, ip]
        L.kg_omega_table_info.argtypes = [C.c_void_p, dp]
        L.kg_omega_table_destroy.argtypes = [C.c_void_p]
        L.kg_omega_table_destroy.restype = None
        L.kg_cross_sections.argtypes = [C.c_void_p, C.c_int, C.c_int, dp, C.c_int, ip, dp, dp, ip]
        L.kg_second_virial.argtypes = [C.c_void_p, C.c_int, C.c_int, dp, dp, dp]
        L.kg_model_set_temperature_dependent.argtypes = [C.c_void_p, C.c_int]
        L.kg_chi_points.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_int, dp, dp, dp, dp, dp, ip]
        L.kg_chi_points_noalloc.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_int, dp, dp, dp, dp, dp, ip]
        L.kg_chi_points_noalloc_max.argtypes = []
        L.kg_omega_batch.argtypes = [C.c_void_p, C.c_int, dp, C.c_int, ip, ip, dp, dp, dp, ip]
        L.kg_omega_plan_create.argtypes = [C.c_void_p, C.c_int, dp, C.c_int, ip, ip, dp, C.POINTER(C.c_void_p)]
        for n in ("kg_omega_plan_run", "kg_omega_plan_sync", "kg_omega_plan_destroy"):
            getattr(L, n).argtypes = [C.c_void_p]
        L.kg_omega_plan_destroy.restype = None
        L.kg_omega_plan_fetch.argtypes = [C.c_void_p, dp, dp, ip]
        L.kg_omega_plan_omega_dev.restype = C.c_void_p
        L.kg_omega_plan_omega_dev.argtypes = [C.c_void_p]
        L.kg_omega_plan_counters.argtypes = [C.c_void_p, dp]
        L.kg_model_set_stream.argtypes = [C.c_void_p, C.c_void_p]
        L.kg_omega_plan_export_dev.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.kg_transfer_lengths_batch.argtypes = [C.c_void_p, C.c_int, dp, dp, dp, dp, ip]
        L.kg_measure_fp64_peak.argtypes = [C.c_int, dp, dp]
        L.kg_omega_plan_create_masked.argtypes = [C.c_void_p, C.c_int, dp, C.c_int, ip, ip, dp, C.POINTER(C.c_uint8), C.POINTER(C.c_void_p)]
        L.kg_tcache_create.argtypes = [C.c_void_p, C.c_void_p, C.c_int, dp, dp, C.POINTER(C.c_void_p)]
        L.kg_last_timings.argtypes = [C.c_void_p, dp]
        L.kg_tcache_destroy.argtypes = [C.c_void_p]
        L.kg_tcache_destroy.restype = None
        L.kg_tcache_lr.argtypes = [C.c_void_p, ip, ip]
        L.kg_tcache_fetch.argtypes = [C.c_void_p, dp, dp, dp, C.POINTER(C.c_uint8)]
        L.kg_transport_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, dp, dp, dp, C.c_void_p, C.c_void_p, C.c_void_p, ip]
        _LIB = L
    return _LIB


# ---- page-locked host arrays (kg_host_alloc): freed blocks are kept for reuse (cudaHostAlloc costs ~1 ms per few MB)
_PINNED_FREE = {}          # nbytes -> [ptr, ...]
_PINNED_CACHED = [0]
_PINNED_CACHE_LIMIT = 2 << 30


class _PinnedBlock:
    """Owner of one cudaHostAlloc block; arrays made by pinned_empty (and their views) keep it alive."""

    def __init__(self, nbytes):
        self.nbytes = (max(1, nbytes) + 65535) & ~65535
        free = _PINNED_FREE.get(self.nbytes)
        if free:
            self.ptr = free.pop()
            _PINNED_CACHED[0] -= self.nbytes
        else:
            self.ptr = lib().kg_host_alloc(self.nbytes)
            if not self.ptr:
                raise MemoryError(f"kg_host_alloc({self.nbytes}) failed (no CUDA device?)")
        self.buf = (C.c_char * self.nbytes).from_address(self.ptr)

    def __del__(self):
        try:
            if _PINNED_CACHED[0] + self.nbytes <= _PINNED_CACHE_LIMIT:
                _PINNED_FREE.setdefault(self.nbytes, []).append(self.ptr)
                _PINNED_CACHED[0] += self.nbytes
            else:
                lib().kg_host_free(self.ptr)
        except Exception:
            pass


def pinned_empty(shape, dtype=float):
    """numpy array in page-locked host memory: H2D / D2H of the batch calls run as asynchronous DMA from / into it."""
    import numpy as np

    class PinnedArray(np.ndarray):
        _kg_block = None

        def __array_finalize__(self, obj):
            if obj is not None:
                self._kg_block = getattr(obj, "_kg_block", None)

    global _PinnedArray
    if _PinnedArray is None:
        _PinnedArray = PinnedArray
    dt = np.dtype(dtype)
    n = int(np.prod(shape))
    blk = _PinnedBlock(n * dt.itemsize)
    arr = np.frombuffer(blk.buf, dtype=dt, count=n).reshape(shape).view(_PinnedArray)
    arr._kg_block = blk
    return arr


_PinnedArray = None


def pinned_copy(a, dtype=float):
    import numpy as np
    a = np.asarray(a, dtype)
    out = pinned_empty(a.shape, dtype)
    out[...] = a
    return out
