"""ctypes binding of csrc/libmcb200.so (include/mcb200.h). No fallback: a missing library or a
failing call raises."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(CSRC, "libmcb200.so")
_LIB = None
LAUNCH_CLASSES = ("gemm", "udt", "rdivp", "sweep_spatial", "fused", "measure", "misc", "ising", "advance",
                  "greens")


class McbError(RuntimeError):
    pass


class DqmcParams(C.Structure):
    _fields_ = [("n_sites", C.c_int32), ("n_slices", C.c_int32), ("safe_mult", C.c_int32),
                ("n_chains", C.c_int32), ("delta_tau", C.c_double), ("U", C.c_double),
                ("seed", C.c_uint64), ("chain_offset", C.c_uint32), ("device_id", C.c_int32),
                ("check_propagation", C.c_int32), ("general_path", C.c_int32)]


class DqmcState(C.Structure):
    _fields_ = [("current_slice", C.c_int64), ("direction", C.c_int64), ("n_updates", C.c_int64),
                ("n_measurements", C.c_int64)]


class DqmcChainStats(C.Structure):
    _fields_ = [("prop_local", C.c_int64), ("acc_local", C.c_int64),
                ("max_propagation_error", C.c_double), ("n_negative", C.c_int64),
                ("n_nonfinite", C.c_int64)]


_f64 = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_i8 = np.ctypeslib.ndpointer(dtype=np.int8, flags="C_CONTIGUOUS")
_i32 = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_i64 = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_u64 = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")

# name -> (restype, argtypes); every symbol include/mcb200.h declares
SIGNATURES = {
    "mcb_last_error": (C.c_char_p, []),
    "mcb_version": (C.c_char_p, []),
    "mcb_dqmc_create": (C.c_int, [C.POINTER(DqmcParams), C.POINTER(C.c_void_p)]),
    "mcb_dqmc_destroy": (C.c_int, [C.c_void_p]),
    "mcb_dqmc_set_hopping": (C.c_int, [C.c_void_p, _f64, _f64, _f64, _f64]),
    "mcb_dqmc_set_conf": (C.c_int, [C.c_void_p, _i8]),
    "mcb_dqmc_get_conf": (C.c_int, [C.c_void_p, _i8]),
    "mcb_dqmc_build_stack": (C.c_int, [C.c_void_p]),
    "mcb_dqmc_update": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32]),
    "mcb_dqmc_sweep": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32]),
    "mcb_dqmc_get_greens": (C.c_int, [C.c_void_p, _f64, C.c_int32]),
    "mcb_dqmc_calculate_greens_at": (C.c_int, [C.c_void_p, C.c_int32, _f64]),
    "mcb_dqmc_greens_unequal": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, _f64, C.c_int32]),
    "mcb_comm_unique_id": (C.c_int, [C.c_void_p]),
    "mcb_comm_create": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    "mcb_comm_destroy": (C.c_int, [C.c_void_p]),
    "mcb_dqmc_moments_len": (C.c_int32, [C.c_void_p]),
    "mcb_dqmc_allreduce_observables": (C.c_int, [C.c_void_p, C.c_void_p, _f64]),
    "mcb_dqmc_allreduce_susceptibilities": (C.c_int, [C.c_void_p, C.c_void_p, _f64]),
    "mcb_dqmc_cgi_begin": (C.c_int, [C.c_void_p, C.c_int32]),
    "mcb_dqmc_cgi_next": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.c_void_p, C.c_void_p, C.c_void_p]),
    "mcb_dqmc_cgi_device_results": (C.c_int, [C.c_void_p] + [C.POINTER(C.c_void_p)] * 3),
    "mcb_dqmc_resume": (C.c_int, [C.c_void_p, C.c_int64]),
    "mcb_dqmc_conf_chunks": (C.c_int, [C.c_void_p]),
    "mcb_dqmc_set_conf_bits": (C.c_int, [C.c_void_p, _u64]),
    "mcb_dqmc_get_conf_bits": (C.c_int, [C.c_void_p, _u64]),
    "mcb_dqmc_replay": (C.c_int, [C.c_void_p, _u64, C.c_int64]),
    "mcb_dqmc_get_state": (C.c_int, [C.c_void_p, C.POINTER(DqmcState)]),
    "mcb_dqmc_get_stats": (C.c_int, [C.c_void_p, C.POINTER(DqmcChainStats)]),
    "mcb_dqmc_set_measurement_tables": (C.c_int, [C.c_void_p, _i32, C.c_int32, C.c_int32, _f64]),
    "mcb_dqmc_measure": (C.c_int, [C.c_void_p]),
    "mcb_dqmc_reset_observables": (C.c_int, [C.c_void_p]),
    "mcb_dqmc_measure_susceptibilities": (C.c_int, [C.c_void_p, C.c_int32]),
    "mcb_dqmc_schedule_susceptibilities": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    "mcb_dqmc_susceptibility_len": (C.c_int32, [C.c_void_p]),
    "mcb_dqmc_get_susceptibilities": (C.c_int, [C.c_void_p, _f64, C.POINTER(C.c_int64)]),
    "mcb_dqmc_rep_measure_susceptibilities": (C.c_int, [C.c_void_p, C.c_int32]),
    "mcb_dqmc_rep_schedule_susceptibilities": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    "mcb_dqmc_rep_susceptibility_len": (C.c_int32, [C.c_void_p]),
    "mcb_dqmc_rep_get_susceptibilities": (C.c_int, [C.c_void_p, _f64, C.POINTER(C.c_int64)]),
    "mcb_dqmc_observable_len": (C.c_int32, [C.c_void_p]),
    "mcb_dqmc_get_observables": (C.c_int, [C.c_void_p, _f64, C.POINTER(C.c_int64)]),
    "mcb_dqmc_get_greens_sum": (C.c_int, [C.c_void_p, _f64]),
    "mcb_dqmc_profile": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    "mcb_dqmc_get_profile": (C.c_int, [C.c_void_p, _i64, _f64, _i64, _f64]),
    "mcb_dqmc_last_elapsed_ms": (C.c_double, [C.c_void_p]),
    "mcb_dqmc_rep_create": (C.c_int, [C.POINTER(DqmcParams), C.POINTER(C.c_void_p)]),
    "mcb_dqmc_rep_destroy": (C.c_int, [C.c_void_p]),
    "mcb_dqmc_rep_set_hopping": (C.c_int, [C.c_void_p, _f64, _f64, _f64, _f64]),
    "mcb_dqmc_rep_set_conf": (C.c_int, [C.c_void_p, _i8]),
    "mcb_dqmc_rep_get_conf": (C.c_int, [C.c_void_p, _i8]),
    "mcb_dqmc_rep_build_stack": (C.c_int, [C.c_void_p]),
    "mcb_dqmc_rep_update": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32]),
    "mcb_dqmc_rep_sweep": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32]),
    "mcb_dqmc_rep_get_greens": (C.c_int, [C.c_void_p, _f64, C.c_int32]),
    "mcb_dqmc_rep_calculate_greens_at": (C.c_int, [C.c_void_p, C.c_int32, _f64]),
    "mcb_dqmc_rep_get_state": (C.c_int, [C.c_void_p, C.POINTER(DqmcState)]),
    "mcb_dqmc_rep_get_stats": (C.c_int, [C.c_void_p, C.POINTER(DqmcChainStats)]),
    "mcb_dqmc_rep_greens_unequal": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, _f64, C.c_int32]),
    "mcb_dqmc_rep_cgi_begin": (C.c_int, [C.c_void_p, C.c_int32]),
    "mcb_dqmc_rep_cgi_next": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.c_void_p, C.c_void_p, C.c_void_p]),
    "mcb_dqmc_rep_conf_chunks": (C.c_int, [C.c_void_p]),
    "mcb_dqmc_rep_get_conf_bits": (C.c_int, [C.c_void_p, _u64]),
    "mcb_dqmc_rep_set_conf_bits": (C.c_int, [C.c_void_p, _u64]),
    "mcb_dqmc_rep_resume": (C.c_int, [C.c_void_p, C.c_int64]),
    "mcb_dqmc_rep_last_elapsed_ms": (C.c_double, [C.c_void_p]),
    "mcb_dqmc_rep_set_measurement_tables": (C.c_int, [C.c_void_p, _i32, C.c_int32, C.c_int32, _f64]),
    "mcb_dqmc_rep_measure": (C.c_int, [C.c_void_p]),
    "mcb_dqmc_rep_reset_observables": (C.c_int, [C.c_void_p]),
    "mcb_dqmc_rep_observable_len": (C.c_int32, [C.c_void_p]),
    "mcb_dqmc_rep_get_observables": (C.c_int, [C.c_void_p, _f64, C.POINTER(C.c_int64)]),
    "mcb_dqmc_rep_get_greens_sum": (C.c_int, [C.c_void_p, _f64]),
    "mcb_dqmc_rep_replay": (C.c_int, [C.c_void_p, _u64, C.c_int64]),
    "mcb_ising_last_elapsed_ms": (C.c_double, [C.c_void_p]),
    "mcb_ising_launch_count": (C.c_int64, [C.c_void_p]),
    "mcb_linalg_force_general": (C.c_int, [C.c_int32]),
    "mcb_linalg_vmul": (C.c_int, [C.c_int32, C.c_int32, _f64, _f64, _f64, C.c_int32, C.c_int32]),
    "mcb_linalg_udt": (C.c_int, [C.c_int32, C.c_int32, _f64, _f64, _f64, _f64, _i32, C.c_int32,
                                 C.c_int32]),
    "mcb_linalg_udt_noalloc": (C.c_int, [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "mcb_linalg_rdivp": (C.c_int, [C.c_int32, C.c_int32, _f64, _f64, _i32, C.c_int32]),
    "mcb_linalg_calculate_greens": (C.c_int, [C.c_int32, C.c_int32] + [_f64] * 7 + [C.c_int32]),
    "mcb_ising_create": (C.c_int, [C.c_int32, C.c_double, C.c_int32, C.c_uint64, C.c_uint32,
                                   C.c_int32, C.POINTER(C.c_void_p)]),
    "mcb_ising_destroy": (C.c_int, [C.c_void_p]),
    "mcb_ising_set_spins": (C.c_int, [C.c_void_p, _i8]),
    "mcb_ising_get_spins": (C.c_int, [C.c_void_p, _i8]),
    "mcb_ising_spin_chunks": (C.c_int, [C.c_void_p]),
    "mcb_ising_set_spins_bits": (C.c_int, [C.c_void_p, _u64]),
    "mcb_ising_get_spins_bits": (C.c_int, [C.c_void_p, _u64]),
    "mcb_ising_sweep": (C.c_int, [C.c_void_p, C.c_int64]),
    "mcb_ising_measure": (C.c_int, [C.c_void_p, _i64, _i64]),
    "mcb_ising_sweep_count": (C.c_int64, [C.c_void_p]),
}


def build(force=False):
    """Compile libmcb200.so for sm_100a with the committed Makefile (nvcc cross-compiles on CPU)."""
    if force:
        subprocess.run(["make", "-C", CSRC, "clean"], check=True, capture_output=True)
    r = subprocess.run(["make", "-C", CSRC, "-j8"], capture_output=True, text=True)
    if r.returncode != 0:
        raise McbError("building libmcb200.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    return LIB_PATH


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise McbError(f"{LIB_PATH} is missing: run __graft_entry__.build() (there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


def check(rc):
    if rc != 0:
        raise McbError(f"libmcb200 error {rc}: {lib().mcb_last_error().decode()}")
