"""ctypes binding of the C ABI in include/smmc.h (libsmmc.so, built in-tree by csrc/Makefile).

The product has no CPU fallback: importing this module without the built library raises, and every compute
entry point raises `SmmcError` when no sm_100a device is visible.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SMMC_LIB") or os.path.join(_HERE, "lib", "libsmmc.so")  # SMMC_LIB: development override

SMMC_ABI_VERSION = 3
SMMC_BLOCK = 256
SMMC_STATS_HEADER = 6
SMMC_MAX_HIST_BINS = 4096
SMMC_NOALLOC_MAX_N = 1 << 30
KERNEL_SIMULATION = 1
KERNEL_SELECT_CANDIDATES = 2

MODE_BOOTSTRAP_COMPAT = 0
MODE_BOOTSTRAP = 1
MODE_GAUSSIAN = 2
MODE_SOBOL_BOOTSTRAP = 3
MODE_SOBOL_GAUSSIAN = 4
MODES = {
    "bootstrap_compat": MODE_BOOTSTRAP_COMPAT,
    "bootstrap": MODE_BOOTSTRAP,
    "gaussian": MODE_GAUSSIAN,
    "sobol_bootstrap": MODE_SOBOL_BOOTSTRAP,
    "sobol_gaussian": MODE_SOBOL_GAUSSIAN,
}

FLAG_DEVICE_POINTERS = 1
FLAG_NO_SYNC = 2

STATUS_NAMES = {0: "SMMC_OK", 1: "SMMC_ERR_INVALID_ARG", 2: "SMMC_ERR_NO_DEVICE", 3: "SMMC_ERR_CUDA",
                4: "SMMC_ERR_UNSUPPORTED"}


class SmmcConfig(C.Structure):
    _fields_ = [
        ("abi_version", C.c_uint32),
        ("mode", C.c_int32),
        ("n_simulations", C.c_int64),
        ("n_periods", C.c_int32),
        ("n_returns", C.c_int32),
        ("returns", C.c_void_p),
        ("seed", C.c_uint64),
        ("stream_offset", C.c_uint64),
        ("initial_capital", C.c_float),
        ("gauss_mean", C.c_float),
        ("gauss_std", C.c_float),
        ("target", C.c_float),
        ("hist_lo", C.c_float),
        ("hist_hi", C.c_float),
        ("hist_bins", C.c_int32),
        ("device", C.c_int32),
        ("flags", C.c_uint32),
        ("stream", C.c_void_p),
        ("withdraw_fixed", C.c_float),
        ("withdraw_pct", C.c_float),
        ("withdraw_pct_schedule", C.c_void_p),
        ("philox_rounds", C.c_int32),
        ("reserved0", C.c_int32),
    ]


class SmmcResult(C.Structure):
    _fields_ = [
        ("finals", C.c_void_p),
        ("block_means", C.c_void_p),
        ("block_vars", C.c_void_p),
        ("hist", C.c_void_p),
        ("stats_dev", C.c_void_p),
        ("count", C.c_uint64),
        ("below_count", C.c_uint64),
        ("mean", C.c_double),
        ("m2", C.c_double),
        ("device_ms", C.c_float),
        ("gpu_launches", C.c_int32),
    ]


SELECT_MAX_TARGETS = 8
SELECT_BINS = 2048


class SmmcSelectState(C.Structure):
    _fields_ = [
        ("n_targets", C.c_int32),
        ("bits_done", C.c_int32),
        ("rank", C.c_uint64 * SELECT_MAX_TARGETS),
        ("prefix", C.c_uint32 * SELECT_MAX_TARGETS),
    ]


class SmmcPlan(C.Structure):
    _fields_ = [
        ("grid", C.c_int32),
        ("threads", C.c_int32),
        ("table_copies", C.c_int32),
        ("ticket", C.c_int32),
        ("smem_bytes", C.c_uint64),
    ]


class SmmcError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"{STATUS_NAMES.get(status, status)}: {message}")
        self.status = status


# every symbol include/smmc.h declares (tests check the library exports exactly these)
C_ABI_SYMBOLS = [
    "smmc_run", "smmc_run_multi", "smmc_run_select", "smmc_stats_len", "smmc_stats_finalize", "smmc_reduce_mean", "smmc_reduce_mean_device", "smmc_reduce_mean_device_noalloc", "smmc_vector_add", "smmc_vector_add_device",
    "smmc_debug_draws", "smmc_debug_philox", "smmc_debug_philox_rounds", "smmc_debug_set_variant", "smmc_kernel_times", "smmc_kernel_kinds", "smmc_last_error", "smmc_version", "smmc_device_count", "smmc_shutdown",
    "smmc_select_begin", "smmc_select_pass", "smmc_select_advance", "smmc_select_values", "smmc_quantiles",
    "smmc_trajectories", "smmc_read_returns_csv", "smmc_write_data_file", "smmc_plan_launch",
    "smmc_comm_unique_id", "smmc_comm_init_rank", "smmc_comm_adopt", "smmc_comm_destroy", "smmc_comm_info", "smmc_comm_allreduce_count", "smmc_run_rank",
    "smmc_rank_range", "smmc_quantiles_rank",
]
COMM_ID_BYTES = 128

_lib = None


def load():
    """Load libsmmc.so (once).  Raises ImportError with build instructions when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C stock_market_monte_carlo_b200/csrc`). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    P = C.POINTER
    lib.smmc_run.argtypes = [P(SmmcConfig), P(SmmcResult)]
    lib.smmc_run.restype = C.c_int
    lib.smmc_run_multi.argtypes = [P(SmmcConfig), C.c_int, P(SmmcResult)]
    lib.smmc_run_multi.restype = C.c_int
    lib.smmc_stats_len.argtypes = [C.c_int32]
    lib.smmc_stats_len.restype = C.c_int32
    lib.smmc_stats_finalize.argtypes = [C.c_void_p, C.c_int32, C.c_float, P(C.c_uint64), P(C.c_uint64),
                                        P(C.c_double), P(C.c_double), C.c_void_p]
    lib.smmc_stats_finalize.restype = C.c_int
    lib.smmc_run_select.argtypes = [P(SmmcConfig), P(SmmcResult), P(SmmcSelectState), C.c_void_p]
    lib.smmc_run_select.restype = C.c_int
    lib.smmc_vector_add_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, P(C.c_float)]
    lib.smmc_vector_add_device.restype = C.c_int
    lib.smmc_reduce_mean.argtypes = [C.c_void_p, C.c_int64, C.c_int32, P(C.c_float)]
    lib.smmc_reduce_mean.restype = C.c_int
    lib.smmc_reduce_mean_device.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, P(C.c_float), P(C.c_float)]
    lib.smmc_reduce_mean_device.restype = C.c_int
    lib.smmc_reduce_mean_device_noalloc.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, P(C.c_float)]
    lib.smmc_reduce_mean_device_noalloc.restype = C.c_int
    lib.smmc_vector_add.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32]
    lib.smmc_vector_add.restype = C.c_int
    lib.smmc_debug_draws.argtypes = [P(SmmcConfig), C.c_int32, C.c_void_p]
    lib.smmc_debug_draws.restype = C.c_int
    lib.smmc_debug_philox.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
    lib.smmc_debug_philox.restype = C.c_int
    lib.smmc_debug_philox_rounds.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
    lib.smmc_debug_philox_rounds.restype = C.c_int
    lib.smmc_debug_set_variant.argtypes = [C.c_char_p, C.c_int32]
    lib.smmc_debug_set_variant.restype = C.c_int
    lib.smmc_kernel_times.argtypes = [C.c_int32, C.c_void_p, C.c_int32, P(C.c_int32)]
    lib.smmc_kernel_times.restype = C.c_int
    lib.smmc_kernel_kinds.argtypes = [C.c_int32, C.c_void_p, C.c_int32, P(C.c_int32)]
    lib.smmc_kernel_kinds.restype = C.c_int
    lib.smmc_select_begin.argtypes = [P(SmmcSelectState), C.c_int32, C.c_void_p]
    lib.smmc_select_begin.restype = C.c_int
    lib.smmc_select_pass.argtypes = [P(SmmcConfig), P(SmmcSelectState), C.c_void_p]
    lib.smmc_select_pass.restype = C.c_int
    lib.smmc_select_advance.argtypes = [P(SmmcSelectState), C.c_void_p]
    lib.smmc_select_advance.restype = C.c_int
    lib.smmc_select_values.argtypes = [P(SmmcSelectState), C.c_void_p]
    lib.smmc_select_values.restype = C.c_int
    lib.smmc_quantiles.argtypes = [P(SmmcConfig), C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, P(C.c_float)]
    lib.smmc_quantiles.restype = C.c_int
    lib.smmc_trajectories.argtypes = [P(SmmcConfig), C.c_int32, C.c_void_p, C.c_void_p]
    lib.smmc_trajectories.restype = C.c_int
    lib.smmc_read_returns_csv.argtypes = [C.c_char_p, C.c_void_p, C.c_int32, P(C.c_int32)]
    lib.smmc_read_returns_csv.restype = C.c_int
    lib.smmc_write_data_file.argtypes = [C.c_char_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32]
    lib.smmc_write_data_file.restype = C.c_int
    lib.smmc_plan_launch.argtypes = [P(SmmcConfig), C.c_int32, C.c_int32, C.c_uint64, P(SmmcPlan)]
    lib.smmc_plan_launch.restype = C.c_int
    lib.smmc_comm_unique_id.argtypes = [C.c_void_p]
    lib.smmc_comm_unique_id.restype = C.c_int
    lib.smmc_comm_init_rank.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, P(C.c_void_p)]
    lib.smmc_comm_init_rank.restype = C.c_int
    lib.smmc_comm_adopt.argtypes = [C.c_void_p, C.c_int32, P(C.c_void_p)]
    lib.smmc_comm_adopt.restype = C.c_int
    lib.smmc_comm_destroy.argtypes = [C.c_void_p]
    lib.smmc_comm_destroy.restype = C.c_int
    lib.smmc_comm_info.argtypes = [C.c_void_p, P(C.c_int32), P(C.c_int32), P(C.c_int32), P(C.c_int32)]
    lib.smmc_comm_info.restype = C.c_int
    lib.smmc_comm_allreduce_count.argtypes = [C.c_void_p, P(C.c_uint64)]
    lib.smmc_comm_allreduce_count.restype = C.c_int
    lib.smmc_run_rank.argtypes = [P(SmmcConfig), P(SmmcResult), C.c_void_p]
    lib.smmc_run_rank.restype = C.c_int
    lib.smmc_rank_range.argtypes = [C.c_int64, C.c_int32, C.c_int32, P(C.c_int64), P(C.c_int64)]
    lib.smmc_rank_range.restype = C.c_int
    lib.smmc_quantiles_rank.argtypes = [P(SmmcConfig), C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, P(C.c_float)]
    lib.smmc_quantiles_rank.restype = C.c_int
    lib.smmc_last_error.restype = C.c_char_p
    lib.smmc_version.restype = C.c_char_p
    lib.smmc_device_count.restype = C.c_int32
    lib.smmc_shutdown.restype = None
    _lib = lib
    return lib


def check(status):
    if status != 0:
        raise SmmcError(status, load().smmc_last_error().decode())
