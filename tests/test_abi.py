"""The drop-in boundary: libsmmc.so loads, exports every symbol include/smmc.h declares plus the reference's five
C++ entry points under their Itanium-mangled names, the ctypes mirrors match the C structs, and
the library fails loudly without a device (no CPU fallback).  No compute happens here."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "smmc.h")

# reference include/stock_market_monte_carlo/simulations.h:69-86 and gpu.h:1-2, mangled (probed with nm on the
# reference's own objects)
REFERENCE_CXX_SYMBOLS = [
    "_Z18mc_simulations_gpuRSt6atomicIlElifRSt6vectorIfSaIfEES5_i",
    "_Z30mc_simulations_gpu_reduceBlockRSt6atomicIlElifRSt6vectorIfSaIfEES5_S5_i",
    "_Z15reduce_mean_gpuRSt6vectorIfSaIfEEl",
    "_Z10vector_addPfS_S_i",
    "_Z14vector_add_gpuPfS_S_i",
]


def _declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(smmc_[a-z_0-9]+)\s*\(", text)))


def _exported(S):
    from stock_market_monte_carlo_b200 import _native
    out = subprocess.run(["nm", "-D", "--defined-only", _native.LIB_PATH], capture_output=True, text=True, check=True)
    return {line.split()[-1] for line in out.stdout.splitlines() if line.strip()}


def test_library_exports_every_declared_symbol(S):
    from stock_market_monte_carlo_b200 import _native
    declared = _declared_functions()
    assert declared == sorted(_native.C_ABI_SYMBOLS)
    exported = _exported(S)
    missing = [s for s in declared if s not in exported]
    assert not missing, missing
    lib = S.load()
    for s in declared:
        getattr(lib, s)


def test_reference_cxx_entry_points_exported(S):
    exported = _exported(S)
    missing = [s for s in REFERENCE_CXX_SYMBOLS if s not in exported]
    assert not missing, missing


def test_struct_layout_matches_header(S, tmp_path):
    """compile a C probe against include/smmc.h and compare sizeof/offsetof with the ctypes mirrors"""
    from stock_market_monte_carlo_b200 import _native as N
    fields_cfg = [f[0] for f in N.SmmcConfig._fields_]
    fields_res = [f[0] for f in N.SmmcResult._fields_]
    src = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', "int main(void){",
           'printf("%zu %zu\\n", sizeof(smmc_config), sizeof(smmc_result));']
    for f in fields_cfg:
        src.append(f'printf("c {f} %zu\\n", offsetof(smmc_config, {f}));')
    for f in fields_res:
        src.append(f'printf("r {f} %zu\\n", offsetof(smmc_result, {f}));')
    src.append('printf("k %d %d %d %d\\n", SMMC_ABI_VERSION, SMMC_BLOCK, SMMC_STATS_HEADER, SMMC_MAX_HIST_BINS);')
    src.append("return 0;}")
    c = tmp_path / "probe.c"
    c.write_text("\n".join(src))
    exe = tmp_path / "probe"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-o", str(exe), str(c)], check=True)
    lines = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines()
    assert lines[0].split() == [str(C.sizeof(N.SmmcConfig)), str(C.sizeof(N.SmmcResult))]
    for line in lines[1:-1]:
        kind, name, off = line.split()
        struct = N.SmmcConfig if kind == "c" else N.SmmcResult
        assert getattr(struct, name).offset == int(off), line
    assert lines[-1].split()[1:] == [str(N.SMMC_ABI_VERSION), str(N.SMMC_BLOCK), str(N.SMMC_STATS_HEADER),
                                     str(N.SMMC_MAX_HIST_BINS)]


def test_header_is_plain_c(tmp_path):
    c = tmp_path / "plain.c"
    c.write_text(f'#include "{HEADER}"\nint main(void) {{ return smmc_stats_len(0) - 8; }}\n')
    subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-c", "-o", str(tmp_path / "plain.o"), str(c)],
                   check=True)


def test_cxx_header_compiles_like_the_reference_callers(tmp_path):
    """a caller written against the reference's signatures compiles against our drop-in header"""
    cpp = tmp_path / "caller.cpp"
    cpp.write_text(f'''
#include "{ROOT}/include/stock_market_monte_carlo/simulations.h"
#include "{ROOT}/include/stock_market_monte_carlo/gpu.h"
int main() {{
  std::atomic<long> n{{0}};
  std::vector<float> returns{{1.f, -2.f}}, totals, means, variances;
  if (false) {{
    mc_simulations_gpu(n, 100L, 360, 1000.f, returns, totals, 1);
    mc_simulations_gpu_reduceBlock(n, 100L, 360, 1000.f, returns, means, variances, 1);
    (void)reduce_mean_gpu(totals, 10L);
    vector_add_gpu(nullptr, nullptr, nullptr, 0);
  }}
  return 0;
}}''')
    from stock_market_monte_carlo_b200 import _native
    libdir = os.path.dirname(_native.LIB_PATH)
    subprocess.run(["g++", "-std=c++17", "-Wall", "-o", str(tmp_path / "caller"), str(cpp), f"-L{libdir}", "-lsmmc",
                    f"-Wl,-rpath,{libdir}"], check=True)
    subprocess.run([str(tmp_path / "caller")], check=True)


def test_no_device_fails_loudly(S):
    lib = S.load()
    if lib.smmc_device_count() > 0:
        pytest.skip("a device is visible")
    with pytest.raises(S.SmmcError) as e:
        S.simulate("gaussian", 1000, 12)
    assert e.value.status == 2 and "no CPU fallback" in str(e.value)
    with pytest.raises(S.SmmcError):
        S.reduce_mean_gpu([1.0, 2.0], 2)


def test_validation_precedes_device_probe(S, table):
    """argument errors are reported as such with or without a GPU (reference behaviour for bad input is UB/abort;
    the C ABI returns status codes instead, include/smmc.h)"""
    bad = [
        dict(mode="bootstrap", n_simulations=10, n_periods=10),                      # no table
        dict(mode="gaussian", n_simulations=-1, n_periods=10),
        dict(mode="gaussian", n_simulations=10, n_periods=-1),
        dict(mode="gaussian", n_simulations=10, n_periods=10, stream_offset=100),    # not a multiple of 256
        dict(mode="gaussian", n_simulations=10, n_periods=10, hist_bins=5000),
        dict(mode="gaussian", n_simulations=10, n_periods=10, hist_bins=10, hist_range=(5.0, 5.0)),
        dict(mode="sobol_gaussian", n_simulations=10, n_periods=1112),
        dict(mode="bootstrap_compat", n_simulations=2 ** 31, n_periods=1, returns=table),
        dict(mode="sobol_bootstrap", n_simulations=2 ** 32, n_periods=1, returns=table),
        dict(mode=17, n_simulations=10, n_periods=10),
    ]
    for kw in bad:
        with pytest.raises(S.SmmcError) as e:
            S.simulate(**kw)
        assert e.value.status == 1, kw


def test_round2_fields_are_validated_before_the_device_probe(S, table):
    """ABI 3: philox_rounds in {0, 7, 10}; non-finite scalars and table entries are argument errors, not NaN results."""
    bad = [
        dict(mode="gaussian", n_simulations=10, n_periods=10, philox_rounds=8),
        dict(mode="bootstrap", n_simulations=10, n_periods=10, returns=table, philox_rounds=-1),
        dict(mode="gaussian", n_simulations=10, n_periods=10, gauss_mean=float("nan")),
        dict(mode="gaussian", n_simulations=10, n_periods=10, gauss_std=float("inf")),
        dict(mode="gaussian", n_simulations=10, n_periods=10, initial_capital=float("nan")),
        dict(mode="gaussian", n_simulations=10, n_periods=10, hist_bins=10, hist_range=(0.0, float("inf"))),
        dict(mode="gaussian", n_simulations=2 ** 38 + 256, n_periods=1),
    ]
    nan_table = table.copy()
    nan_table[17] = np.nan
    bad.append(dict(mode="bootstrap", n_simulations=10, n_periods=10, returns=nan_table))
    for kw in bad:
        with pytest.raises(S.SmmcError) as e:
            S.simulate(**kw)
        assert e.value.status == 1, kw


def test_comm_entry_points_validate_and_report(S):
    """smmc_comm_*: null / out-of-range arguments are INVALID_ARG; destroying nothing is fine."""
    lib = S.load()
    n, r, d, v = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
    assert lib.smmc_comm_info(None, C.byref(n), C.byref(r), C.byref(d), C.byref(v)) == 1
    assert lib.smmc_comm_unique_id(None) == 1
    comm = C.c_void_p()
    uid = C.create_string_buffer(128)  # SMMC_COMM_ID_BYTES
    assert lib.smmc_comm_init_rank(uid, 0, 0, 0, C.byref(comm)) == 1   # n_ranks 0
    assert lib.smmc_comm_init_rank(uid, 2, 2, 0, C.byref(comm)) == 1   # rank out of range
    assert lib.smmc_comm_init_rank(None, 1, 0, 0, C.byref(comm)) == 1  # no id
    assert lib.smmc_comm_destroy(None) == 0


def test_new_entry_points_fail_loudly_without_a_device(S, table):
    """quantiles / select passes / trajectories are device work: no CPU fallback either"""
    lib = S.load()
    if lib.smmc_device_count() > 0:
        pytest.skip("a device is visible")
    for call in (lambda: S.quantiles("gaussian", 1000, 12),
                 lambda: S.trajectories("bootstrap", 10, 12, returns=table),
                 lambda: S.SelectState([0]).run_pass(S.make_config("gaussian", 1000, 12)[0]),
                 lambda: S.reduce_mean_device(0x1000, 10),
                 lambda: S.vector_add_device(0x1000, 0x2000, 0x3000, 10),
                 lambda: S.simulate("gaussian", 1000, 12, select=S.SelectState([0]))):
        with pytest.raises(S.SmmcError) as e:
            call()
        assert e.value.status == 2 and "no CPU fallback" in str(e.value)


def test_new_entry_points_validate_before_probing(S, table):
    bad = [
        lambda: S.quantiles("bootstrap", 100, 12),                               # no table
        lambda: S.quantiles("gaussian", 100, 12, ranks=list(range(9))),          # > SMMC_SELECT_MAX_TARGETS
        lambda: S.quantiles("gaussian", 100, 12, ranks=[]),
        lambda: S.trajectories("sobol_gaussian", 4, 2000),
        lambda: S.trajectories("gaussian", 4, 12, stream_offset=3),
        lambda: S.plan_launch("gaussian", 100, 12, select_targets=9),
        lambda: S.plan_launch("gaussian", 100, 12, n_sms=0),
        lambda: S.simulate("gaussian", 1000, -1, select=S.SelectState([0])),
        lambda: S.vector_add_device(0, 0x1000, 0x1000, 10),
        lambda: S.vector_add_device(0x1000, 0x1002, 0x1000, 10),
        lambda: S.reduce_mean_device(0, 10),          # null pointer
        lambda: S.reduce_mean_device(0x1000, 0),      # n <= 0
        lambda: S.reduce_mean_device(0x1002, 10),     # not 4-byte aligned
    ]
    for call in bad:
        with pytest.raises(S.SmmcError) as e:
            call()
        assert e.value.status == 1
