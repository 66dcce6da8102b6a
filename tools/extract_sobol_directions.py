#!/usr/bin/env python
"""Extract the Joe-Kuo direction numbers vendored in the reference (include/sobol/sobol.cpp) as DATA.

Reads the `poly[1111]` initialiser and the `v[i][j] = m;` initial direction integers of `i4_sobol`
(reference include/sobol/sobol.cpp:340-453 and :472-13709), runs the Bratley-Fox recurrence
(the algorithm of :13776-13795, restated here) out to 32 columns, and writes

    stock_market_monte_carlo_b200/data/sobol_joe_kuo_1111x32.bin   uint32 little-endian, [1111][32],
        V[d][k] = m_{d,k} << (31 - k)   (binary fraction with bit 31 = first digit)

No reference source text is copied: only the published Joe & Kuo (2003) numbers are read out.
When oracle/_ref/libref_sobol.so exists the result is validated against the compiled reference
(`i4_sobol` through its float output, which is exact for <= 24 significant bits).
Run in the build container (needs /root/reference):  python tools/extract_sobol_directions.py
"""
import ctypes
import os
import re
import sys

import numpy as np

REF = os.environ.get("SMMC_REFERENCE", "/root/reference")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "stock_market_monte_carlo_b200", "data", "sobol_joe_kuo_1111x32.bin")
DIMS, COLS = 1111, 32


def parse():
    src = open(os.path.join(REF, "include", "sobol", "sobol.cpp")).read()
    # first generator in the file is i4_sobol; its tables end where i4_sobol_generate begins
    body = src[src.index("void i4_sobol"):src.index("float *i4_sobol_generate")]
    poly_txt = re.search(r"static int poly\[DIM_MAX2\]\s*=\s*\{([^}]*)\}", body).group(1)
    poly = [int(t) for t in re.findall(r"\d+", poly_txt)]
    assert len(poly) == DIMS, len(poly)
    init = {}
    for i, j, val in re.findall(r"v\[(\d+)\]\[(\d+)\]\s*=\s*(\d+)\s*;", body):
        init[(int(i), int(j))] = int(val)
    return poly, init


def directions(poly, init):
    m_tab = np.zeros((DIMS, COLS), dtype=np.uint64)
    m_tab[0, :] = 1  # first dimension: van der Corput
    for d in range(1, DIMS):
        deg = poly[d].bit_length() - 1
        # coefficient bits below the leading one, most significant first (includ[0..deg-1])
        includ = [(poly[d] >> (deg - 1 - k)) & 1 for k in range(deg)]
        for j in range(deg):
            m_tab[d, j] = init[(d, j)]
        for j in range(deg, COLS):
            newv = int(m_tab[d, j - deg])
            l = 1
            for k in range(deg):
                l *= 2
                if includ[k]:
                    newv ^= l * int(m_tab[d, j - k - 1])
            m_tab[d, j] = newv
    v = np.zeros((DIMS, COLS), dtype=np.uint32)
    for k in range(COLS):
        assert int(m_tab[:, k].max()) < (1 << (k + 1))
        v[:, k] = (m_tab[:, k] << np.uint64(31 - k)).astype(np.uint32)
    return v


def sobol_points(v, p):
    g = p ^ (p >> 1)
    x = np.zeros(v.shape[0], dtype=np.uint32)
    k = 0
    while g >> k:
        if (g >> k) & 1:
            x ^= v[:, k]
        k += 1
    return x


def validate(v):
    so = os.path.join(ROOT, "oracle", "_ref", "libref_sobol.so")
    if not os.path.exists(so):
        print("oracle/_ref/libref_sobol.so absent: skipping validation against the compiled reference")
        return
    lib = ctypes.CDLL(so)
    lib.ref_i4_sobol.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_float)]
    quasi = (ctypes.c_float * DIMS)()
    bad = 0
    pts = list(range(0, 300)) + [2 ** k for k in range(8, 20)] + [2 ** k + 12345 for k in range(14, 20)]
    for p in sorted(set(pts)):
        seed = ctypes.c_int(p)
        lib.ref_i4_sobol(DIMS, ctypes.byref(seed), quasi)
        x = sobol_points(v, p)
        # reference integer lastq is 30-bit (LOG_MAX 30): x32 >> 2; float output = lastq * 2^-30 (RN to 24 bits)
        mine = (x >> np.uint32(2)).astype(np.float32) * np.float32(2.0 ** -30)
        ref = np.frombuffer(quasi, dtype=np.float32)
        bad += int((mine != ref).sum())
    print(f"validated {len(set(pts))} points x {DIMS} dims against compiled i4_sobol: {bad} mismatches")
    if bad:
        sys.exit(1)


if __name__ == "__main__":
    poly, init = parse()
    v = directions(poly, init)
    validate(v)
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    v.astype("<u4").tofile(OUT)
    print("wrote", OUT, v.shape, "first rows:", [hex(int(t)) for t in v[1, :4]])
