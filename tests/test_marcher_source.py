"""The device Mandelbulb routines (rt_scene.cuh, block [marcher-core-begin]..[marcher-core-end]) compiled
for the HOST — every __d*_rn / __fma_rn intrinsic mapped to the single IEEE operation it stands for —
must return the same bits as the oracle's canonical routines.  This keeps the two independent
statements of the map in step without a GPU; the GPU itself is checked by the strict parity tests
of the `raymarching` scene (tests/test_gpu_parity.py)."""
import ctypes as C
import os
import subprocess

import numpy as np

import oracle_binding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SHIM = r'''
#include <cmath>
#include <cstdint>
#include <cstring>
#define __device__
#define __forceinline__ inline
#define __noinline__
struct V3 { double x, y, z; };
static inline double __dmul_rn(double a, double b) { return a * b; }
static inline double __dadd_rn(double a, double b) { return a + b; }
static inline double __ddiv_rn(double a, double b) { return a / b; }
static inline double __dsqrt_rn(double a) { return std::sqrt(a); }
static inline double __fma_rn(double a, double b, double c) { return std::fma(a, b, c); }
static inline double __longlong_as_double(long long v) { double d; std::memcpy(&d, &v, 8); return d; }
static inline long long __double_as_longlong(double d) { long long v; std::memcpy(&v, &d, 8); return v; }
%s
extern "C" int dev_iterate(const double* xyz, long n, int iterations, double* r, double* dr, int* inside) {
    for (long i = 0; i < n; ++i) { V3 p{xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]}; inside[i] = marcher_iterate(p, iterations, r[i], dr[i]) ? 1 : 0; }
    return 0;
}
extern "C" int dev_log(const double* x, long n, double* out) { for (long i = 0; i < n; ++i) out[i] = marcher_log(x[i]); return 0; }
'''


def _build(tmp_path):
    src = open(os.path.join(ROOT, "raysnail_b200", "csrc", "device", "rt_scene.cuh")).read()
    block = src[src.index("// [marcher-core-begin]"):src.index("// [marcher-core-end]")].replace("#pragma unroll", "")
    cpp = tmp_path / "marcher_host.cpp"
    cpp.write_text(SHIM % block)
    so = tmp_path / "marcher_host.so"
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", str(cpp), "-o", str(so)])
    return C.CDLL(str(so))


def test_device_source_on_host_matches_the_oracle_bit_for_bit(tmp_path):
    dev = _build(tmp_path)
    orc = oracle_binding.load()
    rng = np.random.default_rng(11)
    n = 200000
    pts = rng.uniform(-1.3, 1.3, size=(n, 3))
    pts[:8] = [[0, 0, 0], [0, 0, 0.65], [0, 0, -1.0], [2, 0, 0], [0.5, 0, 0], [0, 0.5, 0], [-0.0, 0.0, 0.3], [1e-200, 0, 0.2]]
    out = {}
    for name, fn in (("dev", dev.dev_iterate), ("orc", orc.rso_marcher_iterate)):
        r, dr, ins = np.zeros(n), np.zeros(n), np.zeros(n, np.int32)
        fn(C.c_void_p(pts.ctypes.data), C.c_long(n), C.c_int(100), C.c_void_p(r.ctypes.data), C.c_void_p(dr.ctypes.data), C.c_void_p(ins.ctypes.data))
        out[name] = (r, dr, ins)
    assert 0.05 < out["orc"][2].mean() < 0.5            # a real mix of bounded and escaping orbits
    assert np.array_equal(out["dev"][2], out["orc"][2])
    assert np.array_equal(out["dev"][0].view(np.uint64), out["orc"][0].view(np.uint64))
    assert np.array_equal(out["dev"][1].view(np.uint64), out["orc"][1].view(np.uint64))
    # the fixed-series logarithm: same bits, and within 2 ulp of libm on (0, inf)
    x = np.concatenate([rng.uniform(0, 4096, 100000), 10.0 ** rng.uniform(-300, 300, 100000), [0.0, 1.0, 5e-324, 1e-310, np.inf]])
    a, b = np.zeros(x.size), np.zeros(x.size)
    dev.dev_log(C.c_void_p(x.ctypes.data), C.c_long(x.size), C.c_void_p(a.ctypes.data))
    orc.rso_marcher_log(C.c_void_p(x.ctypes.data), C.c_long(x.size), C.c_void_p(b.ctypes.data))
    assert np.array_equal(a.view(np.uint64), b.view(np.uint64))
    fin = np.isfinite(x) & (x > 0)
    ref = np.log(x[fin])
    err = np.abs(b[fin] - ref)
    assert (err <= 2 * np.spacing(np.abs(ref)) + 1e-300).all()
    assert b[x == 0.0][0] == -np.inf and b[-1] == np.inf
