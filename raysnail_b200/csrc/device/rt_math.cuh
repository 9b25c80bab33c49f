// Device math for the sm_100a path tracer: f64 vectors, ONB, Philox4x32-10 streams, samplers.
//
// Geometry is f64 because the reference is (src/prelude/vec3.rs:14-19) and its scenes put
// r=1000..10000 spheres next to r=0.2 ones; B200 issues FP64 at half the FP32 rate, so this
// is affordable.  Colours/throughput are f32 like the reference's Color (prelude/color.rs).
// The fma placement of dot/length_squared/Ray::at follows vec3.rs:152-155,177-179 and
// ray.rs:21-31 so hit distances agree with the CPU oracle to rounding.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

// RS_CHECKED=1 (make VARIANT=_chk TUNE=-DRS_CHECKED=1): the library's own memcheck.  Every indexed store / gather of the
// wave state and every stack push goes through RS_IN_BOUNDS; a failed check is counted on the device, the access is
// skipped, and the render call that follows returns RS_ERR_CUDA with the count (compute-sanitizer is not available on the
// GPU pool this was developed on).  In the default build the macro is the constant `true`: no code.
#ifndef RS_CHECKED
#define RS_CHECKED 0
#endif
#if RS_CHECKED && defined(RS_DEFINE_VIOLATIONS)
namespace rsd { __device__ unsigned long long rs_violations = 0ull; }
#define RS_IN_BOUNDS(cond) ((cond) ? true : (atomicAdd(&rsd::rs_violations, 1ull), false))
#elif RS_CHECKED
#define RS_IN_BOUNDS(cond) (cond)  // (translation units without device entry points)
#else
#define RS_IN_BOUNDS(cond) true
#endif

namespace rsd {

#define RS_PI 3.14159265358979323846
#define RS_F64_MAX 1.7976931348623157e308
#define RS_INV_PI 0.31830988618379067154

struct V3 {
    double x, y, z;
};

__device__ __forceinline__ V3 v3(double x, double y, double z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ V3 operator*(V3 a, double s) { return v3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator*(double s, V3 a) { return v3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ double dot(V3 a, V3 b) { return fma(a.z, b.z, fma(a.x, b.x, a.y * b.y)); }
__device__ __forceinline__ double length_squared(V3 a) { return fma(a.z, a.z, fma(a.x, a.x, a.y * a.y)); }
__device__ __forceinline__ double length(V3 a) { return sqrt(length_squared(a)); }
__device__ __forceinline__ V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
// vec3.rs:486-500: `v / s` is `v * (1/s)`
__device__ __forceinline__ V3 vdiv(V3 a, double s) { return a * (1.0 / s); }
__device__ __forceinline__ V3 unit(V3 a) { return vdiv(a, length(a)); }
// Shade-side variants (scatter directions / weights are compared with the reference at 1e-5,
// not bit for bit): one rsqrt instead of sqrt + reciprocal.  Never used where an object id or a
// primary ray is decided (raygen, extend).
__device__ __forceinline__ V3 unit_fast(V3 a) { return a * rsqrt(length_squared(a)); }
__device__ __forceinline__ double comp(V3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
__device__ __forceinline__ void set_comp(V3& a, int i, double v) { if (i == 0) a.x = v; else if (i == 1) a.y = v; else a.z = v; }

struct Ray {
    V3 o, d;
    double time;
};
// ray.rs:21-31
__device__ __forceinline__ V3 ray_at(const Ray& r, double t) { return v3(fma(r.d.x, t, r.o.x), fma(r.d.y, t, r.o.y), fma(r.d.z, t, r.o.z)); }

// onb.rs
struct ONB {
    V3 u, v, w;
};
__device__ __forceinline__ ONB onb_build(V3 n) {
    ONB o;
    o.w = unit(n);
    V3 uc = cross(v3(0.0, 1.0, 0.0), o.w);
    if (length_squared(uc) < 0.00000001) o.u = unit(cross(v3(1.0, 0.0, 0.0), o.w));
    else o.u = unit(uc);
    o.v = cross(o.w, o.u);
    return o;
}
__device__ __forceinline__ ONB onb_build_fast(V3 n) {  // shade only, see unit_fast
    ONB o;
    o.w = unit_fast(n);
    V3 uc = cross(v3(0.0, 1.0, 0.0), o.w);
    if (length_squared(uc) < 0.00000001) o.u = unit_fast(cross(v3(1.0, 0.0, 0.0), o.w));
    else o.u = unit_fast(uc);
    o.v = cross(o.w, o.u);
    return o;
}
__device__ __forceinline__ V3 onb_local(const ONB& o, V3 a) {
    return v3(o.u.x * a.x + o.v.x * a.y + o.w.x * a.z, o.u.y * a.x + o.v.y * a.y + o.w.y * a.z, o.u.z * a.x + o.v.z * a.y + o.w.z * a.z);
}

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11).  Stream layout (DESIGN.md "RNG streams"):
//   key = (seed_lo, seed_hi);  counter = (pixel, sample, depth, purpose | call << 8)
//   uniform = (word >> 8) * 2^-24  (exact in f32 and f64)
// ---------------------------------------------------------------------------------------------
enum Purpose : uint32_t { P_PIXEL = 0, P_LENS = 1, P_SCATTER = 2, P_MIXED = 3, P_LIGHT = 4, P_BRDF = 5, P_MEDIUM = 6 };

// The ten round keys key + r * (0x9E3779B9, 0xBB67AE85) depend on the frame's seed alone, so the host computes them
// (FrameParams::rk, make_frame) and a round reads its key straight from the kernel-parameter constant bank as an operand
// of the XOR: two additions per round less than bumping the key in registers (3.6 % of the bounce-0 extend's
// instructions went there; RS_TUNE_PHILOX_RK=0 builds that form for A/B).  Same integers either way.
#ifndef RS_TUNE_PHILOX_RK
#define RS_TUNE_PHILOX_RK 1
#endif
struct PhiloxKeys {
    uint2 rk[10];
};
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, const PhiloxKeys& K) {
#if RS_TUNE_PHILOX_RK
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ K.rk[r].x, lo1, hi0 ^ c.w ^ K.rk[r].y, lo0);
    }
#else
    uint2 k = K.rk[0];
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u;
        k.y += 0xBB67AE85u;
    }
#endif
    return c;
}

__device__ __forceinline__ double u01(uint32_t w) { return (double)(w >> 8) * (1.0 / 16777216.0); }

struct RngCtx {
    const PhiloxKeys* key;  // FrameParams::rk of the launch (kernel-parameter space)
    uint32_t pixel, sample, depth;
};
__device__ __forceinline__ uint4 rng_words(const RngCtx& r, uint32_t purpose, uint32_t call) {
    return philox4x32_10(make_uint4(r.pixel, r.sample, r.depth, purpose | (call << 8)), *r.key);
}
// rejection loops: iteration k reads word pair (k & 1) of call base + (k >> 1)
__device__ __forceinline__ void rng_pair(const RngCtx& r, uint32_t purpose, uint32_t base, uint32_t k, double& a, double& b) {
    uint4 w = rng_words(r, purpose, base + (k >> 1));
    if (k & 1) { a = u01(w.z); b = u01(w.w); }
    else { a = u01(w.x); b = u01(w.y); }
}

// vec3.rs:100-111
__device__ __forceinline__ V3 cosine_direction(double r1, double r2) {
    double q2 = sqrt(r2);
    double s, c;
    sincospi(2.0 * r1, &s, &c);  // phi = 2 pi r1; sincospi reduces the argument exactly
    return v3(c * q2, s * q2, sqrt(1.0 - r2));
}
// vec3.rs:115-126
__device__ __forceinline__ V3 cosine_direction_exponent(double exponent, double r1, double xi2) {
    double r2 = pow(xi2, 1.0 / (exponent + 1.0));
    double sin_theta = sqrt(1.0 - r2 * r2);
    double s, c;
    sincospi(2.0 * r1, &s, &c);  // phi = 2 pi r1
    return v3(c * sin_theta, s * sin_theta, r2);
}

}  // namespace rsd
