// Device-side scene: flat buffers + the intersection / containment / sampling routines for
// every Hittable of the reference, textures and material helpers.  Used by the extend, shade
// and probe kernels (kernels.cu).
//
// Object expression trees (CSG, TfFacade, ConstantMedium) are flattened at upload into post-order
// programs that one loop evaluates (see "Object programs" below).  RS_MAX_TREE_DEPTH bounds the
// nesting (operand stack); the host rejects deeper scenes at upload.
#pragma once

#include "../../../include/raysnail_b200.h"
#include "rt_math.cuh"

namespace rsd {

#define RS_MAX_TREE_DEPTH 8

// 64-byte BVH node, two children with their boxes (f32, padded outward on the host):
//   bx = (c0.lo.x, c0.hi.x, c0.lo.y, c0.hi.y)   by = (c1.lo.x, c1.hi.x, c1.lo.y, c1.hi.y)
//   bz = (c0.lo.z, c0.hi.z, c1.lo.z, c1.hi.z)   ch = (child0, child1, -, -)
// child >= 0: inner node index; child < 0: leaf holding root object ~child.
struct __align__(64) BvhNode {
    float4 bx, by, bz;
    int4 ch;
};

// 128-byte QBVH node (the world accelerator): four children, boxes as SoA float4s (component
// k = child k; f32, padded outward on the host), loaded with seven 128-bit ld.global.nc.
// child >= 0: inner node index; child < 0: leaf ~(object << 2 | kind).  Unused slots hold a
// degenerate box at +FLT_MAX (lo = hi), which no ray can hit (t_near = +inf or t_far = -inf).
struct __align__(128) BvhNode4 {
    float4 lox, hix, loy, hiy, loz, hiz;
    int4 child;
    int4 pad;
};

// pre-multiplied TransformStack of one TfFacade node: 3x4 forward, 3x4 inverse (row-major)
struct __align__(16) XformPre {
    double fwd[12];
    double inv[12];
};

// ROOT_FLAT: a rect, a box, a quadric (generic tier only), or a TfFacade around one of them: extend finds their distance with inline code
// (rect_t / box_t) instead of walking the expression tree; shade rebuilds the record like a generic root
// RS_TUNE_QUADRIC_FLAT = 1: the flat-tier kernels (MODE 2 / TIER 1) hold the quadric code too, so a scene of
// spheres, rects, boxes and quadrics needs no generic-tier kernel; 0: quadric roots force the generic tier
#ifndef RS_TUNE_QUADRIC_FLAT
#define RS_TUNE_QUADRIC_FLAT 0
#endif
#define RS_QUADRIC_MODE (RS_TUNE_QUADRIC_FLAT ? 2 : 3)
enum RootKind : int { ROOT_SPHERE = 0, ROOT_TRIANGLE = 1, ROOT_GENERIC = 2, ROOT_FLAT = 3 };

// Everything shade reads of a (resolved) material in ONE record, so that the dependent chain
// material -> texture -> nested texture (three round trips to L1/L2 in a latency-bound kernel) is a
// single level for the common cases.  Built at upload from rs_material + rs_texture:
//   tex_mode 0: the colour is c0 (a solid texture; the constant colour of Dielectric / Isotropic)
//   tex_mode 1: a checker of two solid textures: c0 = `even`, c1 = `odd`, `scale`
//   tex_mode 2: anything else: texture_color(texture)
#define RS_MATHOT_TYPE(m) ((m).type_flags & 0xff)
#define RS_MATHOT_TEX_MODE(m) (((m).type_flags >> 8) & 3)
#define RS_MATHOT_HAS_CURVE(m) (((m).type_flags >> 16) & 1)
#define RS_MATHOT_NEEDS_UV(m) (((m).type_flags >> 17) & 1)
struct __align__(16) MatHot {
    float4 c0, c1;
    double a, b;           // rs_material.a / .b
    double phong_factor;   // CommonMaterialSettings
    double scale;          // checker scale
    int type_flags;        // type | tex_mode << 8 | has_curve << 16 | needs_uv << 17
    int texture;
    int phong_exponent;
    int pad;
};

// object programs (see "Object programs" below)
struct __align__(16) ObjInstr {
    int op, node, xf, aux;
};
struct __align__(16) CsgInfo {
    int c0_off, c0_len, c1_off, c1_len;  // `contains` programs of child0 / child1 in the space of the CSG node
};
struct ObjPrograms {
    const ObjInstr* prog;
    const int2* cprog;
    const CsgInfo* csg;
    const int2* root_prog;  // per root: (first instruction, count); count 0 = not a ROOT_GENERIC object
};

struct DScene {
    const rs_node* nodes;
    const XformPre* xf;          // indexed by rs_node.first_xform (rewritten at upload)
    const BvhNode4* bvh;         // 4-wide QBVH (when selected)
    const BvhNode* bvh2;         // binary tree (default)
    int bvh_root;                // >= 0 node, < 0 single leaf ref
    int n_roots;
    uint32_t bvh2_bytes;         // size of the binary tree and of the sphere table (16-byte multiples): what the
    uint32_t sphere_bytes;       // staged extend kernels copy into shared memory (0 = not staged)
    const float4* root_list;     // objects searched as a list before the tree: (lo.xyz, bits of the leaf ref), (hi.xyz, 1 = never dropped by distance)
    int any_gate;                // some root is gated by its reference bbox (aabb_hit_f64 needs 1 / direction)
    int n_list;                  // unbounded objects; in small worlds every object (then bvh_root is the sentinel: no tree)
    int n_list_lockstep;         // entries [0, n_list_lockstep) are spheres / flat objects, the expression trees follow (closest_hit)
    const int* root_node;        // node index or ~triangle
    const int* root_kind;        // RootKind
    const double4* root_sphere;  // ROOT_SPHERE: (cx, cy, cz, r)
    const double* root_inv_r;    // ROOT_SPHERE: 1.0 / r (the IEEE quotient, computed on the host)
    const int* root_material;    // ROOT_SPHERE / ROOT_TRIANGLE: the object's material (-1 = None)
    const MatHot* mat_hot;       // one per material
    int all_spheres;             // every root is a ROOT_SPHERE
    const rs_node* light_nodes;  // copy of nodes[lights[k]]: the lean shade kernel samples plain sphere / rect lights
    const double* root_bbox;     // 6 per root (reference bbox, for the gate)
    const unsigned char* root_gate;
    ObjPrograms obj;             // programs of the ROOT_GENERIC objects (expression trees)
    const double2* tri_geom;     // RS_TRI_STRIDE per triangle (load_triangle)
    const double* tri_normals;   // 9 per triangle
    const int* tri_material;
    const int* lights;
    int n_lights;
    const rs_material* materials;
    const rs_texture* textures;
    const unsigned char* texels;
    const int* perlin_perm;
    const double* perlin_values;
    int default_material;
    float4 bg0, bg1;
};

struct HitRec {
    V3 point, normal;
    double t1, t2, u, v;
    int material;  // -1 = None
    bool outside;
};

// [start, end) like std::ops::Range<f64>::contains
__device__ __forceinline__ bool in_range(double t, double lo, double hi) { return t >= lo && t < hi; }

// aabb.rs:20-38 in f64 (the gate for roots whose reference bbox does not contain them).  `inv` = 1.0 / direction per
// axis, the quotient the reference computes inside the test: it depends on the ray alone, so the caller divides
// once per ray, not once per object (three f64 divisions cost about what the rest of the test costs).
__device__ __forceinline__ bool aabb_hit_f64(const double* b, const Ray& r, const V3& inv, double lo, double hi) {
    double t_min = lo, t_max = hi;
    {
        double t0 = (b[0] - r.o.x) * inv.x, t1 = (b[3] - r.o.x) * inv.x;
        if (inv.x < 0.0) { double s = t0; t0 = t1; t1 = s; }
        t_min = fmax(t_min, t0); t_max = fmin(t_max, t1);
        if (t_max <= t_min) return false;
    }
    {
        double t0 = (b[1] - r.o.y) * inv.y, t1 = (b[4] - r.o.y) * inv.y;
        if (inv.y < 0.0) { double s = t0; t0 = t1; t1 = s; }
        t_min = fmax(t_min, t0); t_max = fmin(t_max, t1);
        if (t_max <= t_min) return false;
    }
    {
        double t0 = (b[2] - r.o.z) * inv.z, t1 = (b[5] - r.o.z) * inv.z;
        if (inv.z < 0.0) { double s = t0; t0 = t1; t1 = s; }
        t_min = fmax(t_min, t0); t_max = fmin(t_max, t1);
        if (t_max <= t_min) return false;
    }
    return true;
}

// ---------------------------------------------------------------------------------------------
// Primitives
// ---------------------------------------------------------------------------------------------
// sphere.rs:83-109 — returns the accepted root (t1) and the far root (t2)
__device__ __forceinline__ bool sphere_roots(V3 c, double r2, const Ray& ray, double lo, double hi, double& t1, double& t2) {
    V3 l = ray.o - c;
    double half_b = dot(ray.d, l);
    double a = length_squared(ray.d);
    double cc = length_squared(l) - r2;
    double delta = half_b * half_b - a * cc;
    if (delta < 0.0) return false;
    double sq = sqrt(delta);
    double ta = (-half_b - sq) / a;
    double tb = (-half_b + sq) / a;
    if (in_range(ta, lo, hi)) { t1 = ta; t2 = tb; return true; }
    if (in_range(tb, lo, hi)) { t1 = tb; t2 = tb; return true; }
    return false;
}

// root-level variant: only the accepted root is needed (the far root t2 matters to CSG only)
__device__ __forceinline__ bool sphere_nearest(V3 c, double r2, const Ray& ray, double lo, double hi, double& t) {
    V3 l = ray.o - c;
    double half_b = dot(ray.d, l);
    double a = length_squared(ray.d);
    double cc = length_squared(l) - r2;
    double delta = half_b * half_b - a * cc;
    if (delta < 0.0) return false;
    double sq = sqrt(delta);
    double ta = (-half_b - sq) / a;
    if (in_range(ta, lo, hi)) { t = ta; return true; }
    double tb = (-half_b + sq) / a;
    if (in_range(tb, lo, hi)) { t = tb; return true; }
    return false;
}

// sphere.rs:64-71
__device__ __forceinline__ void sphere_uv(V3 center, V3 point, double& u, double& v) {
    V3 p = unit(point - center);
    double phi = atan2(-p.z, p.x);
    double theta = asin(p.y);
    u = phi / 2.0 / RS_PI + 0.5;
    v = theta / RS_PI + 0.5;
}

// HitRecord::new for a sphere (hit.rs:32-53, sphere.rs:56-58: static centre)
__device__ __forceinline__ void sphere_record(const double* p, int material, const Ray& ray, double t1, double t2, bool want_uv, HitRec& h) {
    V3 c = v3(p[0], p[1], p[2]);
    h.point = ray_at(ray, t1);
    V3 n = (h.point - c) * p[7];  // p[7] = 1.0 / r, filled in at upload (vec3.rs:486-500: v / s = v * (1/s))
    h.outside = dot(ray.d, n) < 0.0;
    h.normal = h.outside ? n : -n;
    h.material = material;
    h.t1 = t1;
    h.t2 = t2;
    h.u = 0.0;
    h.v = 0.0;
    if (want_uv) sphere_uv(c, h.point, h.u, h.v);
}

// FULL = false (all `hit` routines below): only what the closest-hit search and the CSG rules
// consult is filled in — t1, t2, point, material — not the normal, uv and `outside` (a sphere's
// uv alone is an atan2 and an asin per candidate hit).  extend searches with FULL = false; shade
// rebuilds the record of the winner with FULL = true from the same distances.
__device__ __forceinline__ void lean_record(const Ray& ray, double t1, double t2, int material, HitRec& h) {
    h.point = ray_at(ray, t1);
    h.t1 = t1;
    h.t2 = t2;
    h.material = material;
}

template <bool FULL>
__device__ __forceinline__ bool hit_sphere_node(const rs_node& n, const Ray& ray, double lo, double hi, HitRec& h) {
    V3 c = v3(n.p[0], n.p[1], n.p[2]);
    if (n.p[4] != 0.0 || n.p[5] != 0.0 || n.p[6] != 0.0) c = c + v3(n.p[4], n.p[5], n.p[6]) * ray.time;  // center_at
    double t1, t2;
    if (!sphere_roots(c, n.p[3] * n.p[3], ray, lo, hi, t1, t2)) return false;
    if (FULL) sphere_record(n.p, n.material, ray, t1, t2, true, h);
    else lean_record(ray, t1, t2, n.material, h);
    return true;
}

__device__ __forceinline__ void rect_axes(int kind, int& a, int& b, int& c) {
    if (kind == RS_RECT_XY) { a = 0; b = 1; c = 2; }
    else if (kind == RS_RECT_XZ) { a = 0; b = 2; c = 1; }
    else { a = 1; b = 2; c = 0; }
}

// ~20-bit reciprocal (MUFU.RCP64H): one instruction where an IEEE division is ~35; for conservative pre-tests only
__device__ __forceinline__ double rcp_rough(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    return r;
}
#ifndef RS_TUNE_RECT_QUICK
#define RS_TUNE_RECT_QUICK 0  // MEASURED 15-20 % SLOWER on every scene with boxes (profiles/r02f_ab_rectquick.txt), see rect_t
#endif

// rect.rs:102-120, the distance only (what the closest-hit search of extend needs).
// A box is six of these and a ray hits two of them at most, so most calls end in `return false` — and the reference's
// first step is an f64 division.  The rejects that can be decided WITHOUT it come first (RS_TUNE_RECT_QUICK):
//  - numerator and denominator of opposite sign: the quotient is negative (or -0), below any lower limit > 0 — exact;
//  - the plane distance taken with a 20-bit reciprocal lies outside [lo, hi), or puts the hit point outside the
//    rectangle, by more than 2^-16 relative — 16 times the error of the estimate, so the exact test would refuse as well.
// Every condition is written so that a NaN or an infinity falls through; whatever is not refused runs the reference's
// own sequence (divide, then compare) unchanged, and an accepted hit is therefore the reference's bit for bit.
// Parity-green (75 GPU tests) but SLOWER, in extend and in shade alike: quadric.sdl 1.08 vs 1.30 G samples/s, csg.sdl
// 1.49 vs 1.83, Cornell 1.83 vs 2.16, example.sdl 4.08 vs 4.53 — the ~14 extra f64 operations and compares per face and
// the registers they hold (spill stores 456 -> 604 B in the generic extend) cost more than the division they avoid on
// four faces of six.  Off by default; the division stays.
__device__ __forceinline__ bool rect_t(int kind, double k, double a0, double a1, double b0, double b1, const Ray& ray, double lo, double hi, double& t1) {
    int ax0, ax1, ax2;
    rect_axes(kind, ax0, ax1, ax2);
    double dk = comp(ray.d, ax2);
#if RS_TUNE_RECT_QUICK
    {
        const double n = k - comp(ray.o, ax2);
        const double big = 1e300, eps = 1.52587890625e-05;  // 2^-16
        if (lo > 0.0 && fabs(n) < big && fabs(dk) < big && ((n < 0.0 && dk > 0.0) || (n > 0.0 && dk < 0.0))) return false;
        if (fabs(dk) > 1e-300 && fabs(dk) < big && fabs(n) < big) {
            const double te = n * rcp_rough(dk);  // |te - n / dk| <= 2^-20 |te|
            if (fabs(te) < big) {
                const double slack = fabs(te) * eps;
                if (te + slack < lo || te - slack >= hi) return false;
                const double pa = te * comp(ray.d, ax0), pb = te * comp(ray.d, ax1);
                const double ea = pa + comp(ray.o, ax0), eb = pb + comp(ray.o, ax1);
                // (2^-40 of the origin covers the one-ulp difference between this sum and the reference's fused one)
                const double tiny = 9.094947017729282e-13;  // 2^-40
                const double sa = fma(fabs(pa), eps, fabs(comp(ray.o, ax0)) * tiny), sb = fma(fabs(pb), eps, fabs(comp(ray.o, ax1)) * tiny);
                if (ea + sa < a0 || ea - sa > a1 || eb + sb < b0 || eb - sb > b1) return false;
            }
        }
    }
#endif
    t1 = (k - comp(ray.o, ax2)) / dk;
    if (!in_range(t1, lo, hi)) return false;
    double a = fma(t1, comp(ray.d, ax0), comp(ray.o, ax0));
    if (a < a0 || a > a1) return false;
    double b = fma(t1, comp(ray.d, ax1), comp(ray.o, ax1));
    if (b < b0 || b > b1) return false;
    return true;
}

// box.rs:125-150, the distance only: the nearer of the first two faces hit (order back, front, left,
// right, down, up; see hit_box)
__device__ __forceinline__ bool box_t(const double* p, const Ray& ray, double lo, double hi, double& t) {
    const double mnx = p[0], mny = p[1], mnz = p[2], mxx = p[3], mxy = p[4], mxz = p[5];
    double first = 0.0, second = 0.0, cur;
    int count = 0;
#define RS_BOX_FACE_T(kind, k, a0, a1, b0, b1)                                  \
    if (count < 2 && rect_t(kind, k, a0, a1, b0, b1, ray, lo, hi, cur)) {       \
        if (count == 0) first = cur; else second = cur;                         \
        ++count;                                                                \
    }
    RS_BOX_FACE_T(RS_RECT_XY, mnz, mnx, mxx, mny, mxy)
    RS_BOX_FACE_T(RS_RECT_XY, mxz, mnx, mxx, mny, mxy)
    RS_BOX_FACE_T(RS_RECT_YZ, mnx, mny, mxy, mnz, mxz)
    RS_BOX_FACE_T(RS_RECT_YZ, mxx, mny, mxy, mnz, mxz)
    RS_BOX_FACE_T(RS_RECT_XZ, mny, mnx, mxx, mnz, mxz)
    RS_BOX_FACE_T(RS_RECT_XZ, mxy, mnx, mxx, mnz, mxz)
#undef RS_BOX_FACE_T
    if (count == 0) return false;
    t = (count == 1 || first < second) ? first : second;
    return true;
}

// HitRecord::new for a rect hit at distance t1 (rect.rs:102-120, hit.rs:32-53)
__device__ __forceinline__ void rect_record(int kind, double a0, double a1, double b0, double b1, int material, const Ray& ray, double t1, HitRec& h) {
    int ax0, ax1, ax2;
    rect_axes(kind, ax0, ax1, ax2);
    h.point = ray_at(ray, t1);
    h.outside = comp(ray.d, ax2) < 0.0;  // dot(d, +axis) < 0
    V3 n = v3(0.0, 0.0, 0.0);
    set_comp(n, ax2, h.outside ? 1.0 : -1.0);
    h.normal = n;
    h.material = material;
    h.t1 = t1;
    h.t2 = RS_F64_MAX;
    h.u = (comp(h.point, ax0) - a0) / (a1 - a0);
    h.v = (comp(h.point, ax1) - b0) / (b1 - b0);
}

template <bool FULL = true>
__device__ __forceinline__ bool hit_rect(int kind, double k, double a0, double a1, double b0, double b1, int material,
                                         const Ray& ray, double lo, double hi, HitRec& h) {
    double t1;
    if (!rect_t(kind, k, a0, a1, b0, b1, ray, lo, hi, t1)) return false;
    if (FULL) rect_record(kind, a0, a1, b0, b1, material, ray, t1, h);
    else lean_record(ray, t1, RS_F64_MAX, material, h);
    return true;
}

// box.rs:125-150: six faces in the order back, front, left, right, down, up; the first two that
// are hit count.  One record is built, for the face that is returned.
template <bool FULL = true>
static __device__ __noinline__ bool hit_box(const rs_node& n, const Ray& ray, double lo, double hi, HitRec& h) {
    const double mnx = n.p[0], mny = n.p[1], mnz = n.p[2], mxx = n.p[3], mxy = n.p[4], mxz = n.p[5];
    double t_first = 0.0, t_second = 0.0, cur;
    int f_first = 0, f_second = 0, count = 0;
#define RS_BOX_FACE(face, kind, k, a0, a1, b0, b1)                              \
    if (count < 2 && rect_t(kind, k, a0, a1, b0, b1, ray, lo, hi, cur)) {       \
        if (count == 0) { t_first = cur; f_first = face; }                      \
        else { t_second = cur; f_second = face; }                               \
        ++count;                                                                \
    }
    RS_BOX_FACE(0, RS_RECT_XY, mnz, mnx, mxx, mny, mxy)
    RS_BOX_FACE(1, RS_RECT_XY, mxz, mnx, mxx, mny, mxy)
    RS_BOX_FACE(2, RS_RECT_YZ, mnx, mny, mxy, mnz, mxz)
    RS_BOX_FACE(3, RS_RECT_YZ, mxx, mny, mxy, mnz, mxz)
    RS_BOX_FACE(4, RS_RECT_XZ, mny, mnx, mxx, mnz, mxz)
    RS_BOX_FACE(5, RS_RECT_XZ, mxy, mnx, mxx, mnz, mxz)
#undef RS_BOX_FACE
    if (count == 0) return false;
    // two hits: the nearer one with HitRecord::with_normal -> outside = true, t2 = the other face
    const bool take_first = count == 1 || t_first < t_second;
    const int face = take_first ? f_first : f_second;
    const double t1 = take_first ? t_first : t_second;
    if (!FULL) {
        lean_record(ray, t1, count == 2 ? (take_first ? t_second : t_first) : RS_F64_MAX, n.material, h);
        return true;
    }
    if (face < 2) rect_record(RS_RECT_XY, mnx, mxx, mny, mxy, n.material, ray, t1, h);
    else if (face < 4) rect_record(RS_RECT_YZ, mny, mxy, mnz, mxz, n.material, ray, t1, h);
    else rect_record(RS_RECT_XZ, mnx, mxx, mnz, mxz, n.material, ray, t1, h);
    if (count == 2) {
        h.t2 = take_first ? t_second : t_first;
        h.outside = true;
    }
    return true;
}

// quadric.rs:67-100
__device__ __forceinline__ V3 quadric_normal(const double* q, V3 p) {
    double x = 2.0 * q[0] * p.x + q[1] * p.y + q[2] * p.z + q[3];
    double y = q[1] * p.x + 2.0 * q[4] * p.y + q[5] * p.z + q[6];
    double z = q[2] * p.x + q[5] * p.y + 2.0 * q[7] * p.z + q[8];
    V3 r = v3(x, y, z);
    double len = length(r);
    if (len == 0.0) return v3(1.0, 0.0, 0.0);
    return vdiv(r, len);
}

// quadric.rs:112-165, the roots only (what the closest-hit search of extend needs for a quadric that is a
// world object itself, bare or under one TfFacade)
__device__ __forceinline__ bool quadric_t(const double* q, const Ray& ray, double lo, double hi, double& t1) {
    const double qa = q[0], qb = q[1], qc = q[2], qd = q[3], qe = q[4], qf = q[5], qg = q[6], qh = q[7], qi = q[8], qj = q[9];
    const double xo = ray.o.x, yo = ray.o.y, zo = ray.o.z, xd = ray.d.x, yd = ray.d.y, zd = ray.d.z;
    double a = xd * (qa * xd + qb * yd + qc * zd) + yd * (qe * yd + qf * zd) + zd * qh * zd;
    double b = xd * (qa * xo + 0.5 * (qb * yo + qc * zo + qd)) + yd * (qe * yo + 0.5 * (qb * xo + qf * zo + qg)) +
               zd * (qh * zo + 0.5 * (qc * xo + qf * yo + qi));
    double c = xo * (qa * xo + qb * yo + qc * zo + qd) + yo * (qe * yo + qf * zo + qg) + zo * (qh * zo + qi) + qj;
    if (a == 0.0) {
        if (b == 0.0) return false;
        t1 = -0.5 * c / b;
        return in_range(t1, lo, hi);
    }
    double d = b * b - a * c;
    if (d <= 0.0) return false;
    double dr = sqrt(d);
    t1 = (-b - dr) / a;
    if (in_range(t1, lo, hi)) return true;
    t1 = (-b + dr) / a;
    return in_range(t1, lo, hi);
}

// quadric.rs:112-182
template <bool FULL = true>
static __device__ __noinline__ bool hit_quadric(const rs_node& n, const Ray& ray, double lo, double hi, HitRec& h) {
    const double qa = n.p[0], qb = n.p[1], qc = n.p[2], qd = n.p[3], qe = n.p[4], qf = n.p[5], qg = n.p[6], qh = n.p[7], qi = n.p[8], qj = n.p[9];
    const double xo = ray.o.x, yo = ray.o.y, zo = ray.o.z, xd = ray.d.x, yd = ray.d.y, zd = ray.d.z;
    double a = xd * (qa * xd + qb * yd + qc * zd) + yd * (qe * yd + qf * zd) + zd * qh * zd;
    double b = xd * (qa * xo + 0.5 * (qb * yo + qc * zo + qd)) + yd * (qe * yo + 0.5 * (qb * xo + qf * zo + qg)) +
               zd * (qh * zo + 0.5 * (qc * xo + qf * yo + qi));
    double c = xo * (qa * xo + qb * yo + qc * zo + qd) + yo * (qe * yo + qf * zo + qg) + zo * (qh * zo + qi) + qj;
    double t1, t2;
    if (a == 0.0) {
        if (b == 0.0) return false;
        t1 = -0.5 * c / b;
        t2 = RS_F64_MAX;
        if (!in_range(t1, lo, hi)) return false;
    } else {
        double d = b * b - a * c;
        if (d <= 0.0) return false;
        double dr = sqrt(d);
        double ta = (-b - dr) / a, tb = (-b + dr) / a;
        if (in_range(ta, lo, hi)) { t1 = ta; t2 = tb; }
        else if (in_range(tb, lo, hi)) { t1 = tb; t2 = RS_F64_MAX; }
        else return false;
    }
    if (!FULL) {
        lean_record(ray, t1, t2, n.material, h);
        return true;
    }
    h.point = ray_at(ray, t1);
    V3 nn = quadric_normal(n.p, h.point);
    h.outside = dot(ray.d, nn) < 0.0;
    h.normal = h.outside ? nn : -nn;
    h.material = n.material;
    h.t1 = t1;
    h.t2 = t2;
    h.u = 0.0;
    h.v = 0.0;
    return true;
}

// One triangle: p0.xyz, (a,b,c) = p0 - p1, (d,e,f) = p0 - p2 (triangle_mesh.rs:51-56) and, in the word after them,
// the world object the triangle is (its root id).  Default: 80-byte records read as five 128-bit ld.global.nc.
// RS_TUNE_TRI256=1: 96-byte records, 32-byte aligned, read as three 256-bit loads (sm_100 LDG.E.256) — MEASURED
// 0.9 % SLOWER on the 1 M-triangle mesh (profiles/r02b_ab_tri256.txt: 1267-1276 vs 1283-1284 M samples/s; the node
// loads gain from 256-bit requests, the triangle records lose more to their 20 % larger footprint), kept as an option.
#ifndef RS_TUNE_TRI256
#define RS_TUNE_TRI256 0
#endif
#define RS_TRI_STRIDE (RS_TUNE_TRI256 ? 6 : 5)  // double2 per triangle
#define RS_TRI_DOUBLES (2 * RS_TRI_STRIDE)
struct TriGeom {
    double p0x, p0y, p0z, a, b, c, d, e, f;
    int root;
};
__device__ __forceinline__ TriGeom load_triangle(const double2* tri_geom, int tri) {
    const double2* g = tri_geom + (size_t)RS_TRI_STRIDE * (size_t)tri;
    TriGeom T;
#if RS_TUNE_TRI256
    uint32_t w[24];
#pragma unroll
    for (int k = 0; k < 3; ++k)
        asm("ld.global.nc.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
            : "=r"(w[8 * k]), "=r"(w[8 * k + 1]), "=r"(w[8 * k + 2]), "=r"(w[8 * k + 3]), "=r"(w[8 * k + 4]), "=r"(w[8 * k + 5]), "=r"(w[8 * k + 6]), "=r"(w[8 * k + 7])
            : "l"(g + 2 * k));
#define RS_TRI_D(i) __hiloint2double((int)w[2 * (i) + 1], (int)w[2 * (i)])
    T.p0x = RS_TRI_D(0); T.p0y = RS_TRI_D(1); T.p0z = RS_TRI_D(2);
    T.a = RS_TRI_D(3); T.b = RS_TRI_D(4); T.c = RS_TRI_D(5);
    T.d = RS_TRI_D(6); T.e = RS_TRI_D(7); T.f = RS_TRI_D(8);
    T.root = (int)w[18];
#undef RS_TRI_D
#else
    const double2 q0 = __ldg(g), q1 = __ldg(g + 1), q2 = __ldg(g + 2), q3 = __ldg(g + 3), q4 = __ldg(g + 4);
    T.p0x = q0.x; T.p0y = q0.y; T.p0z = q1.x;
    T.a = q1.y; T.b = q2.x; T.c = q2.y;
    T.d = q3.x; T.e = q3.y; T.f = q4.x;
    T.root = __double2loint(q4.y);
#endif
    return T;
}

// triangle_mesh.rs:85-131.  Returns t, beta, gamma.
// QUICK = true (extend; most tested triangles are missed) first rejects on the SIGNS and MAGNITUDES of the
// numerators, without a division: for finite, nonzero operands fl(n / d) < 0 <=> n, d differ in sign and
// fl(n / d) >= 1 <=> |n| >= |d| (two doubles n < d differ by at least half an ulp of 1 in ratio, and such a
// quotient never rounds up to 1), so the outcome is the reference's.  Whatever is not rejected that way runs the
// reference's own sequence (divide, then compare) unchanged.
template <bool QUICK = false>
__device__ __forceinline__ bool triangle_test(const TriGeom& T, const Ray& ray, double lo, double hi, double& t, double& beta, double& gamma) {
    const double a = T.a, b = T.b, c = T.c, d = T.d, e = T.e, f = T.f;
    const double gg = ray.d.x, hh = ray.d.y, ii = ray.d.z;
    const double j = T.p0x - ray.o.x, k = T.p0y - ray.o.y, l = T.p0z - ray.o.z;
    double eihf = e * ii - hh * f, gfdi = gg * f - d * ii, dheg = d * hh - e * gg;
    double denom = a * eihf + b * gfdi + c * dheg;
    const double nb = j * eihf + k * gfdi + l * dheg;
    const bool plain = denom != 0.0 && fabs(denom) <= RS_F64_MAX;  // (false for NaN too)
    if (QUICK && plain && nb != 0.0 && fabs(nb) <= RS_F64_MAX && ((nb < 0.0) != (denom < 0.0) || fabs(nb) >= fabs(denom))) return false;
    double akjb = a * k - j * b, jcal = j * c - a * l, blkc = b * l - k * c;
    const double ng = ii * akjb + hh * jcal + gg * blkc;
    if (QUICK && plain && fabs(ng) <= RS_F64_MAX && (ng == 0.0 || (ng < 0.0) != (denom < 0.0))) return false;
    beta = nb / denom;
    if (beta < 0.0 || beta >= 1.0) return false;
    gamma = ng / denom;
    if (gamma <= 0.0 || beta + gamma >= 1.0) return false;
    t = -(f * akjb + e * jcal + d * blkc) / denom;
    return t >= lo && t <= hi;  // inclusive on both ends (triangle_mesh.rs:114)
}

__device__ __forceinline__ void triangle_record(const DScene& S, int tri, const Ray& ray, double t, double beta, double gamma, HitRec& h) {
    const double* nn = S.tri_normals + 9 * (size_t)tri;
    V3 n0 = v3(nn[0], nn[1], nn[2]), n1 = v3(nn[3], nn[4], nn[5]), n2 = v3(nn[6], nn[7], nn[8]);
    h.normal = n0 * (1.0 - beta - gamma) + n1 * beta + n2 * gamma;  // not normalised, not face-forwarded
    h.point = ray_at(ray, t);
    h.material = S.tri_material[tri];
    h.t1 = t;
    h.t2 = RS_F64_MAX;
    h.u = 0.0;
    h.v = 0.0;
    h.outside = true;
}

// `Hittable::random` of the two primitives that act as lights in the shipped scenes
// sphere.rs:149-164: quarter disk of radius 1 about the centre
__device__ __forceinline__ V3 sphere_random(const rs_node& n, V3 origin, const RngCtx& rng) {
    V3 c = v3(n.p[0], n.p[1], n.p[2]);
    ONB uvw = onb_build_fast(c - origin);
    for (uint32_t k = 0;; ++k) {
        double a, b;
        rng_pair(rng, P_LIGHT, 0, k, a, b);
        V3 uv = uvw.u * a + uvw.v * b;
        if (length_squared(uv) < 1.0) return (uv + c) - origin;
    }
}
// rect.rs:141-153: xz only, returns origin - root
__device__ __forceinline__ V3 rect_random(const rs_node& n, V3 origin, const RngCtx& rng) {
    uint4 w = rng_words(rng, P_LIGHT, 0);
    V3 root = v3(n.p[1] + u01(w.x) * (n.p[2] - n.p[1]), n.p[0], n.p[3] + u01(w.y) * (n.p[4] - n.p[3]));
    return origin - root;
}

// ---------------------------------------------------------------------------------------------
// raymarching.rs: the power-8 Mandelbulb about the origin (RayMarcher)
// ---------------------------------------------------------------------------------------------
// [marcher-core-begin] (tests/test_marcher_source.py compiles this block for the host)
// ln(x) by IEEE-exact steps only (every operation individually rounded, fma where written), so
// that the host-side checker's copy of this routine returns the same bits: x = m * 2^e,
// m in [sqrt(1/2), sqrt(2)), s = (m-1)/(m+1), ln m = 2 atanh(s) as a fixed series up to s^23.
__device__ __forceinline__ double marcher_log(double x) {
    if (x != x || x < 0.0) return __longlong_as_double(0x7ff8000000000000LL);
    if (x == 0.0) return __longlong_as_double(0xfff0000000000000LL);
    if (x == __longlong_as_double(0x7ff0000000000000LL)) return x;
    long long bits = __double_as_longlong(x);
    int e = 0;
    if ((bits >> 52) == 0) {  // subnormal: scale by 2^54 (exact)
        x = __dmul_rn(x, 18014398509481984.0);
        bits = __double_as_longlong(x);
        e = -54;
    }
    e += (int)(bits >> 52) - 1023;
    double m = __longlong_as_double((bits & 0x000FFFFFFFFFFFFFLL) | 0x3FF0000000000000LL);  // [1, 2)
    if (m > 1.4142135623730951) { m = __dmul_rn(m, 0.5); e += 1; }
    const double s = __ddiv_rn(__dadd_rn(m, -1.0), __dadd_rn(m, 1.0));
    const double s2 = __dmul_rn(s, s);
    double p = 2.0 / 23.0;
    p = __fma_rn(p, s2, 2.0 / 21.0);
    p = __fma_rn(p, s2, 2.0 / 19.0);
    p = __fma_rn(p, s2, 2.0 / 17.0);
    p = __fma_rn(p, s2, 2.0 / 15.0);
    p = __fma_rn(p, s2, 2.0 / 13.0);
    p = __fma_rn(p, s2, 2.0 / 11.0);
    p = __fma_rn(p, s2, 2.0 / 9.0);
    p = __fma_rn(p, s2, 2.0 / 7.0);
    p = __fma_rn(p, s2, 2.0 / 5.0);
    p = __fma_rn(p, s2, 2.0 / 3.0);
    p = __fma_rn(p, s2, 2.0);
    const double lm = __dmul_rn(s, p);
    const double LN2_HI = 6.93147180369123816490e-01, LN2_LO = 1.90821492927058770002e-10;
    const double de = (double)e;
    return __fma_rn(de, LN2_HI, __fma_rn(de, LN2_LO, lm));
}

// raymarching.rs:202-241 in canonical form: the reference evaluates the triplex power through the
// platform libm (sqrt, atan2, powf, sin, cos); orbits near the surface are chaotic, so two libms move
// ~1 % of the hit distances by whole search-grid steps.  Here — and bit for bit in the CPU checker —
// the same map is written with IEEE-exact operations only: cos/sin of theta and phi as coordinate
// ratios, 8-fold angles by three complex squarings, r^8 and r^56 by repeated multiplication.  No
// FMA contraction (the reference has no mul_add here).  `dr` is computed from the ALREADY-raised
// r, as in the reference.
static __device__ __noinline__ bool marcher_iterate(V3 p, int iterations, double& r_out, double& dr_out) {
    double x = 0.0, y = 0.0, z = 0.0, r = 0.0, dr = 0.0;
    bool inside = true;
    for (int i = 0; i < iterations; ++i) {
        const double rho2 = __dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y));
        const double r2 = __dadd_rn(rho2, __dmul_rn(z, z));
        const double rr = __dsqrt_rn(r2), rho = __dsqrt_rn(rho2);
        double ct = 1.0, st = 0.0, cp = 1.0, sp = 0.0;  // atan2(0, +0) = 0
        if (rr != 0.0) { ct = __ddiv_rn(z, rr); st = __ddiv_rn(rho, rr); }
        if (rho != 0.0) { cp = __ddiv_rn(x, rho); sp = __ddiv_rn(y, rho); }
#pragma unroll
        for (int k = 0; k < 3; ++k) {  // angle doubling: theta, phi -> 8 theta, 8 phi
            const double c2 = __dadd_rn(__dmul_rn(ct, ct), -__dmul_rn(st, st)), s2 = __dmul_rn(__dmul_rn(2.0, ct), st);
            ct = c2; st = s2;
            const double d2 = __dadd_rn(__dmul_rn(cp, cp), -__dmul_rn(sp, sp)), t2 = __dmul_rn(__dmul_rn(2.0, cp), sp);
            cp = d2; sp = t2;
        }
        const double r4 = __dmul_rn(r2, r2);
        r = __dmul_rn(r4, r4);  // |v|^8
        const double r16 = __dmul_rn(r, r), r32 = __dmul_rn(r16, r16), r56 = __dmul_rn(__dmul_rn(r32, r16), r);
        dr = __dadd_rn(__dmul_rn(__dmul_rn(r56, 8.0), dr), 1.0);
        x = __dadd_rn(__dmul_rn(__dmul_rn(r, st), cp), p.x);
        y = __dadd_rn(__dmul_rn(__dmul_rn(r, st), sp), p.y);
        z = __dadd_rn(__dmul_rn(r, ct), p.z);
        if (__dadd_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)), __dmul_rn(z, z)) > 8.0) { inside = false; break; }
    }
    r_out = r;
    dr_out = dr;
    return inside;
}
// [marcher-core-end]
__device__ __forceinline__ bool marcher_inside(V3 p, int iterations) {  // :187-190
    double r, dr;
    return marcher_iterate(p, iterations, r, dr);
}
__device__ __forceinline__ double marcher_distance(V3 p, int iterations) {  // :193-197 (NaN when r == 0)
    double r, dr;
    marcher_iterate(p, iterations, r, dr);
    return __ddiv_rn(__dmul_rn(__dmul_rn(0.5, marcher_log(r)), r), dr);
}
// :83-96 (always 100 iterations)
__device__ __forceinline__ V3 marcher_normal(V3 pos) {
    const double d = 0.01;
    const double gx = __dadd_rn(marcher_distance(v3(__dadd_rn(pos.x, d), pos.y, pos.z), 100), -marcher_distance(v3(__dadd_rn(pos.x, -d), pos.y, pos.z), 100));
    const double gy = __dadd_rn(marcher_distance(v3(pos.x, __dadd_rn(pos.y, d), pos.z), 100), -marcher_distance(v3(pos.x, __dadd_rn(pos.y, -d), pos.z), 100));
    const double gz = __dadd_rn(marcher_distance(v3(pos.x, pos.y, __dadd_rn(pos.z, d)), 100), -marcher_distance(v3(pos.x, pos.y, __dadd_rn(pos.z, -d)), 100));
    const V3 g = v3(gx, gy, gz);
    return v3(__dmul_rn(g.x, __drcp_rn(length(g))), __dmul_rn(g.y, __drcp_rn(length(g))), __dmul_rn(g.z, __drcp_rn(length(g))));
}
// Vector steps with every operation rounded on its own.  Plain `a * b + c` compiles to PTX mul.f64 /
// add.f64 WITHOUT a rounding modifier, which ptxas is free to fuse into an FMA (and does, e.g.
// `v + direction * 0.05`); the orbit tests amplify that last-bit difference into another hit point.
__device__ __forceinline__ V3 add_rn(V3 a, V3 b) { return v3(__dadd_rn(a.x, b.x), __dadd_rn(a.y, b.y), __dadd_rn(a.z, b.z)); }
__device__ __forceinline__ V3 sub_rn(V3 a, V3 b) { return v3(__dadd_rn(a.x, -b.x), __dadd_rn(a.y, -b.y), __dadd_rn(a.z, -b.z)); }
__device__ __forceinline__ V3 mul_rn(V3 a, double s) { return v3(__dmul_rn(a.x, s), __dmul_rn(a.y, s), __dmul_rn(a.z, s)); }

// :113-165 (+ search_surface :59-76, binary_search_surface :43-57).  `hi`: the reference ignores
// limit.end here; the canonical rule (shared with the CPU checker) drops a hit at or beyond it.
template <bool FULL = true>
static __device__ __noinline__ bool hit_marcher(const rs_node& n, const Ray& ray, double lo, double hi, HitRec& h) {
    const int iterations = (int)n.p[0];
    const double len = length(ray.d);
    const V3 direction = mul_rn(ray.d, __drcp_rn(len));
    V3 current = ray.o;
    int steps = 0;
    double best = 1000000.0, est = 0.0;
    while (steps < 1000 && est < __dadd_rn(best, 1.0)) {
        ++steps;
        est = marcher_distance(current, iterations);
        if (est != est) est = 0.1;
        if (est < 1.3) {
            const V3 df = mul_rn(direction, 0.05);
            V3 v = current;
            bool found = false;
            for (int i = 0; i < 200; ++i) {
                if (marcher_inside(v, iterations)) { found = true; break; }
                v = add_rn(v, df);
            }
            if (!found) return false;
            V3 outside = sub_rn(v, df), inside = v;
            for (int depth = 8; depth > 0; --depth) {
                V3 mid = mul_rn(add_rn(outside, inside), 0.5);
                if (marcher_inside(mid, 100)) inside = mid; else outside = mid;
            }
            const double dist = length(sub_rn(outside, ray.o));
            if (!(dist > lo)) return false;
            const double t1 = __ddiv_rn(dist, len);
            if (!(t1 < hi)) return false;
            if (!FULL) { lean_record(ray, t1, t1, n.material, h); return true; }
            h.point = ray_at(ray, t1);
            V3 nn = marcher_normal(h.point);
            h.outside = dot(ray.d, nn) < 0.0;
            h.normal = h.outside ? nn : -nn;
            h.material = n.material;
            h.t1 = t1;
            h.t2 = t1;
            sphere_uv(v3(0.0, 0.0, 0.0), h.point, h.u, h.v);  // :102-111
            return true;
        }
        if (est < best) best = est;
        current = add_rn(current, mul_rn(mul_rn(direction, est), 0.05));  // (direction * est) * 0.05
    }
    return false;
}
// vec3.rs:91-96 (`RayMarcher::random`, raymarching.rs:188-190)
__device__ __forceinline__ V3 marcher_random(const RngCtx& rng) {
    uint4 w = rng_words(rng, P_LIGHT, 0);
    double a = 0.0 + u01(w.x) * (2.0 * RS_PI - 0.0);
    double z = -1.0 + u01(w.y) * (1.0 - -1.0);
    double r = sqrt(1.0 - z * z);
    return v3(r * cos(a), r * sin(a), z);
}

// ---------------------------------------------------------------------------------------------
// Expression trees
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ V3 xf_apply(const double* m, V3 p, double w) {
    return v3(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3] * w, m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7] * w,
              m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11] * w);
}

// ---------------------------------------------------------------------------------------------
// Object programs: a world object that is an expression tree (CSG, TfFacade, ConstantMedium, any nesting)
// is flattened at upload into a post-order program, evaluated by ONE loop with a small operand stack —
// no recursion, no per-depth copies of the code.  What makes this possible:
//   * a TfFacade keeps the ray parameter (tf_facade.rs:41-56: the child is hit with the transformed ray,
//     only the hit POINT is mapped back), so the transforms above a primitive compose into one 3x4
//     world -> leaf matrix per LEAF instruction, and a hit is fully described by (t1, t2, material, source);
//   * the points CSG rules ask `contains` about are ray points: p = ray.at(t) (difference.rs:68-75,
//     intersection.rs:66-83), evaluated through the composed world -> leaf matrices as well.
// The record of the winning hit is rebuilt by shade from (source instruction, flavor, t1): see object_record.
// ---------------------------------------------------------------------------------------------
enum ObjOp : int {
    OP_LEAF = 0,       // node = primitive, xf = composed world -> leaf transform (-1: none): push its hit under the current limits
    OP_JMISS = 1,      // if the top of the stack is a miss, skip `aux` instructions (second operand + combiner: a miss stays a miss)
    OP_DIFF = 2,       // node = Difference, aux = its CsgInfo: pop minus, pop plus, push the result (difference.rs:57-107)
    OP_INTER = 3,      // node = Intersection (intersection.rs:58-96)
    OP_MED_BEGIN = 4,  // ConstantMedium (constant.rs:42-80): the boundary is hit over (-inf, +inf) ...
    OP_MED_MID = 5,    // ... then again from t1 + 0.0001 (the boundary's program follows twice); aux = skip when the first search misses
    OP_MED_END = 6,    // node = the medium, xf = its space: free-flight sample, push the result
    OP_BVH_BEGIN = 7,  // node = a BVH used as an object (bvh.rs:173-192), xf = its space: AABB::hit on its bbox; on a miss push a miss and skip `aux` instructions (past its OP_BVH_END)
    OP_BVH_END = 9,    // pop right, pop left, push the closer one (right only when strictly closer: bvh.rs:181-189)
    OP_BVH_GATE = 11,  // node = a BVH node INSIDE a BVH object: AABB::hit on its bbox; on a miss skip `aux` instructions (its whole subtree), push nothing
    OP_MISS = 10,      // push a miss (a child that is None)
};
enum ObjFlavor : int {
    FLAVOR_LEAF = 0,       // HitRecord::new of the primitive at `src` (hit.rs:32-53)
    FLAVOR_SYNTHETIC = 1,  // Difference's inner surface: normal = -minus.normal(p), outside = true (difference.rs:86-103); src = the OP_DIFF
    FLAVOR_MEDIUM = 2,     // ConstantMedium: normal (1,0,0), outside = false (constant.rs:72-79); src = the OP_MED_END
};
// `contains` programs (boolean post-order): x = node | op << 28, y = xf
#define RS_C_LEAF 0
#define RS_C_AND 1      // Intersection::contains
#define RS_C_ANDNOT 2   // Difference::contains: plus && !minus
#define RS_C_FALSE 3    // rect, medium
#define RS_EVAL_STACK 9  // operands: a tree of depth RS_MAX_TREE_DEPTH needs at most depth + 1

struct ObjHit {
    double t1, t2;
    int mat;   // -1 = None
    int src;   // index of the instruction the record is rebuilt from
    int flavor;
    bool hit;
};

__device__ __forceinline__ Ray xf_ray(const double* inv, const Ray& ray) {
    Ray r;
    r.o = xf_apply(inv, ray.o, 1.0);
    r.d = xf_apply(inv, ray.d, 0.0);
    r.time = ray.time;
    return r;
}

// box.rs:125-150, distances only: t1 = the nearer of the first two faces hit (order back, front, left, right, down,
// up), t2 = the other one or f64::MAX
__device__ __forceinline__ bool box_t2(const double* p, const Ray& ray, double lo, double hi, double& t1, double& t2) {
    const double mnx = p[0], mny = p[1], mnz = p[2], mxx = p[3], mxy = p[4], mxz = p[5];
    double first = 0.0, second = 0.0, cur;
    int count = 0;
#define RS_BOX_FACE_T(kind, k, a0, a1, b0, b1)                                  \
    if (count < 2 && rect_t(kind, k, a0, a1, b0, b1, ray, lo, hi, cur)) {       \
        if (count == 0) first = cur; else second = cur;                         \
        ++count;                                                                \
    }
    RS_BOX_FACE_T(RS_RECT_XY, mnz, mnx, mxx, mny, mxy)
    RS_BOX_FACE_T(RS_RECT_XY, mxz, mnx, mxx, mny, mxy)
    RS_BOX_FACE_T(RS_RECT_YZ, mnx, mny, mxy, mnz, mxz)
    RS_BOX_FACE_T(RS_RECT_YZ, mxx, mny, mxy, mnz, mxz)
    RS_BOX_FACE_T(RS_RECT_XZ, mny, mnx, mxx, mnz, mxz)
    RS_BOX_FACE_T(RS_RECT_XZ, mxy, mnx, mxx, mnz, mxz)
#undef RS_BOX_FACE_T
    if (count == 0) return false;
    const bool take_first = count == 1 || first < second;
    t1 = take_first ? first : second;
    t2 = count == 2 ? (take_first ? second : first) : RS_F64_MAX;
    return true;
}

// quadric.rs:112-182, distances only
__device__ __forceinline__ bool quadric_t2(const double* q, const Ray& ray, double lo, double hi, double& t1, double& t2) {
    const double qa = q[0], qb = q[1], qc = q[2], qd = q[3], qe = q[4], qf = q[5], qg = q[6], qh = q[7], qi = q[8], qj = q[9];
    const double xo = ray.o.x, yo = ray.o.y, zo = ray.o.z, xd = ray.d.x, yd = ray.d.y, zd = ray.d.z;
    double a = xd * (qa * xd + qb * yd + qc * zd) + yd * (qe * yd + qf * zd) + zd * qh * zd;
    double b = xd * (qa * xo + 0.5 * (qb * yo + qc * zo + qd)) + yd * (qe * yo + 0.5 * (qb * xo + qf * zo + qg)) +
               zd * (qh * zo + 0.5 * (qc * xo + qf * yo + qi));
    double c = xo * (qa * xo + qb * yo + qc * zo + qd) + yo * (qe * yo + qf * zo + qg) + zo * (qh * zo + qi) + qj;
    if (a == 0.0) {
        if (b == 0.0) return false;
        t1 = -0.5 * c / b;
        t2 = RS_F64_MAX;
        return in_range(t1, lo, hi);
    }
    double d = b * b - a * c;
    if (d <= 0.0) return false;
    double dr = sqrt(d);
    double ta = (-b - dr) / a, tb = (-b + dr) / a;
    if (in_range(ta, lo, hi)) { t1 = ta; t2 = tb; return true; }
    if (in_range(tb, lo, hi)) { t1 = tb; t2 = RS_F64_MAX; return true; }
    return false;
}

// `Hittable::hit` of a primitive, distances only
static __device__ __noinline__ bool leaf_hit(const rs_node& n, const Ray& ray, double lo, double hi, double& t1, double& t2) {
    switch (n.type) {
        case RS_NODE_SPHERE: {
            V3 c = v3(n.p[0], n.p[1], n.p[2]);
            if (n.p[4] != 0.0 || n.p[5] != 0.0 || n.p[6] != 0.0) c = c + v3(n.p[4], n.p[5], n.p[6]) * ray.time;  // center_at
            return sphere_roots(c, n.p[3] * n.p[3], ray, lo, hi, t1, t2);
        }
        case RS_NODE_BOX: return box_t2(n.p, ray, lo, hi, t1, t2);
        case RS_NODE_RECT: t2 = RS_F64_MAX; return rect_t(n.rect_axis, n.p[0], n.p[1], n.p[2], n.p[3], n.p[4], ray, lo, hi, t1);
        case RS_NODE_QUADRIC: return quadric_t2(n.p, ray, lo, hi, t1, t2);
        case RS_NODE_RAYMARCHER: {
            HitRec h;
            if (!hit_marcher<false>(n, ray, lo, hi, h)) return false;
            t1 = h.t1; t2 = h.t2;
            return true;
        }
        default: return false;
    }
}

// `Hittable::contains` of a primitive
__device__ __forceinline__ bool leaf_contains(const rs_node& n, V3 p) {
    switch (n.type) {
        case RS_NODE_SPHERE: {  // sphere.rs:111-115
            V3 r = v3(n.p[0], n.p[1], n.p[2]) - p;
            return length_squared(r) < n.p[3] * n.p[3];
        }
        case RS_NODE_BOX:  // box.rs:152-157
            return p.x >= n.p[0] && p.x <= n.p[3] && p.y >= n.p[1] && p.y <= n.p[4] && p.z >= n.p[2] && p.z <= n.p[5];
        case RS_NODE_QUADRIC: {  // quadric.rs:184-189
            const double* q = n.p;
            return (p.x * (q[0] * p.x + q[1] * p.y + q[3]) + p.y * (q[4] * p.y + q[5] * p.z + q[6]) + p.z * (q[7] * p.z + q[2] * p.x + q[8]) + q[9]) <= 0.0;
        }
        case RS_NODE_RAYMARCHER: return marcher_inside(p, (int)n.p[0]);  // raymarching.rs:167-170
        default: return false;  // rect
    }
}

// `contains` of a subtree for the world-space point p
__device__ __forceinline__ bool contains_prog(const rs_node* nodes, const XformPre* xf, const int2* cprog, int off, int len, V3 p) {
    uint32_t bits = 0;  // boolean stack, top = bit 0
    for (int k = off; k < off + len; ++k) {
        const int2 c = __ldg(cprog + k);
        const int op = (int)((uint32_t)c.x >> 28);
        if (op == RS_C_LEAF) {
            const V3 q = c.y >= 0 ? xf_apply(xf[c.y].inv, p, 1.0) : p;
            bits = (bits << 1) | (leaf_contains(nodes[c.x & 0x0fffffff], q) ? 1u : 0u);
        } else if (op == RS_C_FALSE) {
            bits <<= 1;
        } else {
            const uint32_t b = bits & 1u, a = (bits >> 1) & 1u;
            bits = ((bits >> 2) << 1) | (op == RS_C_AND ? (a & b) : (a & (b ^ 1u)));
        }
    }
    return bits & 1u;
}

// `Hittable::hit` of a world object through its program.  Limits [lo, hi) as for any hit call.
static __device__ __noinline__ bool eval_object(const rs_node* nodes, const XformPre* xf, const ObjPrograms P, int first, int count, const Ray& ray,
                                                double lo, double hi, const RngCtx& rng, ObjHit& out) {
    ObjHit stack[RS_EVAL_STACK];
    int sp = 0;
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    double lo0 = lo, hi0 = hi, med_t1 = 0.0;
    int cur_xi = -1;  // the space `r` is in: consecutive operands usually share it (the operands of one CSG node)
    Ray r = ray;
    for (int pc = first; pc < first + count; ++pc) {
        const int4 raw = __ldg(reinterpret_cast<const int4*>(P.prog + pc));
        const int op = raw.x, ni = raw.y, xi = raw.z, aux = raw.w;
        if (op == OP_LEAF) {
            const rs_node& n = nodes[ni];
            if (xi != cur_xi) {
                r = xi >= 0 ? xf_ray(xf[xi].inv, ray) : ray;
                cur_xi = xi;
            }
            ObjHit e;
            e.hit = leaf_hit(n, r, lo, hi, e.t1, e.t2);
            e.mat = n.material;
            e.src = pc;
            e.flavor = FLAVOR_LEAF;
            if (RS_IN_BOUNDS(sp < RS_EVAL_STACK)) stack[sp++] = e;  // (rs_upload_scene rejects programs that need more operands)
        } else if (op == OP_JMISS) {
            if (!stack[sp - 1].hit) pc += aux;
        } else if (op == OP_DIFF) {  // difference.rs:57-107
            const ObjHit hm = stack[--sp];
            const ObjHit hp = stack[sp - 1];
            const rs_node& n = nodes[ni];
            ObjHit res = hp;
            if (!hp.hit) {
                res.hit = false;
            } else if (!hm.hit) {
                // only the plus object is hit: its record as it is
            } else if (hp.t1 < hm.t1) {
                const CsgInfo ci = P.csg[aux];
                if (contains_prog(nodes, xf, P.cprog, ci.c1_off, ci.c1_len, ray_at(ray, hp.t1))) res.hit = false;
                else if (res.mat < 0) res.mat = n.material;
            } else if (hm.t2 < hp.t1) {
                if (res.mat < 0) res.mat = n.material;
            } else if (hm.t2 < hp.t2) {
                // the ray leaves the minus object inside the plus object: the cavity's surface
                const rs_node& m = nodes[n.child1];
                const bool primitive = m.type == RS_NODE_SPHERE || m.type == RS_NODE_BOX || m.type == RS_NODE_RECT || m.type == RS_NODE_QUADRIC ||
                                       m.type == RS_NODE_RAYMARCHER;  // only these implement normal() / material() (hit.rs:82-93)
                res.t1 = hm.t2;
                res.t2 = hp.t2;
                res.mat = (primitive && m.material >= 0) ? m.material : n.material;
                res.src = pc;
                res.flavor = FLAVOR_SYNTHETIC;
            } else {
                res.hit = false;
            }
            stack[sp - 1] = res;
        } else if (op == OP_INTER) {  // intersection.rs:58-96
            const ObjHit h2 = stack[--sp];
            const ObjHit h1 = stack[sp - 1];
            const rs_node& n = nodes[ni];
            ObjHit res = h1;
            if (!h1.hit || !h2.hit) {
                res.hit = false;
            } else {
                const CsgInfo ci = P.csg[aux];
                const bool first_is_1 = h1.t1 < h2.t1;
                const ObjHit& near = first_is_1 ? h1 : h2;
                const ObjHit& far = first_is_1 ? h2 : h1;
                // contains(the other object, the hit point), nearest hit first
                const int far_off = first_is_1 ? ci.c1_off : ci.c0_off, far_len = first_is_1 ? ci.c1_len : ci.c0_len;
                const int near_off = first_is_1 ? ci.c0_off : ci.c1_off, near_len = first_is_1 ? ci.c0_len : ci.c1_len;
                if (contains_prog(nodes, xf, P.cprog, far_off, far_len, ray_at(ray, near.t1))) res = near;
                else if (contains_prog(nodes, xf, P.cprog, near_off, near_len, ray_at(ray, far.t1))) res = far;
                else res.hit = false;
                if (res.hit && res.mat < 0) res.mat = n.material;
            }
            stack[sp - 1] = res;
        } else if (op == OP_BVH_BEGIN) {  // bvh.rs:174-177
            const rs_node& n = nodes[ni];
            if (xi != cur_xi) {
                r = xi >= 0 ? xf_ray(xf[xi].inv, ray) : ray;
                cur_xi = xi;
            }
            const V3 inv = v3(1.0 / r.d.x, 1.0 / r.d.y, 1.0 / r.d.z);
            if (n.p[6] == 0.0 || !aabb_hit_f64(n.p, r, inv, lo, hi)) {
                ObjHit e;
                e.hit = false; e.t1 = 0.0; e.t2 = 0.0; e.mat = -1; e.src = pc; e.flavor = FLAVOR_LEAF;
                if (RS_IN_BOUNDS(sp < RS_EVAL_STACK)) stack[sp++] = e;
                pc += aux;
            }
        } else if (op == OP_BVH_GATE) {  // bvh.rs:174-177 for the nodes below the top of a BVH object (see ProgramBuilder::emit)
            const rs_node& n = nodes[ni];
            if (xi != cur_xi) {
                r = xi >= 0 ? xf_ray(xf[xi].inv, ray) : ray;
                cur_xi = xi;
            }
            const V3 inv = v3(1.0 / r.d.x, 1.0 / r.d.y, 1.0 / r.d.z);
            if (n.p[6] == 0.0 || !aabb_hit_f64(n.p, r, inv, lo, hi)) pc += aux;
        } else if (op == OP_BVH_END) {
            // bvh.rs:181-189 searches the right child over start..left.t1 and prefers its hit.  Here both children are
            // searched over the node's own limits and the right one wins when strictly closer — the same winner, and (as
            // for the objects of the world) what a Box member reports does not depend on its sibling
            const ObjHit hr = stack[--sp];
            const ObjHit hl = stack[sp - 1];
            if (hr.hit && (!hl.hit || hr.t1 < hl.t1)) stack[sp - 1] = hr;
        } else if (op == OP_MISS) {
            ObjHit e;
            e.hit = false; e.t1 = 0.0; e.t2 = 0.0; e.mat = -1; e.src = pc; e.flavor = FLAVOR_LEAF;
            if (RS_IN_BOUNDS(sp < RS_EVAL_STACK)) stack[sp++] = e;
        } else if (op == OP_MED_BEGIN) {  // constant.rs:45
            lo0 = lo; hi0 = hi;
            lo = -inf; hi = inf;
        } else if (op == OP_MED_MID) {  // constant.rs:46-49
            const ObjHit r1 = stack[sp - 1];
            if (!r1.hit) {
                lo = lo0; hi = hi0;
                pc += aux;  // past the OP_MED_END: the miss stays as the result
            } else {
                --sp;
                med_t1 = r1.t1;
                lo = r1.t1 + 0.0001; hi = inf;
            }
        } else {  // OP_MED_END: constant.rs:50-79
            const ObjHit r2 = stack[--sp];
            const rs_node& n = nodes[ni];
            lo = lo0; hi = hi0;
            ObjHit res;
            res.hit = false;
            res.t1 = 0.0; res.t2 = 0.0; res.mat = n.material; res.src = pc; res.flavor = FLAVOR_MEDIUM;
            if (r2.hit) {
                double a = med_t1, b = r2.t1;
                if (a < lo) a = lo;
                if (b > hi) b = hi;
                if (a < b) {
                    if (a < 0.0) a = 0.0;
                    if (xi != cur_xi) {
                        r = xi >= 0 ? xf_ray(xf[xi].inv, ray) : ray;
                        cur_xi = xi;
                    }
                    const double length_per_unit = length(r.d);
                    const double distance_inside = (b - a) * length_per_unit;
                    const uint4 w = rng_words(rng, P_MEDIUM, (uint32_t)ni);
                    const double hit_distance = (-1.0 / n.p[0]) * log(u01(w.x));
                    if (!(hit_distance > distance_inside)) {
                        res.hit = true;
                        res.t1 = a + hit_distance / length_per_unit;
                        res.t2 = res.t1;
                    }
                }
            }
            if (RS_IN_BOUNDS(sp < RS_EVAL_STACK)) stack[sp++] = res;
        }
    }
    out = stack[0];
    return sp > 0 && out.hit;
}

// `Hittable::normal` of a primitive (hit.rs:82-87; only primitives implement it)
__device__ __forceinline__ V3 leaf_normal(const rs_node& m, V3 p) {
    V3 nn = v3(0.0, 0.0, 0.0);
    if (m.type == RS_NODE_SPHERE) nn = vdiv(p - v3(m.p[0], m.p[1], m.p[2]), m.p[3]);
    else if (m.type == RS_NODE_BOX) nn = v3(0.0, 1.0, 0.0);  // box.rs:107-109
    else if (m.type == RS_NODE_RECT) { int a, b, c; rect_axes(m.rect_axis, a, b, c); set_comp(nn, c, 1.0); }
    else if (m.type == RS_NODE_QUADRIC) nn = quadric_normal(m.p, p);
    else if (m.type == RS_NODE_RAYMARCHER) nn = marcher_normal(p);
    return nn;
}

// The full HitRecord of the hit (src, flavor, t1) that eval_object reported for this ray.
static __device__ __noinline__ void object_record(const rs_node* nodes, const XformPre* xf, const ObjInstr* prog, int src, int flavor, double t1,
                                                  const Ray& ray, bool want_uv, HitRec& h) {
    const int4 raw = __ldg(reinterpret_cast<const int4*>(prog + src));
    const rs_node& n = nodes[raw.y];
    const XformPre* x = raw.z >= 0 ? xf + raw.z : nullptr;
    const Ray r = x ? xf_ray(x->inv, ray) : ray;
    h.t1 = t1; h.t2 = t1; h.u = 0.0; h.v = 0.0; h.material = n.material;
    if (flavor == FLAVOR_LEAF) {
        // the primitive, in its own space, over the root limits [tmin, +inf) it was searched with
        const double inf = __longlong_as_double(0x7ff0000000000000LL);
        bool ok;
        switch (n.type) {
            case RS_NODE_SPHERE: {
                h.point = ray_at(r, t1);
                V3 nn = (h.point - v3(n.p[0], n.p[1], n.p[2])) * n.p[7];  // static centre (sphere.rs:56-58); p[7] = 1 / r
                h.outside = dot(r.d, nn) < 0.0;
                h.normal = h.outside ? nn : -nn;
                if (want_uv) sphere_uv(v3(n.p[0], n.p[1], n.p[2]), h.point, h.u, h.v);
                ok = true;
                break;
            }
            case RS_NODE_BOX: ok = hit_box<true>(n, r, 0.0001, inf, h); break;
            case RS_NODE_RECT: ok = hit_rect<true>(n.rect_axis, n.p[0], n.p[1], n.p[2], n.p[3], n.p[4], n.material, r, 0.0001, inf, h); break;
            case RS_NODE_QUADRIC: ok = hit_quadric<true>(n, r, 0.0001, inf, h); break;
            default: {  // RayMarcher: the record follows from the distance, without marching again
                h.point = ray_at(r, t1);
                V3 nn = marcher_normal(h.point);
                h.outside = dot(r.d, nn) < 0.0;
                h.normal = h.outside ? nn : -nn;
                if (want_uv) sphere_uv(v3(0.0, 0.0, 0.0), h.point, h.u, h.v);
                ok = true;
                break;
            }
        }
        if (!ok || h.t1 != t1) {
            // the primitive was searched under other limits (inside a ConstantMedium's boundary search never ends
            // here; a CSG operand always uses the root limits): keep the distance, generic record
            h.point = ray_at(r, t1);
            V3 nn = leaf_normal(n, h.point);
            h.outside = dot(r.d, nn) < 0.0;
            h.normal = h.outside ? nn : -nn;
            h.t1 = t1; h.t2 = t1;
        }
    } else if (flavor == FLAVOR_SYNTHETIC) {  // difference.rs:86-103; n = the Difference
        const rs_node& m = nodes[n.child1];
        h.point = ray_at(r, t1);
        h.normal = -leaf_normal(m, h.point);
        h.outside = true;
    } else {  // FLAVOR_MEDIUM (constant.rs:72-79)
        h.point = ray_at(r, t1);
        h.normal = v3(1.0, 0.0, 0.0);
        h.outside = false;
    }
    if (x) h.point = xf_apply(x->fwd, h.point, 1.0);  // tf_facade.rs:52-55: only the point goes back to world space
}

// `Hittable::random` of a light (HittableList::random picks it, list.rs:49-52)
static __device__ __noinline__ V3 light_random(const rs_node* nodes, const XformPre* xf, int ni, V3 origin, const RngCtx& rng) {
    for (int guard = 0; guard <= RS_MAX_TREE_DEPTH; ++guard) {
        const rs_node& n = nodes[ni];
        switch (n.type) {
            case RS_NODE_SPHERE: return sphere_random(n, origin, rng);
            case RS_NODE_RECT: return rect_random(n, origin, rng);
            case RS_NODE_QUADRIC: return -origin;  // quadric.rs:202-205
            case RS_NODE_RAYMARCHER: return marcher_random(rng);
            case RS_NODE_DIFFERENCE:
            case RS_NODE_INTERSECTION: ni = n.child0; break;  // difference.rs:118-120, intersection.rs:120-122
            case RS_NODE_TRANSFORM:  // tf_facade.rs:89-93: the origin goes to object space, the result is not mapped back
                origin = xf_apply(xf[n.first_xform].inv, origin, 1.0);
                ni = n.child0;
                break;
            default: return v3(1.0, 0.0, 0.0);  // box, medium
        }
    }
    return v3(1.0, 0.0, 0.0);
}

// ---------------------------------------------------------------------------------------------
// Textures
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t f64_as_u32_sat(double d) {  // Rust `as u32`
    if (!(d > 0.0)) return 0u;
    if (d >= 4294967295.0) return 4294967295u;
    return (uint32_t)d;
}

static __device__ __noinline__ double perlin_noise(const DScene& S, const rs_texture& t, V3 p) {  // noise.rs:100-137,156-189
    const int pc = t.point_count;
    const long long mask = pc - 1;
    const int* perm = S.perlin_perm + t.perm_offset;
    const double* val = S.perlin_values + t.value_offset;
    if (t.perlin_smooth == 0) {
        int i = (int)(((long long)(4.0 * p.x)) & mask), j = (int)(((long long)(4.0 * p.y)) & mask), k = (int)(((long long)(4.0 * p.z)) & mask);
        int idx = perm[i] ^ perm[pc + j] ^ perm[2 * pc + k];
        return t.perlin_vector ? val[3 * idx] : val[idx];
    }
    long long i = (long long)floor(p.x), j = (long long)floor(p.y), k = (long long)floor(p.z);
    double u = p.x - (double)i, v = p.y - (double)j, w = p.z - (double)k;
    double uu = u, vv = v, ww = w;
    if (t.perlin_smooth == 2) {
        uu = u * u * (3.0 - 2.0 * u);
        vv = v * v * (3.0 - 2.0 * v);
        ww = w * w * (3.0 - 2.0 * w);
    }
    double sum_i = 0.0;
    for (int di = 0; di < 2; ++di) {
        double sum_j = 0.0;
        for (int dj = 0; dj < 2; ++dj) {
            double sum_k = 0.0;
            for (int dk = 0; dk < 2; ++dk) {
                int xi = (int)((i + di) & mask), yi = (int)((j + dj) & mask), zi = (int)((k + dk) & mask);
                int idx = perm[xi] ^ perm[pc + yi] ^ perm[2 * pc + zi];
                double lat;
                if (t.perlin_vector) lat = dot(v3(val[3 * idx], val[3 * idx + 1], val[3 * idx + 2]), v3(u - (double)di, v - (double)dj, w - (double)dk));
                else lat = val[idx];
                sum_k += fma((double)di, uu, (double)(1 - di) * (1.0 - uu)) * fma((double)dj, vv, (double)(1 - dj) * (1.0 - vv)) *
                         fma((double)dk, ww, (double)(1 - dk) * (1.0 - ww)) * lat;
            }
            sum_j += sum_k;
        }
        sum_i += sum_j;
    }
    return sum_i;
}

static __device__ __noinline__ float4 texture_color(const DScene& S, int ti, double u, double v, V3 p) {
    // Checker nests arbitrary textures (checker.rs: generic T1, T2): iterate instead of recursing
    for (int guard = 0; guard < 16; ++guard) {
        const rs_texture& t = S.textures[ti];
        if (t.type == RS_TEX_SOLID) return make_float4(t.color[0], t.color[1], t.color[2], t.color[3]);
        if (t.type == RS_TEX_CHECKER) {  // checker.rs:22-29: sign of sin(s x) sin(s y) sin(s z)
            // only the sign is used: sin(a) < 0 <=> floor(a / pi) is odd, and the product is
            // negative iff an odd number of factors is; a factor of exactly 0 makes it `even`
            const double ax = t.scale * p.x, ay = t.scale * p.y, az = t.scale * p.z;
            const double inv_pi = 0.31830988618379067154;
            const long long par = __double2ll_rd(ax * inv_pi) ^ __double2ll_rd(ay * inv_pi) ^ __double2ll_rd(az * inv_pi);
            const bool negative = (par & 1) && ax != 0.0 && ay != 0.0 && az != 0.0;
            ti = negative ? t.odd : t.even;
            continue;
        }
        if (t.type == RS_TEX_IMAGE) {  // image.rs:36-49
            double vv = 1.0 - v;
            uint32_t px = f64_as_u32_sat(u * (double)t.width), py = f64_as_u32_sat(vv * (double)t.height);
            if (px >= (uint32_t)t.width) px = (uint32_t)t.width - 1;
            if (py >= (uint32_t)t.height) py = (uint32_t)t.height - 1;
            const unsigned char* c = S.texels + t.texel_offset + 4 * ((size_t)py * t.width + px);
            return make_float4((float)c[0] / 255.0f, (float)c[1] / 255.0f, (float)c[2] / 255.0f, 1.0f);
        }
        // RS_TEX_PERLIN (noise.rs:192-210)
        double f;
        if (t.perlin_kind == 0) {
            f = perlin_noise(S, t, p * t.scale);
            if (t.perlin_vector) f = 0.5 * (f + 1.0);
        } else {
            V3 q = p;
            double weight = 1.0, sum = 0.0;
            for (int i = 0; i < t.perlin_depth; ++i) {
                sum += weight * perlin_noise(S, t, q);
                weight *= 0.5;
                q = q * 2.0;
            }
            f = fabs(sum);
            if (t.perlin_kind == 2) f = (sin(fma(t.scale, p.z, 10.0 * f)) + 1.0) * 0.5;
        }
        float ff = (float)f;  // Color * f64 (color.rs:99-108): each channel times `rhs as f32`
        return make_float4(1.0f * ff, 1.0f * ff, 1.0f * ff, 1.0f);
    }
    return make_float4(0.f, 0.f, 0.f, 1.f);
}

// Texture::color of a material through its flattened record
__device__ __forceinline__ float4 material_color(const DScene& S, const MatHot& m, double u, double v, V3 p) {
    const int mode = RS_MATHOT_TEX_MODE(m);
    if (mode == 0) return m.c0;
    if (mode == 1) {  // checker.rs:22-29, see texture_color
        const double ax = m.scale * p.x, ay = m.scale * p.y, az = m.scale * p.z;
        const double inv_pi = 0.31830988618379067154;
        const long long par = __double2ll_rd(ax * inv_pi) ^ __double2ll_rd(ay * inv_pi) ^ __double2ll_rd(az * inv_pi);
        const bool negative = (par & 1) && ax != 0.0 && ay != 0.0 && az != 0.0;
        return negative ? m.c1 : m.c0;
    }
    return texture_color(S, m.texture, u, v, p);
}

__device__ __forceinline__ bool texture_needs_uv(const DScene& S, int ti) {
    // only image lookups read (u, v); a checker may nest one: resolved at upload into the
    // `reserved` word of the device copy
    return S.textures[ti].reserved != 0;
}

// color.rs:50-58 + world background closures
__device__ __forceinline__ float4 background(const DScene& S, V3 dir) {
    double slide = (dir.y + 1.0) * 0.5;
    float a = (float)fmin(fmax(slide, 0.0), 1.0);
    float b = 1.0f - a;
    // two products and a sum, each rounded (Rust never fuses; left to the compiler, which product of `x*b + y*a` ends
    // up inside the FMA depends on the kernel the function is inlined into)
    return make_float4(__fadd_rn(__fmul_rn(S.bg0.x, b), __fmul_rn(S.bg1.x, a)), __fadd_rn(__fmul_rn(S.bg0.y, b), __fmul_rn(S.bg1.y, a)),
                       __fadd_rn(__fmul_rn(S.bg0.z, b), __fmul_rn(S.bg1.z, a)), 1.0f);
}

}  // namespace rsd
