// ORACLE — TEST INFRASTRUCTURE ONLY.  Never linked into, imported by, or called from the
// product path (raysnail_b200/).  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load liboracle.so.
//
// CPU restatement (C++17, FP64 geometry / f32 colour, recursion and dynamic dispatch like the
// original) of the reference's render hot path.  Every function cites the reference file:line
// it follows (paths relative to /root/reference/src).
//
// PARITY PINNING STATUS: "parity unpinned" by the reference's own tests — the reference has a
// single, stale unit test (hittable/transform/transform.rs:187-206: rotate_by_y_axis(pi/2)
// maps (0,0,1) to (1,0,0) +-1e-10 and round-trips), no golden vectors, and it cannot be built
// here (no Rust toolchain).  That one known answer is checked in
// tests/test_oracle_kat.py together with self-made known-answer tests.
//
// Deliberate, documented differences from the reference (all forced by determinism):
//  * RNG: the reference draws from XorShiftRng seeded by thread_rng plus hidden thread_rng()
//    calls (prelude/random.rs:109-145, material/dielectric.rs:72, material/mixed_material.rs:44,
//    hittable/medium/constant.rs:62).  Here every draw site reads a counter-based
//    Philox4x32-10 stream keyed by (seed) and indexed by (pixel, sample, depth, purpose, call),
//    the same definition the CUDA backend uses (DESIGN.md "RNG streams").
//  * BVH split axis: `Random::range(0..2)` (collection/bvh.rs:91) becomes a hash of the range.
//  * Every world object is its own single-object BVH leaf (the `count == 1` arm,
//    collection/bvh.rs:62-69), whose bbox test uses [limit.start, +inf) so that the result
//    does not depend on traversal order when a reference bbox does not contain its object
//    (quadric +-100 box, Intersection::bbox z quirk).
//  * Exact ties in t between two world objects: the reference keeps whichever its (randomly
//    built) BVH visits first (collection/bvh.rs:184-191); here the lower object id wins.
//  * Limit seen by a world object: the reference hands every BVH leaf the limit as shrunk by the hits found
//    before it in ITS (random) visiting order (collection/bvh.rs:180-187).  For a Box the far face may then
//    fall outside the limit (one-face record with the real `outside` instead of the two-face record with
//    `outside = true`, box.rs:131-146), and a Difference / Intersection can return another surface or none.
//    Canonical, order-independent rule used by oracle and GPU alike: a world object is evaluated over
//    [limit.start, +inf) — as if it were the first one visited — and its hit competes with t1.
//  * `assert!`s that abort the reference (pdf.rs:38 etc., box.rs:129) do not abort here.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <thread>
#include <vector>

#include "../include/raysnail_b200.h"

namespace {

constexpr double PI = 3.14159265358979323846;
constexpr double F64_MAX = std::numeric_limits<double>::max();
constexpr double INF = std::numeric_limits<double>::infinity();

// ---------------------------------------------------------------------------------------------
// prelude/vec3.rs
// ---------------------------------------------------------------------------------------------
struct Vec3 {
    double x = 0, y = 0, z = 0;
    Vec3() = default;
    Vec3(double a, double b, double c) : x(a), y(b), z(c) {}
    double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    double& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
};
inline Vec3 operator+(const Vec3& a, const Vec3& b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline Vec3 operator-(const Vec3& a, const Vec3& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec3 operator-(const Vec3& a) { return {-a.x, -a.y, -a.z}; }
inline Vec3 operator*(const Vec3& a, double s) { return {a.x * s, a.y * s, a.z * s}; }
inline Vec3 operator*(double s, const Vec3& a) { return a * s; }
inline Vec3 operator*(const Vec3& a, const Vec3& b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline Vec3 operator/(const Vec3& a, double s) { return a * (1.0 / s); }  // vec3.rs:486-500
inline double length_squared(const Vec3& v) { return std::fma(v.z, v.z, std::fma(v.x, v.x, v.y * v.y)); }  // :152-155
inline double length(const Vec3& v) { return std::sqrt(length_squared(v)); }
inline double dot(const Vec3& a, const Vec3& b) { return std::fma(a.z, b.z, std::fma(a.x, b.x, a.y * b.y)); }  // :177-179
inline Vec3 cross(const Vec3& a, const Vec3& b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline Vec3 unit(const Vec3& v) { return v / length(v); }  // :192-194

// prelude/color.rs
struct Color {
    float r = 0, g = 0, b = 0, a = 0;
    Color() = default;
    Color(float r_, float g_, float b_, float a_) : r(r_), g(g_), b(b_), a(a_) {}
};
inline Vec3 operator*(const Color& c, const Vec3& v) { return {v.x * (double)c.r, v.y * (double)c.g, v.z * (double)c.b}; }  // vec3.rs:386-396
inline Vec3 to_vec3(const Color& c) { return {(double)c.r, (double)c.g, (double)c.b}; }                                      // vec3.rs:580-584
// color.rs:50-58
inline Color gradient(const Color& c0, const Color& c1, double slide) {
    float a = (float)std::fmin(std::fmax(slide, 0.0), 1.0);
    float b = 1.0f - a;
    return Color(c0.r * b + c1.r * a, c0.g * b + c1.g * a, c0.b * b + c1.b * a, 1.0f);
}
// color.rs:99-108: Color * f64
inline Color scale_color(const Color& c, double s) { return Color(c.r * (float)s, c.g * (float)s, c.b * (float)s, c.a); }

// prelude/ray.rs
struct Ray {
    Vec3 origin, direction;
    double departure_time = 0;
    Vec3 at(double t) const {  // ray.rs:21-31
        return {std::fma(direction.x, t, origin.x), std::fma(direction.y, t, origin.y), std::fma(direction.z, t, origin.z)};
    }
};

struct Range {  // std::ops::Range<f64>: [start, end)
    double start, end;
    bool contains(double v) const { return v >= start && v < end; }
};

// prelude/aabb.rs
struct AABB {
    Vec3 min, max;
    bool hit(const Ray& ray, const Range& limit) const {  // aabb.rs:20-38
        double t_min = limit.start, t_max = limit.end;
        for (int i = 0; i < 3; ++i) {
            double inv = 1.0 / ray.direction[i];
            double t0 = (min[i] - ray.origin[i]) * inv;
            double t1 = (max[i] - ray.origin[i]) * inv;
            if (inv < 0.0) std::swap(t0, t1);
            t_min = std::fmax(t_min, t0);  // f64::max: NaN loses, like fmax
            t_max = std::fmin(t_max, t1);
            if (t_max <= t_min) return false;
        }
        return true;
    }
};
inline AABB unite(const AABB& a, const AABB& b) {  // aabb.rs:58-66
    return {{std::fmin(a.min.x, b.min.x), std::fmin(a.min.y, b.min.y), std::fmin(a.min.z, b.min.z)},
            {std::fmax(a.max.x, b.max.x), std::fmax(a.max.y, b.max.y), std::fmax(a.max.z, b.max.z)}};
}

// prelude/onb.rs
struct ONB {
    Vec3 axis[3];
    Vec3 local(const Vec3& a) const {  // onb.rs:11-22
        return {axis[0].x * a.x + axis[1].x * a.y + axis[2].x * a.z, axis[0].y * a.x + axis[1].y * a.y + axis[2].y * a.z,
                axis[0].z * a.x + axis[1].z * a.y + axis[2].z * a.z};
    }
    static ONB build_from(const Vec3& n) {  // onb.rs:24-40
        Vec3 w = unit(n);
        Vec3 uc = cross(Vec3(0.0, 1.0, 0.0), w);
        Vec3 u = length_squared(uc) < 0.00000001 ? unit(cross(Vec3(1.0, 0.0, 0.0), w)) : unit(uc);
        Vec3 v = cross(w, u);
        return ONB{{u, v, w}};
    }
};

// ---------------------------------------------------------------------------------------------
// RNG: Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11),
// replaces prelude/random.rs:109-145.  See DESIGN.md "RNG streams" for the stream layout.
// ---------------------------------------------------------------------------------------------
inline void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

enum Purpose : uint32_t { P_PIXEL = 0, P_LENS = 1, P_SCATTER = 2, P_MIXED = 3, P_LIGHT = 4, P_BRDF = 5, P_MEDIUM = 6 };

inline double u01(uint32_t w) { return (double)(w >> 8) * (1.0 / 16777216.0); }

// The per-path RNG context handed down the recursion instead of `&mut FastRng`.
struct Rng {
    uint32_t key[2];
    uint32_t pixel, sample, depth;
    void words(uint32_t purpose, uint32_t call, uint32_t out[4]) const {
        uint32_t ctr[4] = {pixel, sample, depth, purpose | (call << 8)};
        philox4x32_10(ctr, key, out);
    }
    // rejection loops: iteration k reads word pair (k&1) of call (base + k>>1)
    void pair(uint32_t purpose, uint32_t base_call, uint32_t k, double& a, double& b) const {
        uint32_t w[4];
        words(purpose, base_call + (k >> 1), w);
        a = u01(w[2 * (k & 1)]);
        b = u01(w[2 * (k & 1) + 1]);
    }
};

// The lobe samplers of ReflectionPdf / BlinnPhongPdf (pdf.rs:124-140, 195-209) loop until the
// sample lies above the surface.  When a ray hits the BACK of a shading normal (triangle normals
// are not face-forwarded, triangle_mesh.rs:115-116) the reflected lobe can lie entirely below
// it and the reference's `loop {}` never ends.  Oracle and GPU give up after this many
// candidates and keep the last one.
constexpr uint32_t REJECTION_CAP = 256;

// vec3.rs:100-111
inline Vec3 cosine_direction(double r1, double r2) {
    double q2 = std::sqrt(r2);
    double phi = 2.0 * PI * r1;
    return {std::cos(phi) * q2, std::sin(phi) * q2, std::sqrt(1.0 - r2)};
}
// vec3.rs:115-126
inline Vec3 cosine_direction_exponent(double exponent, double r1, double xi2) {
    double r2 = std::pow(xi2, 1.0 / (exponent + 1.0));
    double sin_theta = std::sqrt(1.0 - r2 * r2);
    double phi = 2.0 * PI * r1;
    return {std::cos(phi) * sin_theta, std::sin(phi) * sin_theta, r2};
}

// ---------------------------------------------------------------------------------------------
// texture/*.rs
// ---------------------------------------------------------------------------------------------
struct Scene;

struct Texture {
    virtual ~Texture() = default;
    virtual Color color(double u, double v, const Vec3& p) const = 0;
};
struct SolidTex : Texture {
    Color c;
    Color color(double, double, const Vec3&) const override { return c; }  // color.rs:61-65
};
struct CheckerTex : Texture {  // checker.rs:22-29
    const Texture *odd, *even;
    double scale;
    Color color(double u, double v, const Vec3& p) const override {
        double value = std::sin(scale * p.x) * std::sin(scale * p.y) * std::sin(scale * p.z);
        return value < 0.0 ? odd->color(u, v, p) : even->color(u, v, p);
    }
};
struct ImageTex : Texture {  // image.rs:36-49
    int w, h;
    const uint8_t* texels;
    Color color(double u, double v, const Vec3&) const override {
        v = 1.0 - v;
        auto as_u32 = [](double d) -> uint32_t {  // Rust `as u32`: saturating, NaN -> 0
            if (!(d > 0.0)) return 0u;
            if (d >= 4294967295.0) return 4294967295u;
            return (uint32_t)d;
        };
        uint32_t px = as_u32(u * (double)w), py = as_u32(v * (double)h);
        if (px >= (uint32_t)w) px = (uint32_t)w - 1;
        if (py >= (uint32_t)h) py = (uint32_t)h - 1;
        const uint8_t* t = texels + 4 * ((size_t)py * w + px);
        return Color((float)t[0] / 255.0f, (float)t[1] / 255.0f, (float)t[2] / 255.0f, 1.0f);
    }
};
struct PerlinTex : Texture {  // noise.rs
    int kind, depth, smooth, vector, point_count;
    double scale;
    const int32_t* perm;    // perm_x | perm_y | perm_z
    const double* values;   // point_count * (vector ? 3 : 1)

    double lattice(int idx, double wx, double wy, double wz) const {
        if (vector) return dot(Vec3(values[3 * idx], values[3 * idx + 1], values[3 * idx + 2]), Vec3(wx, wy, wz));
        return values[idx];
    }
    double noise(const Vec3& p) const {  // noise.rs:100-137
        const int64_t mask = point_count - 1;
        if (smooth == 0) {
            auto as_isize = [](double d) -> int64_t {
                if (d != d) return 0;
                if (d >= 9.2e18) return INT64_MAX;
                if (d <= -9.2e18) return INT64_MIN;
                return (int64_t)d;
            };
            int i = (int)(as_isize(4.0 * p.x) & mask), j = (int)(as_isize(4.0 * p.y) & mask), k = (int)(as_isize(4.0 * p.z) & mask);
            int idx = perm[i] ^ perm[point_count + j] ^ perm[2 * point_count + k];
            return vector ? values[3 * idx] : values[idx];
        }
        double fx = std::floor(p.x), fy = std::floor(p.y), fz = std::floor(p.z);
        int64_t i = (int64_t)fx, j = (int64_t)fy, k = (int64_t)fz;
        double u = p.x - (double)i, v = p.y - (double)j, w = p.z - (double)k;
        // interpolate (noise.rs:156-189)
        double uu = u, vv = v, ww = w;
        if (smooth == 2) {
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
                    int idx = perm[xi] ^ perm[point_count + yi] ^ perm[2 * point_count + zi];
                    double term = std::fma((double)di, uu, (double)(1 - di) * (1.0 - uu)) *
                                  std::fma((double)dj, vv, (double)(1 - dj) * (1.0 - vv)) *
                                  std::fma((double)dk, ww, (double)(1 - dk) * (1.0 - ww)) *
                                  lattice(idx, u - (double)di, v - (double)dj, w - (double)dk);
                    sum_k += term;
                }
                sum_j += sum_k;
            }
            sum_i += sum_j;
        }
        return sum_i;
    }
    double turbulence(const Vec3& point, int d) const {  // noise.rs:139-152
        Vec3 p = point;
        double weight = 1.0, sum = 0.0;
        for (int i = 0; i < d; ++i) {
            sum += weight * noise(p);
            weight *= 0.5;
            p = Vec3(p.x * 2.0, p.y * 2.0, p.z * 2.0);
        }
        return std::fabs(sum);
    }
    Color color(double, double, const Vec3& point) const override {  // noise.rs:192-210
        double f;
        if (kind == 0) {
            f = noise(scale * point);
            if (vector) f = 0.5 * (f + 1.0);
        } else if (kind == 1) {
            f = turbulence(point, depth);
        } else {
            double n = turbulence(point, depth);
            f = (std::sin(std::fma(scale, point.z, 10.0 * n)) + 1.0) * 0.5;
        }
        return scale_color(Color(1.f, 1.f, 1.f, 1.f), f);
    }
};

// ---------------------------------------------------------------------------------------------
// prelude/pdf.rs
// ---------------------------------------------------------------------------------------------
struct PDF {
    virtual ~PDF() = default;
    virtual double value(const Vec3& direction) const = 0;
    virtual Vec3 generate(const Rng& rng) const = 0;
};
struct CosinePdf : PDF {  // pdf.rs:20-48
    ONB onb;
    explicit CosinePdf(const Vec3& n) : onb(ONB::build_from(n)) {}
    double value(const Vec3& d) const override {
        double c = dot(d, onb.axis[2]);
        return c < 0.0 ? 0.0 : c / PI;
    }
    Vec3 generate(const Rng& rng) const override {
        uint32_t w[4];
        rng.words(P_BRDF, 0, w);
        return onb.local(cosine_direction(u01(w[0]), u01(w[1])));
    }
};
struct ReflectionPdf : PDF {  // pdf.rs:86-141
    ONB onb_normal, onb_reflected;
    double exponent;
    ReflectionPdf(const Vec3& r_in, const Vec3& normal, double e) : exponent(e) {
        Vec3 reflected = r_in - (normal * (2.0 * dot(r_in, normal)));  // vec3.rs:171-173
        onb_reflected = ONB::build_from(reflected);
        onb_normal = ONB::build_from(normal);
    }
    double value(const Vec3& d) const override {
        double v = dot(d, onb_reflected.axis[2]) / PI;
        return v < 0.0 ? 0.0 : v;
    }
    Vec3 generate(const Rng& rng) const override {
        for (uint32_t k = 0;; ++k) {
            double r1, xi2;
            rng.pair(P_BRDF, 0, k, r1, xi2);
            Vec3 direction = onb_reflected.local(cosine_direction_exponent(exponent, r1, xi2));
            if (dot(direction, onb_normal.axis[2]) > 0.0) return direction;
            if (k + 1 >= REJECTION_CAP) return direction;  // the reference spins on (see REJECTION_CAP)
        }
    }
};
struct BlinnPhongPdf : PDF {  // pdf.rs:145-210
    Vec3 r_in;
    ONB onb_normal, onb_reflected;
    double k_specular, exponent;
    BlinnPhongPdf(const Vec3& r, const Vec3& normal, double ks, double e) : r_in(r), k_specular(ks), exponent(e) {
        Vec3 reflected = r_in - (normal * (2.0 * dot(r_in, normal)));
        onb_reflected = ONB::build_from(reflected);
        onb_normal = ONB::build_from(normal);
    }
    double value(const Vec3& d) const override {
        double cosine = dot(d, onb_normal.axis[2]);
        Vec3 random_normal = unit((-r_in) + d);
        double cosine_specular = std::fmax(dot(random_normal, onb_normal.axis[2]), 0.0);
        double normal_pdf = (exponent + 1.0) / (2.0 * PI) * std::pow(cosine_specular, exponent);
        return std::fmax(cosine / PI, 0.0) * (1.0 - k_specular) + normal_pdf / (4.0 * dot(r_in * -1.0, random_normal)) * k_specular;
    }
    Vec3 generate(const Rng& rng) const override {
        uint32_t w[4];
        rng.words(P_BRDF, 0, w);
        if (u01(w[0]) < k_specular) {
            for (uint32_t k = 0;; ++k) {
                double r1, xi2;
                rng.pair(P_BRDF, 1, k, r1, xi2);
                Vec3 direction = onb_reflected.local(cosine_direction_exponent(exponent, r1, xi2));
                if (dot(direction, onb_normal.axis[2]) > 0.0) return direction;
                if (k + 1 >= REJECTION_CAP) return direction;
            }
        }
        return onb_normal.local(cosine_direction(u01(w[1]), u01(w[2])));
    }
};
struct SpherePdf : PDF {  // pdf.rs:214-233
    double value(const Vec3&) const override { return 1.0 / (4.0 * PI); }
    Vec3 generate(const Rng& rng) const override {  // vec3.rs:91-96
        uint32_t w[4];
        rng.words(P_BRDF, 0, w);
        double a = 0.0 + u01(w[0]) * (2.0 * PI - 0.0);
        double z = -1.0 + u01(w[1]) * (1.0 - -1.0);
        double r = std::sqrt(1.0 - z * z);
        return {r * std::cos(a), r * std::sin(a), z};
    }
};

// ---------------------------------------------------------------------------------------------
// material/*.rs
// ---------------------------------------------------------------------------------------------
struct Material;

struct HitRecord {  // hittable/hit.rs:10-20
    Vec3 point, normal;
    const Material* material = nullptr;
    double t1 = 0, t2 = 0, u = 0, v = 0;
    bool outside = false;
};

struct ScatterRecord {  // material/mod.rs:27-32
    Color color;
    bool has_ray = false;
    Ray ray;
    std::unique_ptr<PDF> pdf;
    bool skip_pdf = false;
};

struct Settings {
    double phong_factor = 0.0;
    int phong_exponent = 1;
};

struct Material {
    int index = -1;
    Settings own_settings;
    virtual ~Material() = default;
    virtual bool scatter(const Ray&, const HitRecord&, const Rng&, int /*mix_level*/, ScatterRecord&) const { return false; }
    virtual bool emitted(double, double, const Vec3&, Vec3&) const { return false; }
    virtual Settings settings() const { return own_settings; }
};

// material/mod.rs:75-81, metal.rs:13-19
inline Ray reflect(const Ray& ray, const HitRecord& hit) {
    Vec3 d = ray.direction - ((2.0 * dot(ray.direction, hit.normal)) * hit.normal);
    return Ray{hit.point, d, ray.departure_time};
}

struct LambertianMat : Material {  // lambertian.rs:39-50
    const Texture* texture;
    bool scatter(const Ray&, const HitRecord& hit, const Rng&, int, ScatterRecord& s) const override {
        s.color = texture->color(hit.u, hit.v, hit.point);
        s.has_ray = false;
        s.pdf = std::make_unique<CosinePdf>(hit.normal);
        s.skip_pdf = false;
        return true;
    }
};
struct MetalMat : Material {  // metal.rs:104-118
    const Texture* texture;
    bool scatter(const Ray& ray, const HitRecord& hit, const Rng&, int, ScatterRecord& s) const override {
        s.color = texture->color(hit.u, hit.v, hit.point);
        Ray reflected = reflect(ray, hit);
        if (dot(reflected.direction, hit.normal) > 0.0) {
            s.has_ray = true;
            s.ray = reflected;
            s.pdf = std::make_unique<CosinePdf>(hit.normal);
            s.skip_pdf = true;
            return true;
        }
        return false;
    }
};
struct DiffuseMetalMat : Material {  // metal.rs:54-68
    const Texture* texture;
    double exponent;
    bool scatter(const Ray& ray, const HitRecord& hit, const Rng&, int, ScatterRecord& s) const override {
        s.color = texture->color(hit.u, hit.v, hit.point);
        Ray reflected = reflect(ray, hit);
        if (dot(reflected.direction, hit.normal) > 0.0) {
            s.has_ray = true;
            s.ray = reflected;
            s.pdf = std::make_unique<ReflectionPdf>(ray.direction, hit.normal, exponent);
            s.skip_pdf = false;
            return true;
        }
        return false;
    }
};
struct DielectricMat : Material {  // dielectric.rs
    Color color;
    double enter_refractive, outer_refractive;
    bool has_curve;
    static double glass_reflect_prob(double cos_theta, double refractive) {  // :20-24
        double r0 = (1.0 - refractive) / (1.0 + refractive);
        r0 = r0 * r0;
        double x = 1.0 - cos_theta;
        double x5 = x * x * x * x * x;  // f64::powi(5)
        return std::fma(1.0 - r0, x5, r0);
    }
    bool refract(const Ray& ray, const HitRecord& hit, const Rng& rng, Ray& out) const {  // :55-80
        double cos_theta = dot(-ray.direction, hit.normal);
        double sin_theta = std::sqrt(1.0 - cos_theta * cos_theta);
        double refractive = hit.outside ? enter_refractive : outer_refractive;
        if (refractive * sin_theta > 1.0) return false;
        double reflect_prob = has_curve ? glass_reflect_prob(cos_theta, refractive) : 0.0;
        uint32_t w[4];
        rng.words(P_SCATTER, 0, w);
        if (u01(w[3]) < reflect_prob) return false;
        Vec3 r_parallel = refractive * (ray.direction + cos_theta * hit.normal);
        Vec3 r_perpendicular = -std::sqrt(1.0 - length_squared(r_parallel)) * hit.normal;
        out = Ray{hit.point, r_parallel + r_perpendicular, ray.departure_time};
        return true;
    }
    bool scatter(const Ray& ray, const HitRecord& hit, const Rng& rng, int, ScatterRecord& s) const override {  // :83-93
        Ray r;
        if (!refract(ray, hit, rng, r)) r = reflect(ray, hit);
        s.color = color;
        s.has_ray = true;
        s.ray = r;
        s.pdf = std::make_unique<CosinePdf>(hit.normal);
        s.skip_pdf = true;
        return true;
    }
};
struct DiffuseLightMat : Material {  // light.rs:33-35
    const Texture* texture;
    double multiplier;
    bool emitted(double u, double v, const Vec3& p, Vec3& out) const override {
        out = to_vec3(texture->color(u, v, p)) * multiplier;
        return true;
    }
};
struct IsotropicMat : Material {  // isotropic.rs:26-33
    Color color;
    bool scatter(const Ray&, const HitRecord&, const Rng&, int, ScatterRecord& s) const override {
        s.color = color;
        s.has_ray = false;
        s.pdf = std::make_unique<SpherePdf>();
        s.skip_pdf = false;
        return true;
    }
};
struct BlinnPhongMat : Material {  // blinn_phong.rs:32-42
    const Texture* texture;
    double k_specular, exponent;
    bool scatter(const Ray& ray, const HitRecord& hit, const Rng&, int, ScatterRecord& s) const override {
        s.color = texture->color(hit.u, hit.v, hit.point);
        s.has_ray = false;
        s.pdf = std::make_unique<BlinnPhongPdf>(ray.direction, hit.normal, k_specular, exponent);
        s.skip_pdf = false;
        return true;
    }
};
struct MixedMat : Material {  // mixed_material.rs:43-58
    const Material *m1, *m2;
    double probability_1;
    bool scatter(const Ray& ray, const HitRecord& hit, const Rng& rng, int level, ScatterRecord& s) const override {
        uint32_t w[4];
        uint32_t word;
        if (level == 0) { rng.words(P_SCATTER, 0, w); word = w[2]; }
        else { rng.words(P_MIXED, (uint32_t)level, w); word = w[0]; }
        if ((double)word < (double)UINT32_MAX * probability_1) return m1->scatter(ray, hit, rng, level + 1, s);
        return m2->scatter(ray, hit, rng, level + 1, s);
    }
    Settings settings() const override { return m1->settings(); }
};

// ---------------------------------------------------------------------------------------------
// hittable/**
// ---------------------------------------------------------------------------------------------
struct Hittable {
    const Material* mat = nullptr;  // Option<Arc<dyn Material>>
    int node_index = -1;
    virtual ~Hittable() = default;
    virtual bool hit(const Ray& ray, const Range& limit, const Rng& rng, HitRecord& out) const = 0;
    virtual bool contains(const Vec3& p) const = 0;
    virtual AABB bbox(const Range& time) const = 0;
    virtual Vec3 random(const Vec3& origin, const Rng& rng) const = 0;
    // The reference's default normal()/uv() panic ("should not be called directly", hit.rs:82-99);
    // `ok` reports that instead of aborting.
    virtual Vec3 normal(const Vec3&, bool& ok) const { ok = false; return Vec3(0, 0, 0); }
    virtual void uv(const Vec3&, double& u, double& v) const { u = 0; v = 0; }
    virtual const Material* material(bool& ok) const { ok = true; return mat; }
};

// hit.rs:32-53
template <class G>
inline HitRecord new_hit_record(const Ray& ray, const G& object, double t1, double t2) {
    HitRecord h;
    h.point = ray.at(t1);
    bool ok;
    Vec3 normal = object.normal(h.point, ok);
    h.outside = dot(ray.direction, normal) < 0.0;
    if (!h.outside) normal = -normal;
    h.normal = normal;
    h.material = object.mat;
    object.uv(h.point, h.u, h.v);
    h.t1 = t1;
    h.t2 = t2;
    return h;
}
// hit.rs:55-67
inline HitRecord with_normal(const Vec3& point, const Vec3& normal, const Material* m, double u, double v, double t1, double t2) {
    HitRecord h;
    h.point = point; h.normal = normal; h.material = m; h.u = u; h.v = v; h.t1 = t1; h.t2 = t2; h.outside = true;
    return h;
}

struct SphereObj : Hittable {  // geometry/sphere.rs
    Vec3 center, speed;
    double radius, radius_squared;
    Vec3 center_at(double t) const { return center + speed * t; }
    Vec3 normal(const Vec3& p, bool& ok) const override { ok = true; return (p - center) / radius; }  // :56-58 (static centre)
    void uv(const Vec3& point, double& u, double& v) const override {  // :64-71
        Vec3 p = unit(point - center);
        double phi = std::atan2(-p.z, p.x);
        double theta = std::asin(p.y);
        u = phi / 2.0 / PI + 0.5;
        v = theta / PI + 0.5;
    }
    bool hit(const Ray& ray, const Range& limit, const Rng&, HitRecord& out) const override {  // :83-109
        Vec3 current_center = center_at(ray.departure_time);
        Vec3 l = ray.origin - current_center;
        double half_b = dot(ray.direction, l);
        double a = length_squared(ray.direction);
        double c = length_squared(l) - radius_squared;
        double delta = half_b * half_b - a * c;
        if (delta < 0.0) return false;
        double sq = std::sqrt(delta);
        double t1 = (-half_b - sq) / a;
        double t2 = (-half_b + sq) / a;
        if (limit.contains(t1)) { out = new_hit_record(ray, *this, t1, t2); return true; }
        if (limit.contains(t2)) { out = new_hit_record(ray, *this, t2, t2); return true; }
        return false;
    }
    bool contains(const Vec3& p) const override {  // :111-115
        Vec3 r = center - p;
        return length_squared(r) < radius * radius;
    }
    AABB bbox(const Range& time) const override {  // :117-142
        Vec3 rv(radius, radius, radius);
        if (speed.x == 0.0 && speed.y == 0.0 && speed.z == 0.0) return {center - rv, center + rv};
        AABB s{center_at(time.start) - rv, center_at(time.start) + rv};
        AABB e{center_at(time.end) - rv, center_at(time.end) + rv};
        return unite(s, e);
    }
    Vec3 random(const Vec3& origin, const Rng& rng) const override {  // :149-164
        Vec3 direction = center - origin;
        ONB uvw = ONB::build_from(direction);
        for (uint32_t k = 0;; ++k) {
            double a, b;
            rng.pair(P_LIGHT, 0, k, a, b);
            Vec3 u = uvw.axis[0] * a;
            Vec3 v = uvw.axis[1] * b;
            Vec3 uv = u + v;
            if (length_squared(uv) < 1.0) return (uv + center) - origin;
        }
    }
};

struct RectObj : Hittable {  // geometry/rect.rs
    int ax0, ax1, ax2;
    double k, a0, a1, b0, b1, a_len, b_len;
    Vec3 normal(const Vec3&, bool& ok) const override { ok = true; Vec3 n; n[ax2] = 1.0; return n; }  // :85-89
    void uv(const Vec3& p, double& u, double& v) const override { u = (p[ax0] - a0) / a_len; v = (p[ax1] - b0) / b_len; }  // :95-100
    bool hit(const Ray& ray, const Range& limit, const Rng&, HitRecord& out) const override {  // :102-120
        double t1 = (k - ray.origin[ax2]) / ray.direction[ax2];
        if (!limit.contains(t1)) return false;
        double a = std::fma(t1, ray.direction[ax0], ray.origin[ax0]);
        if (a < a0 || a > a1) return false;
        double b = std::fma(t1, ray.direction[ax1], ray.origin[ax1]);
        if (b < b0 || b > b1) return false;
        out = new_hit_record(ray, *this, t1, F64_MAX);
        return true;
    }
    bool contains(const Vec3&) const override { return false; }
    AABB bbox(const Range&) const override {  // :127-139
        Vec3 p0, p1;
        p0[ax0] = a0; p0[ax1] = b0; p0[ax2] = k - 0.0001;
        p1[ax0] = a1; p1[ax1] = b1; p1[ax2] = k + 0.0001;
        return {p0, p1};
    }
    Vec3 random(const Vec3& origin, const Rng& rng) const override {  // :141-153 (xz only, origin - root)
        uint32_t w[4];
        rng.words(P_LIGHT, 0, w);
        Vec3 root(0.0, k, 0.0);
        root.x = a0 + u01(w[0]) * (a1 - a0);
        root.z = b0 + u01(w[1]) * (b1 - b0);
        return origin - root;
    }
};

struct BoxObj : Hittable {  // geometry/box.rs
    Vec3 pmin, pmax;
    RectObj faces[6];
    void init(const Vec3& mn, const Vec3& mx) {  // :49-111 face order: back front left right down up
        pmin = mn; pmax = mx;
        auto mk = [&](RectObj& r, int a0x, int a1x, int a2x, double k, double a0, double a1, double b0, double b1) {
            r.ax0 = a0x; r.ax1 = a1x; r.ax2 = a2x; r.k = k; r.a0 = a0; r.a1 = a1; r.b0 = b0; r.b1 = b1;
            r.a_len = a1 - a0; r.b_len = b1 - b0; r.mat = mat;
        };
        mk(faces[0], 0, 1, 2, mn.z, mn.x, mx.x, mn.y, mx.y);
        mk(faces[1], 0, 1, 2, mx.z, mn.x, mx.x, mn.y, mx.y);
        mk(faces[2], 1, 2, 0, mn.x, mn.y, mx.y, mn.z, mx.z);
        mk(faces[3], 1, 2, 0, mx.x, mn.y, mx.y, mn.z, mx.z);
        mk(faces[4], 0, 2, 1, mn.y, mn.x, mx.x, mn.z, mx.z);
        mk(faces[5], 0, 2, 1, mx.y, mn.x, mx.x, mn.z, mx.z);
    }
    Vec3 normal(const Vec3&, bool& ok) const override { ok = true; return Vec3(0.0, 1.0, 0.0); }  // :118-120
    bool hit(const Ray& ray, const Range& limit, const Rng& rng, HitRecord& out) const override {  // :125-150
        HitRecord hits[6];
        int n = 0;
        for (int i = 0; i < 6; ++i)
            if (faces[i].hit(ray, limit, rng, hits[n])) ++n;
        // box.rs:129 asserts n < 3 (aborts on an exact edge/corner hit); here the first two count
        if (n == 1) { out = hits[0]; return true; }
        if (n >= 2) {
            const HitRecord &h1 = hits[0], &h2 = hits[1];
            if (h1.t1 < h2.t1) out = with_normal(h1.point, h1.normal, h1.material, h1.u, h1.v, h1.t1, h2.t1);
            else out = with_normal(h2.point, h2.normal, h2.material, h2.u, h2.v, h2.t1, h1.t1);
            return true;
        }
        return false;
    }
    bool contains(const Vec3& p) const override {  // :152-157
        return p.x >= pmin.x && p.x <= pmax.x && p.y >= pmin.y && p.y <= pmax.y && p.z >= pmin.z && p.z <= pmax.z;
    }
    AABB bbox(const Range& t) const override {  // :160-162 via list.rs:70-83
        AABB r = faces[0].bbox(t);
        for (int i = 1; i < 6; ++i) r = unite(r, faces[i].bbox(t));
        return r;
    }
    Vec3 random(const Vec3&, const Rng&) const override { return Vec3(1.0, 0.0, 0.0); }
};

struct QuadricObj : Hittable {  // geometry/quadric.rs
    double qa, qb, qc, qd, qe, qf, qg, qh, qi, qj;
    Vec3 normal(const Vec3& p, bool& ok) const override {  // :67-100
        ok = true;
        double x = 2.0 * qa * p.x + qb * p.y + qc * p.z + qd;
        double y = qb * p.x + 2.0 * qe * p.y + qf * p.z + qg;
        double z = qc * p.x + qf * p.y + 2.0 * qh * p.z + qi;
        Vec3 r(x, y, z);
        double len = length(r);
        if (len == 0.0) return Vec3(1.0, 0.0, 0.0);
        return r / len;
    }
    bool hit(const Ray& ray, const Range& limit, const Rng&, HitRecord& out) const override {  // :112-182
        double xo = ray.origin.x, yo = ray.origin.y, zo = ray.origin.z;
        double xd = ray.direction.x, yd = ray.direction.y, zd = ray.direction.z;
        double a = xd * (qa * xd + qb * yd + qc * zd) + yd * (qe * yd + qf * zd) + zd * qh * zd;
        double b = xd * (qa * xo + 0.5 * (qb * yo + qc * zo + qd)) + yd * (qe * yo + 0.5 * (qb * xo + qf * zo + qg)) +
                   zd * (qh * zo + 0.5 * (qc * xo + qf * yo + qi));
        double c = xo * (qa * xo + qb * yo + qc * zo + qd) + yo * (qe * yo + qf * zo + qg) + zo * (qh * zo + qi) + qj;
        if (a == 0.0) {
            if (b == 0.0) return false;
            double t1 = -0.5 * c / b;
            if (limit.contains(t1)) { out = new_hit_record(ray, *this, t1, F64_MAX); return true; }
        } else {
            double d = b * b - a * c;
            if (d <= 0.0) return false;
            double dr = std::sqrt(d);
            double t1 = (-b - dr) / a;
            double t2 = (-b + dr) / a;
            if (limit.contains(t1)) { out = new_hit_record(ray, *this, t1, t2); return true; }
            if (limit.contains(t2)) { out = new_hit_record(ray, *this, t2, F64_MAX); return true; }
        }
        return false;
    }
    bool contains(const Vec3& p) const override {  // :184-189
        return (p.x * (qa * p.x + qb * p.y + qd) + p.y * (qe * p.y + qf * p.z + qg) + p.z * (qh * p.z + qc * p.x + qi) + qj) <= 0.0;
    }
    AABB bbox(const Range&) const override { return {Vec3(-100, -100, -100), Vec3(100, 100, 100)}; }  // :191-196
    Vec3 random(const Vec3& origin, const Rng&) const override { return -origin; }                       // :202-205
};

// geometry/raymarching.rs: the power-8 Mandelbulb about the origin.
//
// Two evaluations of the SAME map (raymarching.rs:202-241):
//   * mode 1, "libm": the literal restatement — sqrt / atan2 / powf / sin / cos / ln through the
//     platform libm, exactly as the Rust source calls them;
//   * mode 0, "canonical" (default; what the GPU computes bit for bit): the same triplex power
//     written with IEEE-exact operations only — cos/sin of theta and phi as coordinate ratios,
//     the 8-fold angles by three complex squarings, r^8 and r^56 by repeated multiplication, ln by
//     a fixed atanh series.
// Why: the surface is found by testing orbit boundedness on a 0.05/256 grid along the ray, and
// the orbits near the surface are chaotic (a last-bit difference grows ~8x per iteration), so two
// libms that round ONE call differently move ~1 % of the hit distances by whole grid steps
// (measured: glibc vs CUDA libm, 40 of 3881 rays).  The Rust reference has the same spread between
// platforms (f64::powf/atan2/sin/cos/ln are the platform's libm).  The canonical form differs from
// the literal one by the same few ulp per call and is reproducible everywhere;
// tests/test_oracle_kat.py::test_raymarcher_canonical_vs_libm pins it to the literal form.
namespace marcher {
int g_mode = 0;

// ln(x) by IEEE-exact steps only: x = m * 2^e, m in [sqrt(1/2), sqrt(2)), s = (m-1)/(m+1),
// ln m = 2 atanh(s) = s * (2 + s^2 * (2/3 + s^2 * (2/5 + ...))) up to s^23 (|s| <= 0.1716).
inline double det_log(double x) {
    if (x != x || x < 0.0) return std::numeric_limits<double>::quiet_NaN();
    if (x == 0.0) return -std::numeric_limits<double>::infinity();
    if (x == std::numeric_limits<double>::infinity()) return x;
    uint64_t bits;
    std::memcpy(&bits, &x, 8);
    int e = 0;
    if ((bits >> 52) == 0) {  // subnormal: scale by 2^54 (exact)
        x *= 18014398509481984.0;
        std::memcpy(&bits, &x, 8);
        e = -54;
    }
    e += (int)(bits >> 52) - 1023;
    bits = (bits & 0x000FFFFFFFFFFFFFull) | 0x3FF0000000000000ull;
    double m;
    std::memcpy(&m, &bits, 8);  // [1, 2)
    if (m > 1.4142135623730951) { m = m * 0.5; e += 1; }
    double s = (m - 1.0) / (m + 1.0);
    double s2 = s * s;
    double p = 2.0 / 23.0;
    p = std::fma(p, s2, 2.0 / 21.0);
    p = std::fma(p, s2, 2.0 / 19.0);
    p = std::fma(p, s2, 2.0 / 17.0);
    p = std::fma(p, s2, 2.0 / 15.0);
    p = std::fma(p, s2, 2.0 / 13.0);
    p = std::fma(p, s2, 2.0 / 11.0);
    p = std::fma(p, s2, 2.0 / 9.0);
    p = std::fma(p, s2, 2.0 / 7.0);
    p = std::fma(p, s2, 2.0 / 5.0);
    p = std::fma(p, s2, 2.0 / 3.0);
    p = std::fma(p, s2, 2.0);
    double lm = s * p;
    const double LN2_HI = 6.93147180369123816490e-01, LN2_LO = 1.90821492927058770002e-10;
    double de = (double)e;
    return std::fma(de, LN2_HI, std::fma(de, LN2_LO, lm));
}

// :202-241, literal.  Note `dr` is computed from the ALREADY-raised r (`r = r.powf(power)` comes first).
inline void iterate_libm(const Vec3& p, int iterations, double& r, double& dr, bool& inside) {
    double x = 0.0, y = 0.0, z = 0.0;
    const double power = 8.0;
    r = 0.0; dr = 0.0;
    for (int i = 0; i < iterations; ++i) {
        r = std::sqrt(x * x + y * y + z * z);
        double theta = std::atan2(std::sqrt(x * x + y * y), z);
        double phi = std::atan2(y, x);
        r = std::pow(r, power);
        theta *= power;
        phi *= power;
        dr = std::pow(r, power - 1.0) * power * dr + 1.0;
        x = r * std::sin(theta) * std::cos(phi);
        y = r * std::sin(theta) * std::sin(phi);
        z = r * std::cos(theta);
        x += p.x; y += p.y; z += p.z;
        if (x * x + y * y + z * z > 8.0) { inside = false; return; }
    }
    inside = true;
}
// the same map, canonical form (see above)
inline void iterate(const Vec3& p, int iterations, double& r, double& dr, bool& inside) {
    if (g_mode == 1) { iterate_libm(p, iterations, r, dr, inside); return; }
    double x = 0.0, y = 0.0, z = 0.0;
    r = 0.0; dr = 0.0;
    for (int i = 0; i < iterations; ++i) {
        double rho2 = x * x + y * y;
        double r2 = rho2 + z * z;
        double rr = std::sqrt(r2), rho = std::sqrt(rho2);
        double ct = 1.0, st = 0.0, cp = 1.0, sp = 0.0;  // atan2(0, +0) = 0
        if (rr != 0.0) { ct = z / rr; st = rho / rr; }
        if (rho != 0.0) { cp = x / rho; sp = y / rho; }
        for (int k = 0; k < 3; ++k) {  // angle doubling: theta, phi -> 8 theta, 8 phi
            double c2 = ct * ct - st * st, s2 = (2.0 * ct) * st;
            ct = c2; st = s2;
            double d2 = cp * cp - sp * sp, t2 = (2.0 * cp) * sp;
            cp = d2; sp = t2;
        }
        double r4 = r2 * r2;
        r = r4 * r4;  // |v|^8
        double r16 = r * r, r32 = r16 * r16, r56 = (r32 * r16) * r;
        dr = r56 * 8.0 * dr + 1.0;
        x = r * st * cp + p.x;
        y = r * st * sp + p.y;
        z = r * ct + p.z;
        if (x * x + y * y + z * z > 8.0) { inside = false; return; }
    }
    inside = true;
}
inline bool is_inside(const Vec3& p, int iterations) {  // :187-190
    double r, dr; bool in;
    iterate(p, iterations, r, dr, in);
    return in;
}
inline double distance_est(const Vec3& p, int iterations) {  // :193-197 (NaN when r == 0: ln(0) * 0)
    double r, dr; bool in;
    iterate(p, iterations, r, dr, in);
    return 0.5 * (g_mode == 1 ? std::log(r) : det_log(r)) * r / dr;
}
}  // namespace marcher

struct RayMarcherObj : Hittable {
    int iterations = 100;
    Vec3 normal(const Vec3& pos, bool& ok) const override {  // :83-96 (always 100 iterations)
        ok = true;
        const double d = 0.01;
        Vec3 xd(d, 0.0, 0.0), yd(0.0, d, 0.0), zd(0.0, 0.0, d);
        return unit(Vec3(marcher::distance_est(pos + xd, 100) - marcher::distance_est(pos - xd, 100),
                         marcher::distance_est(pos + yd, 100) - marcher::distance_est(pos - yd, 100),
                         marcher::distance_est(pos + zd, 100) - marcher::distance_est(pos - zd, 100)));
    }
    void uv(const Vec3& point, double& u, double& v) const override {  // :102-111
        Vec3 p = unit(point - Vec3(0.0, 0.0, 0.0));
        double phi = std::atan2(-p.z, p.x);
        double theta = std::asin(p.y);
        u = phi / 2.0 / PI + 0.5;
        v = theta / PI + 0.5;
    }
    // :43-57 (the recursion unrolled: `depth` halvings, is_inside always with 100 iterations)
    static Vec3 binary_search_surface(Vec3 outside, Vec3 inside, int depth) {
        for (; depth > 0; --depth) {
            Vec3 mid = (outside + inside) * 0.5;
            if (marcher::is_inside(mid, 100)) inside = mid; else outside = mid;
        }
        return outside;
    }
    bool search_surface(const Vec3& p, const Vec3& direction, Vec3& found) const {  // :59-76
        Vec3 df = direction * 0.05;
        Vec3 v = p;
        for (int i = 0; i < 200; ++i) {
            if (marcher::is_inside(v, iterations)) { found = binary_search_surface(v - df, v, 8); return true; }
            v = v + df;
        }
        return false;
    }
    // :113-165.  The reference only checks `length > limit.start`; `limit.end` is ignored there, so with
    // its randomly built BVH a farther marcher hit can replace a closer hit of a sibling
    // (bvh.rs:184-191).  Canonical rule here (oracle and GPU): a hit at or beyond limit.end is no hit.
    bool hit(const Ray& ray, const Range& limit, const Rng&, HitRecord& out) const override {
        double len = length(ray.direction);
        Vec3 direction = ray.direction * (1.0 / len);
        Vec3 current = ray.origin;
        int steps = 0;
        double best = 1000000.0, est = 0.0;
        while (steps < 1000 && est < best + 1.0) {
            ++steps;
            est = marcher::distance_est(current, iterations);
            if (est != est) est = 0.1;
            if (est < 1.3) {
                Vec3 p;
                if (search_surface(current, direction, p)) {
                    double dist = length(p - ray.origin);
                    if (dist > limit.start) {
                        double t1 = dist / len;
                        if (!(t1 < limit.end)) return false;
                        out = new_hit_record(ray, *this, t1, t1);
                        return true;
                    }
                }
                return false;
            }
            if (est < best) best = est;
            current = current + direction * est * 0.05;
        }
        return false;
    }
    bool contains(const Vec3& p) const override { return marcher::is_inside(p, iterations); }  // :167-170
    AABB bbox(const Range&) const override { return {Vec3(-1.3, -1.3, -1.3), Vec3(1.3, 1.3, 1.3)}; }  // :172-181
    Vec3 random(const Vec3&, const Rng& rng) const override {  // :188-190 -> vec3.rs:91-96
        uint32_t w[4];
        rng.words(P_LIGHT, 0, w);
        double a = 0.0 + u01(w[0]) * (2.0 * PI - 0.0);
        double z = -1.0 + u01(w[1]) * (1.0 - -1.0);
        double r = std::sqrt(1.0 - z * z);
        return {r * std::cos(a), r * std::sin(a), z};
    }
};

struct TriangleObj : Hittable {  // geometry/triangle_mesh.rs:17-149
    Vec3 p0, n0, n1, n2;
    double a, b, c, d, e, f;
    AABB box;
    void init(const Vec3& q0, const Vec3& q1, const Vec3& q2) {  // :42-60
        box = {{std::fmin(std::fmin(q0.x, q1.x), q2.x), std::fmin(std::fmin(q0.y, q1.y), q2.y), std::fmin(std::fmin(q0.z, q1.z), q2.z)},
               {std::fmax(std::fmax(q0.x, q1.x), q2.x), std::fmax(std::fmax(q0.y, q1.y), q2.y), std::fmax(std::fmax(q0.z, q1.z), q2.z)}};
        p0 = q0;
        a = q0.x - q1.x; b = q0.y - q1.y; c = q0.z - q1.z;
        d = q0.x - q2.x; e = q0.y - q2.y; f = q0.z - q2.z;
    }
    bool hit(const Ray& ray, const Range& limit, const Rng&, HitRecord& out) const override {  // :85-131
        double g = ray.direction.x, h = ray.direction.y, i = ray.direction.z;
        double j = p0.x - ray.origin.x, k = p0.y - ray.origin.y, l = p0.z - ray.origin.z;
        double eihf = e * i - h * f, gfdi = g * f - d * i, dheg = d * h - e * g;
        double denom = a * eihf + b * gfdi + c * dheg;
        double beta = (j * eihf + k * gfdi + l * dheg) / denom;
        if (beta < 0.0 || beta >= 1.0) return false;
        double akjb = a * k - j * b, jcal = j * c - a * l, blkc = b * l - k * c;
        double gamma = (i * akjb + h * jcal + g * blkc) / denom;
        if (gamma <= 0.0 || beta + gamma >= 1.0) return false;
        double t = -(f * akjb + e * jcal + d * blkc) / denom;
        if (t >= limit.start && t <= limit.end) {
            Vec3 normal = n0 * (1.0 - beta - gamma) + n1 * beta + n2 * gamma;
            out = with_normal(ray.at(t), normal, mat, 0.0, 0.0, t, F64_MAX);
            return true;
        }
        return false;
    }
    bool contains(const Vec3&) const override { return false; }
    AABB bbox(const Range&) const override { return box; }
    Vec3 random(const Vec3& origin, const Rng&) const override { return origin - p0; }
};

struct DifferenceObj : Hittable {  // csg/difference.rs
    const Hittable *plus, *minus;
    bool hit(const Ray& ray, const Range& limit, const Rng& rng, HitRecord& out) const override {  // :57-107
        HitRecord hp, hm;
        bool has_p = plus->hit(ray, limit, rng, hp);
        bool has_m = minus->hit(ray, limit, rng, hm);
        if (has_p) {
            if (has_m) {
                if (hp.t1 < hm.t1) {
                    if (!minus->contains(hp.point)) {
                        out = hp;
                        if (!out.material) out.material = mat;
                        return true;
                    }
                } else {
                    if (hm.t2 < hp.t1) {
                        out = hp;
                        if (!out.material) out.material = mat;
                        return true;
                    } else if (hm.t2 < hp.t2) {
                        Vec3 p = ray.at(hm.t2);
                        bool ok_n, ok_m;
                        Vec3 n = minus->normal(p, ok_n);
                        const Material* mm = minus->material(ok_m);
                        out = with_normal(p, -n, mm, 0.0, 0.0, hm.t2, hp.t2);
                        if (!out.material) out.material = mat;
                        return true;
                    }
                }
            } else {
                out = hp;
                return true;
            }
        }
        return false;
    }
    bool contains(const Vec3& p) const override { return plus->contains(p) && !minus->contains(p); }  // :110-112
    AABB bbox(const Range& t) const override { return plus->bbox(t); }                                   // :114-116
    Vec3 random(const Vec3& o, const Rng& rng) const override { return plus->random(o, rng); }
    const Material* material(bool& ok) const override { ok = false; return nullptr; }  // panics in the reference
};

struct IntersectionObj : Hittable {  // csg/intersection.rs
    const Hittable *o1, *o2;
    bool hit(const Ray& ray, const Range& limit, const Rng& rng, HitRecord& out) const override {  // :58-96
        HitRecord h1, h2;
        bool has1 = o1->hit(ray, limit, rng, h1);
        bool has2 = o2->hit(ray, limit, rng, h2);
        if (has1 && has2) {
            const HitRecord* hits[2];
            const Hittable* objs[2];
            if (h1.t1 < h2.t1) { hits[0] = &h1; hits[1] = &h2; objs[0] = o1; objs[1] = o2; }
            else { hits[0] = &h2; hits[1] = &h1; objs[0] = o2; objs[1] = o1; }
            if (objs[1]->contains(hits[0]->point)) {
                out = *hits[0];
                if (!out.material) out.material = mat;
                return true;
            } else if (objs[0]->contains(hits[1]->point)) {
                out = *hits[1];
                if (!out.material) out.material = mat;
                return true;
            }
        }
        return false;
    }
    bool contains(const Vec3& p) const override { return o1->contains(p) && o2->contains(p); }  // :98-100
    AABB bbox(const Range& t) const override {                                                   // :102-118 (z quirk kept)
        AABB b1 = o1->bbox(t), b2 = o2->bbox(t);
        return {{std::fmax(b1.min.x, b2.min.x), std::fmax(b1.min.y, b2.min.y), std::fmax(b1.min.y, b2.min.z)},
                {std::fmin(b1.max.x, b2.max.x), std::fmin(b1.max.y, b2.max.y), std::fmin(b1.max.z, b2.max.z)}};
    }
    Vec3 random(const Vec3& o, const Rng& rng) const override { return o1->random(o, rng); }
    const Material* material(bool& ok) const override { ok = false; return nullptr; }
};

struct BvhObj : Hittable {  // collection/bvh.rs:173-208 — a BVH used as one object
    bool has_box = false;
    AABB box;
    const Hittable *left = nullptr, *right = nullptr;
    bool hit(const Ray& ray, const Range& limit, const Rng& rng, HitRecord& out) const override {  // :174-190
        if (!has_box || !box.hit(ray, limit)) return false;
        HitRecord hl, hr;
        // The reference searches the right child over start..left.t1 and prefers its hit (:181-189).  Canonical form, as
        // for the objects of the world (header): both children over the node's own limit, the right one wins when strictly
        // closer — the same winner, and what a Box member reports (which faces count, box.rs:125-150) does not depend on
        // its sibling.
        const bool a = left && left->hit(ray, limit, rng, hl);
        const bool b = right && right->hit(ray, limit, rng, hr);
        if (b && (!a || hr.t1 < hl.t1)) out = hr;
        else if (a) out = hl;
        return a || b;
    }
    bool contains(const Vec3&) const override { return false; }  // :194-200 unimplemented!(): refused at upload as a CSG operand
    AABB bbox(const Range&) const override { return box; }
    Vec3 random(const Vec3&, const Rng&) const override { return Vec3(1.0, 0.0, 0.0); }  // :206-208
    const Material* material(bool& ok) const override { ok = false; return nullptr; }
};

struct TransformObj : Hittable {  // transform/tf_facade.rs + transform.rs:133-157
    const Hittable* object;
    std::vector<rs_xform> stack;
    static Vec3 apply(const double* m, const Vec3& p, double w) {  // vecmath::row_mat4_transform, rows 0..2
        return {m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3] * w, m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7] * w,
                m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11] * w};
    }
    Vec3 forward(const Vec3& pos, double w) const {
        Vec3 r = pos;
        for (const rs_xform& t : stack) r = apply(t.m, r, w);
        return r;
    }
    Vec3 inverse(const Vec3& pos, double w) const {
        Vec3 r = pos;
        for (auto it = stack.rbegin(); it != stack.rend(); ++it) r = apply(it->inv, r, w);
        return r;
    }
    bool hit(const Ray& ray_in, const Range& limit, const Rng& rng, HitRecord& out) const override {  // tf_facade.rs:41-56
        Ray ray{inverse(ray_in.origin, 1.0), inverse(ray_in.direction, 0.0), ray_in.departure_time};
        if (!object->hit(ray, limit, rng, out)) return false;
        out.point = forward(out.point, 1.0);
        return true;
    }
    bool contains(const Vec3& p) const override { return object->contains(inverse(p, 1.0)); }  // :90-93
    AABB bbox(const Range& t) const override {                                                  // :59-88
        AABB bb = object->bbox(t);
        Vec3 pmin(INF, INF, INF), pmax(-INF, -INF, -INF);
        for (int i = 0; i < 2; ++i)
            for (int j = 0; j < 2; ++j)
                for (int k = 0; k < 2; ++k) {
                    double x = std::fma((double)i, bb.max.x, (double)(1 - i) * bb.min.x);
                    double y = std::fma((double)j, bb.max.y, (double)(1 - j) * bb.min.y);
                    double z = std::fma((double)k, bb.max.z, (double)(1 - k) * bb.min.z);
                    Vec3 p = forward(Vec3(x, y, z), 1.0);
                    for (int c = 0; c < 3; ++c) { pmin[c] = std::fmin(pmin[c], p[c]); pmax[c] = std::fmax(pmax[c], p[c]); }
                }
        return {pmin, pmax};
    }
    Vec3 random(const Vec3& origin, const Rng& rng) const override { return object->random(inverse(origin, 1.0), rng); }  // :95-98
    const Material* material(bool& ok) const override { ok = false; return nullptr; }
};

struct MediumObj : Hittable {  // medium/constant.rs
    const Hittable* boundary;
    double density, neg_inv_density;
    bool hit(const Ray& ray, const Range& limit, const Rng& rng, HitRecord& out) const override {  // :42-80
        HitRecord rec1, rec2;
        if (!boundary->hit(ray, Range{-INF, INF}, rng, rec1)) return false;
        if (!boundary->hit(ray, Range{rec1.t1 + 0.0001, INF}, rng, rec2)) return false;
        if (rec1.t1 < limit.start) rec1.t1 = limit.start;
        if (rec2.t1 > limit.end) rec2.t1 = limit.end;
        if (rec1.t1 >= rec2.t1) return false;
        if (rec1.t1 < 0.0) rec1.t1 = 0.0;
        double length_per_unit = length(ray.direction);
        double distance_inside = (rec2.t1 - rec1.t1) * length_per_unit;
        uint32_t w[4];
        rng.words(P_MEDIUM, (uint32_t)node_index, w);
        double hit_distance = neg_inv_density * std::log(u01(w[0]));
        if (hit_distance > distance_inside) return false;
        double t = rec1.t1 + hit_distance / length_per_unit;
        out = HitRecord();
        out.point = ray.at(t);
        out.normal = Vec3(1.0, 0.0, 0.0);
        out.material = mat;
        out.t1 = t; out.t2 = t; out.u = 0; out.v = 0; out.outside = false;
        return true;
    }
    bool contains(const Vec3&) const override { return false; }  // unimplemented!() in the reference
    AABB bbox(const Range& t) const override { return boundary->bbox(t); }
    Vec3 random(const Vec3&, const Rng&) const override { return Vec3(1.0, 0.0, 0.0); }
};

inline double next_up(double x) {
    if (!(x < INF) || !(x > 0.0)) return x;
    uint64_t b;
    std::memcpy(&b, &x, 8);
    b += 1;
    std::memcpy(&x, &b, 8);
    return x;
}

// collection/bvh.rs
struct BVHNode : Hittable {
    AABB box;
    const Hittable* left = nullptr;
    const Hittable* right = nullptr;
    bool gate_leaf = false;  // single-object leaf: bbox tested over [start, +inf) (see header)
    int leaf_id = -1;
    bool hit(const Ray& ray, const Range& limit, const Rng& rng, HitRecord& out) const override {  // :173-192
        if (gate_leaf) {
            if (!box.hit(ray, Range{limit.start, INF})) return false;
            // The object is evaluated over [limit.start, +inf) and its hit is accepted if it lies inside the
            // (shrunk) limit.  For spheres, rects, quadrics, triangles and media that is what evaluating under
            // the shrunk limit returns anyway; for a Box (which of its faces count, hence `outside` of the
            // two-face record, box.rs:125-150) and for CSG objects the reference's answer depends on how far
            // the limit had shrunk when its randomly built tree reached the object — canonical rule: as if
            // the object were visited first (see the header).
            if (!left->hit(ray, Range{limit.start, INF}, rng, out)) return false;
            if (!(out.t1 < limit.end)) return false;
            last_leaf_hit(leaf_id);
            return true;
        }
        if (!box.hit(ray, limit)) return false;
        HitRecord hl, hr;
        int id_left = -1;
        bool has_l = left && left->hit(ray, limit, rng, hl);
        if (has_l) id_left = take_leaf_hit();
        // bvh.rs:184-187 shrinks the right limit to `..left.t1`, so on an EXACT tie whichever
        // subtree is visited first wins — and the reference's tree order is random
        // (bvh.rs:91).  Canonical, order-independent tie rule used by oracle and GPU alike:
        // the limit admits t == left.t1 and the lower object id wins.
        Range right_limit{limit.start, has_l ? next_up(hl.t1) : limit.end};
        bool has_r = right && right->hit(ray, right_limit, rng, hr);
        if (has_r && has_l) {
            int id_right = take_leaf_hit();
            bool right_wins = hr.t1 < hl.t1 || (hr.t1 == hl.t1 && id_right < id_left);
            if (right_wins) { out = hr; last_leaf_hit(id_right); }
            else { out = hl; last_leaf_hit(id_left); }
            return true;
        }
        if (has_r) { out = hr; return true; }  // hit_right.or(hit_left)
        if (has_l) { out = hl; last_leaf_hit(id_left); return true; }
        return false;
    }
    bool contains(const Vec3&) const override { return false; }
    AABB bbox(const Range&) const override { return box; }
    Vec3 random(const Vec3&, const Rng&) const override { return Vec3(1.0, 0.0, 0.0); }
    // thread-local side channel carrying the object id of the record being returned
    static int& slot() { static thread_local int id = -1; return id; }
    static void last_leaf_hit(int id) { slot() = id; }
    static int take_leaf_hit() { return slot(); }
};

// ---------------------------------------------------------------------------------------------
// Scene: rs_scene_desc -> object graph (the inverse of the host's flatten)
// ---------------------------------------------------------------------------------------------
struct Scene {
    std::vector<std::unique_ptr<Texture>> textures;
    std::vector<std::unique_ptr<Material>> materials;
    std::vector<std::unique_ptr<Hittable>> nodes;
    std::vector<std::unique_ptr<TriangleObj>> triangles;
    std::vector<const Hittable*> roots;
    std::vector<const Hittable*> lights;
    std::vector<std::unique_ptr<BVHNode>> bvh_nodes;
    const BVHNode* bvh = nullptr;
    const Material* default_material = nullptr;
    Color bg0, bg1;
    std::vector<AABB> root_boxes;

    bool build(const rs_scene_desc& d, bool use_desc_bbox);
    const BVHNode* build_bvh(std::vector<int>& ids, size_t lo, size_t hi);

    Color background(const Ray& ray) const { return gradient(bg0, bg1, (ray.direction.y + 1.0) * 0.5); }
    // list.rs:49-52 + random.rs:138-140
    Vec3 lights_random(const Vec3& origin, const Rng& rng) const {
        uint32_t w[4];
        rng.words(P_SCATTER, 0, w);
        size_t idx = (size_t)w[1] % lights.size();
        return lights[idx]->random(origin, rng);
    }
    bool hit(const Ray& ray, const Range& limit, const Rng& rng, HitRecord& out, int& obj_id) const {  // world.rs:63-65
        if (!bvh) return false;
        if (!bvh->hit(ray, limit, rng, out)) return false;
        obj_id = BVHNode::take_leaf_hit();
        return true;
    }
};

bool Scene::build(const rs_scene_desc& d, bool use_desc_bbox) {
    bg0 = Color(d.bg0[0], d.bg0[1], d.bg0[2], d.bg0[3]);
    bg1 = Color(d.bg1[0], d.bg1[1], d.bg1[2], d.bg1[3]);
    // textures (children have smaller indices: flatten is post-order)
    for (int i = 0; i < d.n_textures; ++i) {
        const rs_texture& t = d.textures[i];
        switch (t.type) {
            case RS_TEX_SOLID: { auto p = std::make_unique<SolidTex>(); p->c = Color(t.color[0], t.color[1], t.color[2], t.color[3]); textures.push_back(std::move(p)); break; }
            case RS_TEX_CHECKER: {
                if (t.odd < 0 || t.odd >= i || t.even < 0 || t.even >= i) return false;
                auto p = std::make_unique<CheckerTex>(); p->odd = textures[t.odd].get(); p->even = textures[t.even].get(); p->scale = t.scale; textures.push_back(std::move(p)); break;
            }
            case RS_TEX_IMAGE: { auto p = std::make_unique<ImageTex>(); p->w = t.width; p->h = t.height; p->texels = d.texels + t.texel_offset; textures.push_back(std::move(p)); break; }
            case RS_TEX_PERLIN: {
                auto p = std::make_unique<PerlinTex>();
                p->kind = t.perlin_kind; p->depth = t.perlin_depth; p->smooth = t.perlin_smooth; p->vector = t.perlin_vector;
                p->point_count = t.point_count; p->scale = t.scale; p->perm = d.perlin_perm + t.perm_offset; p->values = d.perlin_values + t.value_offset;
                textures.push_back(std::move(p)); break;
            }
            default: return false;
        }
    }
    auto tex = [&](int id) -> const Texture* { return (id >= 0 && id < (int)textures.size()) ? textures[id].get() : nullptr; };
    for (int i = 0; i < d.n_materials; ++i) {
        const rs_material& m = d.materials[i];
        std::unique_ptr<Material> out;
        switch (m.type) {
            case RS_MAT_LAMBERTIAN: { auto p = std::make_unique<LambertianMat>(); p->texture = tex(m.texture); out = std::move(p); break; }
            case RS_MAT_METAL: { auto p = std::make_unique<MetalMat>(); p->texture = tex(m.texture); out = std::move(p); break; }
            case RS_MAT_DIFFUSE_METAL: { auto p = std::make_unique<DiffuseMetalMat>(); p->texture = tex(m.texture); p->exponent = m.a; out = std::move(p); break; }
            case RS_MAT_DIELECTRIC: {
                auto p = std::make_unique<DielectricMat>();
                p->color = Color(m.color[0], m.color[1], m.color[2], m.color[3]);
                p->outer_refractive = m.a; p->enter_refractive = 1.0 / m.a;  // dielectric.rs:39-40
                p->has_curve = m.has_curve != 0; out = std::move(p); break;
            }
            case RS_MAT_DIFFUSE_LIGHT: { auto p = std::make_unique<DiffuseLightMat>(); p->texture = tex(m.texture); p->multiplier = m.a; out = std::move(p); break; }
            case RS_MAT_ISOTROPIC: { auto p = std::make_unique<IsotropicMat>(); p->color = Color(m.color[0], m.color[1], m.color[2], m.color[3]); out = std::move(p); break; }
            case RS_MAT_BLINN_PHONG: { auto p = std::make_unique<BlinnPhongMat>(); p->texture = tex(m.texture); p->k_specular = m.a; p->exponent = m.b; out = std::move(p); break; }
            case RS_MAT_MIXED: {
                if (m.mat1 < 0 || m.mat1 >= i || m.mat2 < 0 || m.mat2 >= i) return false;
                auto p = std::make_unique<MixedMat>(); p->m1 = materials[m.mat1].get(); p->m2 = materials[m.mat2].get(); p->probability_1 = m.a; out = std::move(p); break;
            }
            default: return false;
        }
        out->index = i;
        out->own_settings.phong_factor = m.phong_factor;
        out->own_settings.phong_exponent = m.phong_exponent;
        materials.push_back(std::move(out));
    }
    auto mat = [&](int id) -> const Material* { return (id >= 0 && id < (int)materials.size()) ? materials[id].get() : nullptr; };
    default_material = mat(d.default_material);
    for (int i = 0; i < d.n_nodes; ++i) {
        const rs_node& n = d.nodes[i];
        std::unique_ptr<Hittable> out;
        auto child = [&](int c) -> const Hittable* { return (c >= 0 && c < i) ? nodes[c].get() : nullptr; };
        switch (n.type) {
            case RS_NODE_SPHERE: {
                auto p = std::make_unique<SphereObj>();
                p->center = Vec3(n.p[0], n.p[1], n.p[2]); p->radius = n.p[3]; p->radius_squared = n.p[3] * n.p[3];
                p->speed = Vec3(n.p[4], n.p[5], n.p[6]); out = std::move(p); break;
            }
            case RS_NODE_BOX: { auto p = std::make_unique<BoxObj>(); p->mat = mat(n.material); p->init(Vec3(n.p[0], n.p[1], n.p[2]), Vec3(n.p[3], n.p[4], n.p[5])); out = std::move(p); break; }
            case RS_NODE_RECT: {
                auto p = std::make_unique<RectObj>();
                if (n.rect_axis == RS_RECT_XY) { p->ax0 = 0; p->ax1 = 1; p->ax2 = 2; }
                else if (n.rect_axis == RS_RECT_XZ) { p->ax0 = 0; p->ax1 = 2; p->ax2 = 1; }
                else { p->ax0 = 1; p->ax1 = 2; p->ax2 = 0; }
                p->k = n.p[0]; p->a0 = n.p[1]; p->a1 = n.p[2]; p->b0 = n.p[3]; p->b1 = n.p[4]; p->a_len = p->a1 - p->a0; p->b_len = p->b1 - p->b0;
                out = std::move(p); break;
            }
            case RS_NODE_QUADRIC: {
                auto p = std::make_unique<QuadricObj>();
                p->qa = n.p[0]; p->qb = n.p[1]; p->qc = n.p[2]; p->qd = n.p[3]; p->qe = n.p[4]; p->qf = n.p[5]; p->qg = n.p[6]; p->qh = n.p[7]; p->qi = n.p[8]; p->qj = n.p[9];
                out = std::move(p); break;
            }
            case RS_NODE_DIFFERENCE: { auto p = std::make_unique<DifferenceObj>(); p->plus = child(n.child0); p->minus = child(n.child1); if (!p->plus || !p->minus) return false; out = std::move(p); break; }
            case RS_NODE_INTERSECTION: { auto p = std::make_unique<IntersectionObj>(); p->o1 = child(n.child0); p->o2 = child(n.child1); if (!p->o1 || !p->o2) return false; out = std::move(p); break; }
            case RS_NODE_TRANSFORM: {
                auto p = std::make_unique<TransformObj>(); p->object = child(n.child0); if (!p->object) return false;
                if (n.first_xform < 0 || n.first_xform + n.n_xform > d.n_xforms) return false;
                p->stack.assign(d.xforms + n.first_xform, d.xforms + n.first_xform + n.n_xform); out = std::move(p); break;
            }
            case RS_NODE_MEDIUM: {
                auto p = std::make_unique<MediumObj>(); p->boundary = child(n.child0); if (!p->boundary) return false;
                p->density = n.p[0]; p->neg_inv_density = -1.0 / n.p[0]; out = std::move(p); break;
            }
            case RS_NODE_RAYMARCHER: { auto p = std::make_unique<RayMarcherObj>(); p->iterations = (int)n.p[0]; out = std::move(p); break; }
            case RS_NODE_BVH: {
                auto p = std::make_unique<BvhObj>();
                p->has_box = n.p[6] != 0.0;
                p->box = {Vec3(n.p[0], n.p[1], n.p[2]), Vec3(n.p[3], n.p[4], n.p[5])};
                if (n.child0 >= i || n.child1 >= i) return false;
                p->left = n.child0 >= 0 ? child(n.child0) : nullptr;
                p->right = n.child1 >= 0 ? child(n.child1) : nullptr;
                out = std::move(p); break;
            }
            default: return false;
        }
        out->mat = mat(n.material);
        out->node_index = i;
        nodes.push_back(std::move(out));
    }
    for (int i = 0; i < d.n_tris; ++i) {
        auto t = std::make_unique<TriangleObj>();
        const double* v = d.tri_verts + 9 * (size_t)i;
        const double* nn = d.tri_normals + 9 * (size_t)i;
        t->init(Vec3(v[0], v[1], v[2]), Vec3(v[3], v[4], v[5]), Vec3(v[6], v[7], v[8]));
        t->n0 = Vec3(nn[0], nn[1], nn[2]); t->n1 = Vec3(nn[3], nn[4], nn[5]); t->n2 = Vec3(nn[6], nn[7], nn[8]);
        t->mat = mat(d.tri_material[i]);
        triangles.push_back(std::move(t));
    }
    for (int i = 0; i < d.n_roots; ++i) {
        int r = d.roots[i];
        const Hittable* h = r >= 0 ? (r < d.n_nodes ? nodes[r].get() : nullptr) : (~r < d.n_tris ? triangles[~r].get() : nullptr);
        if (!h) return false;
        roots.push_back(h);
        if (use_desc_bbox) {
            const double* b = d.root_bbox + 6 * (size_t)i;
            root_boxes.push_back({Vec3(b[0], b[1], b[2]), Vec3(b[3], b[4], b[5])});
        } else {
            root_boxes.push_back(h->bbox(Range{0.0, 0.0}));
        }
    }
    for (int i = 0; i < d.n_lights; ++i) {
        int r = d.lights[i];
        if (r < 0 || r >= d.n_nodes) return false;
        lights.push_back(nodes[r].get());
    }
    std::vector<int> ids(roots.size());
    for (size_t i = 0; i < ids.size(); ++i) ids[i] = (int)i;
    bvh = ids.empty() ? nullptr : build_bvh(ids, 0, ids.size());
    return true;
}

// bvh.rs:56-112.  Leaves are single-object gated nodes (see header); `count == 2` therefore
// becomes an inner node over two leaves.
const BVHNode* Scene::build_bvh(std::vector<int>& ids, size_t lo, size_t hi) {
    size_t count = hi - lo;
    auto node = std::make_unique<BVHNode>();
    BVHNode* n = node.get();
    bvh_nodes.push_back(std::move(node));
    if (count == 1) {
        n->gate_leaf = true;
        n->leaf_id = ids[lo];
        n->left = roots[ids[lo]];
        n->box = root_boxes[ids[lo]];
        return n;
    }
    if (count > 2) {
        // `Random::range(0..2)`: x or y, never z (bvh.rs:91); deterministic stand-in
        uint64_t hsh = (uint64_t)lo * 0x9E3779B97F4A7C15ull ^ (uint64_t)hi * 0xC2B2AE3D27D4EB4Full;
        hsh ^= hsh >> 29;
        int axis = (int)((hsh * 0xBF58476D1CE4E5B9ull >> 40) & 1);
        std::stable_sort(ids.begin() + lo, ids.begin() + hi, [&](int a, int b) { return root_boxes[a].min[axis] < root_boxes[b].min[axis]; });
    }
    size_t mid = count == 2 ? lo + 1 : lo + count / 2;
    const BVHNode* l = build_bvh(ids, lo, mid);
    const BVHNode* r = build_bvh(ids, mid, hi);
    n->left = l;
    n->right = r;
    n->box = unite(l->box, r->box);
    return n;
}

// ---------------------------------------------------------------------------------------------
// camera.rs
// ---------------------------------------------------------------------------------------------
struct Camera {
    Vec3 origin, lb, horizontal_full, vertical_full, horizontal_unit, vertical_unit;
    double aperture, shutter_speed;
    int width, height;
    explicit Camera(const rs_camera& c)
        : origin(c.origin[0], c.origin[1], c.origin[2]), lb(c.lb[0], c.lb[1], c.lb[2]),
          horizontal_full(c.horizontal_full[0], c.horizontal_full[1], c.horizontal_full[2]),
          vertical_full(c.vertical_full[0], c.vertical_full[1], c.vertical_full[2]),
          horizontal_unit(c.horizontal_unit[0], c.horizontal_unit[1], c.horizontal_unit[2]),
          vertical_unit(c.vertical_unit[0], c.vertical_unit[1], c.vertical_unit[2]),
          aperture(c.aperture), shutter_speed(c.shutter_speed), width(c.width), height(c.height) {}

    // camera.rs:77-85 + vec3.rs:140-147
    Ray ray(double u, double v, const Rng& rng, double time_xi) const {
        Vec3 p;
        for (uint32_t k = 0;; ++k) {
            double a, b;
            rng.pair(P_LENS, 0, k, a, b);
            p = Vec3(-1.0 + a * (1.0 - -1.0), -1.0 + b * (1.0 - -1.0), 0.0);
            if (length_squared(p) < 1.0) break;
        }
        Vec3 rd = aperture / 2.0 * p;
        Vec3 offset = horizontal_unit * rd.x + vertical_unit * rd.y;
        Vec3 o = origin + offset;
        Vec3 direction = lb + u * horizontal_full + v * vertical_full - o;
        return Ray{o, unit(direction), shutter_speed * time_xi};
    }
};

// camera.rs:94-100
double phong_highlight(const Vec3& dir_to_light, const Vec3& ray_dir, const Vec3& normal, int exponent, double factor) {
    Vec3 reflected = dir_to_light - 2.0 * dot(dir_to_light, normal) * normal;
    double base = std::fmax(dot(reflected, -ray_dir), 0.0);
    double specular = std::pow(base, (double)exponent);  // f64::powi
    return specular * factor;
}

struct Counters {
    uint64_t segments = 0;
};

struct ScatterProbe {  // filled for the first bounce when non-null
    int kind = 0;
    Vec3 next_o, next_d, weight, emitted;
};

// camera.rs:156-255
// pdf.rs:237-269.  `value` is what the reference's type computes today: the `objects.pdf_value` call is
// commented out there (no Hittable has a pdf_value), it evaluates a CosinePdf about the stored normal.
struct HittablePdf : PDF {
    const Scene* world;
    Vec3 origin, normal;
    HittablePdf(const Scene* w, const Vec3& o, const Vec3& n) : world(w), origin(o), normal(n) {}
    double value(const Vec3& direction) const override { return CosinePdf(normal).value(direction); }
    Vec3 generate(const Rng& rng) const override { return unit(world->lights_random(origin, rng)); }
};
// pdf.rs:271-307; the selector draw is word 0 of the P_SCATTER call (the draw of camera.rs:198)
struct MixturePdf : PDF {
    const PDF *p0, *p1;
    MixturePdf(const PDF* a, const PDF* b) : p0(a), p1(b) {}
    double value(const Vec3& direction) const override { return 0.5 * p0->value(direction) + 0.5 * p1->value(direction); }
    Vec3 generate(const Rng& rng) const override {
        uint32_t w[4];
        rng.words(P_SCATTER, 0, w);
        return u01(w[0]) < 0.5 ? p0->generate(rng) : p1->generate(rng);
    }
};

Vec3 ray_color(const Ray& ray, const Scene& world, int depth, Rng rng, Counters& cnt, ScatterProbe* probe, uint32_t flags) {
    if (depth == 0) return Vec3();
    rng.depth += 1;  // shading depth index of the hit that ends this segment
    HitRecord hit;
    int obj_id = -1;
    cnt.segments += 1;
    if (world.hit(ray, Range{0.0001, INF}, rng, hit, obj_id)) {
        const Material* material = hit.material ? hit.material : world.default_material;
        Vec3 emitted;
        if (!material->emitted(hit.u, hit.v, hit.point, emitted)) emitted = Vec3();
        ScatterRecord srec;
        if (material->scatter(ray, hit, rng, 0, srec)) {
            if (srec.skip_pdf) {
                if (srec.has_ray) {
                    if (probe) { probe->kind = 2; probe->next_o = srec.ray.origin; probe->next_d = srec.ray.direction; probe->weight = to_vec3(srec.color); probe->emitted = emitted; return Vec3(); }
                    return emitted + srec.color * ray_color(srec.ray, world, depth - 1, rng, cnt, nullptr, flags);
                }
                return emitted + srec.color * Vec3(1.0, 1.0, 1.0);
            }
            double light_multi = 1.0;
            double pdf_val;
            Ray scattered;
            uint32_t w[4];
            rng.words(P_SCATTER, 0, w);
            bool light_branch = u01(w[0]) < 0.5;
            if (flags & RS_FLAG_MIS) {
                // the integrator the reference keeps commented out (camera.rs:188-192, 222-225): direction from
                // MixturePdf(HittablePdf(lights), srec.pdf), density = the mixture's value, where the second
                // HittablePdf is built with `normal = scattered.direction`; the ray starts at the hit point
                HittablePdf light_pdf(&world, hit.point, hit.normal);
                MixturePdf mixture(&light_pdf, srec.pdf.get());
                Vec3 scatter_direction = mixture.generate(rng);
                scattered = Ray{hit.point, scatter_direction, ray.departure_time};
                HittablePdf light_pdf2(&world, hit.point, scattered.direction);
                MixturePdf mixture2(&light_pdf2, srec.pdf.get());
                pdf_val = mixture2.value(scattered.direction);
            } else if (light_branch) {
                pdf_val = 0.3183098861837907;
                Vec3 dir_to_light = unit(world.lights_random(hit.point, rng));
                Settings settings = material->settings();
                if (settings.phong_factor > 0.0)
                    light_multi += phong_highlight(-dir_to_light, ray.direction, hit.normal, settings.phong_exponent, settings.phong_factor);
                Vec3 start = ray.at(hit.t1 - 0.0002);
                scattered = Ray{start, dir_to_light, ray.departure_time};
            } else {
                Vec3 scatter_direction = srec.pdf->generate(rng);
                pdf_val = srec.pdf->value(scatter_direction);
                scattered = Ray{hit.point, scatter_direction, ray.departure_time};
            }
            if (pdf_val <= 0.0 || pdf_val != pdf_val) pdf_val = 1e-5;
            double scattering_pdf_val = srec.pdf->value(scattered.direction);
            double pdf_multiplicator = scattering_pdf_val / pdf_val;
            if (probe) {
                probe->kind = light_branch ? 3 : 4;
                probe->next_o = scattered.origin; probe->next_d = scattered.direction;
                probe->weight = (srec.color * (light_multi * Vec3(1.0, 1.0, 1.0))) * pdf_multiplicator;
                probe->emitted = emitted;
                return Vec3();
            }
            Vec3 sample_color = light_multi * ray_color(scattered, world, depth - 1, rng, cnt, nullptr, flags);
            Vec3 color_from_scatter = (srec.color * sample_color) * pdf_multiplicator;
            return emitted + color_from_scatter;
        }
        if (probe) { probe->kind = 1; probe->emitted = emitted; }
        return emitted;
    }
    Vec3 bg = to_vec3(world.background(ray));
    if (probe) { probe->kind = 0; probe->emitted = bg; }
    return bg;
}

// painter.rs:133-139, 167-179 — one stratified sample of one pixel
Ray sample_ray(const Camera& cam, const rs_render_params& p, int row, int column, int s_i, int s_j, Rng& rng) {
    rng.pixel = (uint32_t)(row * cam.width + column);
    rng.sample = (uint32_t)(s_j * p.sqrt_spp + s_i);
    rng.depth = 0;
    uint32_t w[4];
    rng.words(P_PIXEL, 0, w);
    double jx = u01(w[0]), jy = u01(w[1]), jt = u01(w[2]);
    if (p.flags & RS_FLAG_PIXEL_CENTRE) { jx = 0.5; jy = 0.5; jt = 0.0; }
    double xo = (double)column + ((double)s_i + jx) / (double)p.sqrt_spp;
    double yo = (double)row + ((double)s_j + jy) / (double)p.sqrt_spp;
    double h = (double)cam.height;
    double u = xo / (double)cam.width;
    double v = (h - 1.0 - yo) / h;
    if (p.flags & RS_FLAG_PIXEL_CENTRE) {
        Camera c2 = cam;
        c2.aperture = 0.0;
        return c2.ray(u, v, rng, jt);
    }
    return cam.ray(u, v, rng, jt);
}

}  // namespace

// =============================================================================================
// C API
// =============================================================================================
extern "C" {

struct rso_world {
    Scene scene;
};

rso_world* rso_world_create(const rs_scene_desc* desc, int use_desc_bbox) {
    auto w = std::make_unique<rso_world>();
    if (!w->scene.build(*desc, use_desc_bbox != 0)) return nullptr;
    return w.release();
}
void rso_world_free(rso_world* w) { delete w; }

// The oracle's own `Hittable::bbox` of every world object (6 doubles each).
int rso_root_bbox(const rso_world* w, double time_start, double time_end, double* out) {
    for (size_t i = 0; i < w->scene.roots.size(); ++i) {
        AABB b = w->scene.roots[i]->bbox(Range{time_start, time_end});
        double* o = out + 6 * i;
        o[0] = b.min.x; o[1] = b.min.y; o[2] = b.min.z; o[3] = b.max.x; o[4] = b.max.y; o[5] = b.max.z;
    }
    return 0;
}

// 0 = canonical Mandelbulb arithmetic (default), 1 = literal libm restatement (see namespace marcher)
int rso_set_marcher_mode(int mode) { marcher::g_mode = mode == 1 ? 1 : 0; return 0; }

// test hooks: the canonical orbit routine and logarithm on arrays of n points / values
int rso_marcher_iterate(const double* xyz, int64_t n, int iterations, double* r_out, double* dr_out, int32_t* inside_out) {
    for (int64_t i = 0; i < n; ++i) {
        bool in;
        marcher::iterate(Vec3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]), iterations, r_out[i], dr_out[i], in);
        inside_out[i] = in ? 1 : 0;
    }
    return 0;
}
int rso_marcher_log(const double* x, int64_t n, double* out) {
    for (int64_t i = 0; i < n; ++i) out[i] = marcher::det_log(x[i]);
    return 0;
}

// test hook: Texture::color of texture `tex` at n points (u, v, p.xyz as 5 doubles each) -> n x rgba f32
int rso_texture_color(const rso_world* w, int32_t tex, const double* uvp, int64_t n, float* out) {
    if (tex < 0 || tex >= (int)w->scene.textures.size()) return RS_ERR_INVALID_ARGUMENT;
    for (int64_t i = 0; i < n; ++i) {
        const double* q = uvp + 5 * i;
        Color c = w->scene.textures[tex]->color(q[0], q[1], Vec3(q[2], q[3], q[4]));
        out[4 * i] = c.r; out[4 * i + 1] = c.g; out[4 * i + 2] = c.b; out[4 * i + 3] = c.a;
    }
    return RS_OK;
}

int rso_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    philox4x32_10(ctr, key, out);
    return 0;
}

// Painter::draw (painter.rs:256-336): T threads, thread i renders rows i, i+T, ...
int rso_render(const rso_world* w, const rs_camera* cam_in, const rs_render_params* p, const uint8_t* mask, float* out,
               int n_threads, rs_stats* stats) {
    if (w->scene.lights.empty()) return RS_ERR_INVALID_ARGUMENT;  // the reference panics (irange % 0)
    Camera cam(*cam_in);
    const int W = cam.width, H = cam.height;
    int T = n_threads > 0 ? n_threads : (int)std::thread::hardware_concurrency() + 1;  // painter.rs:321-325
    std::atomic<uint64_t> total_segments{0}, total_samples{0};
    auto t0 = std::chrono::steady_clock::now();
    auto render_rows = [&](int start) {
        Counters cnt;
        uint64_t samples = 0;
        Rng rng;
        rng.key[0] = (uint32_t)p->seed; rng.key[1] = (uint32_t)(p->seed >> 32);
        for (int row = start; row < H; row += T) {
            for (int col = 0; col < W; ++col) {
                float* px = out + 4 * ((size_t)row * W + col);
                if (mask && !mask[(size_t)row * W + col]) { px[0] = px[1] = px[2] = px[3] = 0.f; continue; }
                Vec3 color_vec;
                for (int s_j = p->stratum_row_begin; s_j < p->stratum_row_end; ++s_j)
                    for (int s_i = 0; s_i < p->sqrt_spp; ++s_i) {
                        Ray r = sample_ray(cam, *p, row, col, s_i, s_j, rng);
                        color_vec = color_vec + ray_color(r, w->scene, p->depth, rng, cnt, nullptr, p->flags);
                        ++samples;
                    }
                // Vec3::into_color (vec3.rs:227-240)
                double n = (double)((p->stratum_row_end - p->stratum_row_begin) * p->sqrt_spp);
                double cx = color_vec.x / n, cy = color_vec.y / n, cz = color_vec.z / n;
                if (p->gamma) { cx = std::sqrt(cx); cy = std::sqrt(cy); cz = std::sqrt(cz); }
                px[0] = (float)cx; px[1] = (float)cy; px[2] = (float)cz; px[3] = 1.0f;
            }
        }
        total_segments += cnt.segments;
        total_samples += samples;
    };
    std::vector<std::thread> threads;
    for (int i = 0; i < T; ++i) threads.emplace_back(render_rows, i);
    for (auto& t : threads) t.join();
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        stats->n_samples = total_samples;
        stats->n_segments = total_segments;
        stats->ms_total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    return RS_OK;
}

int rso_debug_primary(const rso_world* w, const rs_camera* cam_in, const rs_render_params* p, int32_t sample_index,
                      int32_t* obj_id, double* t, double* normal3, double* point3, int32_t* front, int32_t* material,
                      double* ray_od6) {
    Camera cam(*cam_in);
    Rng rng;
    rng.key[0] = (uint32_t)p->seed; rng.key[1] = (uint32_t)(p->seed >> 32);
    const int s_j = sample_index / p->sqrt_spp, s_i = sample_index % p->sqrt_spp;
    for (int row = 0; row < cam.height; ++row)
        for (int col = 0; col < cam.width; ++col) {
            size_t i = (size_t)row * cam.width + col;
            Ray r = sample_ray(cam, *p, row, col, s_i, s_j, rng);
            Rng hr = rng;
            hr.depth = 1;
            HitRecord hit;
            int id = -1;
            bool has = w->scene.hit(r, Range{0.0001, INF}, hr, hit, id);
            if (ray_od6) { double* o = ray_od6 + 6 * i; o[0] = r.origin.x; o[1] = r.origin.y; o[2] = r.origin.z; o[3] = r.direction.x; o[4] = r.direction.y; o[5] = r.direction.z; }
            if (obj_id) obj_id[i] = has ? id : -1;
            if (t) t[i] = has ? hit.t1 : 0.0;
            if (normal3) { normal3[3 * i] = has ? hit.normal.x : 0; normal3[3 * i + 1] = has ? hit.normal.y : 0; normal3[3 * i + 2] = has ? hit.normal.z : 0; }
            if (point3) { point3[3 * i] = has ? hit.point.x : 0; point3[3 * i + 1] = has ? hit.point.y : 0; point3[3 * i + 2] = has ? hit.point.z : 0; }
            if (front) front[i] = has ? (hit.outside ? 1 : 0) : 0;
            if (material) material[i] = has ? (hit.material ? hit.material->index : w->scene.default_material->index) : -1;
        }
    return RS_OK;
}

int rso_debug_scatter(const rso_world* w, const rs_camera* cam_in, const rs_render_params* p, int32_t sample_index,
                      int32_t* kind, double* next_od6, double* weight3, double* emitted3) {
    if (w->scene.lights.empty()) return RS_ERR_INVALID_ARGUMENT;
    Camera cam(*cam_in);
    Rng rng;
    rng.key[0] = (uint32_t)p->seed; rng.key[1] = (uint32_t)(p->seed >> 32);
    const int s_j = sample_index / p->sqrt_spp, s_i = sample_index % p->sqrt_spp;
    Counters cnt;
    for (int row = 0; row < cam.height; ++row)
        for (int col = 0; col < cam.width; ++col) {
            size_t i = (size_t)row * cam.width + col;
            Ray r = sample_ray(cam, *p, row, col, s_i, s_j, rng);
            ScatterProbe probe;
            ray_color(r, w->scene, p->depth, rng, cnt, &probe, p->flags);
            if (kind) kind[i] = probe.kind;
            if (next_od6) { double* o = next_od6 + 6 * i; o[0] = probe.next_o.x; o[1] = probe.next_o.y; o[2] = probe.next_o.z; o[3] = probe.next_d.x; o[4] = probe.next_d.y; o[5] = probe.next_d.z; }
            if (weight3) { weight3[3 * i] = probe.weight.x; weight3[3 * i + 1] = probe.weight.y; weight3[3 * i + 2] = probe.weight.z; }
            if (emitted3) { emitted3[3 * i] = probe.emitted.x; emitted3[3 * i + 1] = probe.emitted.y; emitted3[3 * i + 2] = probe.emitted.z; }
        }
    return RS_OK;
}

// src/bin/raysnail.rs:176-208
int rso_combine_passes(const float* old_px, const float* new_px, float* out, int64_t n, float pass) {
    for (int64_t i = 0; i < n; ++i) {
        const float* o = old_px + 4 * i;
        const float* nw = new_px + 4 * i;
        float* r = out + 4 * i;
        if (nw[0] == 0.0f && nw[1] == 0.0f && nw[2] == 0.0f && nw[3] == 0.0f) {
            r[0] = o[0]; r[1] = o[1]; r[2] = o[2]; r[3] = o[3];
        } else {
            float d = pass + 1.0f;
            r[0] = (o[0] * pass + nw[0]) / d; r[1] = (o[1] * pass + nw[1]) / d;
            r[2] = (o[2] * pass + nw[2]) / d; r[3] = (o[3] * pass + nw[3]) / d;
        }
    }
    return RS_OK;
}

// src/bin/raysnail.rs:138-173 (quirk != 0 keeps `let x = y as i32;`)
int rso_noise_map(const float* px, float* noise, int32_t width, int32_t height, int32_t quirk) {
    for (int y0 = 0; y0 < height; ++y0)
        for (int x0 = 0; x0 < width; ++x0) {
            const float* def = px + 4 * ((size_t)y0 * width + x0);
            float diff = 0.0f;
            int x = quirk ? y0 : x0, y = y0;
            for (int yy = y - 2; yy < y + 3; ++yy)
                for (int xx = x - 2; xx < x + 3; ++xx) {
                    const float* q = (xx >= 0 && yy >= 0 && xx < width && yy < height) ? px + 4 * ((size_t)yy * width + xx) : def;
                    float rd = def[0] - q[0], gd = def[1] - q[1], bd = def[2] - q[2];
                    diff += rd * rd + gd * gd + bd * bd;
                }
            noise[(size_t)y0 * width + x0] = diff;
        }
    return RS_OK;
}

// src/bin/raysnail.rs:379-427: the passes loop.  redo_mode as RS_REDO_* of include/raysnail_b200.h; pass k
// renders with key seed + k * 0x9E3779B97F4A7C15 (the reference's RNG is freshly seeded per pass).
int rso_render_passes(const rso_world* w, const rs_camera* cam, const rs_render_params* p, int32_t n_passes, int32_t redo_mode,
                      float threshold, int n_threads, float* out, int64_t* redo_counts) {
    const size_t np = (size_t)cam->width * cam->height;
    std::vector<float> old_px(4 * np), new_px(4 * np), comb(4 * np), noise(np);
    std::vector<uint8_t> redo(np, 1);
    for (size_t i = 0; i < np; ++i) { old_px[4 * i] = 0.f; old_px[4 * i + 1] = 0.f; old_px[4 * i + 2] = 0.f; old_px[4 * i + 3] = 1.f; }  // :318-326
    for (int pass = 0; pass < n_passes; ++pass) {
        rs_render_params pp = *p;
        pp.seed = p->seed + (uint64_t)(int64_t)pass * 0x9E3779B97F4A7C15ull;
        // :369-372: the controller holds a clone of the initial all-ones map; the opt-in modes use the live one
        const uint8_t* mask = (redo_mode != RS_REDO_REFERENCE && pass > 0) ? redo.data() : nullptr;
        int rc = rso_render(w, cam, &pp, mask, new_px.data(), n_threads, nullptr);
        if (rc != RS_OK) return rc;
        rso_combine_passes(old_px.data(), new_px.data(), comb.data(), (int64_t)np, (float)pass);  // :388
        rso_noise_map(comb.data(), noise.data(), cam->width, cam->height, redo_mode == RS_REDO_NOISE_FIXED ? 0 : 1);
        int64_t count = 0;
        for (size_t i = 0; i < np; ++i) { redo[i] = noise[i] >= threshold ? 1 : 0; count += redo[i]; }  // :408-423
        if (redo_counts) redo_counts[pass] = count;
        old_px.swap(comb);  // :426
    }
    std::memcpy(out, old_px.data(), 4 * np * sizeof(float));
    return RS_OK;
}

// src/bin/raysnail.rs:437-441 + prelude/mod.rs:24-39
int rso_quantize_rgb8(const float* px, uint8_t* rgb, int64_t n) {
    for (int64_t i = 0; i < n; ++i)
        for (int c = 0; c < 3; ++c) {
            double v = (double)px[4 * i + c];
            double cl = (0.0 > v) ? 0.0 : (v > 1.0 ? 1.0 : v);  // clamp(): NaN falls through unchanged
            double s = cl * 255.5;
            rgb[3 * i + c] = (s != s) ? 0 : (uint8_t)(s < 0.0 ? 0.0 : (s > 255.0 ? 255.0 : s));  // `as u8` saturates, NaN -> 0
        }
    return RS_OK;
}

}  // extern "C"
