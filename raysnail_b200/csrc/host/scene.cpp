// Host-side scene graph -> flat rs_scene_desc.  See scene.hpp for the mirrored interfaces.
#include "scene.hpp"

#include <algorithm>
#include <cstring>
#include <fstream>
#include <limits>
#include <sstream>
#include <stdexcept>

namespace raysnail {

// ---------------------------------------------------------------------------------------------
// Transform (transform.rs:16-110).  m is row-major; inverse via the general cofactor expansion
// (what vecmath::mat4_inv does: adjugate times reciprocal determinant).
// ---------------------------------------------------------------------------------------------
static void mat4_id(double* m) {
    for (int i = 0; i < 16; ++i) m[i] = (i % 5 == 0) ? 1.0 : 0.0;
}

static void mat4_inv(const double* m, double* out) {
    double inv[16];
    inv[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] + m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
    inv[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] - m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
    inv[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] + m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
    inv[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] - m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
    inv[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] - m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
    inv[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] + m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
    inv[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] - m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
    inv[13] = m[0] * m[9] * m[14] - m[0] * m[10] * m[13] - m[8] * m[1] * m[14] + m[8] * m[2] * m[13] + m[12] * m[1] * m[10] - m[12] * m[2] * m[9];
    inv[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] + m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
    inv[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] - m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
    inv[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] + m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
    inv[14] = -m[0] * m[5] * m[14] + m[0] * m[6] * m[13] + m[4] * m[1] * m[14] - m[4] * m[2] * m[13] - m[12] * m[1] * m[6] + m[12] * m[2] * m[5];
    inv[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] - m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
    inv[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] + m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
    inv[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] - m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
    inv[15] = m[0] * m[5] * m[10] - m[0] * m[6] * m[9] - m[4] * m[1] * m[10] + m[4] * m[2] * m[9] + m[8] * m[1] * m[6] - m[8] * m[2] * m[5];
    double det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
    double inv_det = 1.0 / det;
    for (int i = 0; i < 16; ++i) out[i] = inv[i] * inv_det;
}

static Transform finish_transform(const double* m) {
    Transform t;
    std::memcpy(t.matrix, m, sizeof(t.matrix));
    mat4_inv(m, t.inverse);
    return t;
}

Transform Transform::translate(const Vec3& t) {
    double m[16];
    mat4_id(m);
    m[0 * 4 + 3] = t.x;
    m[1 * 4 + 3] = t.y;
    m[2 * 4 + 3] = t.z;
    return finish_transform(m);
}
Transform Transform::rotate_by_x_axis(double theta) {
    double s = std::sin(theta), c = std::cos(theta), m[16];
    mat4_id(m);
    m[1 * 4 + 1] = c;
    m[1 * 4 + 2] = s;
    m[2 * 4 + 1] = -s;
    m[2 * 4 + 2] = c;
    return finish_transform(m);
}
Transform Transform::rotate_by_y_axis(double theta) {
    double s = std::sin(theta), c = std::cos(theta), m[16];
    mat4_id(m);
    m[0 * 4 + 0] = c;
    m[0 * 4 + 2] = s;
    m[2 * 4 + 0] = -s;
    m[2 * 4 + 2] = c;
    return finish_transform(m);
}
Transform Transform::rotate_by_z_axis(double theta) {
    double s = std::sin(theta), c = std::cos(theta), m[16];
    mat4_id(m);
    m[0 * 4 + 0] = c;
    m[0 * 4 + 1] = s;
    m[1 * 4 + 0] = -s;
    m[1 * 4 + 1] = c;
    return finish_transform(m);
}
Transform Transform::scale(const Vec3& s) {
    double m[16];
    mat4_id(m);
    m[0] = s.x;
    m[5] = s.y;
    m[10] = s.z;
    m[15] = 1.0;
    return finish_transform(m);
}

// row_mat4_transform: out[r] = m[r][0]*v0 + m[r][1]*v1 + m[r][2]*v2 + m[r][3]*v3
static Vec3 row_mat4_transform3(const double* m, const Vec3& p, double w) {
    return Vec3(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3] * w,
                m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7] * w,
                m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11] * w);
}

Vec3 TransformStack::forward(const Vec3& pos, double w) const {
    Vec3 r = pos;
    for (const Transform& t : stack) r = row_mat4_transform3(t.matrix, r, w);
    return r;
}
Vec3 TransformStack::inverse(const Vec3& pos, double w) const {
    Vec3 r = pos;
    for (auto it = stack.rbegin(); it != stack.rend(); ++it) r = row_mat4_transform3(it->inverse, r, w);
    return r;
}

// ---------------------------------------------------------------------------------------------
// Textures
// ---------------------------------------------------------------------------------------------
static rs_texture blank_texture(int type) {
    rs_texture t;
    std::memset(&t, 0, sizeof(t));
    t.type = type;
    t.odd = t.even = -1;
    return t;
}

template <class F>
static int cached(std::map<const void*, int>& cache, const void* key, F&& make) {
    auto it = cache.find(key);
    if (it != cache.end()) return it->second;
    int id = make();
    cache[key] = id;
    return id;
}

int SolidColor::flatten(FlatScene& fs) const {
    return cached(fs.texture_cache, this, [&] {
        rs_texture t = blank_texture(RS_TEX_SOLID);
        t.color[0] = color.r; t.color[1] = color.g; t.color[2] = color.b; t.color[3] = color.a;
        fs.textures.push_back(t);
        return (int)fs.textures.size() - 1;
    });
}

int Checker::flatten(FlatScene& fs) const {
    return cached(fs.texture_cache, this, [&] {
        int o = odd->flatten(fs), e = even->flatten(fs);
        rs_texture t = blank_texture(RS_TEX_CHECKER);
        t.odd = o; t.even = e; t.scale = scale;
        fs.textures.push_back(t);
        return (int)fs.textures.size() - 1;
    });
}

int Image::flatten(FlatScene& fs) const {
    return cached(fs.texture_cache, this, [&] {
        rs_texture t = blank_texture(RS_TEX_IMAGE);
        t.width = width; t.height = height;
        t.texel_offset = (int32_t)fs.texels.size();
        fs.texels.insert(fs.texels.end(), rgba.begin(), rgba.end());
        fs.textures.push_back(t);
        return (int)fs.textures.size() - 1;
    });
}

static uint64_t splitmix64(uint64_t& s) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

Perlin::Perlin(int pc, bool vec, uint64_t seed) : vector(vec), point_count(pc) {
    if (pc <= 0 || (pc & (pc - 1)) != 0) throw std::invalid_argument("Perlin point_count must be a power of two");
    uint64_t s = seed;
    auto gen = [&]() { return (double)(splitmix64(s) >> 11) * (1.0 / 9007199254740992.0); };
    const double PI = 3.14159265358979323846;
    for (int i = 0; i < pc; ++i) {
        if (vec) {  // Vec3::random_unit (vec3.rs:91-96)
            double a = gen() * (2.0 * PI), z = -1.0 + gen() * 2.0, r = std::sqrt(1.0 - z * z);
            random_values.push_back(r * std::cos(a));
            random_values.push_back(r * std::sin(a));
            random_values.push_back(z);
        } else {
            random_values.push_back(gen());
        }
    }
    for (int axis = 0; axis < 3; ++axis) {
        std::vector<int32_t> p(pc);
        for (int i = 0; i < pc; ++i) p[i] = i;
        for (int i = pc - 1; i > 0; --i) std::swap(p[i], p[(size_t)(splitmix64(s) % (uint64_t)(i + 1))]);
        perm.insert(perm.end(), p.begin(), p.end());
    }
}

int Perlin::flatten(FlatScene& fs) const {
    return cached(fs.texture_cache, this, [&] {
        rs_texture t = blank_texture(RS_TEX_PERLIN);
        t.perlin_kind = kind; t.perlin_depth = depth; t.perlin_smooth = (int)smooth_type;
        t.perlin_vector = vector ? 1 : 0; t.point_count = point_count; t.scale = scale_;
        t.perm_offset = (int32_t)fs.perlin_perm.size();
        t.value_offset = (int32_t)fs.perlin_values.size();
        fs.perlin_perm.insert(fs.perlin_perm.end(), perm.begin(), perm.end());
        fs.perlin_values.insert(fs.perlin_values.end(), random_values.begin(), random_values.end());
        fs.textures.push_back(t);
        return (int)fs.textures.size() - 1;
    });
}

// ---------------------------------------------------------------------------------------------
// Materials
// ---------------------------------------------------------------------------------------------
static rs_material blank_material(int type, const CommonMaterialSettings& s) {
    rs_material m;
    std::memset(&m, 0, sizeof(m));
    m.type = type;
    m.texture = m.mat1 = m.mat2 = -1;
    m.phong_factor = s.phong_factor;
    m.phong_exponent = s.phong_exponent;
    return m;
}

static int push_material(FlatScene& fs, const rs_material& m) {
    fs.materials.push_back(m);
    return (int)fs.materials.size() - 1;
}

int Lambertian::flatten(FlatScene& fs) const {
    return cached(fs.material_cache, this, [&] {
        rs_material m = blank_material(RS_MAT_LAMBERTIAN, settings);
        m.texture = texture->flatten(fs);
        return push_material(fs, m);
    });
}
int Metal::flatten(FlatScene& fs) const {
    return cached(fs.material_cache, this, [&] {
        rs_material m = blank_material(RS_MAT_METAL, settings);
        m.texture = texture->flatten(fs);
        return push_material(fs, m);
    });
}
int DiffuseMetal::flatten(FlatScene& fs) const {
    return cached(fs.material_cache, this, [&] {
        rs_material m = blank_material(RS_MAT_DIFFUSE_METAL, settings);
        m.texture = texture->flatten(fs);
        m.a = exponent;
        return push_material(fs, m);
    });
}
int Dielectric::flatten(FlatScene& fs) const {
    return cached(fs.material_cache, this, [&] {
        rs_material m = blank_material(RS_MAT_DIELECTRIC, settings);
        m.a = refractive;        // outer_refractive
        m.b = 1.0 / refractive;  // enter_refractive (dielectric.rs:39-40)
        m.has_curve = glass_curve ? 1 : 0;
        m.color[0] = color.r; m.color[1] = color.g; m.color[2] = color.b; m.color[3] = color.a;
        return push_material(fs, m);
    });
}
int DiffuseLight::flatten(FlatScene& fs) const {
    return cached(fs.material_cache, this, [&] {
        rs_material m = blank_material(RS_MAT_DIFFUSE_LIGHT, settings);
        m.texture = texture->flatten(fs);
        m.a = multiplier_;
        return push_material(fs, m);
    });
}
int Isotropic::flatten(FlatScene& fs) const {
    return cached(fs.material_cache, this, [&] {
        rs_material m = blank_material(RS_MAT_ISOTROPIC, settings);
        m.color[0] = color.r; m.color[1] = color.g; m.color[2] = color.b; m.color[3] = color.a;
        return push_material(fs, m);
    });
}
int BlinnPhong::flatten(FlatScene& fs) const {
    return cached(fs.material_cache, this, [&] {
        rs_material m = blank_material(RS_MAT_BLINN_PHONG, settings);
        m.texture = texture->flatten(fs);
        m.a = k_specular;
        m.b = exponent;
        return push_material(fs, m);
    });
}
int MixedMaterial::flatten(FlatScene& fs) const {
    return cached(fs.material_cache, this, [&] {
        int m1 = material_1->flatten(fs), m2 = material_2->flatten(fs);
        // MixedMaterial::settings() returns material_1's settings (mixed_material.rs:56-58)
        rs_material m = blank_material(RS_MAT_MIXED, effective_settings());
        m.mat1 = m1; m.mat2 = m2; m.a = probability_1;
        return push_material(fs, m);
    });
}

// ---------------------------------------------------------------------------------------------
// Hittables: bbox rules + flatten
// ---------------------------------------------------------------------------------------------
static rs_node blank_node(int type) {
    rs_node n;
    std::memset(&n, 0, sizeof(n));
    n.type = type;
    n.material = n.child0 = n.child1 = -1;
    return n;
}
static int push_node(FlatScene& fs, const rs_node& n) {
    fs.nodes.push_back(n);
    return (int)fs.nodes.size() - 1;
}

// sphere.rs:117-142
AABB Sphere::bbox(const TimeRange& tr) const {
    Vec3 rv(radius, radius, radius);
    if (speed.x == 0.0 && speed.y == 0.0 && speed.z == 0.0) return AABB(center - rv, center + rv);
    AABB start(center_at(tr.start) - rv, center_at(tr.start) + rv);
    AABB end(center_at(tr.end) - rv, center_at(tr.end) + rv);
    return start | end;
}
int Sphere::flatten(FlatScene& fs) const {
    return cached(fs.node_cache, this, [&] {
        rs_node n = blank_node(RS_NODE_SPHERE);
        n.material = fs.material_index(material);
        n.p[0] = center.x; n.p[1] = center.y; n.p[2] = center.z; n.p[3] = radius;
        n.p[4] = speed.x; n.p[5] = speed.y; n.p[6] = speed.z;
        return push_node(fs, n);
    });
}

AARectMetrics::AARectMetrics(double k_, double a0_, double a1_, double b0_, double b1_)
    : k(k_), a0(a0_), a1(a1_), b0(b0_), b1(b1_) {
    // rect.rs:27-28 asserts; a violated assert aborts the reference (panic = "abort")
    if (!(a0 < a1) || !(b0 < b1)) throw std::invalid_argument("AARectMetrics: need a0 < a1 and b0 < b1");
}

static void rect_axes(int kind, int& a, int& b, int& c) {
    if (kind == RS_RECT_XY) { a = 0; b = 1; c = 2; }
    else if (kind == RS_RECT_XZ) { a = 0; b = 2; c = 1; }
    else { a = 1; b = 2; c = 0; }
}

// rect.rs:126-139
AABB AARect::bbox(const TimeRange&) const {
    int a, b, c;
    rect_axes(axis_kind, a, b, c);
    Vec3 p0, p1;
    p0[a] = metrics.a0; p0[b] = metrics.b0; p0[c] = metrics.k - 0.0001;
    p1[a] = metrics.a1; p1[b] = metrics.b1; p1[c] = metrics.k + 0.0001;
    return AABB(p0, p1);
}
int AARect::flatten(FlatScene& fs) const {
    return cached(fs.node_cache, this, [&] {
        rs_node n = blank_node(RS_NODE_RECT);
        n.material = fs.material_index(material);
        n.rect_axis = axis_kind;
        n.p[0] = metrics.k; n.p[1] = metrics.a0; n.p[2] = metrics.a1; n.p[3] = metrics.b0; n.p[4] = metrics.b1;
        return push_node(fs, n);
    });
}

// box.rs:161-163: union of the six face boxes (each padded +-1e-4 on its fixed axis)
AABB Box::bbox(const TimeRange& tr) const {
    const Vec3 &mn = point_min, &mx = point_max;
    AARect faces[6] = {
        AARect::new_xy(AARectMetrics(mn.z, mn.x, mx.x, mn.y, mx.y), nullptr),
        AARect::new_xy(AARectMetrics(mx.z, mn.x, mx.x, mn.y, mx.y), nullptr),
        AARect::new_yz(AARectMetrics(mn.x, mn.y, mx.y, mn.z, mx.z), nullptr),
        AARect::new_yz(AARectMetrics(mx.x, mn.y, mx.y, mn.z, mx.z), nullptr),
        AARect::new_xz(AARectMetrics(mn.y, mn.x, mx.x, mn.z, mx.z), nullptr),
        AARect::new_xz(AARectMetrics(mx.y, mn.x, mx.x, mn.z, mx.z), nullptr),
    };
    AABB r = faces[0].bbox(tr);
    for (int i = 1; i < 6; ++i) r = r | faces[i].bbox(tr);
    return r;
}
int Box::flatten(FlatScene& fs) const {
    return cached(fs.node_cache, this, [&] {
        rs_node n = blank_node(RS_NODE_BOX);
        n.material = fs.material_index(material);
        n.p[0] = point_min.x; n.p[1] = point_min.y; n.p[2] = point_min.z;
        n.p[3] = point_max.x; n.p[4] = point_max.y; n.p[5] = point_max.z;
        return push_node(fs, n);
    });
}

// quadric.rs:191-196
AABB Quadric::bbox(const TimeRange&) const { return AABB(Vec3(-100, -100, -100), Vec3(100, 100, 100)); }
int Quadric::flatten(FlatScene& fs) const {
    return cached(fs.node_cache, this, [&] {
        rs_node n = blank_node(RS_NODE_QUADRIC);
        n.material = fs.material_index(material);
        for (int i = 0; i < 10; ++i) n.p[i] = q[i];
        return push_node(fs, n);
    });
}

// raymarching.rs:172-181
AABB RayMarcher::bbox(const TimeRange&) const { return AABB(Vec3(-1.3, -1.3, -1.3), Vec3(1.3, 1.3, 1.3)); }
int RayMarcher::flatten(FlatScene& fs) const {
    return cached(fs.node_cache, this, [&] {
        rs_node n = blank_node(RS_NODE_RAYMARCHER);
        n.material = fs.material_index(material);
        n.p[0] = (double)iterations;
        return push_node(fs, n);
    });
}

// difference.rs:114-116
AABB Difference::bbox(const TimeRange& tr) const { return plus->bbox(tr); }
int Difference::flatten(FlatScene& fs) const {
    return cached(fs.node_cache, this, [&] {
        int a = plus->flatten(fs), b = minus->flatten(fs);
        if (a < 0 || b < 0) throw std::invalid_argument("triangles cannot be CSG operands");
        rs_node n = blank_node(RS_NODE_DIFFERENCE);
        n.material = fs.material_index(material);
        n.child0 = a; n.child1 = b;
        return push_node(fs, n);
    });
}

// intersection.rs:102-118 — including `min.z = max(b1.min.y, b2.min.z)` as written there
AABB Intersection::bbox(const TimeRange& tr) const {
    AABB b1 = o1->bbox(tr), b2 = o2->bbox(tr);
    Vec3 mn(std::fmax(b1.min.x, b2.min.x), std::fmax(b1.min.y, b2.min.y), std::fmax(b1.min.y, b2.min.z));
    Vec3 mx(std::fmin(b1.max.x, b2.max.x), std::fmin(b1.max.y, b2.max.y), std::fmin(b1.max.z, b2.max.z));
    return AABB(mn, mx);
}
int Intersection::flatten(FlatScene& fs) const {
    return cached(fs.node_cache, this, [&] {
        int a = o1->flatten(fs), b = o2->flatten(fs);
        if (a < 0 || b < 0) throw std::invalid_argument("triangles cannot be CSG operands");
        rs_node n = blank_node(RS_NODE_INTERSECTION);
        n.material = fs.material_index(material);
        n.child0 = a; n.child1 = b;
        return push_node(fs, n);
    });
}

// bvh.rs:47-112
std::shared_ptr<BVH> BVH::build(std::vector<HittablePtr> objects, const TimeRange& tr) {
    if (objects.empty()) return std::make_shared<BVH>();  // Self::default(): no bbox, never hit
    for (const auto& o : objects)
        if (!o) throw std::invalid_argument("BVH: null object");
    return build_range(objects, 0, objects.size(), tr);
}
std::shared_ptr<BVH> BVH::build_range(std::vector<HittablePtr>& objects, size_t begin, size_t end, const TimeRange& tr) {
    auto node = std::make_shared<BVH>();
    const size_t count = end - begin;
    node->has_bbox = true;
    if (count == 1) {
        node->left = objects[begin];
        node->box = node->left->bbox(tr);
    } else if (count == 2) {
        node->left = objects[begin];
        node->right = objects[begin + 1];
        node->box = node->left->bbox(tr) | node->right->bbox(tr);
    } else {
        const int axis = (int)((begin + count) & 1);
        const TimeRange zero{0.0, 0.0};  // cmp_geometry_by compares the boxes at time 0..0 (bvh.rs:29-30)
        auto key = [&](const HittablePtr& o) { const AABB b = o->bbox(zero); return axis == 0 ? b.min.x : b.min.y; };
        std::stable_sort(objects.begin() + (long)begin, objects.begin() + (long)end, [&](const HittablePtr& a, const HittablePtr& b) { return key(a) < key(b); });
        const size_t mid = begin + count / 2;
        auto l = build_range(objects, begin, mid, tr), r = build_range(objects, mid, end, tr);
        node->box = l->box | r->box;
        node->left = l;
        node->right = r;
    }
    return node;
}
AABB BVH::bbox(const TimeRange&) const { return box; }
int BVH::flatten(FlatScene& fs) const {
    return cached(fs.node_cache, this, [&] {
        const int a = left ? left->flatten(fs) : -1, b = right ? right->flatten(fs) : -1;
        if ((left && a < 0) || (right && b < 0)) throw std::invalid_argument("triangles cannot be members of a nested BVH");
        rs_node n = blank_node(RS_NODE_BVH);
        n.child0 = a; n.child1 = b;
        n.p[0] = box.min.x; n.p[1] = box.min.y; n.p[2] = box.min.z;
        n.p[3] = box.max.x; n.p[4] = box.max.y; n.p[5] = box.max.z;
        n.p[6] = has_bbox ? 1.0 : 0.0;
        return push_node(fs, n);
    });
}

// tf_facade.rs:59-88
AABB TfFacade::bbox(const TimeRange& tr) const {
    AABB bb = object->bbox(tr);
    const double inf = std::numeric_limits<double>::infinity();
    Vec3 pmin(inf, inf, inf), pmax(-inf, -inf, -inf);
    for (int i = 0; i < 2; ++i)
        for (int j = 0; j < 2; ++j)
            for (int k = 0; k < 2; ++k) {
                double x = std::fma((double)i, bb.max.x, (double)(1 - i) * bb.min.x);
                double y = std::fma((double)j, bb.max.y, (double)(1 - j) * bb.min.y);
                double z = std::fma((double)k, bb.max.z, (double)(1 - k) * bb.min.z);
                Vec3 p = stack.forward(Vec3(x, y, z), 1.0);
                for (int c = 0; c < 3; ++c) {
                    pmin[c] = std::fmin(pmin[c], p[c]);
                    pmax[c] = std::fmax(pmax[c], p[c]);
                }
            }
    return AABB(pmin, pmax);
}
int TfFacade::flatten(FlatScene& fs) const {
    return cached(fs.node_cache, this, [&] {
        int c = object->flatten(fs);
        if (c < 0) throw std::invalid_argument("triangles cannot be wrapped in TfFacade");
        rs_node n = blank_node(RS_NODE_TRANSFORM);
        n.child0 = c;
        n.first_xform = (int32_t)fs.xforms.size();
        n.n_xform = (int32_t)stack.len();
        for (const Transform& t : stack.stack) {
            rs_xform x;
            std::memcpy(x.m, t.matrix, sizeof(x.m));
            std::memcpy(x.inv, t.inverse, sizeof(x.inv));
            fs.xforms.push_back(x);
        }
        return push_node(fs, n);
    });
}

// constant.rs:88-90
AABB ConstantMedium::bbox(const TimeRange& tr) const { return boundary->bbox(tr); }
int ConstantMedium::flatten(FlatScene& fs) const {
    return cached(fs.node_cache, this, [&] {
        int c = boundary->flatten(fs);
        if (c < 0) throw std::invalid_argument("triangles cannot bound a ConstantMedium");
        rs_node n = blank_node(RS_NODE_MEDIUM);
        n.child0 = c;
        n.material = material->flatten(fs);
        n.p[0] = density;
        return push_node(fs, n);
    });
}

// triangle_mesh.rs:42-45
AABB Triangle::bbox(const TimeRange&) const {
    return AABB(Vec3::new_min(Vec3::new_min(p0, p1), p2), Vec3::new_max(Vec3::new_max(p0, p1), p2));
}
int Triangle::flatten(FlatScene& fs) const {
    int idx = (int)fs.tri_material.size();
    const Vec3* v[3] = {&p0, &p1, &p2};
    const Vec3* nn[3] = {&normal0, &normal1, &normal2};
    for (int i = 0; i < 3; ++i) {
        fs.tri_verts.push_back(v[i]->x); fs.tri_verts.push_back(v[i]->y); fs.tri_verts.push_back(v[i]->z);
        fs.tri_normals.push_back(nn[i]->x); fs.tri_normals.push_back(nn[i]->y); fs.tri_normals.push_back(nn[i]->z);
    }
    fs.tri_material.push_back(fs.material_index(material));
    return ~idx;
}

// triangle_mesh.rs:186-276
TriangleMesh TriangleMesh::from_arrays(const std::vector<float>& positions, const std::vector<float>& normals,
                                       const std::vector<uint32_t>& indices, double scale, const Vec3& offset,
                                       double rotation_angle, int axis, MaterialPtr material) {
    TriangleMesh mesh;
    const double rad = rotation_angle * (3.14159265358979323846 / 180.0);  // f64::to_radians
    const double cos = std::cos(rad), sin = std::sin(rad);
    const size_t n_tri = indices.size() / 3;
    // The reference sizes v_normal by the triangle count but indexes it by vertex id
    // (triangle_mesh.rs:196,227-229) and would panic when a vertex id exceeds it; size it safely.
    std::vector<Vec3> v_normal(std::max(n_tri, positions.size() / 3));
    auto pos = [&](uint32_t i) { return Vec3((double)positions[3 * i], (double)positions[3 * i + 1], (double)positions[3 * i + 2]); };
    mesh.triangles.reserve(n_tri);
    for (size_t i = 0; i < n_tri; ++i) {
        uint32_t i0 = indices[3 * i], i1 = indices[3 * i + 1], i2 = indices[3 * i + 2];
        Vec3 p0 = pos(i0).rotate(axis, cos, sin), p1 = pos(i1).rotate(axis, cos, sin), p2 = pos(i2).rotate(axis, cos, sin);
        if (normals.empty()) {
            Vec3 normal = (p1 - p0).cross(p2 - p0).unit();
            v_normal[i0] = v_normal[i0] + normal;
            v_normal[i1] = v_normal[i1] + normal;
            v_normal[i2] = v_normal[i2] + normal;
        }
        mesh.triangles.emplace_back(p0 * scale + offset, p1 * scale + offset, p2 * scale + offset, material);
    }
    for (size_t i = 0; i < n_tri; ++i) {
        uint32_t id[3] = {indices[3 * i], indices[3 * i + 1], indices[3 * i + 2]};
        if (normals.empty()) {
            mesh.triangles[i].set_normals(v_normal[id[0]].unit(), v_normal[id[1]].unit(), v_normal[id[2]].unit());
        } else {
            Vec3 n[3];
            for (int k = 0; k < 3; ++k)
                n[k] = Vec3((double)normals[3 * id[k]], (double)normals[3 * id[k] + 1], (double)normals[3 * id[k] + 2]).rotate(axis, cos, sin);
            mesh.triangles[i].set_normals(n[0], n[1], n[2]);
        }
    }
    return mesh;
}

TriangleMesh TriangleMesh::load(const std::string& filename, double scale, const Vec3& offset,
                                double rotation_angle, int axis, MaterialPtr material) {
    std::ifstream in(filename);
    if (!in) throw std::runtime_error("Failed to load OBJ file: " + filename);
    std::vector<float> raw_pos, raw_nrm;
    // single_index: one index stream; (v, vn) pairs are de-duplicated into unified vertices
    std::map<std::pair<int, int>, uint32_t> unified;
    std::vector<float> positions, normals;
    std::vector<uint32_t> indices;
    bool any_normal_ref = false;
    std::string line;
    auto resolve = [](int idx, size_t n) { return idx > 0 ? idx - 1 : (int)n + idx; };
    while (std::getline(in, line)) {
        std::istringstream ss(line);
        std::string tag;
        ss >> tag;
        if (tag == "v") {
            float x, y, z; ss >> x >> y >> z;
            raw_pos.insert(raw_pos.end(), {x, y, z});
        } else if (tag == "vn") {
            float x, y, z; ss >> x >> y >> z;
            raw_nrm.insert(raw_nrm.end(), {x, y, z});
        } else if (tag == "f") {
            std::vector<uint32_t> face;
            std::string vert;
            while (ss >> vert) {
                int vi = 0, ni = 0;
                bool has_n = false;
                size_t s1 = vert.find('/');
                vi = std::stoi(vert.substr(0, s1));
                if (s1 != std::string::npos) {
                    size_t s2 = vert.find('/', s1 + 1);
                    if (s2 != std::string::npos && s2 + 1 < vert.size()) { ni = std::stoi(vert.substr(s2 + 1)); has_n = true; }
                }
                int v = resolve(vi, raw_pos.size() / 3);
                int n = has_n ? resolve(ni, raw_nrm.size() / 3) : -1;
                any_normal_ref |= has_n;
                auto key = std::make_pair(v, n);
                auto it = unified.find(key);
                uint32_t id;
                if (it == unified.end()) {
                    id = (uint32_t)(positions.size() / 3);
                    unified[key] = id;
                    positions.insert(positions.end(), {raw_pos[3 * v], raw_pos[3 * v + 1], raw_pos[3 * v + 2]});
                    if (n >= 0) normals.insert(normals.end(), {raw_nrm[3 * n], raw_nrm[3 * n + 1], raw_nrm[3 * n + 2]});
                    else normals.insert(normals.end(), {0.f, 0.f, 0.f});
                } else id = it->second;
                face.push_back(id);
            }
            for (size_t k = 2; k < face.size(); ++k) {  // triangulate: fan
                indices.push_back(face[0]); indices.push_back(face[k - 1]); indices.push_back(face[k]);
            }
        }
    }
    if (!any_normal_ref) normals.clear();
    return from_arrays(positions, normals, indices, scale, offset, rotation_angle, axis, std::move(material));
}

// ---------------------------------------------------------------------------------------------
// Image loading: binary PPM here; PNG in image_io.cpp
// ---------------------------------------------------------------------------------------------
std::shared_ptr<Image> load_png(const std::string& path);  // image_io.cpp

std::shared_ptr<Image> Image::open(const std::string& path) {
    std::ifstream in(path, std::ios::binary);
    if (!in) throw std::runtime_error("cannot open image " + path);
    char magic[2] = {0, 0};
    in.read(magic, 2);
    if (magic[0] == 'P' && magic[1] == '6') {
        int w, h, maxv;
        in >> w >> h >> maxv;
        in.get();
        std::vector<uint8_t> rgb((size_t)w * h * 3), rgba((size_t)w * h * 4);
        in.read((char*)rgb.data(), (std::streamsize)rgb.size());
        for (size_t i = 0; i < (size_t)w * h; ++i) {
            rgba[4 * i] = rgb[3 * i]; rgba[4 * i + 1] = rgb[3 * i + 1]; rgba[4 * i + 2] = rgb[3 * i + 2]; rgba[4 * i + 3] = 255;
        }
        return std::make_shared<Image>(w, h, std::move(rgba));
    }
    in.close();
    return load_png(path);
}

// ---------------------------------------------------------------------------------------------
// FlatScene
// ---------------------------------------------------------------------------------------------
std::unique_ptr<FlatScene> FlatScene::from_world(const World& world) {
    auto fs = std::make_unique<FlatScene>();
    fs->bg0 = world.bg0;
    fs->bg1 = world.bg1;
    // World.default_material: Lambertian over white (world.rs:51)
    auto white = std::make_shared<Lambertian>(std::make_shared<SolidColor>(Color(1.f, 1.f, 1.f, 1.f)));
    fs->default_material = white->flatten(*fs);
    for (const HittablePtr& o : world.list.objects) {
        int id = o->flatten(*fs);
        AABB bb = o->bbox(world.time_range);
        fs->roots.push_back(id);
        fs->root_bbox.insert(fs->root_bbox.end(), {bb.min.x, bb.min.y, bb.min.z, bb.max.x, bb.max.y, bb.max.z});
    }
    for (const HittablePtr& o : world.lights.objects) {
        int id = o->flatten(*fs);
        if (id < 0) throw std::invalid_argument("triangle lights are not supported (Triangle::random is a stub in the reference)");
        fs->lights.push_back(id);
    }
    return fs;
}

rs_scene_desc FlatScene::desc() const {
    rs_scene_desc d;
    std::memset(&d, 0, sizeof(d));
    d.nodes = nodes.data(); d.n_nodes = (int32_t)nodes.size();
    d.roots = roots.data(); d.root_bbox = root_bbox.data(); d.n_roots = (int32_t)roots.size();
    d.lights = lights.data(); d.n_lights = (int32_t)lights.size();
    d.xforms = xforms.data(); d.n_xforms = (int32_t)xforms.size();
    d.tri_verts = tri_verts.data(); d.tri_normals = tri_normals.data(); d.tri_material = tri_material.data();
    d.n_tris = (int32_t)tri_material.size();
    d.materials = materials.data(); d.n_materials = (int32_t)materials.size();
    d.textures = textures.data(); d.n_textures = (int32_t)textures.size();
    d.texels = texels.data(); d.n_texel_bytes = (int64_t)texels.size();
    d.perlin_perm = perlin_perm.data(); d.n_perlin_perm = (int32_t)perlin_perm.size();
    d.perlin_values = perlin_values.data(); d.n_perlin_values = (int32_t)perlin_values.size();
    d.default_material = default_material;
    d.bg0[0] = bg0.r; d.bg0[1] = bg0.g; d.bg0[2] = bg0.b; d.bg0[3] = bg0.a;
    d.bg1[0] = bg1.r; d.bg1[1] = bg1.g; d.bg1[2] = bg1.b; d.bg1[3] = bg1.a;
    return d;
}

// ---------------------------------------------------------------------------------------------
// Camera::new (camera.rs:37-73)
// ---------------------------------------------------------------------------------------------
Camera CameraBuilder::build() const {
    const double theta = fov_ * (3.14159265358979323846 / 180.0);  // f64::to_radians
    const double h = std::tan(theta / 2.0);
    const double viewport_height = 2.0 * h * focus_distance_;
    const double viewport_width = viewport_height * aspect_ratio;
    Vec3 w = (look_at_ - look_from_).unit();
    Vec3 horizontal_unit = w.cross(vup_).unit();
    Vec3 vertical_unit = horizontal_unit.cross(w).unit();
    Vec3 viewport_u = horizontal_unit * viewport_width;
    Vec3 viewport_v = vertical_unit * viewport_height;
    Vec3 lb = look_from_ - viewport_u / 2.0 - viewport_v / 2.0 + w * focus_distance_;
    Camera cam;
    auto put = [](double* d, const Vec3& v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; };
    put(cam.c.origin, look_from_);
    put(cam.c.lb, lb);
    put(cam.c.horizontal_full, viewport_u);
    put(cam.c.vertical_full, viewport_v);
    put(cam.c.horizontal_unit, horizontal_unit);
    put(cam.c.vertical_unit, vertical_unit);
    cam.c.aperture = aperture_;
    cam.c.shutter_speed = shutter_speed_;
    cam.c.width = (int32_t)picture_width;
    cam.c.height = (int32_t)picture_height;
    return cam;
}

}  // namespace raysnail
