// Host-side scene API of raysnail_b200 (C++17).
//
// Mirrors the reference's scene-construction interface (same type names, argument meaning and
// bounding-box rules) so that any scene the reference builds can be built here and flattened
// into the C-ABI `rs_scene_desc` (include/raysnail_b200.h).  This layer holds data and the
// `Hittable::bbox` rules only: no intersection, shading or sampling code lives on the host —
// that is the CUDA backend's job (csrc/device).
//
// Reference interfaces mirrored (all under /root/reference/src):
//   prelude/vec3.rs, prelude/color.rs, prelude/aabb.rs
//   texture/{mod,checker,image,noise}.rs
//   material/{mod,lambertian,metal,dielectric,light,isotropic,blinn_phong,mixed_material}.rs
//   hittable/geometry/{sphere,rect,box,quadric,triangle_mesh}.rs, hittable/csg/*.rs,
//   hittable/transform/{transform,tf_facade}.rs, hittable/medium/constant.rs,
//   hittable/collection/{list,world}.rs, camera.rs (Camera, CameraBuilder)
#pragma once

#include <cmath>
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../../../include/raysnail_b200.h"

namespace raysnail {

struct Vec3 {
    double x = 0, y = 0, z = 0;
    Vec3() = default;
    Vec3(double x_, double y_, double z_) : x(x_), y(y_), z(z_) {}
    double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    double& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    Vec3 operator+(const Vec3& o) const { return {x + o.x, y + o.y, z + o.z}; }
    Vec3 operator-(const Vec3& o) const { return {x - o.x, y - o.y, z - o.z}; }
    Vec3 operator-() const { return {-x, -y, -z}; }
    Vec3 operator*(double s) const { return {x * s, y * s, z * s}; }
    // vec3.rs:486-500: division multiplies by the reciprocal
    Vec3 operator/(double s) const { return *this * (1.0 / s); }
    // vec3.rs:152-155
    double length_squared() const { return std::fma(z, z, std::fma(x, x, y * y)); }
    double length() const { return std::sqrt(length_squared()); }
    // vec3.rs:177-179
    double dot(const Vec3& r) const { return std::fma(z, r.z, std::fma(x, r.x, y * r.y)); }
    Vec3 cross(const Vec3& r) const {
        return {y * r.z - z * r.y, z * r.x - x * r.z, x * r.y - y * r.x};
    }
    Vec3 unit() const { return *this / length(); }
    // vec3.rs:196-214
    Vec3 rotate(int axis, double cos, double sin) const {
        switch (axis) {
            case 0: return {x, y * cos - z * sin, y * sin + z * cos};
            case 1: return {x * cos + z * sin, y, -x * sin + z * cos};
            default: return {x * cos - y * sin, x * sin + y * cos, z};
        }
    }
    static Vec3 new_min(const Vec3& a, const Vec3& b) {
        return {std::fmin(a.x, b.x), std::fmin(a.y, b.y), std::fmin(a.z, b.z)};
    }
    static Vec3 new_max(const Vec3& a, const Vec3& b) {
        return {std::fmax(a.x, b.x), std::fmax(a.y, b.y), std::fmax(a.z, b.z)};
    }
};
using Point3 = Vec3;

struct Color {
    float r = 0, g = 0, b = 0, a = 0;
    Color() = default;
    Color(float r_, float g_, float b_, float a_) : r(r_), g(g_), b(b_), a(a_) {}
    static Color new64(double r, double g, double b, double a) {
        return Color((float)r, (float)g, (float)b, (float)a);
    }
};

// prelude/aabb.rs:7-18, 58-104
struct AABB {
    Vec3 min, max;
    AABB() = default;
    AABB(const Vec3& mn, const Vec3& mx) : min(mn), max(mx) {}
    AABB operator|(const AABB& o) const {
        return AABB(Vec3::new_min(min, o.min), Vec3::new_max(max, o.max));
    }
};

struct TimeRange {
    double start = 0, end = 0;
};

class FlatScene;

// ---------------------------------------------------------------------------------------------
// Textures
// ---------------------------------------------------------------------------------------------
struct Texture {
    virtual ~Texture() = default;
    virtual int flatten(FlatScene& fs) const = 0;
};
using TexturePtr = std::shared_ptr<const Texture>;

struct SolidColor : Texture {  // `impl Texture for Color` (color.rs:61-65)
    Color color;
    explicit SolidColor(Color c) : color(c) {}
    int flatten(FlatScene& fs) const override;
};

struct Checker : Texture {  // checker.rs
    TexturePtr odd, even;
    double scale;
    Checker(TexturePtr o, TexturePtr e, double s) : odd(std::move(o)), even(std::move(e)), scale(s) {}
    int flatten(FlatScene& fs) const override;
};

struct Image : Texture {  // image.rs — texels already decoded to RGBA8, row 0 = top
    int width = 0, height = 0;
    std::vector<uint8_t> rgba;
    Image(int w, int h, std::vector<uint8_t> px) : width(w), height(h), rgba(std::move(px)) {}
    // Loads binary PPM (P6) or PNG (8-bit, non-interlaced RGB/RGBA/grey).  Throws on failure.
    static std::shared_ptr<Image> open(const std::string& path);
    int flatten(FlatScene& fs) const override;
};

enum class SmoothType { None = 0, LinearInterpolate = 1, HermitianCubic = 2 };

struct Perlin : Texture {  // noise.rs
    int kind = 0;          // 0 Normal, 1 Turbulence, 2 Marble
    int depth = 0;
    SmoothType smooth_type = SmoothType::HermitianCubic;
    bool vector = false;
    int point_count = 0;
    double scale_ = 1.0;
    std::vector<double> random_values;  // point_count * (vector ? 3 : 1)
    std::vector<int32_t> perm;          // perm_x | perm_y | perm_z
    // noise.rs:40-71.  The reference fills the tables from FastRng (seeded from thread_rng, so
    // they differ per run); here they come from a seeded SplitMix64 stream.
    Perlin(int point_count, bool vector, uint64_t seed);
    Perlin& scale(double s) { scale_ = s; return *this; }
    Perlin& smooth(SmoothType s) { smooth_type = s; return *this; }
    Perlin& turbulence(int d) { kind = 1; depth = d; return *this; }
    Perlin& marble(int d) { kind = 2; depth = d; return *this; }
    int flatten(FlatScene& fs) const override;
};

// ---------------------------------------------------------------------------------------------
// Materials
// ---------------------------------------------------------------------------------------------
struct CommonMaterialSettings {  // material/mod.rs:41-55
    double phong_factor = 0.0;
    int phong_exponent = 1;
};

struct Material {
    CommonMaterialSettings settings;
    virtual ~Material() = default;
    void set(const CommonMaterialSettings& s) { settings = s; }
    // `Material::settings()` (material/mod.rs:63-65); MixedMaterial forwards to material_1 (mixed_material.rs:56-58)
    virtual CommonMaterialSettings effective_settings() const { return settings; }
    virtual int flatten(FlatScene& fs) const = 0;
};
using MaterialPtr = std::shared_ptr<const Material>;

struct Lambertian : Material {
    TexturePtr texture;
    explicit Lambertian(TexturePtr t) : texture(std::move(t)) {}
    int flatten(FlatScene& fs) const override;
};
struct Metal : Material {
    TexturePtr texture;
    explicit Metal(TexturePtr t) : texture(std::move(t)) {}
    int flatten(FlatScene& fs) const override;
};
struct DiffuseMetal : Material {
    double exponent;
    TexturePtr texture;
    DiffuseMetal(double e, TexturePtr t) : exponent(e), texture(std::move(t)) {}
    int flatten(FlatScene& fs) const override;
};
struct Dielectric : Material {  // dielectric.rs:27-52
    Color color;
    double refractive;
    bool glass_curve = false;
    Dielectric(Color c, double ri) : color(c), refractive(ri) {}
    Dielectric& reflect_curve_glass() { glass_curve = true; return *this; }
    int flatten(FlatScene& fs) const override;
};
struct DiffuseLight : Material {
    TexturePtr texture;
    double multiplier_ = 1.0;
    explicit DiffuseLight(TexturePtr t) : texture(std::move(t)) {}
    DiffuseLight& multiplier(double m) { multiplier_ = m; return *this; }
    int flatten(FlatScene& fs) const override;
};
struct Isotropic : Material {
    Color color;
    explicit Isotropic(Color c) : color(c) {}
    int flatten(FlatScene& fs) const override;
};
struct BlinnPhong : Material {
    double k_specular, exponent;
    TexturePtr texture;
    BlinnPhong(double ks, double e, TexturePtr t) : k_specular(ks), exponent(e), texture(std::move(t)) {}
    int flatten(FlatScene& fs) const override;
};
struct MixedMaterial : Material {
    MaterialPtr material_1, material_2;
    double probability_1;
    MixedMaterial(MaterialPtr m1, MaterialPtr m2, double p)
        : material_1(std::move(m1)), material_2(std::move(m2)), probability_1(p) {}
    CommonMaterialSettings effective_settings() const override { return material_1->effective_settings(); }
    int flatten(FlatScene& fs) const override;
};

// ---------------------------------------------------------------------------------------------
// Transforms (transform.rs)
// ---------------------------------------------------------------------------------------------
struct Transform {
    double matrix[16];
    double inverse[16];
    static Transform translate(const Vec3& t);
    static Transform rotate_by_x_axis(double theta);
    static Transform rotate_by_y_axis(double theta);
    static Transform rotate_by_z_axis(double theta);
    static Transform scale(const Vec3& s);
};

struct TransformStack {
    std::vector<Transform> stack;
    void push(const Transform& t) { stack.push_back(t); }
    size_t len() const { return stack.size(); }
    Vec3 forward(const Vec3& pos, double w) const;  // transform.rs:133-144
    Vec3 inverse(const Vec3& pos, double w) const;  // transform.rs:146-157
};

// ---------------------------------------------------------------------------------------------
// Hittables
// ---------------------------------------------------------------------------------------------
struct Hittable {
    virtual ~Hittable() = default;
    virtual AABB bbox(const TimeRange& tr) const = 0;
    // Appends this object (and its children) to fs.nodes, returns the node index.
    virtual int flatten(FlatScene& fs) const = 0;
};
using HittablePtr = std::shared_ptr<const Hittable>;

struct Sphere : Hittable {
    Point3 center;
    double radius;
    Vec3 speed;
    MaterialPtr material;
    Sphere(Point3 c, double r, MaterialPtr m) : center(c), radius(r), material(std::move(m)) {}
    Sphere& with_speed(const Vec3& s) { speed = s; return *this; }
    Point3 center_at(double t) const { return center + speed * t; }
    AABB bbox(const TimeRange& tr) const override;
    int flatten(FlatScene& fs) const override;
};

struct AARectMetrics {  // rect.rs:13-41
    double k, a0, a1, b0, b1;
    AARectMetrics(double k_, double a0_, double a1_, double b0_, double b1_);
};

struct AARect : Hittable {
    int axis_kind;  // RS_RECT_XY / XZ / YZ
    AARectMetrics metrics;
    MaterialPtr material;
    AARect(int kind, AARectMetrics m, MaterialPtr mat) : axis_kind(kind), metrics(m), material(std::move(mat)) {}
    static AARect new_xy(AARectMetrics m, MaterialPtr mat) { return AARect(RS_RECT_XY, m, std::move(mat)); }
    static AARect new_xz(AARectMetrics m, MaterialPtr mat) { return AARect(RS_RECT_XZ, m, std::move(mat)); }
    static AARect new_yz(AARectMetrics m, MaterialPtr mat) { return AARect(RS_RECT_YZ, m, std::move(mat)); }
    AABB bbox(const TimeRange& tr) const override;
    int flatten(FlatScene& fs) const override;
};

struct Box : Hittable {
    Point3 point_min, point_max;
    MaterialPtr material;
    Box(const Point3& p0, const Point3& p1, MaterialPtr m)
        : point_min(Vec3::new_min(p0, p1)), point_max(Vec3::new_max(p0, p1)), material(std::move(m)) {}
    AABB bbox(const TimeRange& tr) const override;
    int flatten(FlatScene& fs) const override;
};

struct Quadric : Hittable {
    double q[10];  // qa qb qc qd qe qf qg qh qi qj
    MaterialPtr material;
    Quadric(double qa, double qb, double qc, double qd, double qe, double qf, double qg, double qh,
            double qi, double qj, MaterialPtr m)
        : q{qa, qb, qc, qd, qe, qf, qg, qh, qi, qj}, material(std::move(m)) {}
    AABB bbox(const TimeRange& tr) const override;
    int flatten(FlatScene& fs) const override;
};

// geometry/raymarching.rs: the power-8 Mandelbulb about the origin (RayMarcher::new: 100 iterations)
struct RayMarcher : Hittable {
    MaterialPtr material;
    int iterations = 100;
    explicit RayMarcher(MaterialPtr m) : material(std::move(m)) {}
    AABB bbox(const TimeRange& tr) const override;
    int flatten(FlatScene& fs) const override;
};

struct Difference : Hittable {
    HittablePtr plus, minus;
    MaterialPtr material;
    Difference(HittablePtr p, HittablePtr m, MaterialPtr mat)
        : plus(std::move(p)), minus(std::move(m)), material(std::move(mat)) {}
    AABB bbox(const TimeRange& tr) const override;
    int flatten(FlatScene& fs) const override;
};

struct Intersection : Hittable {
    HittablePtr o1, o2;
    MaterialPtr material;
    Intersection(HittablePtr a, HittablePtr b, MaterialPtr mat)
        : o1(std::move(a)), o2(std::move(b)), material(std::move(mat)) {}
    AABB bbox(const TimeRange& tr) const override;
    int flatten(FlatScene& fs) const override;
};

struct TfFacade : Hittable {
    HittablePtr object;
    TransformStack stack;
    TfFacade(HittablePtr o, TransformStack s) : object(std::move(o)), stack(std::move(s)) {}
    AABB bbox(const TimeRange& tr) const override;
    int flatten(FlatScene& fs) const override;
};

struct ConstantMedium : Hittable {  // medium/constant.rs:29-39
    HittablePtr boundary;
    std::shared_ptr<Isotropic> material;
    double density;
    ConstantMedium(HittablePtr b, Color c, double d)
        : boundary(std::move(b)), material(std::make_shared<Isotropic>(c)), density(d) {}
    AABB bbox(const TimeRange& tr) const override;
    int flatten(FlatScene& fs) const override;
};

struct Triangle : Hittable {  // triangle_mesh.rs:17-72
    Vec3 p0, p1, p2;
    Vec3 normal0, normal1, normal2;
    MaterialPtr material;
    Triangle(const Vec3& a, const Vec3& b, const Vec3& c, MaterialPtr m)
        : p0(a), p1(b), p2(c), material(std::move(m)) {}
    void set_normals(const Vec3& n0, const Vec3& n1, const Vec3& n2) { normal0 = n0; normal1 = n1; normal2 = n2; }
    AABB bbox(const TimeRange& tr) const override;
    int flatten(FlatScene& fs) const override;  // returns ~triangle_index
};

struct TriangleMesh {  // triangle_mesh.rs:152-276
    std::vector<Triangle> triangles;
    // Minimal Wavefront reader covering what tobj's {single_index, triangulate} gives the
    // reference: v / vn / f (fan-triangulated).  Throws std::runtime_error on I/O errors.
    static TriangleMesh load(const std::string& filename, double scale, const Vec3& offset,
                             double rotation_angle, int axis, MaterialPtr material);
    // Same post-processing on in-memory arrays (positions/normals as f32 like tobj).
    static TriangleMesh from_arrays(const std::vector<float>& positions, const std::vector<float>& normals,
                                    const std::vector<uint32_t>& indices, double scale, const Vec3& offset,
                                    double rotation_angle, int axis, MaterialPtr material);
};

// collection/bvh.rs — a BVH used as ONE object (bvh.rs:173-192 as a `Hittable`): a world member, the object of a
// TfFacade, the boundary of a ConstantMedium.  (`World::new` builds its own tree over the world list; that one is
// rs_upload_scene's business.)  `BVH::build` follows `BVH::new` / `new_internal` (bvh.rs:47-112): one object -> a node
// with a left child only, two -> both children, more -> stable sort by the bbox minimum on an axis, split in the
// middle.  The reference draws that axis with `Random::range(0..2)` (bvh.rs:91); here it is (start + count) & 1 — any
// axis gives a valid tree, and the tree travels explicitly in the scene description, so oracle and GPU walk the same one.
struct BVH : Hittable {
    bool has_bbox = false;
    AABB box;
    HittablePtr left, right;
    static std::shared_ptr<BVH> build(std::vector<HittablePtr> objects, const TimeRange& tr);
    AABB bbox(const TimeRange& tr) const override;
    int flatten(FlatScene& fs) const override;

private:
    static std::shared_ptr<BVH> build_range(std::vector<HittablePtr>& objects, size_t begin, size_t end, const TimeRange& tr);
};

struct HittableList {  // collection/list.rs
    std::vector<HittablePtr> objects;
    template <class G>
    HittableList& add(G object) {
        objects.push_back(std::make_shared<G>(std::move(object)));
        return *this;
    }
    HittableList& add_ref(HittablePtr o) { objects.push_back(std::move(o)); return *this; }
    void clear() { objects.clear(); }
};

// collection/world.rs:25-53.  The background closure is restricted to the gradient form every
// shipped caller uses: bg0.gradient(bg1, 0.5*(dir.y+1)).
struct World {
    HittableList list;
    HittableList lights;
    Color bg0{1.f, 1.f, 1.f, 1.f}, bg1{0.5f, 0.7f, 1.0f, 1.f};  // default_background (world.rs:20-23)
    TimeRange time_range;
    World(HittableList l, HittableList li, Color c0, Color c1, TimeRange tr)
        : list(std::move(l)), lights(std::move(li)), bg0(c0), bg1(c1), time_range(tr) {}
};

// ---------------------------------------------------------------------------------------------
// FlatScene: owner of the buffers behind an rs_scene_desc.
// ---------------------------------------------------------------------------------------------
class FlatScene {
public:
    std::vector<rs_node> nodes;
    std::vector<int32_t> roots;
    std::vector<double> root_bbox;
    std::vector<int32_t> lights;
    std::vector<rs_xform> xforms;
    std::vector<double> tri_verts, tri_normals;
    std::vector<int32_t> tri_material;
    std::vector<rs_material> materials;
    std::vector<rs_texture> textures;
    std::vector<uint8_t> texels;
    std::vector<int32_t> perlin_perm;
    std::vector<double> perlin_values;
    int32_t default_material = -1;
    Color bg0, bg1;

    // pointer -> index caches so shared Arc<> objects flatten once
    std::map<const void*, int> node_cache, material_cache, texture_cache;

    static std::unique_ptr<FlatScene> from_world(const World& world);
    rs_scene_desc desc() const;

    int material_index(const MaterialPtr& m) { return m ? m->flatten(*this) : -1; }
};

// ---------------------------------------------------------------------------------------------
// Camera (camera.rs:18-91, 299-414)
// ---------------------------------------------------------------------------------------------
struct Camera {
    rs_camera c{};
    double shutter_speed() const { return c.shutter_speed; }
};

struct CameraBuilder {
    Point3 look_from_{0, 0, 0};
    Point3 look_at_{0, 0, -1};
    Vec3 vup_{0, 1, 0};
    double fov_ = 90.0, aperture_ = 0.0, focus_distance_ = 1.0, shutter_speed_ = 0.0;
    size_t picture_width = 400, picture_height = 200;
    double aspect_ratio = 2.0;

    CameraBuilder& look_from(const Point3& p) { look_from_ = p; return *this; }
    CameraBuilder& look_at(const Point3& p) { look_at_ = p; return *this; }
    CameraBuilder& vup(const Vec3& v) { vup_ = v; return *this; }
    CameraBuilder& fov(double f) { fov_ = f; return *this; }
    CameraBuilder& aperture(double a) { aperture_ = a; return *this; }
    CameraBuilder& focus(double d) { focus_distance_ = d; return *this; }
    CameraBuilder& focus_to_look_at() { return focus((look_at_ - look_from_).length()); }
    CameraBuilder& shutter_speed(double d) { shutter_speed_ = d; return *this; }
    CameraBuilder& width(size_t w) {
        picture_width = w;
        aspect_ratio = (double)picture_width / (double)picture_height;
        return *this;
    }
    CameraBuilder& height(size_t h) {
        picture_height = h;
        aspect_ratio = (double)picture_width / (double)picture_height;
        return *this;
    }
    Camera build() const;
};

// ---------------------------------------------------------------------------------------------
// SDL front end (sdl_parser.rs) and the example scene recipes (examples/common/scene.rs)
// ---------------------------------------------------------------------------------------------
struct CameraData {
    Vec3 location, look_at;
    double fov_angle = 60.0;
};
struct LightData {
    Vec3 location;
    Color color;
};
struct SceneData {
    bool has_camera = false;
    CameraData camera;
    HittableList hittables;
    std::vector<LightData> lights;
};

struct SdlParser {
    // Returns false (and fills `error`) on a parse error, like `SdlParser::parse` -> Err.
    static bool parse(const std::string& filename, SceneData& out, std::string& error);
    static bool parse_text(const std::string& text, SceneData& out, std::string& error);
};

// `parse_and_render`'s scene assembly (src/bin/raysnail.rs:327-377): lights become r=12 emissive
// spheres (x1.7) added to both lists; sky gradient; aperture 0.01, focus 10.
std::unique_ptr<World> world_from_scene_data(SceneData& sd, CameraBuilder& cam_out);

namespace recipes {
// examples/common/scene.rs:157-191 — `balls_scene(seed, need_speed=false, checker)`
HittableList balls_scene(uint64_t seed, bool checker);
CameraBuilder balls_scene_camera(bool need_shutter_speed);  // scene.rs:194-208
// examples/rtow_13_1.rs:15-46 — final scene + light sphere + sky
std::unique_ptr<World> rtow_final_world(uint64_t seed, CameraBuilder& cam_out);
// examples/common/scene.rs:211-334 — cornell_box_scene(carton, carton_rotation, smoke)
std::unique_ptr<World> cornell_box_world(bool carton, bool carton_rotation, bool smoke, CameraBuilder& cam_out);
// scene shape of examples/preview_sdl2.rs:452-525 with a procedural mesh instead of dragon.obj
std::unique_ptr<World> mesh_test_world(int grid_n, bool blinn_phong, CameraBuilder& cam_out);
std::unique_ptr<World> all_feature_world(uint64_t seed, CameraBuilder& cam_out);  // scene.rs:336-468
// examples/preview_sdl2.rs:338-407 — `render_raymarching_test` (Mandelbulb over a checker ground)
std::unique_ptr<World> raymarching_test_world(CameraBuilder& cam_out);
}  // namespace recipes

}  // namespace raysnail
