// extern "C" wrappers over the C++ host layer (include/raysnail_b200_host.h).
#include <cstring>
#include <stdexcept>
#include <string>

#include "../../../include/raysnail_b200_host.h"
#include "scene.hpp"
#include "seed_random.hpp"

namespace raysnail {
bool write_png_rgb8(const std::string& path, const uint8_t* rgb, int w, int h);
}

using namespace raysnail;

struct rs_host_scene {
    std::vector<std::shared_ptr<Texture>> textures;
    std::vector<std::shared_ptr<Material>> materials;
    std::vector<std::shared_ptr<Hittable>> objects;
    HittableList list, lights;
    Color bg0{1.f, 1.f, 1.f, 1.f}, bg1{0.5f, 0.7f, 1.0f, 1.f};
    TimeRange time_range;
    CameraBuilder camera;
    std::unique_ptr<World> world;  // set for whole-scene constructors
    std::unique_ptr<FlatScene> flat;
    rs_scene_desc desc;
};

static thread_local std::string g_host_error;

const char* rs_host_last_error(void) { return g_host_error.c_str(); }

template <class F>
static int guarded(F&& f) {
    try {
        return f();
    } catch (const std::exception& e) {
        g_host_error = e.what();
        return -1;
    } catch (...) {
        g_host_error = "unknown error";
        return -1;
    }
}

static int fail(const std::string& msg) {
    g_host_error = msg;
    return -1;
}

int rs_host_scene_from_sdl_file(const char* path, rs_host_scene** out) {
    return guarded([&] {
        SceneData sd;
        std::string err;
        if (!SdlParser::parse(path, sd, err)) { g_host_error = err; return (int)RS_ERR_PARSE; }
        if (!sd.has_camera) { g_host_error = "scene has no camera"; return (int)RS_ERR_PARSE; }
        auto s = std::make_unique<rs_host_scene>();
        s->world = world_from_scene_data(sd, s->camera);
        *out = s.release();
        return (int)RS_OK;
    });
}

int rs_host_scene_from_sdl_text(const char* text, rs_host_scene** out) {
    return guarded([&] {
        SceneData sd;
        std::string err;
        if (!SdlParser::parse_text(text, sd, err)) { g_host_error = err; return (int)RS_ERR_PARSE; }
        if (!sd.has_camera) { g_host_error = "scene has no camera"; return (int)RS_ERR_PARSE; }
        auto s = std::make_unique<rs_host_scene>();
        s->world = world_from_scene_data(sd, s->camera);
        *out = s.release();
        return (int)RS_OK;
    });
}

int rs_host_scene_builtin(const char* name, uint64_t seed, int32_t iparam, rs_host_scene** out) {
    return guarded([&] {
        auto s = std::make_unique<rs_host_scene>();
        std::string n = name;
        if (n == "rtow_final") s->world = recipes::rtow_final_world(seed, s->camera);
        else if (n == "cornell") s->world = recipes::cornell_box_world(true, true, false, s->camera);
        else if (n == "cornell_smoke") s->world = recipes::cornell_box_world(true, true, true, s->camera);
        else if (n == "cornell_plain") s->world = recipes::cornell_box_world(true, false, false, s->camera);
        else if (n == "all_feature") s->world = recipes::all_feature_world(seed, s->camera);
        else if (n == "raymarching") s->world = recipes::raymarching_test_world(s->camera);
        else if (n == "mesh") s->world = recipes::mesh_test_world(iparam > 2 ? iparam : 64, false, s->camera);
        else if (n == "mesh_blinn") s->world = recipes::mesh_test_world(iparam > 2 ? iparam : 64, true, s->camera);
        else { g_host_error = "unknown builtin scene: " + n; return (int)RS_ERR_INVALID_ARGUMENT; }
        *out = s.release();
        return (int)RS_OK;
    });
}

void rs_host_scene_free(rs_host_scene* s) { delete s; }

const rs_scene_desc* rs_host_scene_desc(rs_host_scene* s) {
    if (!s) return nullptr;
    int rc = guarded([&] {
        if (!s->world)
            s->world = std::make_unique<World>(s->list, s->lights, s->bg0, s->bg1, s->time_range);
        s->flat = FlatScene::from_world(*s->world);
        s->desc = s->flat->desc();
        return 0;
    });
    return rc == 0 ? &s->desc : nullptr;
}

int rs_host_scene_camera(rs_host_scene* s, int32_t width, int32_t height, rs_camera* out) {
    return guarded([&] {
        CameraBuilder b = s->camera;
        b.width((size_t)width).height((size_t)height);
        *out = b.build().c;
        return (int)RS_OK;
    });
}

int rs_host_camera_build(const double f[3], const double a[3], const double up[3], double fov, double aperture,
                         double focus, double shutter, int32_t width, int32_t height, rs_camera* out) {
    return guarded([&] {
        CameraBuilder b;
        b.look_from(Vec3(f[0], f[1], f[2])).look_at(Vec3(a[0], a[1], a[2])).vup(Vec3(up[0], up[1], up[2]));
        b.fov(fov).aperture(aperture).focus(focus).shutter_speed(shutter).width((size_t)width).height((size_t)height);
        *out = b.build().c;
        return (int)RS_OK;
    });
}

rs_host_scene* rs_host_scene_new(void) { return new rs_host_scene(); }

static void invalidate(rs_host_scene* s) { s->world.reset(); s->flat.reset(); }

static TexturePtr tex_of(rs_host_scene* s, int id) {
    if (id < 0 || id >= (int)s->textures.size()) throw std::invalid_argument("bad texture handle");
    return s->textures[id];
}
static MaterialPtr mat_of(rs_host_scene* s, int id) {
    if (id < 0) return nullptr;
    if (id >= (int)s->materials.size()) throw std::invalid_argument("bad material handle");
    return s->materials[id];
}
static HittablePtr obj_of(rs_host_scene* s, int id) {
    if (id < 0 || id >= (int)s->objects.size()) throw std::invalid_argument("bad object handle");
    return s->objects[id];
}
template <class T>
static int add_tex(rs_host_scene* s, std::shared_ptr<T> t) { s->textures.push_back(std::move(t)); return (int)s->textures.size() - 1; }
template <class T>
static int add_mat(rs_host_scene* s, std::shared_ptr<T> m) { s->materials.push_back(std::move(m)); return (int)s->materials.size() - 1; }
template <class T>
static int add_obj(rs_host_scene* s, std::shared_ptr<T> o) { s->objects.push_back(std::move(o)); return (int)s->objects.size() - 1; }

int rs_host_tex_solid(rs_host_scene* s, float r, float g, float b, float a) {
    return guarded([&] { return add_tex(s, std::make_shared<SolidColor>(Color(r, g, b, a))); });
}
int rs_host_tex_checker(rs_host_scene* s, int odd, int even, double scale) {
    return guarded([&] { return add_tex(s, std::make_shared<Checker>(tex_of(s, odd), tex_of(s, even), scale)); });
}
int rs_host_tex_image(rs_host_scene* s, int32_t w, int32_t h, const uint8_t* rgba) {
    return guarded([&] {
        if (w <= 0 || h <= 0 || !rgba) throw std::invalid_argument("bad image");
        return add_tex(s, std::make_shared<Image>(w, h, std::vector<uint8_t>(rgba, rgba + (size_t)w * h * 4)));
    });
}
int rs_host_tex_image_file(rs_host_scene* s, const char* path) {
    return guarded([&] { return add_tex(s, Image::open(path)); });
}
int rs_host_tex_perlin(rs_host_scene* s, int32_t point_count, int32_t vector, uint64_t seed, int32_t kind,
                       int32_t depth, int32_t smooth, double scale) {
    return guarded([&] {
        auto p = std::make_shared<Perlin>(point_count, vector != 0, seed);
        p->scale(scale).smooth((SmoothType)smooth);
        if (kind == 1) p->turbulence(depth);
        if (kind == 2) p->marble(depth);
        return add_tex(s, p);
    });
}

int rs_host_mat_lambertian(rs_host_scene* s, int tex) {
    return guarded([&] { return add_mat(s, std::make_shared<Lambertian>(tex_of(s, tex))); });
}
int rs_host_mat_metal(rs_host_scene* s, int tex) {
    return guarded([&] { return add_mat(s, std::make_shared<Metal>(tex_of(s, tex))); });
}
int rs_host_mat_diffuse_metal(rs_host_scene* s, double exponent, int tex) {
    return guarded([&] { return add_mat(s, std::make_shared<DiffuseMetal>(exponent, tex_of(s, tex))); });
}
int rs_host_mat_dielectric(rs_host_scene* s, float r, float g, float b, double refractive, int glass_curve) {
    return guarded([&] {
        auto d = std::make_shared<Dielectric>(Color(r, g, b, 1.f), refractive);
        if (glass_curve) d->reflect_curve_glass();
        return add_mat(s, d);
    });
}
int rs_host_mat_diffuse_light(rs_host_scene* s, int tex, double multiplier) {
    return guarded([&] {
        auto l = std::make_shared<DiffuseLight>(tex_of(s, tex));
        l->multiplier(multiplier);
        return add_mat(s, l);
    });
}
int rs_host_mat_isotropic(rs_host_scene* s, float r, float g, float b) {
    return guarded([&] { return add_mat(s, std::make_shared<Isotropic>(Color(r, g, b, 1.f))); });
}
int rs_host_mat_blinn_phong(rs_host_scene* s, double ks, double e, int tex) {
    return guarded([&] { return add_mat(s, std::make_shared<BlinnPhong>(ks, e, tex_of(s, tex))); });
}
int rs_host_mat_mixed(rs_host_scene* s, int m1, int m2, double p) {
    return guarded([&] {
        if (m1 < 0 || m2 < 0) throw std::invalid_argument("mixed material needs two materials");
        return add_mat(s, std::make_shared<MixedMaterial>(mat_of(s, m1), mat_of(s, m2), p));
    });
}
int rs_host_mat_set_phong(rs_host_scene* s, int mat, double factor, int32_t exponent) {
    return guarded([&] {
        if (mat < 0 || mat >= (int)s->materials.size()) throw std::invalid_argument("bad material handle");
        CommonMaterialSettings cs;
        cs.phong_factor = factor;
        cs.phong_exponent = exponent;
        s->materials[mat]->set(cs);
        invalidate(s);
        return 0;
    });
}

int rs_host_obj_sphere(rs_host_scene* s, const double c[3], double r, const double v[3], int mat) {
    return guarded([&] {
        auto sp = std::make_shared<Sphere>(Vec3(c[0], c[1], c[2]), r, mat_of(s, mat));
        if (v) sp->with_speed(Vec3(v[0], v[1], v[2]));
        return add_obj(s, sp);
    });
}
int rs_host_obj_box(rs_host_scene* s, const double p0[3], const double p1[3], int mat) {
    return guarded([&] {
        return add_obj(s, std::make_shared<Box>(Vec3(p0[0], p0[1], p0[2]), Vec3(p1[0], p1[1], p1[2]), mat_of(s, mat)));
    });
}
int rs_host_obj_rect(rs_host_scene* s, int32_t axis, double k, double a0, double a1, double b0, double b1, int mat) {
    return guarded([&] {
        if (axis < 0 || axis > 2) throw std::invalid_argument("bad rect axis");
        return add_obj(s, std::make_shared<AARect>(axis, AARectMetrics(k, a0, a1, b0, b1), mat_of(s, mat)));
    });
}
int rs_host_obj_quadric(rs_host_scene* s, const double q[10], int mat) {
    return guarded([&] {
        return add_obj(s, std::make_shared<Quadric>(q[0], q[1], q[2], q[3], q[4], q[5], q[6], q[7], q[8], q[9], mat_of(s, mat)));
    });
}
int rs_host_obj_raymarcher(rs_host_scene* s, int mat) {
    return guarded([&] { return add_obj(s, std::make_shared<RayMarcher>(mat_of(s, mat))); });
}
int rs_host_obj_difference(rs_host_scene* s, int plus, int minus, int mat) {
    return guarded([&] { return add_obj(s, std::make_shared<Difference>(obj_of(s, plus), obj_of(s, minus), mat_of(s, mat))); });
}
int rs_host_obj_intersection(rs_host_scene* s, int o1, int o2, int mat) {
    return guarded([&] { return add_obj(s, std::make_shared<Intersection>(obj_of(s, o1), obj_of(s, o2), mat_of(s, mat))); });
}
int rs_host_obj_bvh(rs_host_scene* s, int32_t n, const int32_t* objs) {
    return guarded([&] {
        std::vector<HittablePtr> members;
        for (int i = 0; i < n; ++i) members.push_back(obj_of(s, objs[i]));
        return add_obj(s, BVH::build(std::move(members), s->time_range));
    });
}
int rs_host_obj_transform(rs_host_scene* s, int child, int32_t n_ops, const int32_t* ops, const double* args) {
    return guarded([&] {
        TransformStack st;
        for (int i = 0; i < n_ops; ++i) {
            Vec3 v(args[3 * i], args[3 * i + 1], args[3 * i + 2]);
            switch (ops[i]) {
                case 0: st.push(Transform::translate(v)); break;
                case 1: st.push(Transform::rotate_by_x_axis(v.x)); break;
                case 2: st.push(Transform::rotate_by_y_axis(v.x)); break;
                case 3: st.push(Transform::rotate_by_z_axis(v.x)); break;
                case 4: st.push(Transform::scale(v)); break;
                default: throw std::invalid_argument("bad transform op");
            }
        }
        return add_obj(s, std::make_shared<TfFacade>(obj_of(s, child), st));
    });
}
int rs_host_obj_medium(rs_host_scene* s, int boundary, float r, float g, float b, double density) {
    return guarded([&] { return add_obj(s, std::make_shared<ConstantMedium>(obj_of(s, boundary), Color(r, g, b, 1.f), density)); });
}

int rs_host_world_add(rs_host_scene* s, int obj) {
    return guarded([&] { s->list.add_ref(obj_of(s, obj)); invalidate(s); return 0; });
}
int rs_host_world_add_light(rs_host_scene* s, int obj) {
    return guarded([&] { s->lights.add_ref(obj_of(s, obj)); invalidate(s); return 0; });
}
int rs_host_world_add_mesh(rs_host_scene* s, const float* positions, int64_t n_positions, const float* normals,
                           int64_t n_normals, const uint32_t* indices, int64_t n_indices, double scale,
                           const double offset[3], double rotation_angle, int32_t axis, int mat) {
    return guarded([&] {
        std::vector<float> p(positions, positions + n_positions);
        std::vector<float> n;
        if (normals && n_normals > 0) n.assign(normals, normals + n_normals);
        std::vector<uint32_t> idx(indices, indices + n_indices);
        for (uint32_t i : idx)
            if ((int64_t)i * 3 + 2 >= n_positions) throw std::invalid_argument("mesh index out of range");
        TriangleMesh mesh = TriangleMesh::from_arrays(p, n, idx, scale, Vec3(offset[0], offset[1], offset[2]),
                                                      rotation_angle, axis, mat_of(s, mat));
        for (Triangle& t : mesh.triangles) s->list.add(std::move(t));
        invalidate(s);
        return (int)mesh.triangles.size();
    });
}
int rs_host_world_add_obj_file(rs_host_scene* s, const char* path, double scale, const double offset[3],
                               double rotation_angle, int32_t axis, int mat) {
    return guarded([&] {
        TriangleMesh mesh = TriangleMesh::load(path, scale, Vec3(offset[0], offset[1], offset[2]), rotation_angle, axis, mat_of(s, mat));
        for (Triangle& t : mesh.triangles) s->list.add(std::move(t));
        invalidate(s);
        return (int)mesh.triangles.size();
    });
}
int rs_host_world_set_background(rs_host_scene* s, const float c0[4], const float c1[4]) {
    s->bg0 = Color(c0[0], c0[1], c0[2], c0[3]);
    s->bg1 = Color(c1[0], c1[1], c1[2], c1[3]);
    invalidate(s);
    return 0;
}
int rs_host_world_set_time_range(rs_host_scene* s, double t0, double t1) {
    s->time_range = TimeRange{t0, t1};
    invalidate(s);
    return 0;
}
int rs_host_scene_set_camera(rs_host_scene* s, const double f[3], const double a[3], const double up[3], double fov,
                             double aperture, double focus, double shutter) {
    s->camera = CameraBuilder();
    s->camera.look_from(Vec3(f[0], f[1], f[2])).look_at(Vec3(a[0], a[1], a[2])).vup(Vec3(up[0], up[1], up[2]));
    s->camera.fov(fov).aperture(aperture).focus(focus).shutter_speed(shutter);
    return 0;
}

int rs_host_seed_random_u32(uint64_t seed, uint32_t* out, int32_t n) {
    SeedRandom rng(seed);
    for (int i = 0; i < n; ++i) out[i] = rng.next_u32();
    return 0;
}
int rs_host_chacha12_block(const uint32_t key[8], uint64_t counter, uint32_t out[16]) {
    SeedRandom::chacha12_block(key, counter, out);
    return 0;
}
int rs_host_write_png_rgb8(const char* path, const uint8_t* rgb, int32_t w, int32_t h) {
    return write_png_rgb8(path, rgb, w, h) ? 0 : fail(std::string("cannot write ") + path);
}
