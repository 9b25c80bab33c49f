// Scene recipes of the reference's examples, restated for the benchmark configurations.
//   balls_scene / balls_scene_camera : examples/common/scene.rs:23-75,133-208
//   rtow final scene                 : examples/rtow_13_1.rs:15-46
//   cornell_box_scene                : examples/common/scene.rs:211-334
//   mesh test                        : examples/preview_sdl2.rs:452-525 (procedural mesh: the
//                                      reference's objects/dragon.obj is not in its repository)
#include <cmath>
#include <cstdlib>
#include <string>

#include "scene.hpp"
#include "seed_random.hpp"

namespace raysnail {
namespace recipes {

static TexturePtr solid(float r, float g, float b) { return std::make_shared<SolidColor>(Color(r, g, b, 1.f)); }

static bool range_contains(double lo, double hi, double v) { return v >= lo && v < hi; }

// scene.rs:23-75
static void add_small_balls(HittableList& world, SeedRandom& rng, double bounce_height) {
    const double small_ball_radius = 0.2;
    Point3 avoid(0.0, 0.2, 0.0);
    for (int a = -11; a < 11; ++a) {
        for (int b = -11; b < 11; ++b) {
            double cx = std::fma(0.9, rng.normal(), (double)a);
            double cy = 0.2 + rng.normal() * bounce_height;
            double cz = std::fma(0.9, rng.normal(), (double)b);
            Point3 center(cx, cy, cz);
            avoid.x = center.x;
            double ax = std::fabs(center.x);
            if (!(range_contains(0.0, 0.9, ax) || range_contains(3.1, 4.9, ax)) || (center - avoid).length() >= 0.9) {
                double mat = rng.normal();
                if (mat < 0.8) {
                    double r = rng.normal(), g = rng.normal(), bl = rng.normal();
                    auto color = std::make_shared<SolidColor>(Color::new64(r, g, bl, 1.0));
                    world.add(Sphere(center, small_ball_radius, std::make_shared<Lambertian>(color)));
                } else if (mat < 0.95) {
                    double r = rng.range(0.5, 1.0), g = rng.range(0.5, 1.0), bl = rng.range(0.5, 1.0);
                    auto color = std::make_shared<SolidColor>(Color::new64(r, g, bl, 1.0));
                    double fuzz = rng.range(0.0, 0.5);
                    if (fuzz < 0.1) world.add(Sphere(center, small_ball_radius, std::make_shared<Metal>(color)));
                    else world.add(Sphere(center, small_ball_radius, std::make_shared<DiffuseMetal>(fuzz * 1000.0, color)));
                } else {
                    auto glass = std::make_shared<Dielectric>(Color(1.f, 1.f, 1.f, 1.f), 1.5);
                    glass->reflect_curve_glass();
                    world.add(Sphere(center, small_ball_radius, glass));
                }
            }
        }
    }
}

// scene.rs:133-154
static void add_big_balls(HittableList& world) {
    auto glass = std::make_shared<Dielectric>(Color(1.f, 1.f, 1.f, 1.f), 1.5);
    glass->reflect_curve_glass();
    world.add(Sphere(Point3(0.0, 1.0, 0.0), 1.0, glass));
    world.add(Sphere(Point3(-4.0, 1.0, 0.0), 1.0, std::make_shared<Lambertian>(solid(0.4f, 0.2f, 0.1f))));
    world.add(Sphere(Point3(4.0, 1.0, 0.0), 1.0, std::make_shared<Metal>(solid(0.7f, 0.6f, 0.5f))));
}

HittableList balls_scene(uint64_t seed, bool checker) {
    HittableList list;
    if (checker) {
        auto tex = std::make_shared<Checker>(solid(0.3f, 0.3f, 0.3f), solid(0.1f, 0.1f, 0.1f), 10.0);
        list.add(Sphere(Point3(0.0, -1000.0, 0.0), 1000.0, std::make_shared<Lambertian>(tex)));
    } else {
        list.add(Sphere(Point3(0.0, -1000.0, 0.0), 1000.0, std::make_shared<Lambertian>(solid(0.5f, 0.5f, 0.5f))));
    }
    SeedRandom rng(seed);
    add_small_balls(list, rng, 0.9);
    add_big_balls(list);
    return list;
}

CameraBuilder balls_scene_camera(bool need_shutter_speed) {
    CameraBuilder b;
    b.look_from(Point3(13.0, 2.0, 3.0)).look_at(Point3(0.0, 0.0, 0.0)).fov(20.0).aperture(0.02).focus(10.0);
    if (need_shutter_speed) b.shutter_speed(1.0);
    return b;
}

std::unique_ptr<World> rtow_final_world(uint64_t seed, CameraBuilder& cam_out) {
    HittableList world = balls_scene(seed, true);
    cam_out = balls_scene_camera(false);
    HittableList lights;
    auto lm = std::make_shared<DiffuseLight>(solid(1.0f, 0.9f, 0.7f));
    lm->multiplier(1.5);
    auto rs = std::make_shared<Sphere>(Vec3(300.0, 400.0, 100.0), 12.0, lm);
    lights.add_ref(rs);
    world.add_ref(rs);
    return std::make_unique<World>(world, lights, Color(0.3f, 0.4f, 0.5f, 1.0f), Color(0.7f, 0.89f, 1.0f, 1.0f),
                                   TimeRange{0.0, 0.0});
}

std::unique_ptr<World> cornell_box_world(bool carton, bool carton_rotation, bool smoke, CameraBuilder& cam_out) {
    auto red = std::make_shared<Lambertian>(solid(0.65f, 0.05f, 0.05f));
    auto green = std::make_shared<Lambertian>(solid(0.12f, 0.45f, 0.15f));
    auto white = std::make_shared<Lambertian>(solid(0.73f, 0.73f, 0.73f));
    auto light = std::make_shared<DiffuseLight>(solid(1.f, 1.f, 1.f));
    light->multiplier(smoke ? 7.0 : 15.0);

    HittableList objects;
    objects.add(AARect::new_yz(AARectMetrics(555.0, 0.0, 555.0, 0.0, 555.0), green));
    objects.add(AARect::new_yz(AARectMetrics(0.0, 0.0, 555.0, 0.0, 555.0), red));
    objects.add(AARect::new_xz(AARectMetrics(0.0, 0.0, 555.0, 0.0, 555.0), white));
    objects.add(AARect::new_xz(AARectMetrics(555.0, 0.0, 555.0, 0.0, 555.0), white));
    objects.add(AARect::new_xy(AARectMetrics(555.0, 0.0, 555.0, 0.0, 555.0), white));
    HittablePtr light_rect;
    if (smoke) light_rect = std::make_shared<AARect>(AARect::new_xz(AARectMetrics(554.0, 113.0, 443.0, 127.0, 432.0), light));
    else light_rect = std::make_shared<AARect>(AARect::new_xz(AARectMetrics(554.0, 213.0, 343.0, 227.0, 332.0), light));
    objects.add_ref(light_rect);

    if (carton) {
        HittablePtr box1, box2;
        if (carton_rotation) {
            // NB: the reference passes -18.0 / 15.0 straight to rotate_by_y_axis, i.e. RADIANS
            TransformStack s1;
            s1.push(Transform::rotate_by_y_axis(-18.0));
            s1.push(Transform::translate(Vec3(130.0, 0.0, 65.0)));
            box1 = std::make_shared<TfFacade>(std::make_shared<Box>(Point3(0, 0, 0), Point3(165, 165, 165), white), s1);
            TransformStack s2;
            s2.push(Transform::rotate_by_y_axis(15.0));
            s2.push(Transform::translate(Vec3(265.0, 0.0, 295.0)));
            box2 = std::make_shared<TfFacade>(std::make_shared<Box>(Point3(0, 0, 0), Point3(165, 330, 165), white), s2);
        } else {
            box1 = std::make_shared<Box>(Point3(130, 0, 65), Point3(295, 165, 230), white);
            box2 = std::make_shared<Box>(Point3(265, 0, 295), Point3(430, 330, 460), white);
        }
        if (smoke) {
            objects.add_ref(std::make_shared<ConstantMedium>(box1, Color(1.f, 1.f, 1.f, 1.f), 0.01));
            objects.add_ref(std::make_shared<ConstantMedium>(box2, Color(0.f, 0.f, 0.f, 1.f), 0.01));
        } else {
            objects.add_ref(box1);
            objects.add_ref(box2);
        }
    }
    cam_out = CameraBuilder();
    cam_out.fov(40.0).look_from(Point3(278.0, 278.0, -800.0)).look_at(Point3(278.0, 278.0, 0.0));
    // No binary in the reference renders this scene, so its World is not given: the light list
    // is the ceiling rect (book convention) and the background is black (book convention).
    HittableList lights;
    lights.add_ref(light_rect);
    return std::make_unique<World>(objects, lights, Color(0.f, 0.f, 0.f, 1.f), Color(0.f, 0.f, 0.f, 1.f), TimeRange{0.0, 0.0});
}

// A closed, smooth, displaced UV-sphere ("blob") of 2*grid_n*(grid_n-1) triangles standing on
// the ground, with analytic-free (accumulated) vertex normals supplied as `vn`.
std::unique_ptr<World> mesh_test_world(int grid_n, bool blinn_phong, CameraBuilder& cam_out) {
    const double PI = 3.14159265358979323846;
    const int nu = grid_n, nv = grid_n;  // longitude x latitude segments
    std::vector<float> positions, normals;
    std::vector<uint32_t> indices;
    auto radius = [&](double th, double ph) {
        return 5.0 * (1.0 + 0.18 * std::sin(5.0 * ph) * std::sin(4.0 * th) + 0.08 * std::cos(9.0 * th + 2.0 * ph));
    };
    auto point = [&](double th, double ph) {
        double r = radius(th, ph);
        return Vec3(r * std::sin(th) * std::cos(ph), r * std::cos(th) + 5.5, r * std::sin(th) * std::sin(ph));
    };
    for (int j = 0; j <= nv; ++j) {
        double th = PI * (double)j / nv;
        for (int i = 0; i < nu; ++i) {
            double ph = 2.0 * PI * (double)i / nu;
            Vec3 p = point(th, ph);
            // numerical normal from the parametrisation
            const double e = 1e-4;
            Vec3 dth = point(std::fmin(th + e, PI), ph) - point(std::fmax(th - e, 0.0), ph);
            Vec3 dph = point(th, ph + e) - point(th, ph - e);
            Vec3 n = dph.cross(dth);
            if (n.length() < 1e-12) n = Vec3(0.0, j == 0 ? 1.0 : -1.0, 0.0);
            n = n.unit();
            positions.insert(positions.end(), {(float)p.x, (float)p.y, (float)p.z});
            normals.insert(normals.end(), {(float)n.x, (float)n.y, (float)n.z});
        }
    }
    for (int j = 0; j < nv; ++j)
        for (int i = 0; i < nu; ++i) {
            uint32_t a = (uint32_t)(j * nu + i), b = (uint32_t)(j * nu + (i + 1) % nu);
            uint32_t c = (uint32_t)((j + 1) * nu + i), d = (uint32_t)((j + 1) * nu + (i + 1) % nu);
            if (j != 0) { indices.push_back(a); indices.push_back(b); indices.push_back(c); }
            if (j != nv - 1) { indices.push_back(b); indices.push_back(d); indices.push_back(c); }
        }

    MaterialPtr material;
    if (blinn_phong) {
        material = std::make_shared<BlinnPhong>(0.2, 4.0, solid(0.87f, 0.25f, 0.1f));
    } else {
        auto m = std::make_shared<Lambertian>(solid(0.87f, 0.25f, 0.1f));
        m->settings.phong_factor = 4.0;
        m->settings.phong_exponent = 4;
        material = m;
    }
    TriangleMesh mesh = TriangleMesh::from_arrays(positions, normals, indices, 0.2, Vec3(0, 0, 0), 120.0, 1, material);

    cam_out = CameraBuilder();
    cam_out.look_from(Point3(13.0, 1.5, 3.0)).look_at(Point3(0.0, 0.3, 0.0)).fov(20.0).aperture(0.01).focus(10.0);
    HittableList world, lights;
    auto lm = std::make_shared<DiffuseLight>(solid(1.0f, 0.9f, 0.8f));
    lm->multiplier(1.5);
    auto rs = std::make_shared<Sphere>(Vec3(300.0, 400.0, 100.0), 12.0, lm);
    lights.add_ref(rs);
    world.add_ref(rs);
    for (Triangle& t : mesh.triangles) world.add(std::move(t));
    world.add(Sphere(Point3(0.0, -1000.0, 0.0), 1000.0, std::make_shared<Metal>(solid(0.08f, 0.1f, 0.06f))));
    return std::make_unique<World>(world, lights, Color(0.68f, 0.80f, 0.95f, 1.0f), Color(0.2f, 0.4f, 0.7f, 1.0f),
                                   TimeRange{0.0, 0.0});
}

// examples/preview_sdl2.rs:338-407 (`render_raymarching_test`): light sphere, the Mandelbulb with a
// Lambertian that carries Phong settings (4.0, 2), r=1000 DiffuseMetal(800) checker ground.
std::unique_ptr<World> raymarching_test_world(CameraBuilder& cam_out) {
    cam_out = CameraBuilder();
    cam_out.look_from(Point3(13.0, -1.7, 3.0) * 0.7).look_at(Point3(0.0, -0.4, 0.0)).fov(20.0).aperture(0.01).focus(10.0);
    HittableList world, lights;
    auto lm = std::make_shared<DiffuseLight>(solid(1.0f, 0.9f, 0.8f));
    lm->multiplier(1.5);
    auto rs = std::make_shared<Sphere>(Vec3(300.0, 400.0, 100.0), 12.0, lm);
    lights.add_ref(rs);
    world.add_ref(rs);
    auto m = std::make_shared<Lambertian>(solid(0.5f, 0.5f, 0.5f));
    m->settings.phong_factor = 4.0;
    m->settings.phong_exponent = 2;
    world.add(RayMarcher(m));
    auto ground = std::make_shared<Checker>(solid(0.26f, 0.3f, 0.16f), solid(0.1f, 0.1f, 0.1f), 10.0);
    world.add(Sphere(Point3(0.0, -1002.0, 0.0), 1000.0, std::make_shared<DiffuseMetal>(800.0, ground)));
    return std::make_unique<World>(world, lights, Color(0.68f, 0.80f, 0.95f, 1.0f), Color(0.2f, 0.4f, 0.7f, 1.0f),
                                   TimeRange{0.0, 0.0});
}

// examples/common/scene.rs:336-468 (`all_feature_scene`, the book-2 final scene).  The nested
// `BVH::new(boxes1)` object (scene.rs:361) is an accelerator, not geometry: `BVH::hit` returns the
// closest hit of its members, so its 400 boxes are added to the world list directly (the world
// BVH takes its place).  `boxes2` is built but never added by the reference (scene.rs:437-458);
// its 3000 draws are still consumed so that the RNG stream stays aligned.  No binary of the
// reference renders this scene: lights = the ceiling rect, black background (book convention).
// The earth texture is `examples/earth-map.png` relative to the working directory like in the
// reference, else $RAYSNAIL_ASSETS/earth-map.png, else a procedural 1920x960 stand-in.
std::unique_ptr<World> all_feature_world(uint64_t seed, CameraBuilder& cam_out) {
    const TimeRange time_limit{0.0, 1.0};
    const int boxes_per_side = 20;
    SeedRandom rng(seed);
    HittableList objects;
    auto ground = std::make_shared<Lambertian>(solid(0.48f, 0.83f, 0.53f));
    for (int i = 0; i < boxes_per_side; ++i)
        for (int j = 0; j < boxes_per_side; ++j) {
            const double w = 100.0;
            const double x0 = -1000.0 + (double)i * w, z0 = -1000.0 + (double)j * w, y0 = 0.0;
            const double x1 = x0 + w, y1 = rng.range(1.0, 100.0), z1 = z0 + w;
            objects.add(Box(Point3(x0, y0, z0), Point3(x1, y1, z1), ground));
        }
    auto light = std::make_shared<DiffuseLight>(solid(1.f, 1.f, 1.f));
    light->multiplier(7.0);
    HittablePtr light_rect = std::make_shared<AARect>(AARect::new_xz(AARectMetrics(554.0, 123.0, 423.0, 147.0, 412.0), light));
    objects.add_ref(light_rect);

    Sphere moving(Point3(400.0, 400.0, 200.0), 50.0, std::make_shared<Lambertian>(solid(0.7f, 0.3f, 0.1f)));
    moving.with_speed(Vec3(30.0, 0.0, 0.0));
    objects.add(std::move(moving));

    auto glass = [] {
        auto d = std::make_shared<Dielectric>(Color(1.f, 1.f, 1.f, 1.f), 1.5);
        d->reflect_curve_glass();
        return d;
    };
    objects.add(Sphere(Point3(260.0, 150.0, 45.0), 50.0, glass()));
    objects.add(Sphere(Point3(0.0, 150.0, 145.0), 50.0, std::make_shared<Metal>(solid(0.8f, 0.8f, 0.9f))));
    objects.add(Sphere(Point3(360.0, 170.0, 145.0), 70.0, glass()));
    objects.add_ref(std::make_shared<ConstantMedium>(
        std::make_shared<Sphere>(Point3(360.0, 170.0, 145.0), 70.0, std::make_shared<Lambertian>(solid(1.f, 1.f, 1.f))),
        Color(0.2f, 0.4f, 0.9f, 1.f), 0.2));
    objects.add_ref(std::make_shared<ConstantMedium>(std::make_shared<Sphere>(Point3(0.0, 0.0, 0.0), 5000.0, glass()),
                                                     Color(1.f, 1.f, 1.f, 1.f), 0.0001));

    std::shared_ptr<Image> earth;
    for (std::string path : {std::string("examples/earth-map.png"),
                             std::string(std::getenv("RAYSNAIL_ASSETS") ? std::getenv("RAYSNAIL_ASSETS") : ".") + "/earth-map.png"}) {
        try { earth = Image::open(path); break; } catch (...) {}
    }
    if (!earth) {  // procedural stand-in: oceans, two "continents", polar caps
        const int W = 1920, H = 960;
        std::vector<uint8_t> px((size_t)W * H * 4);
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x) {
                const double u = (double)x / W, v = (double)y / H;
                const double land = std::sin(6.2831853 * 2.0 * u + 3.0 * std::sin(9.0 * v)) * std::cos(3.14159265 * (v - 0.5) * 1.7);
                uint8_t r = 20, g = 60, b = 140;
                if (land > 0.35) { r = 60; g = 130; b = 50; }
                if (v < 0.07 || v > 0.93) { r = 235; g = 240; b = 245; }
                uint8_t* q = &px[4 * ((size_t)y * W + x)];
                q[0] = r; q[1] = g; q[2] = b; q[3] = 255;
            }
        earth = std::make_shared<Image>(W, H, std::move(px));
    }
    objects.add(Sphere(Point3(400.0, 200.0, 400.0), 100.0, std::make_shared<Lambertian>(earth)));

    auto tex = std::make_shared<Perlin>(256, true, seed ^ 0x9E3779B97F4A7C15ull);
    tex->scale(0.1).smooth(SmoothType::HermitianCubic);
    objects.add(Sphere(Point3(220.0, 280.0, 300.0), 80.0, std::make_shared<Lambertian>(tex)));

    for (int k = 0; k < 1000; ++k) {  // boxes2: drawn, then dropped (scene.rs:437-458)
        (void)rng.range(0.0, 165.0);
        (void)rng.range(0.0, 165.0);
        (void)rng.range(0.0, 165.0);
    }

    cam_out = CameraBuilder();
    cam_out.look_from(Point3(478.0, 278.0, -600.0)).look_at(Point3(278.0, 278.0, 0.0)).fov(40.0).shutter_speed(1.0);
    HittableList lights;
    lights.add_ref(light_rect);
    return std::make_unique<World>(objects, lights, Color(0.f, 0.f, 0.f, 1.f), Color(0.f, 0.f, 0.f, 1.f), time_limit);
}

}  // namespace recipes
}  // namespace raysnail
