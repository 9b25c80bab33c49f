/*
 * raysnail_b200_host.h — C view of the C++ host layer (csrc/host/scene.hpp).
 *
 * The reference's scene API lives in Rust (Hittable/Material/Texture constructors, SdlParser,
 * CameraBuilder).  No Rust toolchain exists in this environment, so the host side above the
 * hot-path ABI (raysnail_b200.h) is C++ that mirrors those interfaces; this header lets
 * Python (ctypes) drive it.  Every function here runs on the CPU and needs no GPU.
 *
 * Handles are small non-negative ints scoped to one rs_host_scene; negative = error.
 */
#ifndef RAYSNAIL_B200_HOST_H
#define RAYSNAIL_B200_HOST_H

#include "raysnail_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rs_host_scene rs_host_scene;

const char* rs_host_last_error(void); /* thread-local message of the last failing rs_host_* call */

/* ---- whole scenes ------------------------------------------------------------------------*/
/* SdlParser::parse (src/sdl_parser.rs:187-205) + parse_and_render's scene assembly
 * (src/bin/raysnail.rs:327-377). */
int rs_host_scene_from_sdl_file(const char* path, rs_host_scene** out);
int rs_host_scene_from_sdl_text(const char* text, rs_host_scene** out);
/* Example recipes: "rtow_final" (examples/rtow_13_1.rs; seed), "cornell" / "cornell_smoke" /
 * "cornell_plain" (examples/common/scene.rs:211-334), "mesh" / "mesh_blinn"
 * (examples/preview_sdl2.rs:452-525; iparam = grid size, triangles = 2*n*(n-1)). */
int rs_host_scene_builtin(const char* name, uint64_t seed, int32_t iparam, rs_host_scene** out);
void rs_host_scene_free(rs_host_scene* s);

/* Flat description (valid until the scene is freed or modified). */
const rs_scene_desc* rs_host_scene_desc(rs_host_scene* s);
/* The scene's CameraBuilder with width/height applied -> Camera::new (src/camera.rs:37-73). */
int rs_host_scene_camera(rs_host_scene* s, int32_t width, int32_t height, rs_camera* out);

/* CameraBuilder (src/camera.rs:299-414). */
int rs_host_camera_build(const double look_from[3], const double look_at[3], const double vup[3],
                         double fov, double aperture, double focus, double shutter_speed,
                         int32_t width, int32_t height, rs_camera* out);

/* ---- incremental construction (mirrors the Rust constructors) ----------------------------*/
rs_host_scene* rs_host_scene_new(void);
int rs_host_tex_solid(rs_host_scene* s, float r, float g, float b, float a);
int rs_host_tex_checker(rs_host_scene* s, int odd, int even, double scale);
int rs_host_tex_image(rs_host_scene* s, int32_t width, int32_t height, const uint8_t* rgba);
int rs_host_tex_image_file(rs_host_scene* s, const char* path);
/* kind 0 Normal / 1 Turbulence / 2 Marble; smooth 0/1/2 (src/texture/noise.rs) */
int rs_host_tex_perlin(rs_host_scene* s, int32_t point_count, int32_t vector, uint64_t seed,
                       int32_t kind, int32_t depth, int32_t smooth, double scale);

int rs_host_mat_lambertian(rs_host_scene* s, int tex);
int rs_host_mat_metal(rs_host_scene* s, int tex);
int rs_host_mat_diffuse_metal(rs_host_scene* s, double exponent, int tex);
int rs_host_mat_dielectric(rs_host_scene* s, float r, float g, float b, double refractive, int glass_curve);
int rs_host_mat_diffuse_light(rs_host_scene* s, int tex, double multiplier);
int rs_host_mat_isotropic(rs_host_scene* s, float r, float g, float b);
int rs_host_mat_blinn_phong(rs_host_scene* s, double k_specular, double exponent, int tex);
int rs_host_mat_mixed(rs_host_scene* s, int mat1, int mat2, double probability_1);
int rs_host_mat_set_phong(rs_host_scene* s, int mat, double phong_factor, int32_t phong_exponent);

/* mat < 0 = None */
int rs_host_obj_sphere(rs_host_scene* s, const double center[3], double radius, const double speed[3], int mat);
int rs_host_obj_box(rs_host_scene* s, const double p0[3], const double p1[3], int mat);
int rs_host_obj_rect(rs_host_scene* s, int32_t rect_axis, double k, double a0, double a1, double b0, double b1, int mat);
int rs_host_obj_quadric(rs_host_scene* s, const double q[10], int mat);
int rs_host_obj_raymarcher(rs_host_scene* s, int mat); /* RayMarcher::new (raymarching.rs:35-40) */
int rs_host_obj_difference(rs_host_scene* s, int plus, int minus, int mat);
int rs_host_obj_intersection(rs_host_scene* s, int o1, int o2, int mat);
/* `BVH::new(list, time_range)` used as ONE object (bvh.rs:47-112, 173-192); the scene's time range as set so far */
int rs_host_obj_bvh(rs_host_scene* s, int32_t n, const int32_t* objs);
/* ops[i]: 0 translate(v) 1 rotate_x(v[0] rad) 2 rotate_y 3 rotate_z 4 scale(v); args = n*3 doubles */
int rs_host_obj_transform(rs_host_scene* s, int child, int32_t n_ops, const int32_t* ops, const double* args);
int rs_host_obj_medium(rs_host_scene* s, int boundary, float r, float g, float b, double density);

int rs_host_world_add(rs_host_scene* s, int obj);
int rs_host_world_add_light(rs_host_scene* s, int obj);
/* TriangleMesh::load semantics on arrays; every triangle is added to the world list. */
int rs_host_world_add_mesh(rs_host_scene* s, const float* positions, int64_t n_positions,
                           const float* normals, int64_t n_normals, const uint32_t* indices,
                           int64_t n_indices, double scale, const double offset[3],
                           double rotation_angle, int32_t axis, int mat);
int rs_host_world_add_obj_file(rs_host_scene* s, const char* path, double scale, const double offset[3],
                               double rotation_angle, int32_t axis, int mat);
int rs_host_world_set_background(rs_host_scene* s, const float c0[4], const float c1[4]);
int rs_host_world_set_time_range(rs_host_scene* s, double t0, double t1);
int rs_host_scene_set_camera(rs_host_scene* s, const double look_from[3], const double look_at[3],
                             const double vup[3], double fov, double aperture, double focus,
                             double shutter_speed);

/* ---- utilities ---------------------------------------------------------------------------*/
/* Raw StdRng stream of SeedRandom (src/prelude/random.rs:75-105) for the known-answer tests. */
int rs_host_seed_random_u32(uint64_t seed, uint32_t* out, int32_t n);
int rs_host_chacha12_block(const uint32_t key[8], uint64_t counter, uint32_t out[16]);
int rs_host_write_png_rgb8(const char* path, const uint8_t* rgb, int32_t width, int32_t height);

#ifdef __cplusplus
}
#endif
#endif
