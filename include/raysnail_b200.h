/*
 * raysnail_b200.h — C ABI of the B200 (sm_100a) path-tracing backend for Raysnail.
 *
 * This is the drop-in boundary for ONE hot path of the reference renderer: the per-pixel
 * Monte-Carlo sample loop that `TakePhotoSettings::shot_to_target` (reference
 * src/camera.rs:261-287) hands to `Painter::draw` (src/painter.rs:307-336).  A host (the
 * reference's Rust crate through `extern "C"`, or the C++/Python host in this repository)
 * flattens its scene graph into the plain SoA/AoS buffers below, uploads them once, and
 * calls `rs_render*`; the library returns exactly what `shot_to_target` returns: a row-major
 * `[f32;4]` image (src/painter.rs:154-187, src/prelude/vec3.rs:227-240).
 *
 * Rules of the boundary:
 *   - plain C: pointers, sizes, PODs; no exceptions cross it; every call returns an rs_status;
 *   - the caller owns every host buffer; `rs_upload_scene` copies what it needs;
 *   - one rs_ctx is bound to one CUDA device and is used by one host thread at a time;
 *   - there is no CPU fallback: if no CUDA device is usable, rs_create fails.
 *
 * All geometry is f64 (the reference's Vec3 is f64, src/prelude/vec3.rs:14-19); colours are
 * f32 (src/prelude/color.rs:10-16).
 */
#ifndef RAYSNAIL_B200_H
#define RAYSNAIL_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RS_ABI_VERSION 3

typedef enum rs_status {
    RS_OK = 0,
    RS_ERR_INVALID_ARGUMENT = 1,
    RS_ERR_CUDA = 2,
    RS_ERR_NO_DEVICE = 3,
    RS_ERR_NO_SCENE = 4,
    RS_ERR_UNSUPPORTED = 5,
    RS_ERR_OUT_OF_MEMORY = 6,
    RS_ERR_PARSE = 7,
    RS_ERR_IO = 8
} rs_status;

/* ------------------------------------------------------------------------------------------
 * Scene graph nodes.  One node per reference `Hittable` object (src/hittable/hit.rs:80-112).
 * ---------------------------------------------------------------------------------------- */
typedef enum rs_node_type {
    RS_NODE_SPHERE = 0,       /* src/hittable/geometry/sphere.rs   p: cx cy cz r vx vy vz          */
    RS_NODE_BOX = 1,          /* src/hittable/geometry/box.rs      p: minx miny minz maxx maxy maxz */
    RS_NODE_RECT = 2,         /* src/hittable/geometry/rect.rs     p: k a0 a1 b0 b1 ; rect_axis     */
    RS_NODE_QUADRIC = 3,      /* src/hittable/geometry/quadric.rs  p: qa qb qc qd qe qf qg qh qi qj */
    RS_NODE_DIFFERENCE = 4,   /* src/hittable/csg/difference.rs    child0 = plus, child1 = minus    */
    RS_NODE_INTERSECTION = 5, /* src/hittable/csg/intersection.rs  child0 = o1,   child1 = o2       */
    RS_NODE_TRANSFORM = 6,    /* src/hittable/transform/tf_facade.rs child0; xforms[first..first+n) */
    RS_NODE_MEDIUM = 7,       /* src/hittable/medium/constant.rs   child0 = boundary; p[0]=density;
                                 material = the Isotropic material                                  */
    RS_NODE_RAYMARCHER = 8,   /* src/hittable/geometry/raymarching.rs  the power-8 Mandelbulb about the
                                 origin; p[0] = iterations (RayMarcher::new: 100)                  */
    RS_NODE_BVH = 9           /* src/hittable/collection/bvh.rs:173-192 — a `BVH` used as ONE object (a world
                                 member, under a TfFacade, as a medium boundary): child0 = left, child1 = right
                                 (-1 = None), p[0..5] = the node's bbox (min xyz, max xyz), p[6] = 1 when the
                                 bbox is Some.  Its `contains` is `unimplemented!()` in the reference, so a BVH
                                 is refused as a CSG operand (RS_ERR_UNSUPPORTED)                     */
} rs_node_type;

/* rect_axis: which AARect constructor built it (src/hittable/geometry/rect.rs:56-78) */
#define RS_RECT_XY 0 /* axis = (0,1,2) */
#define RS_RECT_XZ 1 /* axis = (0,2,1) */
#define RS_RECT_YZ 2 /* axis = (1,2,0) */

typedef struct rs_node {
    int32_t type;        /* rs_node_type */
    int32_t material;    /* index into materials, -1 = None (Option<Arc<dyn Material>>) */
    int32_t child0;
    int32_t child1;
    int32_t first_xform; /* RS_NODE_TRANSFORM: first entry of its TransformStack */
    int32_t n_xform;     /* RS_NODE_TRANSFORM: number of stack entries           */
    int32_t rect_axis;   /* RS_NODE_RECT */
    int32_t reserved;
    double p[10];
} rs_node;

/* One TransformStack entry: `Transform{matrix, inverse}`, row-major 4x4
 * (src/hittable/transform/transform.rs:8-13). */
typedef struct rs_xform {
    double m[16];
    double inv[16];
} rs_xform;

/* ------------------------------------------------------------------------------------------
 * Materials (src/material/) and textures (src/texture/).
 * ---------------------------------------------------------------------------------------- */
typedef enum rs_material_type {
    RS_MAT_LAMBERTIAN = 0,    /* lambertian.rs    texture                                      */
    RS_MAT_METAL = 1,         /* metal.rs:86-127  texture                                      */
    RS_MAT_DIFFUSE_METAL = 2, /* metal.rs:22-84   texture, a = exponent                        */
    RS_MAT_DIELECTRIC = 3,    /* dielectric.rs    color, a = refractive, has_curve (Glass)     */
    RS_MAT_DIFFUSE_LIGHT = 4, /* light.rs         texture, a = multiplier                      */
    RS_MAT_ISOTROPIC = 5,     /* isotropic.rs     color                                        */
    RS_MAT_BLINN_PHONG = 6,   /* blinn_phong.rs   texture, a = k_specular, b = exponent        */
    RS_MAT_MIXED = 7          /* mixed_material.rs mat1, mat2, a = probability_1               */
} rs_material_type;

typedef struct rs_material {
    int32_t type;           /* rs_material_type */
    int32_t texture;        /* index into textures, -1 if unused */
    int32_t mat1;           /* RS_MAT_MIXED */
    int32_t mat2;           /* RS_MAT_MIXED */
    int32_t has_curve;      /* RS_MAT_DIELECTRIC: 1 = Glass Schlick curve (dielectric.rs:17-25) */
    int32_t phong_exponent; /* CommonMaterialSettings (material/mod.rs:41-55), default 1 */
    double phong_factor;    /* default 0 */
    double a;
    double b;
    float color[4];
} rs_material;

typedef enum rs_texture_type {
    RS_TEX_SOLID = 0,   /* prelude/color.rs:61-65 */
    RS_TEX_CHECKER = 1, /* texture/checker.rs     */
    RS_TEX_IMAGE = 2,   /* texture/image.rs       */
    RS_TEX_PERLIN = 3   /* texture/noise.rs       */
} rs_texture_type;

typedef struct rs_texture {
    int32_t type;          /* rs_texture_type */
    int32_t odd;           /* CHECKER: texture id used when the sign product is < 0 */
    int32_t even;          /* CHECKER */
    int32_t width;         /* IMAGE */
    int32_t height;        /* IMAGE */
    int32_t texel_offset;  /* IMAGE: byte offset into texels (RGBA8, row-major, row 0 = top) */
    int32_t perlin_kind;   /* PERLIN: 0 Normal, 1 Turbulence, 2 Marble (noise.rs:10-15) */
    int32_t perlin_depth;  /* PERLIN: turbulence depth */
    int32_t perlin_smooth; /* PERLIN: 0 None, 1 LinearInterpolate, 2 HermitianCubic */
    int32_t perlin_vector; /* PERLIN: 1 = unit-vector lattice, 0 = float lattice */
    int32_t point_count;   /* PERLIN: power of two */
    int32_t perm_offset;   /* PERLIN: offset (in int32) into perlin_perm: perm_x, perm_y, perm_z */
    int32_t value_offset;  /* PERLIN: offset (in double) into perlin_values: point_count*(vector?3:1) */
    int32_t reserved;
    double scale;          /* CHECKER, PERLIN */
    float color[4];        /* SOLID */
} rs_texture;

/* ------------------------------------------------------------------------------------------
 * The flattened `World` (src/hittable/collection/world.rs:25-53).
 *
 * roots[i] >= 0  : node index of the i-th object of the world's HittableList;
 * roots[i] <  0  : triangle ~roots[i] (src/hittable/geometry/triangle_mesh.rs:17-149).
 * The position i in `roots` is the object id reported by the parity probes.
 * root_bbox[6*i..] = the object's `Hittable::bbox` as the reference computes it
 * (min xyz, max xyz); the host already has it.  An object only counts as hit when the ray
 * passes `AABB::hit` on this box (src/prelude/aabb.rs:20-38): that reproduces the clipping the
 * reference gets from e.g. the fixed +-100 quadric box and the Intersection bbox quirk
 * (src/hittable/csg/intersection.rs:102-118).
 * ---------------------------------------------------------------------------------------- */
typedef struct rs_scene_desc {
    const rs_node* nodes;
    int32_t n_nodes;
    const int32_t* roots;
    const double* root_bbox;
    int32_t n_roots;
    const int32_t* lights; /* node indices of `World.lights` in list order */
    int32_t n_lights;
    const rs_xform* xforms;
    int32_t n_xforms;
    const double* tri_verts;     /* n_tris * 9: p0 p1 p2                     */
    const double* tri_normals;   /* n_tris * 9: normal0 normal1 normal2      */
    const int32_t* tri_material; /* n_tris, -1 = None                        */
    int32_t n_tris;
    const rs_material* materials;
    int32_t n_materials;
    const rs_texture* textures;
    int32_t n_textures;
    const uint8_t* texels;
    int64_t n_texel_bytes;
    const int32_t* perlin_perm;
    int32_t n_perlin_perm;
    const double* perlin_values;
    int32_t n_perlin_values;
    int32_t default_material; /* World.default_material (world.rs:29,51): white Lambertian */
    /* Background closure restricted to the one form every shipped caller uses:
     * bg0.gradient(bg1, 0.5*(dir.y+1))  (raysnail.rs:364-367, rtow_13_1.rs:38-41, world.rs:20-23) */
    float bg0[4];
    float bg1[4];
} rs_scene_desc;

/* `Camera` after `Camera::new` (src/camera.rs:18-73). */
typedef struct rs_camera {
    double origin[3];
    double lb[3];
    double horizontal_full[3];
    double vertical_full[3];
    double horizontal_unit[3];
    double vertical_unit[3];
    double aperture;
    double shutter_speed;
    int32_t width;
    int32_t height;
} rs_camera;

/* `Painter` + `TakePhotoSettings` (src/painter.rs:66-130, src/camera.rs:103-154). */
typedef struct rs_render_params {
    int32_t sqrt_spp;          /* Painter::samples: floor(sqrt(samples_requested)) */
    int32_t depth;             /* TakePhotoSettings::depth, default 8 */
    int32_t gamma;             /* 1 = sqrt per channel (vec3.rs:229-233) */
    int32_t stratum_row_begin; /* render strata rows s_j in [begin, end); whole image: 0..sqrt_spp */
    int32_t stratum_row_end;
    int32_t max_wave_paths;    /* 0 = library default; upper bound on paths in flight */
    uint64_t seed;             /* Philox4x32-10 key */
    uint32_t flags;            /* RS_FLAG_* */
    uint32_t reserved;
} rs_render_params;

#define RS_FLAG_PIXEL_CENTRE 1u  /* probes only: no jitter, no lens, time 0 (u,v at stratum centre) */
#define RS_FLAG_KERNEL_TIMING 2u /* record CUDA events around every kernel; fills rs_stats.ms_* (syncs the stream) */
/* Opt-in integrator: the MixturePdf(HittablePdf(world.lights), srec.pdf) sampling that the reference keeps
 * commented out in ray_color (src/camera.rs:188-192, 222-225; types in src/prelude/pdf.rs:237-307).  The
 * scattered ray starts at the hit point, its direction is MixturePdf::generate, and the sample is weighted
 * by srec.pdf.value / MixturePdf::value.  Default (flag clear) = the live 50/50 branch of camera.rs:194-247. */
#define RS_FLAG_MIS 4u

typedef struct rs_stats {
    uint64_t n_samples;        /* camera paths started */
    uint64_t n_segments;       /* closest-hit queries (ray segments) */
    uint64_t n_launches;       /* kernels launched by the last render call */
    double ms_raygen, ms_extend, ms_shade, ms_accumulate; /* CUDA-event time of last render */
    double ms_total;
} rs_stats;

typedef struct rs_ctx rs_ctx;

/* ---- life cycle --------------------------------------------------------------------------*/
int rs_abi_version(void);
/* sizeof() of the ABI structs as this library was compiled, so a binding (cgo, ctypes, Rust
 * #[repr(C)]) can verify its layout: 0 rs_node, 1 rs_xform, 2 rs_material, 3 rs_texture,
 * 4 rs_scene_desc, 5 rs_camera, 6 rs_render_params, 7 rs_stats; -1 for anything else. */
int rs_abi_struct_size(int which);
/* Bind a context to CUDA device `device`.  Fails (RS_ERR_NO_DEVICE) without a GPU. */
int rs_create(rs_ctx** out, int device);
void rs_destroy(rs_ctx* ctx);
/* Message of the last failed call on this context (or of a failed rs_create when ctx == NULL). */
const char* rs_last_error(const rs_ctx* ctx);

/* Tuning options; set before rs_upload_scene.  RS_OPT_BVH_WIDTH: 2 (default) = binary world BVH of
 * 64-byte nodes, 4 = QBVH of 128-byte nodes (both give identical hits; see DESIGN.md for the
 * measured trade-off).  */
#define RS_OPT_BVH_WIDTH 1
/* RS_OPT_WAVE_STREAMS: n in 1..4 (default 4) = consecutive waves rotate over the caller's stream and
 * n-1 internal ones, so the tail of each persistent kernel overlaps with another wave's work (the image
 * is bit-identical: accumulation stays in wave order); 1 = every kernel on the caller's stream. */
#define RS_OPT_WAVE_STREAMS 2
/* RS_OPT_LIST_MAX: n in 0..64 (default 12, 24 for worlds that hold expression trees) = worlds of at most n objects that hold boxes, rects or expression
 * trees are searched by walking the object list (every lane of a warp tests the same object at the same time)
 * instead of the tree; 0 = always the tree.  Identical hits either way.  Set before rs_upload_scene. */
#define RS_OPT_LIST_MAX 3
/* RS_OPT_WAIT_BUDGET_MS: n >= 0 (default 600000 = 10 min) = how long the wait kernels of rs_wait_counter and
 * rs_multi_render poll for their peers before they give up; 0 = for ever.  A peer whose host failed between two
 * launches never signals: with a budget, rs_multi_render then returns an error, re-bases its counters and the next
 * call works again; a wait queued by rs_wait_counter raises a flag of its context and lets the stream go on (what was
 * queued behind it then runs on incomplete data): ask rs_wait_check.  Choose it well above the longest frame a device renders. */
#define RS_OPT_WAIT_BUDGET_MS 4
/* RS_OPT_INJECT_FAULT (rs_multi_set_option only; for tests): the next rs_multi_render fails on purpose on its last
 * device — 1 = before anything is signalled, 2 = between its two signals — to exercise the recovery path. */
#define RS_OPT_INJECT_FAULT 100
int rs_set_option(rs_ctx* ctx, int option, int value);

/* ---- scene: replaces `World::new` (world.rs:40-53) + `BVH::new` (bvh.rs:47-112) -----------*/
int rs_upload_scene(rs_ctx* ctx, const rs_scene_desc* scene);

/* ---- the hot path: replaces `shot_to_target` (camera.rs:261-287) ---------------------------*/
/* Renders into a host image: out_rgba = W*H*[f32;4]; pixel_mask (nullable, W*H bytes) is the
 * flattened `PixelController::calculate_pixel` (painter.rs:204-210): 0 -> [0,0,0,0]. */
int rs_render(rs_ctx* ctx, const rs_camera* cam, const rs_render_params* params,
              const uint8_t* pixel_mask, float* out_rgba);

/* Device-resident variant used for multi-GPU sharding: ADDS the linear per-pixel radiance sums
 * of strata rows [begin,end) into d_accum (device, W*H float4, xyz = sum, w = sample count).
 * d_pixel_mask: device pointer or NULL.  `stream` is a cudaStream_t (0 = legacy default). */
int rs_render_accumulate(rs_ctx* ctx, const rs_camera* cam, const rs_render_params* params,
                         const uint8_t* d_pixel_mask, float* d_accum, void* stream);

/* `Vec3::into_color` (vec3.rs:227-240) over a summed buffer: out = [sqrt(sum/n)..., 1] or
 * [0,0,0,0] where masked.  All pointers are device pointers. n_samples = sqrt_spp^2. */
int rs_finalize(rs_ctx* ctx, const float* d_accum, const uint8_t* d_pixel_mask, float* d_out_rgba,
                int32_t width, int32_t height, int32_t n_samples, int32_t gamma, void* stream);

/* ---- multi-GPU: fused reduce + finalize over NVLink peer memory ------------------------------
 * The reference has no multi-device path (its only parallelism is painter.rs:275-299's thread
 * pool); these entry points shard ONE `shot_to_target` over the GPUs of a box.  Every process
 * (one per GPU) renders its strata rows with rs_render_accumulate into a buffer obtained from
 * rs_device_alloc, exports it (rs_ipc_export: a 64-byte cudaIpcMemHandle_t that the host side
 * passes to the other processes by any means) and maps the buffers of its peers (rs_ipc_open).
 * rs_finalize_peers then does, in ONE kernel and with no intermediate buffer, what
 * ncclReduce + rs_finalize would do: for pixels [pixel_begin, pixel_end) it sums the n_peers
 * accumulators (peer loads over NVLink, fixed order => deterministic), applies
 * `Vec3::into_color` (vec3.rs:227-240) and stores the pixel into d_out_rgba, which may itself
 * be a peer pointer (the image of the rank that owns the result).  Each rank finalizes its own
 * pixel slice, so the traffic is spread over all NVLink ports (reduce-scatter + gather shape).
 * The caller orders it after the peers' renders (stream sync + a host barrier). */
int rs_device_alloc(rs_ctx* ctx, int64_t bytes, void** d_ptr); /* cudaMalloc + zero fill */
int rs_device_free(rs_ctx* ctx, void* d_ptr);
#define RS_IPC_HANDLE_BYTES 64
int rs_ipc_export(rs_ctx* ctx, const void* d_ptr, uint8_t handle[RS_IPC_HANDLE_BYTES]);
int rs_ipc_open(rs_ctx* ctx, const uint8_t handle[RS_IPC_HANDLE_BYTES], void** d_ptr);
int rs_ipc_close(rs_ctx* ctx, void* d_ptr);
#define RS_MAX_PEERS 16
int rs_finalize_peers(rs_ctx* ctx, const float* const* d_accums, int32_t n_peers,
                      const uint8_t* d_pixel_mask, float* d_out_rgba, int64_t pixel_begin,
                      int64_t pixel_end, int32_t n_samples, int32_t gamma, void* stream);

/* ---- passes logic of the CLI (src/bin/raysnail.rs:157-208, 379-427); device pointers -------*/
/* combine_pixel: new == (0,0,0,0) ? old : (old*pass + new)/(pass+1)   (raysnail.rs:176-193) */
int rs_combine_passes(rs_ctx* ctx, const float* d_old, const float* d_new, float* d_out,
                      int64_t n_pixels, float pass, void* stream);
/* calc_noise (raysnail.rs:157-173), including its `x = y` indexing quirk when quirk != 0. */
int rs_noise_map(rs_ctx* ctx, const float* d_pixels, float* d_noise, int32_t width, int32_t height,
                 int32_t quirk, void* stream);
/* PNG quantisation (raysnail.rs:437-441): (clamp(c,0,1)*255.5) as u8; out = W*H*3 bytes. */
int rs_quantize_rgb8(rs_ctx* ctx, const float* d_pixels, uint8_t* d_rgb, int64_t n_pixels,
                     void* stream);

/* The whole passes loop of the CLI (src/bin/raysnail.rs:379-427) in one call, device-resident between
 * passes: old = [0,0,0,1]; for pass in 0..n_passes { new = shot_to_target(mask); old = combine_pixels(old,
 * new, pass); noise = calc_noise(old); redo = noise >= noise_threshold }.  Pass k renders with the Philox
 * key rs_pass_seed(params->seed, k) (the reference gets fresh RNG state on every shot_to_target call).
 *   RS_REDO_REFERENCE: what the reference does — its RedoController holds a stale clone of the all-ones
 *                      map (raysnail.rs:369-372), so every pass re-renders every pixel; the map is only counted;
 *   RS_REDO_NOISE:     opt-in, the map is used: pass k+1 renders only pixels with noise >= threshold
 *                      (noise computed like the reference, `x = y` indexing quirk included);
 *   RS_REDO_NOISE_FIXED: as RS_REDO_NOISE with the noise window centred on (x, y).
 * The reference's threshold is 0.01 (raysnail.rs:404).  out_rgba: host, W*H*[f32;4] (the final old_pixels);
 * out_rgb8: host, W*H*3 PNG bytes (raysnail.rs:431-441) or NULL; redo_counts: host, n_passes entries
 * ("oversampling N pixels", raysnail.rs:424) or NULL. */
#define RS_REDO_REFERENCE 0
#define RS_REDO_NOISE 1
#define RS_REDO_NOISE_FIXED 2
int rs_render_passes(rs_ctx* ctx, const rs_camera* cam, const rs_render_params* params, int32_t n_passes,
                     int32_t redo_mode, float noise_threshold, float* out_rgba, uint8_t* out_rgb8,
                     int64_t* redo_counts);
/* Philox key of pass k: seed + k * 0x9E3779B97F4A7C15 (wrapping); pass 0 renders with `seed` itself. */
uint64_t rs_pass_seed(uint64_t seed, int32_t pass);

/* ---- the RNG, for host harnesses (replaces FastRng, src/prelude/random.rs:109-145) ---------------
 * Philox4x32-10 (Salmon et al., SC'11) exactly as the kernels evaluate it, so that a host-side harness
 * (the Rust reference with its RNG swapped, or a test) can draw the very numbers a path sees.  Plain CPU
 * integer code, no GPU needed.  Stream layout: key = (seed lo, seed hi), counter = (pixel, sample, depth,
 * purpose | call << 8); uniform = (word >> 8) * 2^-24.  Purposes: RS_RNG_*; `depth` is 0 for the camera
 * ray and b + 1 for everything drawn at the b-th hit of the path. */
#define RS_RNG_PIXEL 0u   /* words 0,1: jitter x,y (painter.rs:169-170); word 2: shutter time (camera.rs:84) */
#define RS_RNG_LENS 1u    /* random_unit_disk rejection loop (vec3.rs:140-147): iteration k = word pair k&1 of call k>>1 */
#define RS_RNG_SCATTER 2u /* word 0: light/BRDF branch (camera.rs:198); 1: light pick (list.rs:49-52); 2: MixedMaterial level 0; 3: Dielectric */
#define RS_RNG_MIXED 3u   /* call = nesting level >= 1 of a MixedMaterial (mixed_material.rs:44), word 0 */
#define RS_RNG_LIGHT 4u   /* Hittable::random of the picked light (sphere.rs:153-163 loop / rect.rs:149-150) */
#define RS_RNG_BRDF 5u    /* PDF::generate (pdf.rs) */
#define RS_RNG_MEDIUM 6u  /* ConstantMedium free flight (constant.rs:62), call = node index of the medium */
int rs_philox4x32_10(const uint32_t counter[4], const uint32_t key[2], uint32_t out[4]);
int rs_philox_uniform(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t depth, uint32_t purpose,
                      uint32_t call, double out4[4]);

/* ---- parity probes (host output buffers, one entry per pixel) ------------------------------*/
/* Primary ray of sample `sample_index` (= s_j*sqrt_spp + s_i) of every pixel: closest hit.
 * obj_id: index into roots or -1; t: hit.t1; normal/point: HitRecord fields; front: hit.outside;
 * material: resolved material index (default material substituted). Any pointer may be NULL. */
int rs_debug_primary(rs_ctx* ctx, const rs_camera* cam, const rs_render_params* params,
                     int32_t sample_index, int32_t* obj_id, double* t, double* normal3,
                     double* point3, int32_t* front, int32_t* material, double* ray_od6);

/* One bounce: the primary hit is shaded once (camera.rs:165-250).
 * kind: 0 miss, 1 no scatter (emitted only), 2 skip_pdf specular, 3 light branch, 4 pdf branch.
 * next_od6: scattered ray origin+direction; weight3: colour factor applied to the recursive
 * ray_color (srec.color * light_multi * pdf_multiplicator); emitted3: radiance added here
 * (emitted, or the background colour on a miss). */
int rs_debug_scatter(rs_ctx* ctx, const rs_camera* cam, const rs_render_params* params,
                     int32_t sample_index, int32_t* kind, double* next_od6, double* weight3,
                     double* emitted3);

/* ---- several devices behind one call (replaces the caller-side fan-out a multi-GPU
 * `Camera::take_photo` — camera.rs:283-330 is the single-device loop — would otherwise need) ------------
 * One process, one host thread per device.  The scene is replicated; the strata rows
 * [stratum_row_begin, stratum_row_end) of `params` are split evenly over the devices, every device
 * sum-reduces and finalizes its slice of the pixels out of all accumulators (NVLink peer memory) into
 * device 0's image.  The devices order themselves with arrival counters in device 0's memory — no host
 * barrier between render and combine.  device_ids may name a device twice (separate contexts on one GPU). */
typedef struct rs_multi rs_multi;
int rs_create_multi(rs_multi** out, const int32_t* device_ids, int32_t n);
void rs_destroy_multi(rs_multi* m);
const char* rs_multi_last_error(const rs_multi* m); /* m may be NULL: error of the last failed rs_create_multi */
int rs_multi_device_count(const rs_multi* m);
int rs_multi_set_option(rs_multi* m, int option, int value);
int rs_multi_upload_scene(rs_multi* m, const rs_scene_desc* scene);
/* out_rgba: HOST, width*height*4 floats (as rs_render).  stats_sum (may be NULL): counters summed over devices. */
int rs_multi_render(rs_multi* m, const rs_camera* cam, const rs_render_params* params, const uint8_t* pixel_mask,
                    float* out_rgba, rs_stats* stats_sum);
/* The counters themselves, for callers that run one PROCESS per device (torchrun): d_counter is a uint32 in
 * memory every rank mapped (rs_device_alloc on one rank + rs_ipc_export / rs_ipc_open on the others).
 * rs_signal: when everything queued on `stream` so far is done and visible system-wide, counter += 1.
 * rs_wait_counter: work queued on `stream` after this call starts once counter >= expect (wrap-safe). */
int rs_signal(rs_ctx* ctx, uint32_t* d_counter, void* stream);
int rs_wait_counter(rs_ctx* ctx, const uint32_t* d_counter, uint32_t expect, void* stream);
/* Synchronises `stream` and reports whether a wait queued by rs_wait_counter on this context has given up since the last
 * check (RS_OPT_WAIT_BUDGET_MS): RS_OK = every wait was satisfied; RS_ERR_CUDA = a peer never signalled and the frame
 * behind that wait is not valid (rs_last_error says so).  Clears the flag.  The multi-process caller's counterpart of the
 * check rs_multi_render makes by itself. */
int rs_wait_check(rs_ctx* ctx, void* stream);

/* ---- the redo-map step without any allocation -------------------------------------------------------
 * No counterpart in the reference (its passes loop works on host Vecs, src/bin/raysnail.rs:404-423); this
 * is the library's own addition for callers that cannot let it allocate device memory: hosts whose own
 * pool already owns the whole card, per-frame loops where the device-wide stall of a cudaMalloc is not
 * acceptable, and multi-GPU code sitting between rs_signal and rs_wait_counter, where a cudaMalloc waits
 * for peer devices that may be spinning and never returns.
 *
 * One call = what rs_render_passes does after combine_pixels: noise = calc_noise(d_pixels) (exactly
 * rs_noise_map, same `quirk`), d_redo_mask[i] = noise[i] >= noise_threshold ? 1 : 0, *d_redo_count = number
 * of ones.  All three pointers are DEVICE pointers the caller owns (W*H float4 in; W*H bytes and one
 * uint64 out).  The noise map in between is kept in a fixed array inside the library, RS_STATIC_MAX_PIXELS
 * floats = 16 MiB, present on every GPU the library is loaded on for the life of the process.
 *   - width*height > RS_STATIC_MAX_PIXELS is refused with RS_ERR_UNSUPPORTED (rs_last_error says by how much);
 *   - the call allocates and frees nothing on the device, and does not synchronise: work is queued on `stream`;
 *   - the first call on each device looks the array's address up (one-time set-up that may also load the
 *     library's device code): make that first call outside of any stream capture;
 *   - the array is shared by every context, stream and thread on that GPU and keeps what the last call wrote:
 *     two calls whose kernels can run at the same time on one GPU (different streams, different contexts)
 *     overwrite each other's noise map.  Order them on one stream, or with events. */
#define RS_STATIC_MAX_PIXELS 4194304 /* 4 Mi pixels (2048 x 2048; a 1920 x 1080 frame is 2 073 600) */
int rs_redo_map_static(rs_ctx* ctx, const float* d_pixels, uint8_t* d_redo_mask, uint64_t* d_redo_count,
                       int32_t width, int32_t height, int32_t quirk, float noise_threshold, void* stream);

int rs_get_stats(const rs_ctx* ctx, rs_stats* out);

#ifdef __cplusplus
}
#endif
#endif /* RAYSNAIL_B200_H */
