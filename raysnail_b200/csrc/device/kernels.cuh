// Shared declarations between kernels.cu (device code) and backend.cu (host orchestration).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "rt_scene.cuh"

namespace rsd {

#define RS_MAX_DEPTH 64   // TakePhotoSettings::depth upper bound supported by the counter block
#define RS_MAX_MIXED_DEPTH 8  // MixedMaterial nesting the kernels resolve (resolve_mixed); deeper scenes are rejected at upload
#define RS_TRAVERSAL_STACK 96  // per-thread QBVH traversal stack entries (3 pushes per level at most)
// Shade work classes: 0 = miss, 1 + rs_material_type for the resolved (non-mixed) material, and —
// lean kernels, where extend already knows the light/BRDF branch draw — 8, 9, 10 = the hits of
// Lambertian / Isotropic / BlinnPhong whose next ray is the light-directed one (camera.rs:198-213),
// so that a warp of shade runs ONE of the two branches instead of both with half its lanes
#define RS_NUM_CLASSES 11
#define RS_COUNTER_STRIDE 32  // uint32 slots between two class counters (128 B)
#define RS_FRONT_CLASS(cls) ((cls) == 1 + RS_MAT_LAMBERTIAN ? 8 : ((cls) == 1 + RS_MAT_ISOTROPIC ? 9 : ((cls) == 1 + RS_MAT_BLINN_PHONG ? 10 : (cls))))
#define RS_BASE_CLASS(k) ((k) < 8 ? (k) : ((k) == 8 ? 1 + RS_MAT_LAMBERTIAN : ((k) == 9 ? 1 + RS_MAT_ISOTROPIC : 1 + RS_MAT_BLINN_PHONG)))

// One path record is 64 contiguous, 64-byte aligned bytes (four 128-bit elements), so a record
// that is GATHERED through a queue index (shade) costs exactly two 32-byte DRAM sectors, and a
// linear sweep (extend) stays fully coalesced:
//   q0 = (o.x, o.y)  q1 = (o.z, d.x)  q2 = (d.y, d.z)  q3 = (bits T.r, T.g, T.b, slot)
// `slot` = sample-major index of the path in its wave; pixel and sample index are derived from it.
// The ray's shutter time lives in a separate array that only scenes with moving spheres touch.
struct PathBuf {
    uint4* rec;    // 4 * capacity
    double* time;  // capacity (used when WaveState::uses_time)
};

// closest-hit record written by extend, read by shade (16 B)
struct __align__(16) HitA {
    double t;
    int root;          // -1 = miss
    uint32_t matinfo;  // resolved material | (settings material << 16)
};

struct WaveCounters {
    // Live paths entering bounce b sit in two stacks of the path buffer: n_rays[b] records grow
    // up from index 0 (camera rays; after a bounce: the rays sampled TOWARDS A LIGHT, camera.rs:198-213)
    // and n_back[b] records grow down from index capacity-1 (BRDF-sampled and specular rays).
    // Light-directed rays are mutually coherent and usually short; keeping them apart from the
    // incoherent ones keeps the warps of extend homogeneous.
    uint32_t n_rays[RS_MAX_DEPTH + 1];
    uint32_t n_back[RS_MAX_DEPTH + 1];
    // every hot counter on its own 128-byte line: atomics on one line serialise in L2, and in bounce 0
    // all warps of extend bump the same two or three buckets
    uint32_t class_count[RS_MAX_DEPTH][RS_NUM_CLASSES * RS_COUNTER_STRIDE];
    uint32_t extend_head[RS_MAX_DEPTH];           // persistent-thread work cursors
    uint32_t shade_head[RS_MAX_DEPTH];
    uint32_t shade_head2[RS_MAX_DEPTH];          // second class set of a bounce (see shade_kernel)
};

struct DeviceStats {
    unsigned long long n_samples;
    unsigned long long n_segments;
};

struct WaveState {
    PathBuf buf[2];
    HitA* hit;
    int2* hit_src;          // ROOT_GENERIC hits: (source instruction, flavor) the record is rebuilt from (object_record)
    uint2* class_queue;     // RS_NUM_CLASSES * capacity entries of (path index, output slot reserved by extend or ~0)
    float4* contrib;        // per slot: radiance of the finished path
    WaveCounters* ctr;
    DeviceStats* stats;
    uint32_t capacity;
#if defined(RS_TUNE_COST_DUMP) && RS_TUNE_COST_DUMP
    int2* dbg_cost;  // profiling builds: per ray index (node visits, leaf tests) of the last extend launch
#endif
    int uses_time;  // the scene holds moving spheres: rays carry their shutter time (sphere.rs:50-53)
};

struct FrameParams {
    rs_camera cam;
    int sqrt_spp;
    int depth;
    int stratum_row_begin;
    uint32_t flags;
    uint2 key;
    PhiloxKeys rk;  // the ten Philox round keys of `key` (rt_math.cuh)
    // wave
    uint32_t pixel0;     // first pixel of the wave
    uint32_t n_pixels;   // P
    uint32_t sample0;    // first local sample index (0 .. S)
    uint32_t n_samples;  // K
    // floor(2^32 / d) for d = n_pixels and d = sqrt_spp: slot -> (pixel, sample) without integer
    // divisions (decode_slot; two divisions per ray in extend and per hit in shade otherwise)
    uint32_t magic_pixels, magic_q;
    const unsigned char* mask;
};

struct DebugOut {  // all device pointers, indexed by pixel; any may be null
    int* obj_id;
    double* t;
    double* normal3;
    double* point3;
    int* front;
    int* material;
    double* ray_od6;
    int* kind;
    double* next_od6;
    double* weight3;
    double* emitted3;
};

inline uint32_t div_magic(uint32_t d) { return d <= 1u ? 0xffffffffu : (uint32_t)(0x100000000ull / d); }

// launchers (kernels.cu)
void launch_raygen(const FrameParams& fp, const WaveState& ws, cudaStream_t s);
void launch_extend(const DScene& S, const FrameParams& fp, const WaveState& ws, int bounce, int mode, int wide, int miss_in_extend, int grid,
                   int stage_bytes, cudaStream_t s);
void launch_shade(const DScene& S, const FrameParams& fp, const WaveState& ws, int bounce, int generic, int set, int grid, const DebugOut* dbg, cudaStream_t s);
void launch_probe_primary(const DScene& S, const FrameParams& fp, const WaveState& ws, const DebugOut& dbg, cudaStream_t s);
void launch_accumulate(const FrameParams& fp, const WaveState& ws, float4* accum, cudaStream_t s);
void launch_finalize(const float4* accum, const unsigned char* mask, float4* out, int64_t n, int n_samples, int gamma, cudaStream_t s);
struct PeerAccums {
    const float4* p[16];  // RS_MAX_PEERS
    int n;
};
void launch_finalize_peers(const PeerAccums& peers, const unsigned char* mask, float4* out, int64_t begin, int64_t end, int n_samples, int gamma,
                           cudaStream_t s);
unsigned long long read_violations();  // RS_CHECKED builds: failed bounds checks since the last call (0 otherwise)
void launch_signal(uint32_t* counter, cudaStream_t s);
// budget_ns: 0 = poll without limit; otherwise give up after that long and set *gave_up
void launch_wait_counter(const uint32_t* counter, uint32_t expect, cudaStream_t s, unsigned long long budget_ns = 0, uint32_t* gave_up = nullptr);
void launch_combine(const float4* old_px, const float4* new_px, float4* out, int64_t n, float pass, cudaStream_t s);
void launch_noise(const float4* px, float* noise, int w, int h, int quirk, cudaStream_t s);
void launch_quantize(const float4* px, unsigned char* rgb, int64_t n, cudaStream_t s);
void launch_redo_mask(const float* noise, unsigned char* mask, int64_t n, float t, unsigned long long* count, cudaStream_t s);
void launch_fill(float4* p, int64_t n, float4 v, cudaStream_t s);
void launch_extend_first(const DScene& S, const FrameParams& fp, const WaveState& ws, int mode, int grid, int stage_bytes, cudaStream_t s);
#define RS_STAGE_MAX_BYTES (100 * 1024)  // tree + sphere table a staged block may hold (two blocks of 512 threads per SM)
// stage_bytes > 0 selects the kernels that keep the tree + sphere table in shared memory; 0 blocks = they do not fit
int extend_first_blocks_per_sm(int mode, int stage_bytes);
int extend_blocks_per_sm(int mode, int wide, int stage_bytes);
// the streaming form of extend (lean modes, binary tree, bounces >= 1): per-lane traversal state, dynamic ray fetch
void launch_extend_stream(const DScene& S, const FrameParams& fp, const WaveState& ws, int bounce, int mode, int miss_in_extend, int grid, cudaStream_t s);
int extend_stream_blocks_per_sm(int mode);
int shade_blocks_per_sm(int generic, int set);
int shade_split_trees();  // tier-2 scenes launch the Lambertian hits separately (RS_TUNE_SPLIT_TREES)

}  // namespace rsd
