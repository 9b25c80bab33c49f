// sm_100a kernels of the wavefront path tracer.
//
//   raygen      replaces Painter::render_pixel's strata loop + Camera::ray
//               (reference src/painter.rs:154-187, src/camera.rs:77-85)
//   extend      replaces World::hit -> BVH::hit -> leaf Hittable::hit
//               (src/hittable/collection/bvh.rs:173-192 and every geometry/*.rs, csg/*.rs ...)
//   shade       replaces the body of TakePhotoSettings::ray_color (src/camera.rs:156-255),
//               Material::scatter/emitted (src/material/*.rs), the PDFs (src/prelude/pdf.rs)
//               and the textures (src/texture/*.rs)
//   accumulate  replaces the f64 `color_vec` sum (painter.rs:163-176)
//   finalize    replaces Vec3::into_color (src/prelude/vec3.rs:227-240)
//   combine / noise / quantize: the CLI's passes logic (src/bin/raysnail.rs:157-208,437-441)
//
// Scheduling: extend and shade are persistent kernels (grid = SMs x resident blocks); each
// warp fetches 32 work items at a time with one atomicAdd on a cursor in global memory.
// extend buckets its hits by resolved material class with __match_any_sync so that every
// warp of shade runs one material's code; shade compacts the surviving paths into the other
// ping-pong buffer with __ballot_sync-aggregated atomics.
#include <cfloat>

#define RS_DEFINE_VIOLATIONS 1
#include "kernels.cuh"

namespace rsd {

#define RS_BLOCK 256
#define RS_TMIN 0.0001
#define RS_STACK 96  // must match RS_TRAVERSAL_STACK in kernels.cuh (the host checks the tree against it)
// relative slack for the f32 slab test: covers the rounding of 1/d, the FMA and the f64->f32
// conversions of the ray; the absolute part (origin rounding) is folded into the padded boxes.
#define RS_SLAB_EPS 9.5367431640625e-07f  // 2^-20

// ---- 256-bit global loads / stores (sm_100: LDG.E.256 / STG.E.256, PTX ld.global.v8.b32) ---------------------------
// A 64-byte BVH node or path record is two requests to L1 instead of four.  The walk of incoherent rays (every lane
// of a warp at another node) is bound by L1 request throughput, not by bytes: ncu, 1 M-triangle mesh, bounce >= 1:
// l1tex throughput 80 % of peak at 44 % hit rate.  RS_TUNE_LDG256=0 builds the 128-bit form for A/B.
#ifndef RS_TUNE_LDG256
#define RS_TUNE_LDG256 1
#endif
struct U8 { uint32_t v[8]; };
__device__ __forceinline__ U8 ldg256_nc(const void* p) {  // read-only data (scene tables): ld.global.nc
    U8 r;
    asm("ld.global.nc.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
        : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3]), "=r"(r.v[4]), "=r"(r.v[5]), "=r"(r.v[6]), "=r"(r.v[7]) : "l"(p));
    return r;
}
__device__ __forceinline__ U8 ldg256(const void* p) {  // data another kernel wrote (path records)
    U8 r;
    asm volatile("ld.global.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
        : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3]), "=r"(r.v[4]), "=r"(r.v[5]), "=r"(r.v[6]), "=r"(r.v[7]) : "l"(p));
    return r;
}
__device__ __forceinline__ void stg256(void* p, uint4 a, uint4 b) {
    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};\n" ::"l"(p), "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w) : "memory");
}
struct PathRec {
    Ray ray;
    float3 T;
    uint32_t slot, pixel, sample;
};

// slot -> (pixel, absolute sample index): the inverse of raygen's numbering
__device__ __forceinline__ void decode_slot(const FrameParams& fp, uint32_t slot, uint32_t& pixel, uint32_t& sample) {
    // x / d = umulhi(x, floor(2^32 / d)) or one more (FrameParams::magic_*)
    uint32_t k_local = __umulhi(slot, fp.magic_pixels), p_local = slot - k_local * fp.n_pixels;
    if (p_local >= fp.n_pixels) { ++k_local; p_local -= fp.n_pixels; }
    pixel = fp.pixel0 + p_local;
    const uint32_t j = fp.sample0 + k_local, q = (uint32_t)fp.sqrt_spp;
    uint32_t row = __umulhi(j, fp.magic_q), col = j - row * q;
    if (col >= q) { ++row; col -= q; }
    sample = ((uint32_t)fp.stratum_row_begin + row) * q + col;
}
__device__ __forceinline__ void unpack_path(const FrameParams& fp, uint4 q0, uint4 q1, uint4 q2, uint4 q3, PathRec& p) {
    p.ray.o = v3(__hiloint2double((int)q0.y, (int)q0.x), __hiloint2double((int)q0.w, (int)q0.z), __hiloint2double((int)q1.y, (int)q1.x));
    p.ray.d = v3(__hiloint2double((int)q1.w, (int)q1.z), __hiloint2double((int)q2.y, (int)q2.x), __hiloint2double((int)q2.w, (int)q2.z));
    p.ray.time = 0.0;
    p.T = make_float3(__uint_as_float(q3.x), __uint_as_float(q3.y), __uint_as_float(q3.z));
    p.slot = q3.w;
    decode_slot(fp, p.slot, p.pixel, p.sample);
}
__device__ __forceinline__ uint4 pack_d2(double a, double b) {
    return make_uint4((uint32_t)__double2loint(a), (uint32_t)__double2hiint(a), (uint32_t)__double2loint(b), (uint32_t)__double2hiint(b));
}
__device__ __forceinline__ void store_path(const WaveState& ws, const PathBuf& B, uint32_t i, const PathRec& p) {
    if (!RS_IN_BOUNDS(i < ws.capacity)) return;
    uint4* r = B.rec + 4 * (size_t)i;
#if RS_TUNE_LDG256
    stg256(r, pack_d2(p.ray.o.x, p.ray.o.y), pack_d2(p.ray.o.z, p.ray.d.x));
    stg256(r + 2, pack_d2(p.ray.d.y, p.ray.d.z), make_uint4(__float_as_uint(p.T.x), __float_as_uint(p.T.y), __float_as_uint(p.T.z), p.slot));
#else
    r[0] = pack_d2(p.ray.o.x, p.ray.o.y);
    r[1] = pack_d2(p.ray.o.z, p.ray.d.x);
    r[2] = pack_d2(p.ray.d.y, p.ray.d.z);
    r[3] = make_uint4(__float_as_uint(p.T.x), __float_as_uint(p.T.y), __float_as_uint(p.T.z), p.slot);
#endif
    if (ws.uses_time) B.time[i] = p.ray.time;
}
__device__ __forceinline__ void load_path(const FrameParams& fp, const WaveState& ws, const PathBuf& B, uint32_t i, PathRec& p) {
    if (!RS_IN_BOUNDS(i < ws.capacity)) i = 0;
    const uint4* r = B.rec + 4 * (size_t)i;
#if RS_TUNE_LDG256
    const U8 a = ldg256(r), b = ldg256(r + 2);
    const uint4 q0 = make_uint4(a.v[0], a.v[1], a.v[2], a.v[3]), q1 = make_uint4(a.v[4], a.v[5], a.v[6], a.v[7]);
    const uint4 q2 = make_uint4(b.v[0], b.v[1], b.v[2], b.v[3]), q3 = make_uint4(b.v[4], b.v[5], b.v[6], b.v[7]);
#else
    const uint4 q0 = r[0], q1 = r[1], q2 = r[2], q3 = r[3];
#endif
    unpack_path(fp, q0, q1, q2, q3, p);
    if (ws.uses_time) p.ray.time = B.time[i];
}
__device__ __forceinline__ void load_ray_only(const FrameParams& fp, const WaveState& ws, const PathBuf& B, uint32_t i, Ray& r, uint32_t& pixel,
                                              uint32_t& sample) {
    if (!RS_IN_BOUNDS(i < ws.capacity)) i = 0;
    const uint4* rec = B.rec + 4 * (size_t)i;
#if RS_TUNE_LDG256
    const U8 a = ldg256(rec), b = ldg256(rec + 2);
    const uint4 q0 = make_uint4(a.v[0], a.v[1], a.v[2], a.v[3]), q1 = make_uint4(a.v[4], a.v[5], a.v[6], a.v[7]);
    const uint4 q2 = make_uint4(b.v[0], b.v[1], b.v[2], b.v[3]);
    const uint32_t slot = b.v[7];
#else
    const uint4 q0 = rec[0], q1 = rec[1], q2 = rec[2];
    const uint32_t slot = rec[3].w;
#endif
    r.o = v3(__hiloint2double((int)q0.y, (int)q0.x), __hiloint2double((int)q0.w, (int)q0.z), __hiloint2double((int)q1.y, (int)q1.x));
    r.d = v3(__hiloint2double((int)q1.w, (int)q1.z), __hiloint2double((int)q2.y, (int)q2.x), __hiloint2double((int)q2.w, (int)q2.z));
    r.time = ws.uses_time ? B.time[i] : 0.0;
    decode_slot(fp, slot, pixel, sample);
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }

// atomicAdd whose result is consumed LATER.  ptxas rewrites an atomic add on a warp-uniform
// address into its own warp-aggregated form (elect + ATOMG + SHFL) — also for inline PTX and
// even when a single lane executes it — and that SHFL waits for the atomic on the spot, which
// exposes the full L2 round trip (the top stall of shade and extend in the first profiles).
// `lane_zero` is a per-lane value that is always 0 at run time but that the compiler cannot
// prove uniform, so the address looks divergent and the instruction stays as written: the
// latency then overlaps with the work between the atomic and the first use of its result.
__device__ __forceinline__ uint32_t atomic_add_async(uint32_t* p, uint32_t v, uint32_t lane_zero) {
    uint32_t old;
    asm volatile("atom.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p + lane_zero), "r"(v) : "memory");
    return old;
}
#define RS_LANE_ZERO(ws) (lane_id() & ((ws).capacity >> 31))  // capacity < 2^31 always

// warp-aggregated append: returns this lane's position; `counter` is bumped once per warp
__device__ __forceinline__ uint32_t warp_append(uint32_t* counter, bool pred, uint32_t lane_zero) {
    uint32_t mask = __ballot_sync(0xffffffffu, pred);
    if (mask == 0) return 0;
    uint32_t leader = __ffs(mask) - 1, base = 0;
    if (lane_id() == leader) base = atomic_add_async(counter, __popc(mask), lane_zero);
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + __popc(mask & ((1u << lane_id()) - 1u));
}

// Asynchronous global->shared staging (LDGSTS, 16 bytes per lane, L2 only).  The persistent
// kernels know the NEXT batch of work items while they process the current one, so its path /
// hit records are copied into shared memory in the background and the DRAM latency of the
// queue reads is off the critical path (the kernels were long-scoreboard bound without it).
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async8(void* smem, const void* gmem) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;\n" ::: "memory"); }

// ---- 1-D bulk copy global -> shared (TMA, `cp.async.bulk`) completing on an mbarrier -----------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");  // visible to the async proxy
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar), "r"(parity)
        : "memory");
}
__device__ __forceinline__ float4 lds_f4(uint32_t a) {
    float4 v;
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];\n" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ int4 lds_i4(uint32_t a) {
    int4 v;
    asm("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];\n" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ double2 lds_d2(uint32_t a) {
    double2 v;
    asm("ld.shared.v2.f64 {%0, %1}, [%2];\n" : "=d"(v.x), "=d"(v.y) : "r"(a));
    return v;
}
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];\n" ::"l"(p)); }

// one 64-byte node of the binary tree: two children with their boxes (BvhNode)
__device__ __forceinline__ void load_node(const BvhNode* n, float4& bx, float4& by, float4& bz, int4& ch) {
#if RS_TUNE_LDG256
    const U8 a = ldg256_nc(n), b = ldg256_nc(reinterpret_cast<const unsigned char*>(n) + 32);
    bx = make_float4(__uint_as_float(a.v[0]), __uint_as_float(a.v[1]), __uint_as_float(a.v[2]), __uint_as_float(a.v[3]));
    by = make_float4(__uint_as_float(a.v[4]), __uint_as_float(a.v[5]), __uint_as_float(a.v[6]), __uint_as_float(a.v[7]));
    bz = make_float4(__uint_as_float(b.v[0]), __uint_as_float(b.v[1]), __uint_as_float(b.v[2]), __uint_as_float(b.v[3]));
    ch = make_int4((int)b.v[4], (int)b.v[5], (int)b.v[6], (int)b.v[7]);
#else
    const float4* np = reinterpret_cast<const float4*>(n);
    bx = __ldg(np), by = __ldg(np + 1), bz = __ldg(np + 2);
    ch = __ldg(reinterpret_cast<const int4*>(np + 3));
#endif
}
// where the staged kernels keep the tree and the sphere table (shared-space addresses)
struct Staged {
    uint32_t nodes, spheres;
};

#ifndef RS_TUNE_FIRST_LATE_FETCH
#define RS_TUNE_FIRST_LATE_FETCH 1
#endif
// threads per block of the lean shade kernel (the generic one uses RS_BLOCK); tunable
#ifndef RS_TUNE_SHADE_BLOCK
#define RS_TUNE_SHADE_BLOCK 256
#endif
#define RS_WARPS (RS_TUNE_SHADE_BLOCK / 32)
// staging slots per lane of shade: the 4 path elements + the hit record; double buffered
#define RS_SHADE_STAGE 5
// per warp: two record buffers + two 32-entry queue-entry buffers (uint2) + two 32-entry buffers
// of reserved output slots + two 32-entry buffers of path indices
#define RS_SHADE_WARP_SMEM (2 * RS_SHADE_STAGE * 32 * 16 + 2 * 32 * 8 + 2 * 32 * 4 + 2 * 32 * 4)
// highest shade tier that runs the staged (cp.async) pipeline; 0 = lean scenes only
#ifndef RS_TUNE_STAGE_TIERS
#define RS_TUNE_STAGE_TIERS 0
#endif
// 1: scenes with expression trees (tier 2) also shade their Lambertian hits in a kernel of their own
#ifndef RS_TUNE_SPLIT_TREES
#define RS_TUNE_SPLIT_TREES 1
#endif
#define RS_SHADE_STAGED(TIER) ((TIER) <= RS_TUNE_STAGE_TIERS)
#define RS_SHADE_SMEM (RS_WARPS * RS_SHADE_WARP_SMEM)

// =============================================================================================
// raygen
// =============================================================================================
__device__ __forceinline__ Ray camera_ray(const FrameParams& fp, int row, int col, int s_i, int s_j, RngCtx& rng) {
    const rs_camera& c = fp.cam;
    uint4 w = rng_words(rng, P_PIXEL, 0);
    double jx = u01(w.x), jy = u01(w.y), jt = u01(w.z);
    const bool centre = (fp.flags & RS_FLAG_PIXEL_CENTRE) != 0;
    if (centre) { jx = 0.5; jy = 0.5; jt = 0.0; }
    // painter.rs:169-172, 133-139
    double xo = (double)col + ((double)s_i + jx) / (double)fp.sqrt_spp;
    double yo = (double)row + ((double)s_j + jy) / (double)fp.sqrt_spp;
    double h = (double)c.height;
    double u = xo / (double)c.width;
    double v = (h - 1.0 - yo) / h;
    // camera.rs:77-85
    double aperture = centre ? 0.0 : c.aperture;
    V3 p = v3(0.0, 0.0, 0.0);
    if (aperture != 0.0) {  // with aperture 0 the disk sample is multiplied by 0: skip the draws
        for (uint32_t k = 0;; ++k) {
            double a, b;
            rng_pair(rng, P_LENS, 0, k, a, b);
            p = v3(-1.0 + a * 2.0, -1.0 + b * 2.0, 0.0);
            if (length_squared(p) < 1.0) break;
        }
    }
    V3 rd = (aperture / 2.0) * p;
    V3 hu = v3(c.horizontal_unit[0], c.horizontal_unit[1], c.horizontal_unit[2]);
    V3 vu = v3(c.vertical_unit[0], c.vertical_unit[1], c.vertical_unit[2]);
    // every product and sum rounded on its own (no FMA contraction): the ray must be the oracle's bit
    // for bit, because objects found by orbit tests (RayMarcher) amplify a last-bit difference
    V3 offset = v3(__dadd_rn(__dmul_rn(hu.x, rd.x), __dmul_rn(vu.x, rd.y)), __dadd_rn(__dmul_rn(hu.y, rd.x), __dmul_rn(vu.y, rd.y)),
                   __dadd_rn(__dmul_rn(hu.z, rd.x), __dmul_rn(vu.z, rd.y)));
    Ray r;
    r.o = v3(c.origin[0], c.origin[1], c.origin[2]) + offset;
    V3 hf = v3(c.horizontal_full[0], c.horizontal_full[1], c.horizontal_full[2]);
    V3 vf = v3(c.vertical_full[0], c.vertical_full[1], c.vertical_full[2]);
    V3 dir = v3(__dadd_rn(__dadd_rn(__dadd_rn(c.lb[0], __dmul_rn(u, hf.x)), __dmul_rn(v, vf.x)), -r.o.x),
                __dadd_rn(__dadd_rn(__dadd_rn(c.lb[1], __dmul_rn(u, hf.y)), __dmul_rn(v, vf.y)), -r.o.y),
                __dadd_rn(__dadd_rn(__dadd_rn(c.lb[2], __dmul_rn(u, hf.z)), __dmul_rn(v, vf.z)), -r.o.z));
    r.d = unit(dir);
    r.time = c.shutter_speed * jt;
    return r;
}

__global__ void __launch_bounds__(RS_BLOCK) raygen_kernel(const __grid_constant__ FrameParams fp, const __grid_constant__ WaveState ws) {
    const uint32_t total = fp.n_pixels * fp.n_samples;
    const uint32_t stride = gridDim.x * blockDim.x;
    // every thread of a warp iterates the same number of times (warp_append is collective)
    for (uint32_t base = blockIdx.x * blockDim.x; base < total; base += stride) {
        uint32_t i = base + threadIdx.x;
        bool live = i < total;
        PathRec p;
        if (live) {
            uint32_t pixel, sample;
            decode_slot(fp, i, pixel, sample);
            ws.contrib[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (fp.mask && fp.mask[pixel] == 0) {
                live = false;
            } else {
                int row = (int)(pixel / (uint32_t)fp.cam.width), col = (int)(pixel % (uint32_t)fp.cam.width);
                uint32_t s_j = __umulhi(sample, fp.magic_q), s_i = sample - s_j * (uint32_t)fp.sqrt_spp;
                if (s_i >= (uint32_t)fp.sqrt_spp) { ++s_j; s_i -= (uint32_t)fp.sqrt_spp; }
                RngCtx rng;
                rng.key = &fp.rk;
                rng.pixel = pixel;
                rng.sample = sample;
                rng.depth = 0;
                p.ray = camera_ray(fp, row, col, (int)s_i, (int)s_j, rng);
                p.T = make_float3(1.f, 1.f, 1.f);
                p.slot = i;
                p.pixel = pixel;
                p.sample = rng.sample;
            }
        }
        // without a pixel mask every path is live and path i simply goes to record i: no compaction,
        // no atomic (its round trip to L2 was the top stall of this kernel)
        uint32_t pos = i;
        if (fp.mask) pos = warp_append(&ws.ctr->n_rays[0], live, RS_LANE_ZERO(ws));
        if (live) store_path(ws, ws.buf[0], pos, p);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(&ws.stats->n_samples, (unsigned long long)total);
        if (!fp.mask) ws.ctr->n_rays[0] = total;
    }
}

// =============================================================================================
// extend
// =============================================================================================
#ifndef RS_TUNE_COST_DUMP
#define RS_TUNE_COST_DUMP 0  // 1 (profiling builds only): extend records per-ray node visits / leaf tests (profiles/cost_dump.py)
#endif
struct Closest {
#if RS_TUNE_COST_DUMP
    int n_nodes, n_leaves;
#endif
    double t;        // current limit end / best t
    int root;
    int material;
    int src, flavor;  // ROOT_GENERIC: what the record is rebuilt from (object_record)
};

// smallest double above x (x > 0); +inf stays +inf
__device__ __forceinline__ double next_up(double x) {
    return x < __longlong_as_double(0x7ff0000000000000LL) ? __longlong_as_double(__double_as_longlong(x) + 1) : x;
}

// Closest-hit update.  The reference shrinks the limit to `..best.t` (bvh.rs:184-187), so on an
// exact tie the object its randomly-built tree visits first wins; the canonical rule here (and
// in the oracle) is order-independent: the limit admits t == best.t and the lower id wins.
// MODE selects what a leaf may hold, so that scenes made of plain spheres (and triangles) get a
// lean kernel without the expression-tree code: 0 = static spheres, 1 = + triangles, 2 = + rects and
// boxes, bare or under one TfFacade (ROOT_FLAT: still no expression-tree code), 3 = everything.
// `ref` is the decoded leaf reference (root << 2 | kind).
template <int MODE, bool STAGE = false>
__device__ __forceinline__ void test_root(const DScene& S, int ref, const Ray& ray, const V3& inv_d, const RngCtx& rng, Closest& best, const Staged st = Staged{0, 0}) {
    const int kind = MODE == 0 ? (int)ROOT_SPHERE : (ref & 3);
    const int root = ref >> 2;
    if (kind == ROOT_SPHERE) {
        double2 c01, c23;
        if constexpr (STAGE) {
            c01 = lds_d2(st.spheres + 32u * (uint32_t)root);
            c23 = lds_d2(st.spheres + 32u * (uint32_t)root + 16u);
        } else {
            const double2* sp = reinterpret_cast<const double2*>(S.root_sphere + root);
            c01 = __ldg(sp), c23 = __ldg(sp + 1);  // two 128-bit ld.global.nc
        }
        double t1;
        if (sphere_nearest(v3(c01.x, c01.y, c23.x), c23.y * c23.y, ray, RS_TMIN, next_up(best.t), t1) && (t1 < best.t || root < best.root)) {
            best.t = t1;
            best.root = root;
            if (MODE >= 2) best.material = -2;  // look it up in root_material
        }
    } else if (MODE >= 1 && kind == ROOT_TRIANGLE) {
        // (a triangle's leaf reference holds its index in the triangle table, the record names the world object)
        const TriGeom T = load_triangle(S.tri_geom, root);
        double t, beta, gamma;
        if (triangle_test<true>(T, ray, RS_TMIN, best.t, t, beta, gamma) && (t < best.t || T.root < best.root)) {
            best.t = t;
            best.root = T.root;
            if (MODE >= 2) best.material = -2;
        }
    } else if (MODE >= 2 && kind == ROOT_FLAT) {
        const rs_node* n = S.nodes + __ldg(&S.root_node[root]);
        Ray r = ray;
        if (n->type == RS_NODE_TRANSFORM) {  // tf_facade.rs:41-56: the ray goes to object space, t is kept
            const XformPre& x = S.xf[n->first_xform];
            r.o = xf_apply(x.inv, ray.o, 1.0);
            r.d = xf_apply(x.inv, ray.d, 0.0);
            n = S.nodes + n->child0;
        }
        // a world object is evaluated over [tmin, +inf) — as if it were the first one visited — and its hit then
        // competes with t1: which faces of a Box count (box.rs:125-150: `outside` of the record), and what a CSG
        // object returns, must not depend on the order in which the tree reaches the objects
        const double inf = __longlong_as_double(0x7ff0000000000000LL);
        const double hi = inf;
        double t;
        bool hit;
        if (MODE >= RS_QUADRIC_MODE && n->type == RS_NODE_QUADRIC) {
            // quadric.rs:191-196: the surface is unbounded, the reference clips it with its +-100 bbox
            if (!aabb_hit_f64(S.root_bbox + 6 * (size_t)root, ray, inv_d, RS_TMIN, inf)) return;
            hit = quadric_t(n->p, r, RS_TMIN, hi, t);
        } else {
            hit = n->type == RS_NODE_BOX ? box_t(n->p, r, RS_TMIN, hi, t)
                                         : rect_t(n->rect_axis, n->p[0], n->p[1], n->p[2], n->p[3], n->p[4], r, RS_TMIN, hi, t);
        }
        if (hit && (t < best.t || (t == best.t && root < best.root))) {
            best.t = t;
            best.root = root;
            best.material = -2;
        }
    } else if (MODE >= 3) {
        const int2 pr = __ldg(&S.obj.root_prog[root]);
        ObjHit h;
        const double hi = __longlong_as_double(0x7ff0000000000000LL);  // see ROOT_FLAT
        // the object only counts when the ray passes AABB::hit on its reference bbox (bvh.rs:173-176 on the leaf's box)
        if (__ldg(&S.root_gate[root]) && !aabb_hit_f64(S.root_bbox + 6 * (size_t)root, ray, inv_d, RS_TMIN, hi)) return;
        const bool hit = eval_object(S.nodes, S.xf, S.obj, pr.x, pr.y, ray, RS_TMIN, hi, rng, h) && (h.t1 < best.t || (h.t1 == best.t && root < best.root));
        if (hit) {
            best.t = h.t1;
            best.root = root;
            best.material = h.mat;
            best.src = h.src;
            best.flavor = h.flavor;
        }
    }
}

__device__ __forceinline__ float up(float x) { return x * (1.0f + RS_SLAB_EPS); }

// Packed f32x2 FMA (sm_100: one FFMA2 issue slot for two fused multiply-adds, each rounded like
// fmaf): the slab test of a node is 12 FMAs = 6 FFMA2.  extend is issue-bound.
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long ra, rb, rc, rd;
    float2 d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
    return d;
}
#ifndef RS_TUNE_FFMA2
#define RS_TUNE_FFMA2 1
#endif

#define RS_BVH_SENTINEL ((int)0x80000000)
// material classes (1 + rs_material_type) whose `scatter` always returns Some
#define RS_ALWAYS_CONTINUES(cls) \
    ((cls) == 1 + RS_MAT_LAMBERTIAN || (cls) == 1 + RS_MAT_DIELECTRIC || (cls) == 1 + RS_MAT_ISOTROPIC || (cls) == 1 + RS_MAT_BLINN_PHONG)
// resident 256-thread blocks per SM the kernels are compiled for (register budget); tunable
// at build time (make TUNE="-DRS_TUNE_EXTEND_LEAN=4 ...")
#ifndef RS_TUNE_EXTEND_LEAN
#define RS_TUNE_EXTEND_LEAN 4
#endif
#ifndef RS_TUNE_EXTEND_GENERIC
#define RS_TUNE_EXTEND_GENERIC 2
#endif
#ifndef RS_TUNE_SHADE_LEAN
#define RS_TUNE_SHADE_LEAN 2
#endif
// work distribution of the persistent kernels: 0 = one atomicAdd on a global cursor per batch of 32
// (issued a batch ahead), 1 = static round robin (batch = warp index + k * warps in the grid)
#ifndef RS_TUNE_EXTEND_STATIC
#define RS_TUNE_EXTEND_STATIC 0
#endif
#ifndef RS_TUNE_SHADE_STATIC
#define RS_TUNE_SHADE_STATIC 0
#endif
// extend reserves the output slots of always-continuing classes for shade (lean kernels)
#ifndef RS_TUNE_RESERVE
#define RS_TUNE_RESERVE 1
#endif
// shade: batches of 32 hits a warp takes per atomicAdd on the work cursor
#ifndef RS_TUNE_SHADE_CHUNK
#define RS_TUNE_SHADE_CHUNK 2
#endif

#ifndef RS_TUNE_SHADE_GENERIC
#define RS_TUNE_SHADE_GENERIC 2
#endif
// shade (lean): percent of the batches dealt statically before the cursor takes over (0 = cursor only)
// resident blocks per SM of the Lambertian-only shade kernel
#ifndef RS_TUNE_SHADE_LAMBERT
#define RS_TUNE_SHADE_LAMBERT 2
#endif
#ifndef RS_TUNE_SHADE_HYBRID
#define RS_TUNE_SHADE_HYBRID 0
#endif
#ifndef RS_TUNE_EXTEND_FLAT
#define RS_TUNE_EXTEND_FLAT 3
#endif
#ifndef RS_TUNE_EXTEND_TRI
#define RS_TUNE_EXTEND_TRI RS_TUNE_EXTEND_LEAN
#endif
#define RS_EXTEND_MIN_BLOCKS(MODE) ((MODE) >= 3 ? RS_TUNE_EXTEND_GENERIC : ((MODE) == 2 ? RS_TUNE_EXTEND_FLAT : ((MODE) == 1 ? RS_TUNE_EXTEND_TRI : RS_TUNE_EXTEND_LEAN)))

// while-while traversal: every lane first walks inner nodes until it holds a leaf (or is done),
// the warp reconverges, then the lanes holding leaves run the (expensive, f64) object test
// together.
// WIDE = false: binary tree of 64-byte nodes (the default: fewest instructions per ray on
// these cache-resident scenes); WIDE = true: the 4-wide QBVH of 128-byte nodes.
// `sstack` (binary tree only): the TOP RS_TUNE_SSTACK entries of the traversal stack live in shared memory, laid out
// [entry][thread] so that a lane owns its banks (a warp's push or pop is conflict-free whatever depth each lane is at);
// deeper entries spill to the local-memory array, which short walks never touch.  nullptr = the whole stack in local memory.
// MEASURED SLOWER in this kernel (profiles/r02d_ab_sstack.txt, two runs each: headline 3.49 vs 3.67 G samples/s, quadric.sdl
// 1.065 vs 1.133, csg.sdl 1.74 vs 1.83, Cornell 2.10 vs 2.16, example.sdl 4.42 vs 4.54, mesh equal): the walk is
// issue-bound and the spill bookkeeping (`lo`, two compares per push / pop) costs more issue slots than the LDL / STL it
// replaces cost L1 bandwidth — a warp's local-memory stack entries of one depth share a 128-byte line, so the lanes of a
// coherent warp hit one line.  Off by default (RS_TUNE_SSTACK_MODES = 0); extend_stream_kernel, whose lanes sit at
// unrelated depths, keeps its shared-memory stack (there it was 22 % of all L1 sector accesses).
#ifndef RS_TUNE_LIST_SKIP
#define RS_TUNE_LIST_SKIP 1  // object list: lanes skip ahead to their own next expression tree (0 = all lanes walk in lockstep)
#endif
#ifndef RS_TUNE_SSTACK
#define RS_TUNE_SSTACK 8  // entries per thread in shared memory (power of two; 0 = local memory only)
#endif
#ifndef RS_TUNE_SSTACK_MODES
#define RS_TUNE_SSTACK_MODES 0  // bit m: extend_kernel<MODE = m> keeps its stack top in shared memory (0xf = all tiers)
#endif
#define RS_SSTACK_MODE(MODE) (((RS_TUNE_SSTACK_MODES) >> (MODE)) & 1)
template <int MODE, bool WIDE, bool STAGE = false>
__device__ __forceinline__ void closest_hit(const DScene& S, const Ray& ray, const RngCtx& rng, Closest& best, const Staged st = Staged{0, 0},
                                            unsigned long long* const sstack = nullptr) {
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    best.t = inf;
    best.root = -1;
    best.material = -1;
#if RS_TUNE_COST_DUMP
    best.n_nodes = 0;
    best.n_leaves = 0;
#endif
    if (S.n_roots == 0) return;
    // f32 copy of the ray for the slab tests
    const float ox = (float)ray.o.x, oy = (float)ray.o.y, oz = (float)ray.o.z;
    float dx = (float)ray.d.x, dy = (float)ray.d.y, dz = (float)ray.d.z;
    const float tiny = 1e-20f;
    dx = fabsf(dx) > tiny ? dx : copysignf(tiny, dx);
    dy = fabsf(dy) > tiny ? dy : copysignf(tiny, dy);
    dz = fabsf(dz) > tiny ? dz : copysignf(tiny, dz);
    const float idx = 1.0f / dx, idy = 1.0f / dy, idz = 1.0f / dz;
    const float oodx = ox * idx, oody = oy * idy, oodz = oz * idz;
    const float tmin_f = (float)RS_TMIN * (1.0f - 2.0f * RS_SLAB_EPS);
    float tmax_f = FLT_MAX;
    const float slack = (1.0f + RS_SLAB_EPS) * (1.0f + RS_SLAB_EPS);
    // 1 / direction in f64 for the bbox gates of the objects (AABB::hit divides per call; the quotient is the ray's)
    V3 inv_d = v3(0.0, 0.0, 0.0);
    if constexpr (MODE >= 2) {
        if (S.any_gate) inv_d = v3(1.0 / ray.d.x, 1.0 / ray.d.y, 1.0 / ray.d.z);
    }

    // The object list, walked before the tree.  It holds
    //  - the objects whose surface has no bound (quadrics): the reference visits them whenever the ray
    //    passes their bbox at all, wherever the hit then lies, so no closest hit may drop them (hi.w = 1);
    //  - in worlds of a few objects (the .sdl scenes, the Cornell box) every object, and there is no tree:
    //    all lanes walk the SAME objects in the same order, so the object test a warp runs in an iteration is
    //    one object's code for all its lanes (a tree hands every lane another object first and the kinds —
    //    sphere, box, expression tree — serialise: 12 of 32 lanes per instruction on bounce 1 of quadric.sdl).
    // Same conservative f32 slab test as a tree node; the tie rule (lower id wins) is order-independent.
    if constexpr (MODE >= 2) {
        const int n_lock = MODE == 4 ? S.n_list_lockstep : S.n_list;
        for (int k = 0; k < n_lock; ++k) {
            const float4 lo = __ldg(S.root_list + 2 * k), hi = __ldg(S.root_list + 2 * k + 1);
            const float x0 = fmaf(lo.x, idx, -oodx), x1 = fmaf(hi.x, idx, -oodx);
            const float y0 = fmaf(lo.y, idy, -oody), y1 = fmaf(hi.y, idy, -oody);
            const float z0 = fmaf(lo.z, idz, -oodz), z1 = fmaf(hi.z, idz, -oodz);
            const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin_f));
            const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), hi.w != 0.f ? FLT_MAX : tmax_f));
            if (tn <= tf * slack) {
#if RS_TUNE_COST_DUMP
                ++best.n_leaves;
#endif
                test_root<MODE>(S, __float_as_int(lo.w), ray, inv_d, rng, best);
                tmax_f = best.t < (double)FLT_MAX ? up(__double2float_ru(best.t)) : FLT_MAX;
            }
        }
        // The expression trees at the end of the list all run ONE piece of code (eval_object), so here every lane
        // skips ahead to ITS next tree whose box it enters and the warp evaluates 32 different trees together: the
        // rounds a warp needs are the most trees any one of its rays enters, not the trees its rays enter between them
        // (bounce >= 1 of sdl/quadric.sdl: four TfFacade(Intersection(quadric, box)), a ray enters one or two).
        if constexpr (MODE == 4) {
            int k = n_lock;
            while (true) {
                int ref = 0;
                bool found = false;
                while (k < S.n_list) {
                    const float4 lo = __ldg(S.root_list + 2 * k), hi = __ldg(S.root_list + 2 * k + 1);
                    ++k;
                    const float x0 = fmaf(lo.x, idx, -oodx), x1 = fmaf(hi.x, idx, -oodx);
                    const float y0 = fmaf(lo.y, idy, -oody), y1 = fmaf(hi.y, idy, -oody);
                    const float z0 = fmaf(lo.z, idz, -oodz), z1 = fmaf(hi.z, idz, -oodz);
                    const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin_f));
                    const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), hi.w != 0.f ? FLT_MAX : tmax_f));
                    if (tn <= tf * slack) {
                        ref = __float_as_int(lo.w);
                        found = true;
                        break;
                    }
                }
                if (!found) break;
#if RS_TUNE_COST_DUMP
                ++best.n_leaves;
#endif
                test_root<MODE>(S, ref, ray, inv_d, rng, best);
                tmax_f = best.t < (double)FLT_MAX ? up(__double2float_ru(best.t)) : FLT_MAX;
            }
        }
    }
    if (S.bvh_root == RS_BVH_SENTINEL) return;

    // stack entries carry the entry distance of the child (its sort key) next to the reference,
    // so entries that fell behind the closest hit found meanwhile are dropped when popped —
    // most stacked entries are leaves, and a leaf costs an f64 object test
    unsigned long long stack[RS_STACK];
    int sp = 0;
    int node = S.bvh_root;  // may itself be a leaf (single object)
    constexpr bool SST = !WIDE && RS_TUNE_SSTACK > 0;
    const bool use_s = SST && sstack != nullptr;  // (uniform per kernel: folded after inlining)
    int lo = 0;  // entries [lo, sp) are in shared memory (slot = index mod RS_TUNE_SSTACK), [0, lo) in local memory
#define RS_SSLOT(k) sstack[((k) & ((RS_TUNE_SSTACK > 0 ? RS_TUNE_SSTACK : 1) - 1)) * RS_BLOCK]
    if (use_s) RS_SSLOT(0) = (unsigned long long)(uint32_t)RS_BVH_SENTINEL;  // key 0: never dropped
    else stack[0] = (unsigned long long)(uint32_t)RS_BVH_SENTINEL;
    sp = 1;
#define RS_PUSH2(e)                                                                        \
    do {                                                                                   \
        if (use_s) {                                                                       \
            if (sp - lo == RS_TUNE_SSTACK) { stack[lo] = RS_SSLOT(lo); ++lo; }             \
            RS_SSLOT(sp) = (e);                                                            \
        } else {                                                                           \
            stack[sp] = (e);                                                               \
        }                                                                                  \
        ++sp;                                                                              \
    } while (false)
#define RS_POP()                                                                   \
    do {                                                                           \
        unsigned long long e_;                                                     \
        --sp;                                                                      \
        if (use_s) {                                                               \
            e_ = RS_SSLOT(sp);                                                     \
            if (sp == lo && lo > 0) { --lo; RS_SSLOT(lo) = stack[lo]; }            \
        } else {                                                                   \
            e_ = stack[sp];                                                        \
        }                                                                          \
        node = (int)(uint32_t)e_;                                                  \
        if (!(__uint_as_float((uint32_t)(e_ >> 32) & ~3u) > tmax_f)) break;        \
    } while (true)
    while (true) {
        while (!WIDE && node >= 0) {
#if RS_TUNE_COST_DUMP
            ++best.n_nodes;
#endif
            // one 64-byte node: two children with their boxes, four 128-bit ld.global.nc
            float4 bx, by, bz;
            int4 ch;
            if constexpr (STAGE) {  // the tree sits in shared memory: four ld.shared.v4
                const uint32_t na = st.nodes + 64u * (uint32_t)node;
                bx = lds_f4(na), by = lds_f4(na + 16u), bz = lds_f4(na + 32u);
                ch = lds_i4(na + 48u);
            } else {
                load_node(S.bvh2 + node, bx, by, bz, ch);
            }
#if RS_TUNE_FFMA2
            const float2 ix2 = make_float2(idx, idx), iy2 = make_float2(idy, idy), iz2 = make_float2(idz, idz);
            const float2 ox2 = make_float2(-oodx, -oodx), oy2 = make_float2(-oody, -oody), oz2 = make_float2(-oodz, -oodz);
            float2 ta = ffma2(make_float2(bx.x, bx.y), ix2, ox2), tb = ffma2(make_float2(bx.z, bx.w), iy2, oy2), tc = ffma2(make_float2(bz.x, bz.y), iz2, oz2);
            const float near0 = fmaxf(fmaxf(fminf(ta.x, ta.y), fminf(tb.x, tb.y)), fmaxf(fminf(tc.x, tc.y), tmin_f));
            const float far0 = fminf(fminf(fmaxf(ta.x, ta.y), fmaxf(tb.x, tb.y)), fminf(fmaxf(tc.x, tc.y), tmax_f));
            ta = ffma2(make_float2(by.x, by.y), ix2, ox2); tb = ffma2(make_float2(by.z, by.w), iy2, oy2); tc = ffma2(make_float2(bz.z, bz.w), iz2, oz2);
            const float near1 = fmaxf(fmaxf(fminf(ta.x, ta.y), fminf(tb.x, tb.y)), fmaxf(fminf(tc.x, tc.y), tmin_f));
            const float far1 = fminf(fminf(fmaxf(ta.x, ta.y), fmaxf(tb.x, tb.y)), fminf(fmaxf(tc.x, tc.y), tmax_f));
#else
            float a0 = fmaf(bx.x, idx, -oodx), a1 = fmaf(bx.y, idx, -oodx);
            float b0 = fmaf(bx.z, idy, -oody), b1 = fmaf(bx.w, idy, -oody);
            float c0 = fmaf(bz.x, idz, -oodz), c1 = fmaf(bz.y, idz, -oodz);
            const float near0 = fmaxf(fmaxf(fminf(a0, a1), fminf(b0, b1)), fmaxf(fminf(c0, c1), tmin_f));
            const float far0 = fminf(fminf(fmaxf(a0, a1), fmaxf(b0, b1)), fminf(fmaxf(c0, c1), tmax_f));
            a0 = fmaf(by.x, idx, -oodx); a1 = fmaf(by.y, idx, -oodx);
            b0 = fmaf(by.z, idy, -oody); b1 = fmaf(by.w, idy, -oody);
            c0 = fmaf(bz.z, idz, -oodz); c1 = fmaf(bz.w, idz, -oodz);
            const float near1 = fmaxf(fmaxf(fminf(a0, a1), fminf(b0, b1)), fmaxf(fminf(c0, c1), tmin_f));
            const float far1 = fminf(fminf(fmaxf(a0, a1), fmaxf(b0, b1)), fminf(fmaxf(c0, c1), tmax_f));
#endif
            const bool h0 = near0 <= far0 * slack, h1 = near1 <= far1 * slack;
            if (h0 && h1) {
                const bool swap = near1 < near0;
                const int second = swap ? ch.x : ch.y;
                const float second_near = swap ? near0 : near1;
                if (sp < RS_STACK) RS_PUSH2(((unsigned long long)__float_as_uint(second_near) << 32) | (unsigned long long)(uint32_t)second);
                else (void)RS_IN_BOUNDS(false);  // a dropped push would lose a subtree: upload bounds the depth so that it cannot happen
                node = swap ? ch.y : ch.x;
            } else if (h0) {
                node = ch.x;
            } else if (h1) {
                node = ch.y;
            } else {
                RS_POP();
            }
        }
        while (WIDE && node >= 0) {
#if RS_TUNE_COST_DUMP
            ++best.n_nodes;
#endif
            // one 128-byte QBVH node: seven 128-bit ld.global.nc
            const float4* np = reinterpret_cast<const float4*>(S.bvh + node);
            const float4 lox = __ldg(np), hix = __ldg(np + 1), loy = __ldg(np + 2), hiy = __ldg(np + 3), loz = __ldg(np + 4), hiz = __ldg(np + 5);
            const int4 ch = __ldg(reinterpret_cast<const int4*>(np + 6));
            // slab test of child k; hits become sortable keys: (bits(t_near) & ~3) | k  (t_near > 0)
#define RS_SLAB4(k, C)                                                                                         \
            uint32_t key##k;                                                                                       \
            {                                                                                                      \
                const float x0 = fmaf(lox.C, idx, -oodx), x1 = fmaf(hix.C, idx, -oodx);                            \
                const float y0 = fmaf(loy.C, idy, -oody), y1 = fmaf(hiy.C, idy, -oody);                            \
                const float z0 = fmaf(loz.C, idz, -oodz), z1 = fmaf(hiz.C, idz, -oodz);                            \
                const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin_f));         \
                const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), tmax_f));         \
                key##k = (tn <= tf * slack) ? ((__float_as_uint(tn) & ~3u) | (uint32_t)k) : 0xffffffffu;           \
            }
            RS_SLAB4(0, x)
            RS_SLAB4(1, y)
            RS_SLAB4(2, z)
            RS_SLAB4(3, w)
#undef RS_SLAB4
            // 5-comparator sorting network: key0 = nearest hit ... misses (0xffffffff) last
#define RS_CSWAP(a, b) { const uint32_t lo_ = min(a, b), hi_ = max(a, b); a = lo_; b = hi_; }
            RS_CSWAP(key0, key1) RS_CSWAP(key2, key3) RS_CSWAP(key0, key2) RS_CSWAP(key1, key3) RS_CSWAP(key1, key2)
#undef RS_CSWAP
#define RS_CHILD(key) (((key) & 3u) == 0u ? ch.x : (((key) & 3u) == 1u ? ch.y : (((key) & 3u) == 2u ? ch.z : ch.w)))
#define RS_PUSH(key) stack[sp++] = ((unsigned long long)(key) << 32) | (unsigned long long)(uint32_t)RS_CHILD(key)
            if (key0 == 0xffffffffu) {
                RS_POP();
            } else {
                // push the farther hits (farthest first) so that the nearest is popped first
                if (key3 != 0xffffffffu && sp < RS_STACK) RS_PUSH(key3);
                if (key2 != 0xffffffffu && sp < RS_STACK) RS_PUSH(key2);
                if (key1 != 0xffffffffu && sp < RS_STACK) RS_PUSH(key1);
                node = RS_CHILD(key0);
            }
#undef RS_PUSH
#undef RS_CHILD
        }
        if (node == RS_BVH_SENTINEL) break;
#if RS_TUNE_COST_DUMP
        ++best.n_leaves;
#endif
        test_root<MODE, STAGE>(S, ~node, ray, inv_d, rng, best, st);
        // shrink the f32 limit (rounded up) so farther boxes are culled
        tmax_f = best.t < (double)FLT_MAX ? up(__double2float_ru(best.t)) : FLT_MAX;
        RS_POP();
    }
#undef RS_POP
#undef RS_PUSH2
#undef RS_SSLOT
}

template <int MODE>
__device__ __forceinline__ int root_material(const DScene& S, const Closest& best) {
    int m = best.material;  // generic objects: the material their expression tree resolved
    if (MODE < 2 || m == -2) m = __ldg(&S.root_material[best.root]);
    return m < 0 ? S.default_material : m;
}

// MixedMaterial::scatter (mixed_material.rs:43-51) resolved ahead of shading, so that the
// class buckets hold concrete materials.  Level 0 uses word 2 of the P_SCATTER call.
__device__ __forceinline__ int resolve_mixed(const DScene& S, int m, const RngCtx& rng) {
    for (uint32_t level = 0; level < RS_MAX_MIXED_DEPTH && RS_MATHOT_TYPE(S.mat_hot[m]) == RS_MAT_MIXED; ++level) {
        const rs_material& mm = S.materials[m];
        uint32_t word = level == 0 ? rng_words(rng, P_SCATTER, 0).z : rng_words(rng, P_MIXED, level).x;
        m = ((double)word < 4294967295.0 * mm.a) ? mm.mat1 : mm.mat2;
    }
    return m;
}

// FIRST = true (bounce 0, no pixel mask): the kernel GENERATES the camera ray of path i itself
// (raygen fused in: Camera::ray + the strata loop of painter.rs:154-187) instead of reading a
// record raygen wrote — one kernel, one 64-byte record read per path less, and a path that ends on
// its first segment (sky, light) never gets a record at all.
// STAGE = true (lean scenes whose binary tree + sphere table fit, rs_upload_scene decides): every persistent
// block first copies the tree and the sphere table into shared memory — two 1-D bulk copies (TMA,
// cp.async.bulk) completing on one mbarrier — and the traversal reads them with ld.shared: a node fetch
// no longer competes with the streaming ray / hit / stack traffic for L1 lines (ncu, headline scene:
// L1 hit 75-78 % on a 61 KB tree; every miss is an L2 round trip on the critical path of the walk).
// MEASURED SLOWER (profiles/r02_ab_stage.txt: headline extend 30.5 vs 29.2 ms at 64 spp; 512 x 2 and 1024 x 1
// blocks, all or only the first bounces staged): the walk is issue- and divergence-bound, an L1 hit costs
// what an LDS costs, and 76 KB per block per launch is a fixed ~10 us the small late-bounce launches cannot
// amortise.  Kept as a build option (RS_TUNE_STAGE_TREE=1 compiles the kernels, RAYSNAIL_B200_STAGE_TREE=0/1 and
// RAYSNAIL_B200_STAGE_BOUNCES=k then choose at run time); the default build does not instantiate them.
#ifndef RS_TUNE_STAGE_TREE
#define RS_TUNE_STAGE_TREE 0
#endif
#ifndef RS_TUNE_STAGE_BLOCK
#define RS_TUNE_STAGE_BLOCK 512  // threads per block of the staged kernels (x 1024 / block per SM)
#endif
template <int MODE, bool WIDE, bool FIRST, bool STAGE = false>
__global__ void __launch_bounds__(STAGE ? RS_TUNE_STAGE_BLOCK : RS_BLOCK, STAGE ? 1024 / RS_TUNE_STAGE_BLOCK : RS_EXTEND_MIN_BLOCKS(MODE))
    extend_kernel(const __grid_constant__ DScene S, const __grid_constant__ FrameParams fp, const __grid_constant__ WaveState ws, const int bounce,
                  const int miss_in_extend) {
    Staged st{0, 0};
    // the hot end of every thread's traversal stack (see closest_hit)
    unsigned long long* sstack = nullptr;
    if constexpr (!WIDE && RS_TUNE_SSTACK > 0 && RS_SSTACK_MODE(MODE)) {
        __shared__ unsigned long long s_stack[(RS_TUNE_SSTACK > 0 ? RS_TUNE_SSTACK : 1) * (STAGE ? RS_TUNE_STAGE_BLOCK : RS_BLOCK)];
        sstack = s_stack + threadIdx.x;
    }
    if constexpr (STAGE) {
        extern __shared__ __align__(128) unsigned char stage_smem[];
        __shared__ __align__(8) unsigned long long stage_bar;
        st.nodes = smem_u32(stage_smem);
        st.spheres = st.nodes + S.bvh2_bytes;
        const uint32_t bar = smem_u32(&stage_bar);
        if (threadIdx.x == 0) {
            mbar_init(bar, 1);
            mbar_expect_tx(bar, S.bvh2_bytes + S.sphere_bytes);
            bulk_g2s(st.nodes, S.bvh2, S.bvh2_bytes, bar);
            bulk_g2s(st.spheres, S.root_sphere, S.sphere_bytes, bar);
        }
        __syncthreads();  // the barrier is initialised for everyone
        mbar_wait(bar, 0);
    }
    // virtual index space: [0, nf32) = the front stack padded to a multiple of 32 (so that no warp
    // mixes the two kinds of rays), [nf32, nf32 + nb) = the back stack, top first
    const uint32_t nf = FIRST ? fp.n_pixels * fp.n_samples : ws.ctr->n_rays[bounce], nb = FIRST ? 0u : ws.ctr->n_back[bounce];
    const uint32_t nf32 = (nf + 31u) & ~31u;
    const uint32_t n = nf32 + nb;
    const uint32_t lane_zero = RS_LANE_ZERO(ws);
    const PathBuf& in = ws.buf[bounce & 1];
    uint32_t* head = &ws.ctr->extend_head[bounce];
    uint32_t* class_count = ws.ctr->class_count[bounce];
    const bool last_bounce = bounce + 1 >= fp.depth;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(&ws.stats->n_segments, (unsigned long long)(nf + nb));
        if (FIRST) atomicAdd(&ws.stats->n_samples, (unsigned long long)nf);
    }
    // the cursor bump for the NEXT batch is issued before the current batch is processed, so the
    // atomic's round trip to L2 overlaps with the traversal instead of heading every batch
#if RS_TUNE_EXTEND_STATIC
    (void)head;
    uint32_t next_base = (blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5)) * 32u;
    const uint32_t base_stride = gridDim.x * (blockDim.x / 32) * 32u;
#else
    uint32_t pend = 0;
    if (lane_id() == 0) pend = atomic_add_async(head, 32u, lane_zero);
#endif
    while (true) {
#if RS_TUNE_EXTEND_STATIC
        const uint32_t base = next_base;
        if (base >= n) break;
        next_base += base_stride;
#else
        const uint32_t base = __shfl_sync(0xffffffffu, pend, 0);
        if (base >= n) break;
        // RS_TUNE_FIRST_LATE_FETCH: the bounce-0 kernels (camera ray generation fused in) spill `pend` at 64
        // registers, so lane 0 waits for the atomic at the spill store; issuing it after the
        // traversal instead keeps it in a register and overlaps it with the epilogue only
        if (!(RS_TUNE_FIRST_LATE_FETCH && FIRST) && lane_id() == 0) pend = atomic_add_async(head, 32u, lane_zero);
#endif
        const uint32_t v = base + lane_id();
        const bool live = v < nf32 ? v < nf : v < n;
        const uint32_t i = v < nf32 ? v : ws.capacity - 1u - (v - nf32);
        int cls = -1;
        uint2* queue_slot = nullptr;
        RngCtx rng_of_lane;
        rng_of_lane.key = &fp.rk;
        rng_of_lane.pixel = 0;
        rng_of_lane.sample = 0;
        rng_of_lane.depth = (uint32_t)bounce + 1;
        if (live) {
            Ray ray;
            RngCtx rng;
            rng.key = &fp.rk;
            if constexpr (FIRST) {
                decode_slot(fp, i, rng.pixel, rng.sample);
                const uint32_t width = (uint32_t)fp.cam.width, row = rng.pixel / width, col = rng.pixel - row * width;
                uint32_t s_j = __umulhi(rng.sample, fp.magic_q), s_i = rng.sample - s_j * (uint32_t)fp.sqrt_spp;
                if (s_i >= (uint32_t)fp.sqrt_spp) { ++s_j; s_i -= (uint32_t)fp.sqrt_spp; }
                rng.depth = 0;
                ray = camera_ray(fp, (int)row, (int)col, (int)s_i, (int)s_j, rng);
            } else {
                load_ray_only(fp, ws, in, i, ray, rng.pixel, rng.sample);
            }
            rng.depth = (uint32_t)bounce + 1;
            if constexpr (RS_TUNE_RESERVE && MODE < 2) rng_of_lane = rng;
            Closest best;
            closest_hit<MODE, WIDE, STAGE>(S, ray, rng, best, st, sstack);
#if RS_TUNE_COST_DUMP
            if (ws.dbg_cost) ws.dbg_cost[i] = make_int2(best.n_nodes, best.n_leaves);
#endif
            HitA h;
            h.t = best.t;
            h.root = best.root;
            h.matinfo = 0;
            cls = 0;
            if (best.root >= 0) {
                int top = root_material<MODE>(S, best);
                int leaf = resolve_mixed(S, top, rng);
                h.matinfo = (uint32_t)leaf | ((uint32_t)top << 16);
                const MatHot& mh = S.mat_hot[leaf];
                const int tf = mh.type_flags;
                cls = 1 + (tf & 0xff);
                if (miss_in_extend && (tf & 0xff) == RS_MAT_DIFFUSE_LIGHT && leaf != top) {
                    // a DiffuseLight reached THROUGH a MixedMaterial: `emitted` is asked of the hit's
                    // own (top) material, and MixedMaterial::emitted is None (mixed_material.rs:53-55,
                    // camera.rs:172-174); the light's `scatter` is None as well (material/mod.rs:59-61), so the
                    // path ends here and adds nothing
                    if constexpr (FIRST) { if (RS_IN_BOUNDS(i < ws.capacity)) ws.contrib[i] = make_float4(0.f, 0.f, 0.f, 0.f); }
                    cls = -1;
                } else if (miss_in_extend && (tf & 0x3ff) == RS_MAT_DIFFUSE_LIGHT) {
                    // a DiffuseLight never scatters (light.rs:33-35): the path ends here with
                    // T * colour * multiplier.  With a solid colour (tex_mode 0) nothing about the
                    // hit point is needed, so it is finished on the spot like a miss — no hit
                    // record, no queue entry, no gather of the path record in shade.
                    const float4 c = mh.c0;
                    const double mul = mh.a;
                    const uint4 q3 = FIRST ? make_uint4(0x3f800000u, 0x3f800000u, 0x3f800000u, i) : in.rec[4 * (size_t)i + 3];  // (T, slot)
                    if (RS_IN_BOUNDS(q3.w < ws.capacity)) ws.contrib[q3.w] = make_float4(__uint_as_float(q3.x) * (float)((double)c.x * mul), __uint_as_float(q3.y) * (float)((double)c.y * mul),
                                                   __uint_as_float(q3.z) * (float)((double)c.z * mul), 0.f);
                    cls = -1;
                } else {
                    if (MODE >= 3 && S.root_kind[best.root] == ROOT_GENERIC && RS_IN_BOUNDS(i < ws.capacity)) ws.hit_src[i] = make_int2(best.src, best.flavor);
                    if (RS_IN_BOUNDS(i < ws.capacity && best.root < S.n_roots)) ws.hit[i] = h;
                }
            } else if (miss_in_extend) {
                // a miss ends the path with `world.background(ray)` (camera.rs:253-254): finish it
                // here instead of sending it through a queue to the shade kernel
                const float4 bg = background(S, ray.d);
                const uint4 q3 = FIRST ? make_uint4(0x3f800000u, 0x3f800000u, 0x3f800000u, i) : in.rec[4 * (size_t)i + 3];  // (T, slot)
                if (RS_IN_BOUNDS(q3.w < ws.capacity)) ws.contrib[q3.w] = make_float4(__uint_as_float(q3.x) * bg.x, __uint_as_float(q3.y) * bg.y, __uint_as_float(q3.z) * bg.z, 0.f);
                cls = -1;
            } else {
                if (RS_IN_BOUNDS(i < ws.capacity)) ws.hit[i] = h;
            }
            if constexpr (FIRST) {
                if (RS_CHECKED && (fp.flags & 0x80000000u) && i < 32u) {  // (checked builds: the negative control, see render_into)
                    PathRec q;
                    q.ray = ray;
                    q.T = make_float3(1.f, 1.f, 1.f);
                    q.slot = i;
                    store_path(ws, in, ws.capacity + i, q);
                }
                if (cls >= 0) {  // the path goes on to shade: now it needs its record (and a zeroed contribution)
                    PathRec p;
                    p.ray = ray;
                    p.T = make_float3(1.f, 1.f, 1.f);
                    p.slot = i;
                    store_path(ws, in, i, p);
                    if (RS_IN_BOUNDS(i < ws.capacity)) ws.contrib[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
        }
#if !RS_TUNE_EXTEND_STATIC
        if (RS_TUNE_FIRST_LATE_FETCH && FIRST && lane_id() == 0) pend = atomic_add_async(head, 32u, lane_zero);
#endif
        // material-coherent bucketing
        const uint32_t active = __ballot_sync(0xffffffffu, live && cls >= 0);
        // Paths of the classes in RS_ALWAYS_CONTINUES always go on after shading (their `scatter`
        // never returns None), and whether the next ray is the light-directed one is decided by the
        // first word of the P_SCATTER draw alone (camera.rs:198).  Their slot in the next bounce's
        // path buffer is therefore reserved HERE and handed to shade with the queue entry: shade
        // (4 warps per scheduler, latency-bound) runs these classes without any atomic or ballot,
        // while this kernel has the warps to hide the two extra round trips to L2.
        constexpr bool RESERVE = RS_TUNE_RESERVE && MODE < 2;  // lean kernels only: the generic one is register-bound and its shade does not gain
        bool reserve = false, res_front = false;
        uint32_t mf = 0, mb = 0, bf = 0, bb = 0, lf = 0, lb = 0;
        if constexpr (RESERVE) {
            if (live && cls >= 0 && !last_bounce && RS_ALWAYS_CONTINUES(cls)) {
                reserve = true;
                res_front = cls != 1 + RS_MAT_DIELECTRIC && u01(rng_words(rng_of_lane, P_SCATTER, 0).x) < 0.5;
            }
            mf = __ballot_sync(0xffffffffu, reserve && res_front);
            mb = __ballot_sync(0xffffffffu, reserve && !res_front);
            lf = mf ? (uint32_t)(__ffs(mf) - 1) : 0u;
            lb = mb ? (uint32_t)(__ffs(mb) - 1) : 0u;
            if (mf && lane_id() == lf) bf = atomic_add_async(&ws.ctr->n_rays[bounce + 1], __popc(mf), lane_zero);
            if (mb && lane_id() == lb) bb = atomic_add_async(&ws.ctr->n_back[bounce + 1], __popc(mb), lane_zero);
        }
        if (live && cls >= 0) {
            const int key = (RESERVE && res_front) ? RS_FRONT_CLASS(cls) : cls;  // (class, branch) bucket
            const uint32_t peers = __match_any_sync(active, key);
            const uint32_t leader = __ffs(peers) - 1;
            uint32_t qbase = 0;
            if (lane_id() == leader) qbase = atomic_add_async(&class_count[key * RS_COUNTER_STRIDE], __popc(peers), lane_zero);
            qbase = __shfl_sync(peers, qbase, leader);
            const uint32_t rank = __popc(peers & ((1u << lane_id()) - 1u));
            (void)RS_IN_BOUNDS(key >= 0 && key < RS_NUM_CLASSES && qbase + rank < ws.capacity);
            uint2* const slot = ws.class_queue + ((size_t)key * ws.capacity + (RS_CHECKED ? min(qbase + rank, ws.capacity - 1u) : qbase + rank));
            if constexpr (!RESERVE) *slot = make_uint2(i, 0xffffffffu);
            else queue_slot = slot;
        }
        if constexpr (RESERVE) {
            bf = __shfl_sync(0xffffffffu, bf, lf);
            bb = __shfl_sync(0xffffffffu, bb, lb);
            if (live && cls >= 0) {
                const uint32_t lt = (1u << lane_id()) - 1u;
                const uint32_t out_pos = !reserve ? 0xffffffffu : (res_front ? bf + __popc(mf & lt) : ws.capacity - 1u - (bb + __popc(mb & lt)));
                *queue_slot = make_uint2(i, out_pos);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// extend, streaming form (triangle-mesh scenes, bounces >= 1; chosen per scene in rs_upload_scene)
// ---------------------------------------------------------------------------------------------
// After the first bounce the rays of a warp need wildly different numbers of node visits on a mesh
// (1 M triangles, bounce 1: median 4, mean 18, 99th percentile 90) and extend_kernel's warp, which
// takes 32 rays and walks them to the end — every round of its while-while loop as long as the
// slowest lane's walk to its next leaf — issues with 6-7 of its 32 lanes.  Here
//  - every lane owns its traversal state, and one loop iteration is ONE node step for the lanes that
//    hold an inner node;
//  - lanes that reached a leaf wait until RS_TUNE_STREAM_LEAF lanes hold one (or no lane walks any
//    more), then test their triangles together;
//  - lanes whose ray is done park (ray, t, object) in a per-warp shared-memory queue and, once
//    RS_TUNE_STREAM_REFILL lanes are idle, take new rays from the warp's pool (32 indices per atomic,
//    fetched one chunk ahead);
//  - the epilogue of a ray (material, miss / light contribution, class bucket, output slot) runs off
//    that queue 32 rays at a time, i.e. with full warps.
// The f64 ray is not kept in registers: a leaf test reloads it from the path record (three 128-bit
// loads that hit L1), which is what lets the per-lane state fit the 64-register budget.
#ifndef RS_TUNE_STREAM_REFILL
#define RS_TUNE_STREAM_REFILL 8
#endif
#ifndef RS_TUNE_STREAM_LEAF
#define RS_TUNE_STREAM_LEAF 8
#endif
#ifndef RS_TUNE_STREAM_BLOCKS
#define RS_TUNE_STREAM_BLOCKS 4
#endif
#ifndef RS_TUNE_STREAM_PREFETCH
#define RS_TUNE_STREAM_PREFETCH 0
#endif
#ifndef RS_TUNE_STREAM_SSTACK
#define RS_TUNE_STREAM_SSTACK 16  // stack entries per thread kept in shared memory (power of two; 8 -> 16: +1.2 % on the 1 M-triangle mesh, profiles/r02j_ab_stream.txt)
#endif
#define RS_STREAM_SMEM ((RS_BLOCK / 32) * 64 * sizeof(uint4) + 6 * RS_BLOCK * sizeof(double) + RS_TUNE_STREAM_SSTACK * RS_BLOCK * sizeof(unsigned long long))
template <int MODE>
__global__ void __launch_bounds__(RS_BLOCK, RS_TUNE_STREAM_BLOCKS) extend_stream_kernel(const __grid_constant__ DScene S, const __grid_constant__ FrameParams fp,
                                                                                        const __grid_constant__ WaveState ws, const int bounce,
                                                                                        const int miss_in_extend) {
    static_assert(MODE < 2, "lean scenes only");
    // shared memory of a block: the queues of finished rays awaiting their epilogue (per warp: 64 x (ray index,
    // object, t)), every thread's f64 ray (what a triangle test needs) and the TOP RS_TUNE_STREAM_SSTACK entries of
    // every thread's traversal stack.  All per-thread arrays are laid out [entry][thread]: a lane owns its banks,
    // so a warp's access is conflict-free whatever entry each lane is at — where the same stack in local memory
    // costs one L1 sector access per lane and push or pop (ncu, 1 M-triangle mesh: 22 % of all L1 sector accesses,
    // on a kernel whose L1 runs at 76 % of peak).
    extern __shared__ __align__(16) unsigned char stream_smem[];
    uint4* const queue = reinterpret_cast<uint4*>(stream_smem) + 64 * (threadIdx.x >> 5);
    double* const sray = reinterpret_cast<double*>(stream_smem + (RS_BLOCK / 32) * 64 * sizeof(uint4)) + threadIdx.x;  // [6][RS_BLOCK]
    unsigned long long* const sstack = reinterpret_cast<unsigned long long*>(stream_smem + (RS_BLOCK / 32) * 64 * sizeof(uint4) + 6 * RS_BLOCK * sizeof(double)) + threadIdx.x;  // [SSTACK][RS_BLOCK]
    const uint32_t nf = ws.ctr->n_rays[bounce], nb = ws.ctr->n_back[bounce];
    const uint32_t nf32 = (nf + 31u) & ~31u;
    const uint32_t n = nf32 + nb;
    const uint32_t lane = lane_id(), lt = (1u << lane) - 1u;
    const uint32_t lane_zero = RS_LANE_ZERO(ws);
    const PathBuf& in = ws.buf[bounce & 1];
    uint32_t* head = &ws.ctr->extend_head[bounce];
    uint32_t* class_count = ws.ctr->class_count[bounce];
    const bool last_bounce = bounce + 1 >= fp.depth;
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&ws.stats->n_segments, (unsigned long long)(nf + nb));
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    const float tmin_f = (float)RS_TMIN * (1.0f - 2.0f * RS_SLAB_EPS);
    const float slack = (1.0f + RS_SLAB_EPS) * (1.0f + RS_SLAB_EPS);

    // per-lane traversal state
    // entries [lo, sp) of the stack live in shared memory (slot = index mod RS_TUNE_STREAM_SSTACK), [0, lo) in local memory
    unsigned long long stack[RS_STACK];
    int sp = 0, lo = 0, node = RS_BVH_SENTINEL;
    bool has_ray = false;
    uint32_t i = 0;
    float idx = 0.f, idy = 0.f, idz = 0.f, oodx = 0.f, oody = 0.f, oodz = 0.f, tmax_f = FLT_MAX;
    Closest best;
    best.t = inf; best.root = -1; best.material = -1;
    // the warp's pool of virtual ray indices [pool_next, pool_end) + the chunk fetched ahead
    uint32_t pool_next = 0, pool_end = 0, pend = 0, queued = 0;
    bool exhausted = false;
    if (lane == 0) pend = atomic_add_async(head, 32u, lane_zero);
    RngCtx no_rng;
    no_rng.key = &fp.rk; no_rng.pixel = 0; no_rng.sample = 0; no_rng.depth = 0;

    // epilogue of up to 32 queued rays (extend_kernel's, for a ray that came from the path buffer)
    auto epilogue = [&](const bool live, const uint4 e) {
        const uint32_t ri = e.x;
        int cls = -1;
        RngCtx rng;
        rng.key = &fp.rk; rng.pixel = 0; rng.sample = 0; rng.depth = (uint32_t)bounce + 1;
        if (live) {
            const uint4 q3 = in.rec[4 * (size_t)ri + 3];  // (T, slot)
            decode_slot(fp, q3.w, rng.pixel, rng.sample);
            Closest b;
            b.t = __hiloint2double((int)e.w, (int)e.z);
            b.root = (int)e.y;
            b.material = -1;
            HitA h;
            h.t = b.t; h.root = b.root; h.matinfo = 0;
            cls = 0;
            if (b.root >= 0) {
                const int top = root_material<MODE>(S, b);
                const int leaf = resolve_mixed(S, top, rng);
                h.matinfo = (uint32_t)leaf | ((uint32_t)top << 16);
                const MatHot& mh = S.mat_hot[leaf];
                const int tf = mh.type_flags;
                cls = 1 + (tf & 0xff);
                if (miss_in_extend && (tf & 0xff) == RS_MAT_DIFFUSE_LIGHT && leaf != top) {
                    cls = -1;  // a light behind a MixedMaterial neither emits nor scatters (see extend_kernel)
                } else if (miss_in_extend && (tf & 0x3ff) == RS_MAT_DIFFUSE_LIGHT) {
                    const float4 c = mh.c0;
                    const double mul = mh.a;
                    if (RS_IN_BOUNDS(q3.w < ws.capacity)) ws.contrib[q3.w] = make_float4(__uint_as_float(q3.x) * (float)((double)c.x * mul), __uint_as_float(q3.y) * (float)((double)c.y * mul),
                                                   __uint_as_float(q3.z) * (float)((double)c.z * mul), 0.f);
                    cls = -1;
                } else {
                    if (RS_IN_BOUNDS(ri < ws.capacity)) ws.hit[ri] = h;
                }
            } else if (miss_in_extend) {
                const uint4 q1 = in.rec[4 * (size_t)ri + 1], q2 = in.rec[4 * (size_t)ri + 2];
                const V3 d = v3(__hiloint2double((int)q1.w, (int)q1.z), __hiloint2double((int)q2.y, (int)q2.x), __hiloint2double((int)q2.w, (int)q2.z));
                const float4 bg = background(S, d);
                if (RS_IN_BOUNDS(q3.w < ws.capacity)) ws.contrib[q3.w] = make_float4(__uint_as_float(q3.x) * bg.x, __uint_as_float(q3.y) * bg.y, __uint_as_float(q3.z) * bg.z, 0.f);
                cls = -1;
            } else {
                if (RS_IN_BOUNDS(ri < ws.capacity)) ws.hit[ri] = h;
            }
        }
        const bool go = live && cls >= 0;
        const uint32_t active = __ballot_sync(0xffffffffu, go);
        constexpr bool RESERVE = RS_TUNE_RESERVE != 0;
        bool reserve = false, res_front = false;
        uint32_t mf = 0, mb = 0, bf = 0, bb = 0, lf = 0, lb = 0;
        if constexpr (RESERVE) {
            if (go && !last_bounce && RS_ALWAYS_CONTINUES(cls)) {
                reserve = true;
                res_front = cls != 1 + RS_MAT_DIELECTRIC && u01(rng_words(rng, P_SCATTER, 0).x) < 0.5;
            }
            mf = __ballot_sync(0xffffffffu, reserve && res_front);
            mb = __ballot_sync(0xffffffffu, reserve && !res_front);
            lf = mf ? (uint32_t)(__ffs(mf) - 1) : 0u;
            lb = mb ? (uint32_t)(__ffs(mb) - 1) : 0u;
            if (mf && lane == lf) bf = atomic_add_async(&ws.ctr->n_rays[bounce + 1], __popc(mf), lane_zero);
            if (mb && lane == lb) bb = atomic_add_async(&ws.ctr->n_back[bounce + 1], __popc(mb), lane_zero);
        }
        uint2* queue_slot = nullptr;
        if (go) {
            const int key = (RESERVE && res_front) ? RS_FRONT_CLASS(cls) : cls;
            const uint32_t peers = __match_any_sync(active, key);
            const uint32_t leader = __ffs(peers) - 1;
            uint32_t qbase = 0;
            if (lane == leader) qbase = atomic_add_async(&class_count[key * RS_COUNTER_STRIDE], __popc(peers), lane_zero);
            qbase = __shfl_sync(peers, qbase, leader);
            (void)RS_IN_BOUNDS(key >= 0 && key < RS_NUM_CLASSES && qbase + __popc(peers & lt) < ws.capacity);
            queue_slot = ws.class_queue + ((size_t)key * ws.capacity + (RS_CHECKED ? min(qbase + (uint32_t)__popc(peers & lt), ws.capacity - 1u) : qbase + __popc(peers & lt)));
            if constexpr (!RESERVE) *queue_slot = make_uint2(ri, 0xffffffffu);
        }
        if constexpr (RESERVE) {
            bf = __shfl_sync(0xffffffffu, bf, lf);
            bb = __shfl_sync(0xffffffffu, bb, lb);
            if (go) {
                const uint32_t out_pos = !reserve ? 0xffffffffu : (res_front ? bf + __popc(mf & lt) : ws.capacity - 1u - (bb + __popc(mb & lt)));
                *queue_slot = make_uint2(ri, out_pos);
            }
        }
    };

#define RS_SSLOT(k) sstack[((k) & (RS_TUNE_STREAM_SSTACK - 1)) * RS_BLOCK]
#define RS_PUSH(e)                                                                 \
    do {                                                                           \
        if (sp - lo == RS_TUNE_STREAM_SSTACK) { stack[lo] = RS_SSLOT(lo); ++lo; }  \
        RS_SSLOT(sp) = (e);                                                        \
        ++sp;                                                                      \
    } while (false)
#define RS_POP()                                                                   \
    do {                                                                           \
        --sp;                                                                      \
        const unsigned long long e_ = RS_SSLOT(sp);                                \
        if (sp == lo && lo > 0) { --lo; RS_SSLOT(lo) = stack[lo]; }                \
        node = (int)(uint32_t)e_;                                                  \
        if (!(__uint_as_float((uint32_t)(e_ >> 32) & ~3u) > tmax_f)) break;        \
    } while (true)
    while (true) {
        // ---- lanes whose ray is done hand it to the queue ----
        {
            const bool fin = has_ray && node == RS_BVH_SENTINEL;
            const uint32_t fm = __ballot_sync(0xffffffffu, fin);
            if (fm) {
                if (fin) {
                    queue[queued + __popc(fm & lt)] = make_uint4(i, (uint32_t)best.root, (uint32_t)__double2loint(best.t), (uint32_t)__double2hiint(best.t));
                    has_ray = false;
                }
                queued += __popc(fm);
                __syncwarp();
                if (queued >= 32u) {
                    queued -= 32u;
                    const uint4 e = queue[queued + lane];
                    __syncwarp();
                    epilogue(true, e);
                }
            }
        }
        // ---- idle lanes take new rays from the pool ----
        const uint32_t idle = __ballot_sync(0xffffffffu, !has_ray);
        if (!exhausted && __popc(idle) >= RS_TUNE_STREAM_REFILL) {
            const uint32_t need = __popc(idle), avail = pool_end - pool_next, rank = __popc(idle & lt);
            uint32_t v;
            if (need > avail) {  // warp-uniform: cross into the chunk fetched ahead
                const uint32_t nextb = __shfl_sync(0xffffffffu, pend, 0);
                v = rank < avail ? pool_next + rank : nextb + (rank - avail);
                pool_next = nextb + (need - avail);
                pool_end = nextb + 32u;
                if (nextb >= n) exhausted = true;
                else if (lane == 0) pend = atomic_add_async(head, 32u, lane_zero);
            } else {
                v = pool_next + rank;
                pool_next += need;
            }
            if (!has_ray && (v < nf32 ? v < nf : v < n)) {
                i = v < nf32 ? v : ws.capacity - 1u - (v - nf32);
                const uint4* rec = in.rec + 4 * (size_t)i;
                const uint4 q0 = rec[0], q1 = rec[1], q2 = rec[2];
                const double o0 = __hiloint2double((int)q0.y, (int)q0.x), o1 = __hiloint2double((int)q0.w, (int)q0.z), o2 = __hiloint2double((int)q1.y, (int)q1.x);
                const double d0 = __hiloint2double((int)q1.w, (int)q1.z), d1 = __hiloint2double((int)q2.y, (int)q2.x), d2 = __hiloint2double((int)q2.w, (int)q2.z);
                sray[0] = o0; sray[RS_BLOCK] = o1; sray[2 * RS_BLOCK] = o2;
                sray[3 * RS_BLOCK] = d0; sray[4 * RS_BLOCK] = d1; sray[5 * RS_BLOCK] = d2;
                const float ox = (float)o0, oy = (float)o1, oz = (float)o2;
                float dx = (float)d0, dy = (float)d1, dz = (float)d2;
                const float tiny = 1e-20f;
                dx = fabsf(dx) > tiny ? dx : copysignf(tiny, dx);
                dy = fabsf(dy) > tiny ? dy : copysignf(tiny, dy);
                dz = fabsf(dz) > tiny ? dz : copysignf(tiny, dz);
                idx = 1.0f / dx; idy = 1.0f / dy; idz = 1.0f / dz;
                oodx = ox * idx; oody = oy * idy; oodz = oz * idz;
                tmax_f = FLT_MAX;
                best.t = inf; best.root = -1;
                has_ray = true;
                RS_SSLOT(0) = (unsigned long long)(uint32_t)RS_BVH_SENTINEL;
                sp = 1;
                lo = 0;
                node = S.n_roots ? S.bvh_root : RS_BVH_SENTINEL;
            }
        }
        if (__ballot_sync(0xffffffffu, has_ray) == 0u) {
            if (exhausted) break;
            continue;  // a chunk of padding only
        }
        // ---- node steps and batched leaf tests until enough lanes are done ----
        while (true) {
            if (node >= 0) {
                float4 bx, by, bz;
                int4 ch;
                load_node(S.bvh2 + node, bx, by, bz, ch);
                const float2 ix2 = make_float2(idx, idx), iy2 = make_float2(idy, idy), iz2 = make_float2(idz, idz);
                const float2 ox2 = make_float2(-oodx, -oodx), oy2 = make_float2(-oody, -oody), oz2 = make_float2(-oodz, -oodz);
                float2 ta = ffma2(make_float2(bx.x, bx.y), ix2, ox2), tb = ffma2(make_float2(bx.z, bx.w), iy2, oy2), tc = ffma2(make_float2(bz.x, bz.y), iz2, oz2);
                const float near0 = fmaxf(fmaxf(fminf(ta.x, ta.y), fminf(tb.x, tb.y)), fmaxf(fminf(tc.x, tc.y), tmin_f));
                const float far0 = fminf(fminf(fmaxf(ta.x, ta.y), fmaxf(tb.x, tb.y)), fminf(fmaxf(tc.x, tc.y), tmax_f));
                ta = ffma2(make_float2(by.x, by.y), ix2, ox2); tb = ffma2(make_float2(by.z, by.w), iy2, oy2); tc = ffma2(make_float2(bz.z, bz.w), iz2, oz2);
                const float near1 = fmaxf(fmaxf(fminf(ta.x, ta.y), fminf(tb.x, tb.y)), fmaxf(fminf(tc.x, tc.y), tmin_f));
                const float far1 = fminf(fminf(fmaxf(ta.x, ta.y), fmaxf(tb.x, tb.y)), fminf(fmaxf(tc.x, tc.y), tmax_f));
                const bool h0 = near0 <= far0 * slack, h1 = near1 <= far1 * slack;
                if (h0 && h1) {
                    const bool swap = near1 < near0;
                    const int second = swap ? ch.x : ch.y;
                    const float second_near = swap ? near0 : near1;
                    if (sp < RS_STACK) RS_PUSH(((unsigned long long)__float_as_uint(second_near) << 32) | (unsigned long long)(uint32_t)second);
                    else (void)RS_IN_BOUNDS(false);
#if RS_TUNE_STREAM_PREFETCH
                    // the walk is bound by the latency of its dependent node fetches (ncu: long_scoreboard 56 % of the
                    // stall samples at 8 warps per scheduler): start the fetch of the postponed child now
                    if (second >= 0) prefetch_l1(S.bvh2 + second);
                    else if (RS_TUNE_STREAM_PREFETCH >= 2 && MODE == 1 && ((~second) & 3) == ROOT_TRIANGLE) prefetch_l1(S.tri_geom + (size_t)RS_TRI_STRIDE * (size_t)((~second) >> 2));
#endif
                    node = swap ? ch.y : ch.x;
                } else if (h0) {
                    node = ch.x;
                } else if (h1) {
                    node = ch.y;
                } else {
                    RS_POP();
                }
            }
            const bool at_leaf = node < 0 && node != RS_BVH_SENTINEL;
            const uint32_t lm = __ballot_sync(0xffffffffu, at_leaf), im = __ballot_sync(0xffffffffu, node >= 0);
            if (lm && (__popc(lm) >= RS_TUNE_STREAM_LEAF || im == 0u)) {
                if (at_leaf) {
                    Ray ray;
                    ray.o = v3(sray[0], sray[RS_BLOCK], sray[2 * RS_BLOCK]);
                    ray.d = v3(sray[3 * RS_BLOCK], sray[4 * RS_BLOCK], sray[5 * RS_BLOCK]);
                    ray.time = 0.0;
                    test_root<MODE>(S, ~node, ray, v3(0.0, 0.0, 0.0), no_rng, best);
                    tmax_f = best.t < (double)FLT_MAX ? up(__double2float_ru(best.t)) : FLT_MAX;
                    RS_POP();
                }
                continue;  // (lanes may have popped another leaf: look again before deciding to leave)
            }
            const uint32_t busy = im | lm;
            if (busy == 0u) break;
            if (!exhausted && 32 - __popc(busy) >= RS_TUNE_STREAM_REFILL) break;
            if (exhausted && __popc(__ballot_sync(0xffffffffu, has_ray && node == RS_BVH_SENTINEL)) + (int)queued >= 32) break;  // a full epilogue is ready
        }
    }
#undef RS_POP
#undef RS_PUSH
#undef RS_SSLOT
    // the rays still queued
    if (queued) {
        const bool live = lane < queued;
        const uint4 e = live ? queue[lane] : make_uint4(0u, 0u, 0u, 0u);
        epilogue(live, e);
    }
}

// =============================================================================================
// shade
// =============================================================================================
// GENERIC = false: the scene holds only static spheres and triangles and its lights are plain
// spheres/rects, so no expression-tree code is compiled into the kernel.
template <bool GENERIC, bool TREES>
__device__ __forceinline__ void rebuild_hit(const DScene& S, const Ray& ray, const HitA& ha, int2 src, bool want_uv, HitRec& h) {
    // world objects are evaluated over [tmin, +inf) (see test_root): the same call returns the same record here
    const double end_then = __longlong_as_double(0x7ff0000000000000LL);
    // static spheres: centre and 1/r straight from the per-root tables (one level of loads
    // instead of root -> node index -> node)
    const int kind = (!GENERIC && S.all_spheres) ? (int)ROOT_SPHERE : __ldg(&S.root_kind[ha.root]);
    if (kind == ROOT_SPHERE) {
        const double2* sp = reinterpret_cast<const double2*>(S.root_sphere + ha.root);
        const double2 c01 = __ldg(sp), c23 = __ldg(sp + 1);
        const double pr[8] = {c01.x, c01.y, c23.x, c23.y, 0.0, 0.0, 0.0, __ldg(&S.root_inv_r[ha.root])};
        sphere_record(pr, -1, ray, ha.t, ha.t, want_uv, h);
        return;
    }
    const int rn = S.root_node[ha.root];
    if (kind == ROOT_TRIANGLE) {
        double t, beta, gamma;
        const double inf = __longlong_as_double(0x7ff0000000000000LL);
        triangle_test(load_triangle(S.tri_geom, ~rn), ray, -inf, inf, t, beta, gamma);
        triangle_record(S, ~rn, ray, ha.t, beta, gamma, h);
    } else if (GENERIC) {
        // same limit as in extend => bit-identical record
        bool ok;
        if (!TREES || kind == ROOT_FLAT) {  // rect / box, bare or under one TfFacade (tf_facade.rs:41-56)
            const rs_node* n = S.nodes + rn;
            const XformPre* x = nullptr;
            Ray r = ray;
            if (n->type == RS_NODE_TRANSFORM) {
                x = S.xf + n->first_xform;
                r.o = xf_apply(x->inv, ray.o, 1.0);
                r.d = xf_apply(x->inv, ray.d, 0.0);
                n = S.nodes + n->child0;
            }
            if ((TREES || RS_TUNE_QUADRIC_FLAT) && n->type == RS_NODE_QUADRIC) ok = hit_quadric<true>(*n, r, RS_TMIN, end_then, h);
            else ok = n->type == RS_NODE_BOX ? hit_box(*n, r, RS_TMIN, end_then, h)
                                             : hit_rect(n->rect_axis, n->p[0], n->p[1], n->p[2], n->p[3], n->p[4], n->material, r, RS_TMIN, end_then, h);
            if (ok && x) h.point = xf_apply(x->fwd, h.point, 1.0);
        } else {
            object_record(S.nodes, S.xf, S.obj.prog, src.x, src.y, ha.t, ray, want_uv, h);
            ok = true;
        }
        if (!ok) {
            h.point = ray_at(ray, ha.t);
            h.normal = v3(0.0, 1.0, 0.0);
            h.t1 = ha.t; h.t2 = ha.t; h.u = 0.0; h.v = 0.0; h.material = -1; h.outside = true;
        }
    }
}

// camera.rs:94-100
__device__ __forceinline__ double phong_highlight(V3 dir_to_light, V3 ray_dir, V3 normal, int exponent, double factor) {
    V3 reflected = dir_to_light - (2.0 * dot(dir_to_light, normal)) * normal;
    double base = fmax(dot(reflected, -ray_dir), 0.0);
    double specular = pow(base, (double)exponent);
    return specular * factor;
}

// The lobe samplers of ReflectionPdf / BlinnPhongPdf (pdf.rs:124-140, 195-209) loop until the
// sample lies above the surface; when a ray hits the BACK of a shading normal (triangle normals
// are not face-forwarded, triangle_mesh.rs:115-116) the lobe points below it and the reference
// spins (forever if the lobe is entirely below).  Oracle and GPU give up after this many
// candidates and keep the last one (the CPU checker uses the same cap).
#define RS_REJECTION_CAP 256u

struct PdfState {  // the `Box<dyn PDF>` of a ScatterRecord, flattened
    int kind;      // 0 Cosine, 1 Reflection, 2 BlinnPhong, 3 Sphere
    V3 n_w;        // onb_normal.axis[2]
    V3 r_w;        // onb_reflected.axis[2]
    V3 r_in;
    double exponent, k_specular;
};

__device__ __forceinline__ double pdf_value(const PdfState& P, V3 dir) {
    switch (P.kind) {
        case 0: {  // pdf.rs:36-43
            double c = dot(dir, P.n_w);
            return c < 0.0 ? 0.0 : c * RS_INV_PI;
        }
        case 1: {  // pdf.rs:114-122
            double v = dot(dir, P.r_w) * RS_INV_PI;
            return v < 0.0 ? 0.0 : v;
        }
        case 2: {  // pdf.rs:177-192
            double cosine = dot(dir, P.n_w);
            V3 random_normal = unit_fast((-P.r_in) + dir);
            double cosine_specular = fmax(dot(random_normal, P.n_w), 0.0);
            double normal_pdf = (P.exponent + 1.0) / (2.0 * RS_PI) * pow(cosine_specular, P.exponent);
            return fmax(cosine * RS_INV_PI, 0.0) * (1.0 - P.k_specular) + normal_pdf / (4.0 * dot(P.r_in * -1.0, random_normal)) * P.k_specular;
        }
        default: return 1.0 / (4.0 * RS_PI);  // pdf.rs:226-228
    }
}

__device__ __noinline__ V3 pdf_generate(const PdfState& P, V3 normal, V3 reflected, const RngCtx& rng) {
    switch (P.kind) {
        case 0: {  // pdf.rs:45-47
            uint4 w = rng_words(rng, P_BRDF, 0);
            ONB onb = onb_build_fast(normal);
            return onb_local(onb, cosine_direction(u01(w.x), u01(w.y)));
        }
        case 1: {  // pdf.rs:124-140
            ONB onb = onb_build_fast(reflected);
            V3 dir = v3(0.0, 0.0, 0.0);
            for (uint32_t k = 0; k < RS_REJECTION_CAP; ++k) {
                double r1, xi2;
                rng_pair(rng, P_BRDF, 0, k, r1, xi2);
                dir = onb_local(onb, cosine_direction_exponent(P.exponent, r1, xi2));
                if (dot(dir, P.n_w) > 0.0) break;
            }
            return dir;
        }
        case 2: {  // pdf.rs:195-209
            uint4 w = rng_words(rng, P_BRDF, 0);
            if (u01(w.x) < P.k_specular) {
                ONB onb = onb_build_fast(reflected);
                V3 dir = v3(0.0, 0.0, 0.0);
                for (uint32_t k = 0; k < RS_REJECTION_CAP; ++k) {
                    double r1, xi2;
                    rng_pair(rng, P_BRDF, 1, k, r1, xi2);
                    dir = onb_local(onb, cosine_direction_exponent(P.exponent, r1, xi2));
                    if (dot(dir, P.n_w) > 0.0) break;
                }
                return dir;
            }
            ONB onb = onb_build_fast(normal);
            return onb_local(onb, cosine_direction(u01(w.y), u01(w.z)));
        }
        default: {  // vec3.rs:91-96
            uint4 w = rng_words(rng, P_BRDF, 0);
            double z = -1.0 + u01(w.y) * 2.0;
            double r = sqrt(1.0 - z * z);
            double s, c;
            sincospi(2.0 * u01(w.x), &s, &c);  // a = U(0, 2 pi)
            return v3(r * c, r * s, z);
        }
    }
}

__device__ __forceinline__ void dbg_store3(double* p, uint32_t i, V3 v) {
    if (p) { p[3 * (size_t)i] = v.x; p[3 * (size_t)i + 1] = v.y; p[3 * (size_t)i + 2] = v.z; }
}

// SET selects the work classes a launch takes: 0 = all of them, 1 = the hits of Lambertian
// surfaces only (most hits of most scenes: a kernel that holds nothing but that one material
// fits its 128 registers without a single spill, which the all-material kernel does not; measured:
// shade -8 %; capping it at 96 / 85 / 64 registers for a third, fourth or fifth resident block
// spills again and is slower than two blocks), 2 = everything else.  Lean scenes launch 1 then 2,
// generic scenes 0.
#define RS_SHADE_IN_SET(SET, c) ((SET) == 0 || (((c) == 1 + RS_MAT_LAMBERTIAN || (c) == RS_FRONT_CLASS(1 + RS_MAT_LAMBERTIAN)) == ((SET) == 1)))
// TIER: 0 = lean scenes (spheres, triangles; staged pipeline), 1 = + ROOT_FLAT objects, lights are plain
// spheres / rects (no expression-tree code), 2 = everything
template <int TIER, int SET>
__global__ void __launch_bounds__(RS_SHADE_STAGED(TIER) ? RS_TUNE_SHADE_BLOCK : RS_BLOCK, TIER ? RS_TUNE_SHADE_GENERIC : (SET == 1 ? RS_TUNE_SHADE_LAMBERT : RS_TUNE_SHADE_LEAN)) shade_kernel(const __grid_constant__ DScene S, const __grid_constant__ FrameParams fp,
                                                            const __grid_constant__ WaveState ws, const int bounce,
                                                            const __grid_constant__ DebugOut dbg, const int has_dbg) {
    constexpr bool GENERIC = TIER >= 1, TREES = TIER >= 2;
    extern __shared__ __align__(16) uint4 s_stage[];
    const PathBuf& in = ws.buf[bounce & 1];
    const PathBuf& out = ws.buf[(bounce + 1) & 1];
    const uint32_t* class_count = ws.ctr->class_count[bounce];
    uint32_t* head = SET == 2 ? &ws.ctr->shade_head2[bounce] : &ws.ctr->shade_head[bounce];
    uint32_t* out_front = &ws.ctr->n_rays[bounce + 1];
    uint32_t* out_back = &ws.ctr->n_back[bounce + 1];
    const bool last_bounce = bounce + 1 >= fp.depth;
    const uint32_t lane = lane_id();
    const uint32_t lane_zero = RS_LANE_ZERO(ws);
    unsigned char* const warp_smem = reinterpret_cast<unsigned char*>(s_stage) + (size_t)(threadIdx.x >> 5) * RS_SHADE_WARP_SMEM;
    uint4* const stage = reinterpret_cast<uint4*>(warp_smem) + lane;
    uint2* const idxq = reinterpret_cast<uint2*>(warp_smem + 2 * RS_SHADE_STAGE * 32 * 16) + lane;                   // [2][32]
    uint32_t* const outq = reinterpret_cast<uint32_t*>(warp_smem + 2 * RS_SHADE_STAGE * 32 * 16 + 2 * 32 * 8) + lane;  // [2][32]
    uint32_t* const pidx = reinterpret_cast<uint32_t*>(warp_smem + 2 * RS_SHADE_STAGE * 32 * 16 + 2 * 32 * 8 + 2 * 32 * 4) + lane;  // [2][32]

    // virtual index space: each class padded to a multiple of 32 so a warp never mixes classes.
    // The table is spread over the lanes of the warp (lane c holds class c: two registers instead
    // of 2 * RS_NUM_CLASSES): padded begins by a warp scan, lane RS_NUM_CLASSES holds the total.
    uint32_t my_count = (lane < RS_NUM_CLASSES && RS_SHADE_IN_SET(SET, lane)) ? class_count[lane * RS_COUNTER_STRIDE] : 0u;
    uint32_t my_begin = (my_count + 31u) & ~31u;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, my_begin, d);
        if ((int)lane >= d) my_begin += v;
    }
    my_begin = __shfl_up_sync(0xffffffffu, my_begin, 1);  // inclusive -> exclusive
    if (lane == 0) my_begin = 0;
    const uint32_t total = __shfl_sync(0xffffffffu, my_begin, RS_NUM_CLASSES);

    // batch -> (class, is this lane's entry real, its address in the class queue)
    auto classify = [&](uint32_t base, int& cls, bool& live, const uint2*& entry) {
        // highest class whose padded range starts at or before `base` (empty classes share their
        // begin with the next one, so the highest index is the non-empty one)
        const uint32_t m = __ballot_sync(0xffffffffu, lane < RS_NUM_CLASSES && my_begin <= base);
        cls = 31 - __clz((int)(m | 1u));
        const uint32_t cb = __shfl_sync(0xffffffffu, my_begin, cls), cn = __shfl_sync(0xffffffffu, my_count, cls);
        const uint32_t qi = base - cb + lane;
        live = base < total && qi < cn;
        entry = ws.class_queue + ((size_t)cls * ws.capacity + qi);
    };
#if RS_TUNE_SHADE_STATIC
    // static round robin (measured slower than the cursor: kept for experiments)
    (void)head;
    uint32_t next_base = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 32u;
    const uint32_t base_stride = gridDim.x * (blockDim.x >> 5) * 32u;
    auto next_batch = [&]() -> uint32_t {
        const uint32_t b = next_base;
        next_base = b < total ? b + base_stride : b;  // saturates: no wrap-around past `total`
        return b;
    };
#else
    // Lean kernel: one atomicAdd on the cursor per RS_TUNE_SHADE_CHUNK batches.  At 128 registers ptxas spills
    // a cursor value that is fetched a batch ahead, and the spill store waits for the atomic on the
    // spot (the top stall of this kernel in the profiles) — so the round trip is paid where the
    // value is needed, but only once per chunk.
    // The generic kernel keeps the cursor value fetched a batch ahead (it has no staging pipeline
    // to feed, and measured faster that way).
    constexpr uint32_t CHUNK = GENERIC ? 1u : (uint32_t)RS_TUNE_SHADE_CHUNK;
    // Hybrid distribution (lean kernel): the first RS_TUNE_SHADE_HYBRID percent of the batches are
    // dealt round robin (batch = warp + k * warps in the grid: no atomic, no round trip to L2 on the
    // critical path of this latency-bound kernel), the rest through the cursor so that the tail
    // still balances.
    constexpr uint32_t HYBRID = GENERIC ? 0u : (uint32_t)RS_TUNE_SHADE_HYBRID;
    const uint32_t n_warps_grid = gridDim.x * (blockDim.x >> 5);
    uint32_t static_left = HYBRID ? (uint32_t)(((unsigned long long)(total >> 5) * HYBRID / 100u) / n_warps_grid) : 0u;
    const uint32_t dynamic_begin = static_left * n_warps_grid * 32u;
    uint32_t static_next = ((threadIdx.x >> 5) * gridDim.x + blockIdx.x) * 32u;
    uint32_t chunk_base = 0, chunk_left = 0, pend = 0;
    if (CHUNK == 1 && lane == 0) pend = atomic_add_async(head, 32u, lane_zero);
    auto next_batch = [&]() -> uint32_t {
        if constexpr (CHUNK == 1) {
            const uint32_t b = __shfl_sync(0xffffffffu, pend, 0);
            if (lane == 0) pend = atomic_add_async(head, 32u, lane_zero);
            return b;
        }
        if constexpr (HYBRID != 0) {
            if (static_left) {
                const uint32_t b = static_next;
                static_next += n_warps_grid * 32u;
                --static_left;
                return b;
            }
        }
        if (chunk_left == 0) {
            uint32_t v = 0;
            if (lane == 0) v = atomic_add_async(head, 32u * CHUNK, lane_zero) + dynamic_begin;
            chunk_base = __shfl_sync(0xffffffffu, v, 0);
            chunk_left = CHUNK;
        }
        const uint32_t b = chunk_base;
        chunk_base += 32u;
        --chunk_left;
        return b;
    };
#endif
    auto stage_item = [&](int buf, uint2 entry) {  // entry = (path index, output slot reserved by extend)
        const uint32_t i = RS_IN_BOUNDS(entry.x < ws.capacity) ? entry.x : 0u;
        outq[buf * 32] = entry.y;
        if constexpr (GENERIC) pidx[buf * 32] = i;
        uint4* s = stage + buf * (RS_SHADE_STAGE * 32);
        const uint4* r = in.rec + 4 * (size_t)i;
        cp_async16(s, r);
        cp_async16(s + 32, r + 1);
        cp_async16(s + 64, r + 2);
        cp_async16(s + 96, r + 3);
        cp_async16(s + 128, ws.hit + i);
    };

    // Lean scenes (STAGED): three-deep software pipeline over the batches a warp fetches (one
    // atomic each), entirely through cp.async so that no register is held across the latency:
    //   batch k   : records staged in shared memory -> shade
    //   batch k+1 : its queue entries arrived in shared memory -> issue the cp.async of its records
    //   batch k+2 : its cursor value arrived -> issue the cp.async of its queue entries
    // Generic scenes keep the plain loads: their kernel is register-bound (expression-tree
    // interpreter) and the extra pipeline state costs more in spills than the latency it hides
    // (measured on the Cornell box).
    constexpr bool STAGED = RS_SHADE_STAGED(TIER);
    uint32_t base0 = 0, base1 = 0;
    int cls0 = 0, cls1 = 0;
    bool live0 = false, live1 = false;
    int buf = 0;
    if constexpr (STAGED) {
        base0 = next_batch();
        base1 = next_batch();
        const uint2* entry;
        classify(base0, cls0, live0, entry);
        if (live0) stage_item(0, *entry);
        classify(base1, cls1, live1, entry);
        if (live1) cp_async8(idxq + 32, entry);
        cp_async_commit();
    }
    while (true) {
        int cls;
        bool live;
        uint32_t i = 0, out_pos = 0xffffffffu;
        uint32_t base2 = 0;
        if constexpr (STAGED) {
            if (base0 >= total) break;
            base2 = next_batch();
            cls = cls0;
            live = live0;
        } else {
            base0 = next_batch();
            if (base0 >= total) break;
            const uint2* entry;
            classify(base0, cls, live, entry);
            if (live) {
                const uint2 e = *entry;
                i = e.x;
                out_pos = e.y;
            }
        }
        cls = RS_BASE_CLASS(cls);  // the (class, branch) bucket -> the material class
        const bool never = cls == 0 || cls == 1 + RS_MAT_DIFFUSE_LIGHT || last_bounce;  // miss / light never continue

        const uint4* s = stage + buf * (RS_SHADE_STAGE * 32);
        int cls2 = 0;
        bool live2 = false;
        if constexpr (STAGED) {
            cp_async_wait_all();  // records of batch k (stage[buf]) and queue entries of k+1 (idxq[buf ^ 1])
            if (live1) stage_item(buf ^ 1, idxq[(buf ^ 1) * 32]);
            const uint2* entry;
            classify(base2, cls2, live2, entry);
            if (live2) cp_async8(idxq + buf * 32, entry);
            cp_async_commit();
        }

        bool survive = false, front = false;
        PathRec p;
        HitA ha;
        RngCtx rng;
        rng.key = &fp.rk;
        rng.pixel = 0;
        rng.sample = 0;
        rng.depth = (uint32_t)bounce + 1;
        if (live) {
            if constexpr (STAGED) {
                const uint4 hh = s[128];
                unpack_path(fp, s[0], s[32], s[64], s[96], p);
                ha.t = __hiloint2double((int)hh.y, (int)hh.x);
                ha.root = (int)hh.z;
                ha.matinfo = hh.w;
                if constexpr (GENERIC) {
                    i = pidx[buf * 32];
                    if (ws.uses_time) p.ray.time = in.time[i];
                }
            } else {
                load_path(fp, ws, in, i, p);
                ha = ws.hit[RS_IN_BOUNDS(i < ws.capacity) ? i : 0u];
            }
            (void)RS_IN_BOUNDS(p.slot < ws.capacity && (cls == 0 || (ha.root >= 0 && ha.root < S.n_roots)));
            rng.pixel = p.pixel;
            rng.sample = p.sample;
        }
        if (live) {
            float3 add = make_float3(0.f, 0.f, 0.f);  // radiance added at this vertex (before T)
            int kind = 0;
            V3 next_o = v3(0, 0, 0), next_d = v3(0, 0, 0);
            double w_r = 0.0, w_g = 0.0, w_b = 0.0;

            if (cls == 0) {
                // miss: world.background(ray) (camera.rs:253-254)
                float4 bg = background(S, p.ray.d);
                add = make_float3(bg.x, bg.y, bg.z);
            } else {
                const MatHot& mat = S.mat_hot[ha.matinfo & 0xffffu];
                const MatHot& smat = S.mat_hot[ha.matinfo >> 16];
                const int mat_type = SET == 1 ? (int)RS_MAT_LAMBERTIAN : RS_MATHOT_TYPE(mat);
                const bool want_uv = RS_MATHOT_NEEDS_UV(mat);
                HitRec h;
                int2 src = make_int2(0, 0);
                if constexpr (TREES) { if (S.root_kind[ha.root] == ROOT_GENERIC) src = ws.hit_src[i]; }
                rebuild_hit<GENERIC, TREES>(S, p.ray, ha, src, want_uv, h);

                const float4 col = material_color(S, mat, h.u, h.v, h.point);

                kind = 1;
                bool scattered = false, specular = false;
                PdfState P;
                P.kind = 0;
                V3 reflected = v3(0, 0, 0);
                switch (mat_type) {
                    case RS_MAT_DIFFUSE_LIGHT: {  // light.rs:33-35: emitted = colour * multiplier, no scatter
                        // `emitted` is asked of the hit's own material (camera.rs:172-174): a light that was
                        // selected THROUGH a MixedMaterial emits nothing (mixed_material.rs:53-55)
                        if ((ha.matinfo >> 16) == (ha.matinfo & 0xffffu))
                            add = make_float3((float)((double)col.x * mat.a), (float)((double)col.y * mat.a), (float)((double)col.z * mat.a));
                        break;
                    }
                    case RS_MAT_LAMBERTIAN: {  // lambertian.rs:39-50
                        scattered = true;
                        P.kind = 0;
                        break;
                    }
                    case RS_MAT_METAL:            // metal.rs:104-118
                    case RS_MAT_DIFFUSE_METAL: {  // metal.rs:54-68
                        reflected = p.ray.d - (2.0 * dot(p.ray.d, h.normal)) * h.normal;
                        if (dot(reflected, h.normal) > 0.0) {
                            scattered = true;
                            if (mat_type == RS_MAT_METAL) {
                                specular = true;
                                next_o = h.point;
                                next_d = reflected;
                            } else {
                                P.kind = 1;
                                P.exponent = mat.a;
                            }
                        }
                        break;
                    }
                    case RS_MAT_DIELECTRIC: {  // dielectric.rs:55-93
                        scattered = true;
                        specular = true;
                        next_o = h.point;
                        double cos_theta = dot(-p.ray.d, h.normal);
                        double sin_theta = sqrt(1.0 - cos_theta * cos_theta);
                        double refractive = h.outside ? mat.b : mat.a;
                        bool refracted = false;
                        if (!(refractive * sin_theta > 1.0)) {
                            double reflect_prob = 0.0;
                            if (RS_MATHOT_HAS_CURVE(mat)) {  // Glass::reflect_prob (dielectric.rs:20-24)
                                double r0 = (1.0 - refractive) / (1.0 + refractive);
                                r0 = r0 * r0;
                                double x = 1.0 - cos_theta;
                                reflect_prob = fma(1.0 - r0, x * x * x * x * x, r0);
                            }
                            const uint4 w = rng_words(rng, P_SCATTER, 0);
                            if (!(u01(w.w) < reflect_prob)) {
                                V3 r_parallel = refractive * (p.ray.d + cos_theta * h.normal);
                                V3 r_perp = (-sqrt(1.0 - length_squared(r_parallel))) * h.normal;
                                next_d = r_parallel + r_perp;
                                refracted = true;
                            }
                        }
                        if (!refracted) next_d = p.ray.d - (2.0 * dot(p.ray.d, h.normal)) * h.normal;
                        break;
                    }
                    case RS_MAT_ISOTROPIC: {  // isotropic.rs:26-33
                        scattered = true;
                        P.kind = 3;
                        break;
                    }
                    case RS_MAT_BLINN_PHONG: {  // blinn_phong.rs:32-42
                        scattered = true;
                        P.kind = 2;
                        P.k_specular = mat.a;
                        P.exponent = mat.b;
                        reflected = p.ray.d - (2.0 * dot(p.ray.d, h.normal)) * h.normal;  // vec3.rs:171-173
                        break;
                    }
                    default: break;
                }

                if (scattered && specular) {
                    // skip_pdf: emitted + srec.color * ray_color(srec.ray) (camera.rs:178-182)
                    kind = 2;
                    w_r = (double)col.x; w_g = (double)col.y; w_b = (double)col.z;
                } else if (scattered) {
                    P.n_w = unit_fast(h.normal);
                    P.r_in = p.ray.d;
                    P.r_w = (P.kind == 1 || P.kind == 2) ? unit_fast(reflected) : v3(0, 0, 0);
                    double light_multi = 1.0, pdf_val;
                    const uint4 w = rng_words(rng, P_SCATTER, 0);
                    const bool mis = (fp.flags & RS_FLAG_MIS) != 0;
                    if (u01(w.x) < 0.5) {  // camera.rs:198-213
                        kind = 3;
                        pdf_val = 0.3183098861837907;
                        const uint32_t light_k = S.n_lights == 1 ? 0u : w.y % (uint32_t)S.n_lights;  // list.rs:49-52
                        V3 to_light;
                        if (TREES) {
                            to_light = light_random(S.nodes, S.xf, S.lights[light_k], h.point, rng);
                        } else {
                            const rs_node& ln = S.light_nodes[light_k];
                            to_light = ln.type == RS_NODE_SPHERE ? sphere_random(ln, h.point, rng) : rect_random(ln, h.point, rng);
                        }
                        V3 dir_to_light = unit_fast(to_light);
                        if (!mis && smat.phong_factor > 0.0)
                            light_multi += phong_highlight(-dir_to_light, p.ray.d, h.normal, smat.phong_exponent, smat.phong_factor);
                        next_o = mis ? h.point : ray_at(p.ray, h.t1 - 0.0002);
                        next_d = dir_to_light;
                    } else {  // camera.rs:214-218
                        kind = 4;
                        if constexpr (SET == 1) {  // CosinePdf::generate (pdf.rs:45-47), inlined: the only PDF of this kernel
                            const uint4 wb = rng_words(rng, P_BRDF, 0);
                            const ONB onb = onb_build_fast(h.normal);
                            next_d = onb_local(onb, cosine_direction(u01(wb.x), u01(wb.y)));
                        } else {
                            next_d = pdf_generate(P, h.normal, reflected, rng);
                        }
                        pdf_val = pdf_value(P, next_d);
                        next_o = h.point;
                    }
                    if (mis) {
                        // RS_FLAG_MIS: the integrator the reference keeps commented out (camera.rs:188-192,
                        // 222-225): density of MixturePdf(HittablePdf, srec.pdf) (pdf.rs:237-307), where
                        // HittablePdf::value is a CosinePdf about ITS normal, and that normal is the sampled direction
                        const double c = dot(next_d, unit_fast(next_d));
                        pdf_val = 0.5 * (c < 0.0 ? 0.0 : c * RS_INV_PI) + 0.5 * pdf_value(P, next_d);
                    }
                    // camera.rs:236-243: a pdf value that is <= 0 or NaN is replaced by 1e-5 and the
                    // (unchanged) scattering value is divided by it: 0, a large negative number or NaN
                    const bool degenerate = pdf_val <= 0.0 || pdf_val != pdf_val;
                    if (degenerate) pdf_val = 1e-5;
                    // BRDF branch with a regular value: the two values are the same number, x / x == 1 exactly
                    const double ratio = (kind == 4 && !mis && !degenerate && pdf_val < RS_F64_MAX) ? 1.0 : pdf_value(P, next_d) / pdf_val;
                    const double mult = light_multi * ratio;
                    w_r = (double)col.x * mult; w_g = (double)col.y * mult; w_b = (double)col.z * mult;
                }
            }

            if (add.x != 0.f || add.y != 0.f || add.z != 0.f) {
                // No material of the reference both emits and scatters (DiffuseLight never scatters,
                // material/light.rs; nothing else overrides `emitted`), so a path adds radiance exactly
                // once — when it ends on a light or leaves the scene: a plain 128-bit store, no RMW.
                if (RS_IN_BOUNDS(p.slot < ws.capacity)) ws.contrib[p.slot] = make_float4(p.T.x * add.x, p.T.y * add.y, p.T.z * add.z, 0.f);
            }
            if (has_dbg) {
                if (dbg.kind) dbg.kind[p.pixel] = kind;
                if (dbg.next_od6) {
                    double* o = dbg.next_od6 + 6 * (size_t)p.pixel;
                    o[0] = next_o.x; o[1] = next_o.y; o[2] = next_o.z; o[3] = next_d.x; o[4] = next_d.y; o[5] = next_d.z;
                }
                dbg_store3(dbg.weight3, p.pixel, v3(w_r, w_g, w_b));
                dbg_store3(dbg.emitted3, p.pixel, v3((double)add.x, (double)add.y, (double)add.z));
            }
            if (kind >= 2 && !last_bounce) {
                survive = true;
                front = kind == 3;
                p.ray.o = next_o;
                p.ray.d = next_d;
                p.T.x *= (float)w_r; p.T.y *= (float)w_g; p.T.z *= (float)w_b;
            }
        }
        if (!never) {
            if constexpr (STAGED) out_pos = outq[buf * 32];
            if (__any_sync(0xffffffffu, live && out_pos != 0xffffffffu)) {
                // the slots were reserved by extend (see there; the same for every hit of the class):
                // no atomic, no ballot
                if (survive) store_path(ws, out, out_pos, p);
            } else {
                // survivors: light-directed rays onto the front stack, the others onto the back stack
                const uint32_t mf = __ballot_sync(0xffffffffu, survive && front), mb = __ballot_sync(0xffffffffu, survive && !front);
                const uint32_t lf = mf ? (uint32_t)(__ffs(mf) - 1) : 0u, lb = mb ? (uint32_t)(__ffs(mb) - 1) : 0u;
                uint32_t bf = 0, bb = 0;
                if (mf && lane == lf) bf = atomic_add_async(out_front, __popc(mf), lane_zero);
                if (mb && lane == lb) bb = atomic_add_async(out_back, __popc(mb), lane_zero);
                bf = __shfl_sync(0xffffffffu, bf, lf);
                bb = __shfl_sync(0xffffffffu, bb, lb);
                const uint32_t lt = (1u << lane) - 1u;
                if (survive) store_path(ws, out, front ? bf + __popc(mf & lt) : ws.capacity - 1u - (bb + __popc(mb & lt)), p);
            }
        }
        if constexpr (STAGED) {  // rotate the pipeline
            base0 = base1; cls0 = cls1; live0 = live1;
            base1 = base2; cls1 = cls2; live1 = live2;
            buf ^= 1;
        }
    }
}

// =============================================================================================
// parity probe: primary hit of every pixel (no shading)
// =============================================================================================
__global__ void __launch_bounds__(RS_BLOCK) probe_primary_kernel(const __grid_constant__ DScene S, const __grid_constant__ FrameParams fp,
                                                                 const __grid_constant__ WaveState ws, const __grid_constant__ DebugOut dbg) {
    const uint32_t n = ws.ctr->n_rays[0];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        PathRec p;
        load_path(fp, ws, ws.buf[0], i, p);
        const HitA ha = ws.hit[i];
        const uint32_t px = p.pixel;
        if (dbg.ray_od6) {
            double* o = dbg.ray_od6 + 6 * (size_t)px;
            o[0] = p.ray.o.x; o[1] = p.ray.o.y; o[2] = p.ray.o.z; o[3] = p.ray.d.x; o[4] = p.ray.d.y; o[5] = p.ray.d.z;
        }
        if (dbg.obj_id) dbg.obj_id[px] = ha.root;
        if (ha.root < 0) {
            if (dbg.t) dbg.t[px] = 0.0;
            dbg_store3(dbg.normal3, px, v3(0, 0, 0));
            dbg_store3(dbg.point3, px, v3(0, 0, 0));
            if (dbg.front) dbg.front[px] = 0;
            if (dbg.material) dbg.material[px] = -1;
            continue;
        }
        RngCtx rng;
        rng.key = &fp.rk; rng.pixel = p.pixel; rng.sample = p.sample; rng.depth = 1;
        HitRec h;
        const int2 src = S.root_kind[ha.root] == ROOT_GENERIC ? ws.hit_src[i] : make_int2(0, 0);
        rebuild_hit<true, true>(S, p.ray, ha, src, true, h);
        if (dbg.t) dbg.t[px] = h.t1;
        dbg_store3(dbg.normal3, px, h.normal);
        dbg_store3(dbg.point3, px, h.point);
        if (dbg.front) dbg.front[px] = h.outside ? 1 : 0;
        if (dbg.material) dbg.material[px] = (int)(ha.matinfo >> 16);
    }
}

// =============================================================================================
// accumulate / finalize / passes
// =============================================================================================
// One thread per pixel of the wave; contrib is sample-major so consecutive threads read
// consecutive float4s (coalesced 128-bit loads), and the read-modify-write of `accum` is a
// float4 too.  No atomics: the sum order is fixed, the image is reproducible.
__global__ void __launch_bounds__(RS_BLOCK) accumulate_kernel(const __grid_constant__ FrameParams fp, const __grid_constant__ WaveState ws,
                                                              float4* __restrict__ accum) {
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < fp.n_pixels; p += gridDim.x * blockDim.x) {
        const uint32_t pixel = fp.pixel0 + p;
        if (fp.mask && fp.mask[pixel] == 0) continue;
        // the reference sums `color_vec` in f64 (painter.rs:163-176): the samples of a wave are summed in
        // f64 and rounded once into the f32 accumulator
        double sx = 0.0, sy = 0.0, sz = 0.0;
        for (uint32_t k = 0; k < fp.n_samples; ++k) {
            const float4 c = ws.contrib[(size_t)k * fp.n_pixels + p];
            sx += (double)c.x; sy += (double)c.y; sz += (double)c.z;
        }
        float4 a = accum[pixel];
        a.x = (float)((double)a.x + sx); a.y = (float)((double)a.y + sy); a.z = (float)((double)a.z + sz); a.w += (float)fp.n_samples;
        accum[pixel] = a;
    }
}

// Vec3::into_color (vec3.rs:227-240) + the [0,0,0,0] of masked pixels (painter.rs:208-210)
__global__ void finalize_kernel(const float4* __restrict__ accum, const unsigned char* __restrict__ mask, float4* __restrict__ out,
                                int64_t n, float inv_n, int gamma) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (mask && mask[i] == 0) { out[i] = make_float4(0.f, 0.f, 0.f, 0.f); continue; }
        float4 a = accum[i];
        float r = a.x * inv_n, g = a.y * inv_n, b = a.z * inv_n;
        if (gamma) { r = sqrtf(r); g = sqrtf(g); b = sqrtf(b); }
        out[i] = make_float4(r, g, b, 1.0f);
    }
}

// Fused multi-GPU reduce + finalize: the sum over the per-GPU accumulators is taken straight
// from peer memory (128-bit loads over NVLink / NVSwitch, one per peer per pixel, all issued
// before the first use) and the finished pixel is stored once — possibly into a peer's image.
// Replaces ncclReduce (which writes a reduced copy that finalize would read again).
__global__ void __launch_bounds__(RS_BLOCK) finalize_peers_kernel(const __grid_constant__ PeerAccums peers, const unsigned char* __restrict__ mask,
                                                                  float4* __restrict__ out, int64_t begin, int64_t end, float inv_n, int gamma) {
    for (int64_t i = begin + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < end; i += (int64_t)gridDim.x * blockDim.x) {
        if (mask && mask[i] == 0) { out[i] = make_float4(0.f, 0.f, 0.f, 0.f); continue; }
        float4 v[16];
#pragma unroll
        for (int k = 0; k < 16; ++k)
            if (k < peers.n) v[k] = __ldcv(peers.p[k] + i);  // volatile-class load: never served from a stale L1 line
        float sx = 0.f, sy = 0.f, sz = 0.f;
#pragma unroll
        for (int k = 0; k < 16; ++k)
            if (k < peers.n) { sx += v[k].x; sy += v[k].y; sz += v[k].z; }
        float r = sx * inv_n, g = sy * inv_n, b = sz * inv_n;
        if (gamma) { r = sqrtf(r); g = sqrtf(g); b = sqrtf(b); }
        out[i] = make_float4(r, g, b, 1.0f);
    }
}

// Device-side arrival counters in peer memory (multi-GPU combine without a host barrier).
// signal: everything this stream did before is visible system-wide, then counter += 1.
__global__ void signal_kernel(uint32_t* counter) {
    __threadfence_system();
    asm volatile("red.release.sys.global.add.u32 [%0], 1;\n" ::"l"(counter) : "memory");
}
// wait: ONE warp polls (ld.acquire.sys) until the counter reaches `expect`; the kernels queued behind it on
// the stream then see what the signalling devices wrote.  (The poll sits in its own one-warp kernel rather than
// at the top of the reduce kernel: a whole grid of spinning blocks could keep co-resident work — another
// context's render on the same device — from ever running.)
// budget_ns > 0 bounds the poll (RS_OPT_WAIT_BUDGET_MS): a peer that never signals — its host failed between two
// launches — must not hang this device for good.  When the budget runs out the kernel raises *gave_up and ends; the
// kernels queued behind it then run on incomplete data, and the host, which checks the flag once the stream has
// drained (rs_multi_render, rs_wait_check), reports the frame as failed.
__global__ void wait_counter_kernel(const uint32_t* counter, uint32_t expect, unsigned long long budget_ns, uint32_t* gave_up) {
    if (threadIdx.x == 0) {
        uint32_t v;
        unsigned long long t0 = 0, now;
        if (budget_ns) asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t0));
        while (true) {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(counter) : "memory");
            if ((int32_t)(v - expect) >= 0) break;
            if (budget_ns) {
                asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(now));
                if (now - t0 > budget_ns) {
                    if (gave_up) asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(gave_up), "r"(1u) : "memory");
                    break;
                }
            }
            __nanosleep(500);
        }
    }
}

// combine_pixel (raysnail.rs:176-193)
__global__ void combine_kernel(const float4* __restrict__ old_px, const float4* __restrict__ new_px, float4* __restrict__ out, int64_t n, float pass) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 o = old_px[i], w = new_px[i];
        if (w.x == 0.0f && w.y == 0.0f && w.z == 0.0f && w.w == 0.0f) { out[i] = o; continue; }
        const float d = pass + 1.0f;
        out[i] = make_float4(__fdiv_rn(__fadd_rn(__fmul_rn(o.x, pass), w.x), d), __fdiv_rn(__fadd_rn(__fmul_rn(o.y, pass), w.y), d),
                             __fdiv_rn(__fadd_rn(__fmul_rn(o.z, pass), w.z), d), __fdiv_rn(__fadd_rn(__fmul_rn(o.w, pass), w.w), d));
    }
}

// calc_noise (raysnail.rs:138-173); quirk keeps its `let x = y as i32;`
__global__ void noise_kernel(const float4* __restrict__ px, float* __restrict__ noise, int w, int h, int quirk) {
    const int64_t n = (int64_t)w * h;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int y0 = (int)(i / w), x0 = (int)(i % w);
        const float4 def = px[i];
        const int x = quirk ? y0 : x0, y = y0;
        float diff = 0.0f;
        for (int yy = y - 2; yy < y + 3; ++yy)
            for (int xx = x - 2; xx < x + 3; ++xx) {
                float4 q = def;
                if (xx >= 0 && yy >= 0 && xx < w && yy < h) q = px[(int64_t)yy * w + xx];
                const float rd = def.x - q.x, gd = def.y - q.y, bd = def.z - q.z;
                diff = __fadd_rn(diff, __fadd_rn(__fadd_rn(__fmul_rn(rd, rd), __fmul_rn(gd, gd)), __fmul_rn(bd, bd)));
            }
        noise[i] = diff;
    }
}

// the redo map of the passes loop (raysnail.rs:408-423): 1 where noise >= t; *count += number of ones
__global__ void __launch_bounds__(RS_BLOCK) redo_mask_kernel(const float* __restrict__ noise, unsigned char* __restrict__ mask, int64_t n, float t,
                                                             unsigned long long* __restrict__ count) {
    unsigned int mine = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const bool redo = noise[i] >= t;
        mask[i] = redo ? 1 : 0;
        mine += redo ? 1u : 0u;
    }
    mine = __reduce_add_sync(0xffffffffu, mine);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(count, (unsigned long long)mine);
}

__global__ void fill_kernel(float4* __restrict__ p, int64_t n, float4 v) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}

// (clamp(c,0,1) * 255.5) as u8 (raysnail.rs:437-439)
__device__ __forceinline__ unsigned char quant(float c) {
    double v = (double)c;
    double cl = (0.0 > v) ? 0.0 : (v > 1.0 ? 1.0 : v);
    double s = cl * 255.5;
    if (s != s) return 0;
    return (unsigned char)(s < 0.0 ? 0.0 : (s > 255.0 ? 255.0 : s));
}
__global__ void quantize_kernel(const float4* __restrict__ px, unsigned char* __restrict__ rgb, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 c = px[i];
        rgb[3 * i] = quant(c.x); rgb[3 * i + 1] = quant(c.y); rgb[3 * i + 2] = quant(c.z);
    }
}

// =============================================================================================
// launchers
// =============================================================================================
static int grid_for(int64_t n) {
    int64_t b = (n + RS_BLOCK - 1) / RS_BLOCK;
    if (b < 1) b = 1;
    if (b > 148 * 16) b = 148 * 16;
    return (int)b;
}

void launch_raygen(const FrameParams& fp, const WaveState& ws, cudaStream_t s) {
    raygen_kernel<<<grid_for((int64_t)fp.n_pixels * fp.n_samples), RS_BLOCK, 0, s>>>(fp, ws);
}
template <int MODE, bool WIDE>
static const void* extend_fn() { return (const void*)extend_kernel<MODE, WIDE, false>; }
static const void* extend_first_fn_for(int mode, int stage) {
#if RS_TUNE_STAGE_TREE
    if (stage) return mode == 0 ? (const void*)extend_kernel<0, false, true, true> : (const void*)extend_kernel<1, false, true, true>;
#endif
    return mode == 0 ? (const void*)extend_kernel<0, false, true> : (mode == 1 ? (const void*)extend_kernel<1, false, true>
                     : (mode == 2 ? (const void*)extend_kernel<2, false, true> : (mode == 3 ? (const void*)extend_kernel<3, false, true> : (const void*)extend_kernel<4, false, true>)));
}
static const void* extend_fn_for(int mode, int wide, int stage) {
#if RS_TUNE_STAGE_TREE
    if (stage) return mode == 0 ? (const void*)extend_kernel<0, false, false, true> : (const void*)extend_kernel<1, false, false, true>;
#endif
    if (wide) return mode == 0 ? extend_fn<0, true>() : (mode == 1 ? extend_fn<1, true>() : (mode == 2 ? extend_fn<2, true>() : extend_fn<3, true>()));
    return mode == 0 ? extend_fn<0, false>() : (mode == 1 ? extend_fn<1, false>() : (mode == 2 ? extend_fn<2, false>() : (mode == 3 ? extend_fn<3, false>() : extend_fn<4, false>())));
}
// stage_bytes > 0: the staged kernels (lean modes, binary tree), with that much dynamic shared memory
void launch_extend_first(const DScene& S, const FrameParams& fp, const WaveState& ws, int mode, int grid, int stage_bytes, cudaStream_t s) {
    int bounce = 0, miss_in_extend = 1;
    void* args[] = {(void*)&S, (void*)&fp, (void*)&ws, (void*)&bounce, (void*)&miss_in_extend};
    cudaLaunchKernel(extend_first_fn_for(mode, stage_bytes > 0), dim3(grid), dim3(stage_bytes > 0 ? RS_TUNE_STAGE_BLOCK : RS_BLOCK), args, (size_t)stage_bytes, s);
}
void launch_extend(const DScene& S, const FrameParams& fp, const WaveState& ws, int bounce, int mode, int wide, int miss_in_extend, int grid,
                   int stage_bytes, cudaStream_t s) {
    void* args[] = {(void*)&S, (void*)&fp, (void*)&ws, (void*)&bounce, (void*)&miss_in_extend};
    cudaLaunchKernel(extend_fn_for(mode, wide, stage_bytes > 0), dim3(grid), dim3(stage_bytes > 0 ? RS_TUNE_STAGE_BLOCK : RS_BLOCK), args, (size_t)stage_bytes, s);
}
// resident blocks per SM of the kernel the scene will use; staged kernels are first given their dynamic shared memory
static int extend_occupancy(const void* fn, int stage_bytes) {
    int n = 0;
    if (stage_bytes > 0 && !RS_TUNE_STAGE_TREE) return 0;  // not compiled in
    if (stage_bytes > 0) {
        if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, stage_bytes) != cudaSuccess) { cudaGetLastError(); return 0; }
        cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    }
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, fn, stage_bytes > 0 ? RS_TUNE_STAGE_BLOCK : RS_BLOCK, (size_t)stage_bytes);
    if (stage_bytes > 0 && n < 1024 / RS_TUNE_STAGE_BLOCK) return 0;  // staging must not cost resident warps
    return n;
}
int extend_first_blocks_per_sm(int mode, int stage_bytes) {
    const int n = extend_occupancy(extend_first_fn_for(mode, stage_bytes > 0), stage_bytes);
    return n > 0 ? n : (stage_bytes > 0 ? 0 : 1);
}
int extend_blocks_per_sm(int mode, int wide, int stage_bytes) {
    const int n = extend_occupancy(extend_fn_for(mode, wide, stage_bytes > 0), stage_bytes);
    return n > 0 ? n : (stage_bytes > 0 ? 0 : 1);
}
static const void* extend_stream_fn(int mode) { return mode == 0 ? (const void*)extend_stream_kernel<0> : (const void*)extend_stream_kernel<1>; }
void launch_extend_stream(const DScene& S, const FrameParams& fp, const WaveState& ws, int bounce, int mode, int miss_in_extend, int grid, cudaStream_t s) {
    void* args[] = {(void*)&S, (void*)&fp, (void*)&ws, (void*)&bounce, (void*)&miss_in_extend};
    cudaLaunchKernel(extend_stream_fn(mode), dim3(grid), dim3(RS_BLOCK), args, RS_STREAM_SMEM, s);
}
int extend_stream_blocks_per_sm(int mode) {
    int n = 0;
    cudaFuncSetAttribute(extend_stream_fn(mode), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RS_STREAM_SMEM);
    cudaFuncSetAttribute(extend_stream_fn(mode), cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, extend_stream_fn(mode), RS_BLOCK, RS_STREAM_SMEM);
    return n > 0 ? n : 1;
}
template <int TIER, int SET>
static void launch_shade_t(const DScene& S, const FrameParams& fp, const WaveState& ws, int bounce, int grid, const DebugOut& d, int has_dbg, cudaStream_t s) {
    if constexpr (RS_SHADE_STAGED(TIER)) shade_kernel<TIER, SET><<<grid, RS_TUNE_SHADE_BLOCK, RS_SHADE_SMEM, s>>>(S, fp, ws, bounce, d, has_dbg);
    else shade_kernel<TIER, SET><<<grid, RS_BLOCK, 0, s>>>(S, fp, ws, bounce, d, has_dbg);
}
template <int TIER, int SET>
static int shade_occupancy_t() {
    int n = 0;
    if constexpr (RS_SHADE_STAGED(TIER)) {
        cudaFuncSetAttribute(shade_kernel<TIER, SET>, cudaFuncAttributeMaxDynamicSharedMemorySize, RS_SHADE_SMEM);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, shade_kernel<TIER, SET>, RS_TUNE_SHADE_BLOCK, RS_SHADE_SMEM);
    } else {
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, shade_kernel<TIER, SET>, RS_BLOCK, 0);
    }
    return n > 0 ? n : 1;
}
#define RS_SHADE_DISPATCH(FN, ...)                                                                   \
    do {                                                                                             \
        if (generic == 2) { if (set == 1) return FN<2, 1>(__VA_ARGS__); if (set == 2) return FN<2, 2>(__VA_ARGS__); return FN<2, 0>(__VA_ARGS__); } \
        if (generic == 1) { if (set == 1) return FN<1, 1>(__VA_ARGS__); if (set == 2) return FN<1, 2>(__VA_ARGS__); return FN<1, 0>(__VA_ARGS__); } \
        if (set == 1) return FN<0, 1>(__VA_ARGS__);                                                  \
        if (set == 2) return FN<0, 2>(__VA_ARGS__);                                                  \
        return FN<0, 0>(__VA_ARGS__);                                                                \
    } while (0)
void launch_shade(const DScene& S, const FrameParams& fp, const WaveState& ws, int bounce, int generic, int set, int grid, const DebugOut* dbg, cudaStream_t s) {
    DebugOut d = {};
    if (dbg) d = *dbg;
    const int has_dbg = dbg ? 1 : 0;
    RS_SHADE_DISPATCH(launch_shade_t, S, fp, ws, bounce, grid, d, has_dbg, s);
}
int shade_split_trees() { return RS_TUNE_SPLIT_TREES; }
void launch_probe_primary(const DScene& S, const FrameParams& fp, const WaveState& ws, const DebugOut& dbg, cudaStream_t s) {
    probe_primary_kernel<<<grid_for((int64_t)fp.n_pixels * fp.n_samples), RS_BLOCK, 0, s>>>(S, fp, ws, dbg);
}
void launch_accumulate(const FrameParams& fp, const WaveState& ws, float4* accum, cudaStream_t s) {
    accumulate_kernel<<<grid_for(fp.n_pixels), RS_BLOCK, 0, s>>>(fp, ws, accum);
}
void launch_finalize(const float4* accum, const unsigned char* mask, float4* out, int64_t n, int n_samples, int gamma, cudaStream_t s) {
    finalize_kernel<<<grid_for(n), RS_BLOCK, 0, s>>>(accum, mask, out, n, 1.0f / (float)n_samples, gamma);
}
void launch_finalize_peers(const PeerAccums& peers, const unsigned char* mask, float4* out, int64_t begin, int64_t end, int n_samples, int gamma,
                           cudaStream_t s) {
    if (end <= begin) return;
    finalize_peers_kernel<<<grid_for(end - begin), RS_BLOCK, 0, s>>>(peers, mask, out, begin, end, 1.0f / (float)n_samples, gamma);
}
unsigned long long read_violations() {
#if RS_CHECKED
    unsigned long long v = 0, zero = 0;
    cudaMemcpyFromSymbol(&v, rs_violations, sizeof(v));  // (synchronises with the default stream; checked builds only)
    cudaMemcpyToSymbol(rs_violations, &zero, sizeof(zero));
    return v;
#else
    return 0ull;
#endif
}

void launch_signal(uint32_t* counter, cudaStream_t s) { signal_kernel<<<1, 1, 0, s>>>(counter); }
void launch_wait_counter(const uint32_t* counter, uint32_t expect, cudaStream_t s, unsigned long long budget_ns, uint32_t* gave_up) {
    wait_counter_kernel<<<1, 32, 0, s>>>(counter, expect, budget_ns, gave_up);
}
void launch_combine(const float4* old_px, const float4* new_px, float4* out, int64_t n, float pass, cudaStream_t s) {
    combine_kernel<<<grid_for(n), RS_BLOCK, 0, s>>>(old_px, new_px, out, n, pass);
}
void launch_noise(const float4* px, float* noise, int w, int h, int quirk, cudaStream_t s) {
    noise_kernel<<<grid_for((int64_t)w * h), RS_BLOCK, 0, s>>>(px, noise, w, h, quirk);
}
void launch_redo_mask(const float* noise, unsigned char* mask, int64_t n, float t, unsigned long long* count, cudaStream_t s) {
    redo_mask_kernel<<<grid_for(n), RS_BLOCK, 0, s>>>(noise, mask, n, t, count);
}
void launch_fill(float4* p, int64_t n, float4 v, cudaStream_t s) { fill_kernel<<<grid_for(n), RS_BLOCK, 0, s>>>(p, n, v); }
void launch_quantize(const float4* px, unsigned char* rgb, int64_t n, cudaStream_t s) {
    quantize_kernel<<<grid_for(n), RS_BLOCK, 0, s>>>(px, rgb, n);
}
int shade_blocks_per_sm(int generic, int set) { RS_SHADE_DISPATCH(shade_occupancy_t); }

}  // namespace rsd
