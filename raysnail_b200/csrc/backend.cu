// Host orchestration of the CUDA backend: the C ABI of include/raysnail_b200.h.
//
//   rs_upload_scene   validates the flat scene, pre-multiplies TransformStacks, builds the
//                     world BVH (binned SAH, replaces BVH::new of the reference,
//                     src/hittable/collection/bvh.rs:47-112 — any topology returns the same
//                     closest hit) and copies everything to HBM.
//   rs_render*        drives the wavefront: raygen -> depth x (extend, shade) -> accumulate
//                     per wave of (pixel range x sample range); replaces Painter::draw
//                     (src/painter.rs:256-336).
#include <algorithm>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/raysnail_b200.h"
#include "device/kernels.cuh"

using namespace rsd;

namespace {

thread_local std::string g_create_error;

struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    ~DevBuf() { release(); }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        bytes = 0;
    }
    cudaError_t ensure(size_t n) {
        if (n <= bytes) return cudaSuccess;
        release();
        cudaError_t e = cudaMalloc(&p, n ? n : 16);
        if (e == cudaSuccess) bytes = n;
        return e;
    }
    template <class T>
    T* as() const { return reinterpret_cast<T*>(p); }
};

struct HostBvhNode {  // binary SAH tree: two children with f64 boxes
    double box[2][6];
    int child[2];
};
// one member of the tree under construction: bounds, centroid and object id
struct BuildPrim {
    double lo[3], hi[3], c[3];
    int root;
    int pad;  // (kept zero: the records are compared bytewise by the tree cache of rs_upload_scene)
};
struct HostBvhNode4 {  // collapsed 4-wide tree (what the device traverses)
    double box[4][6];
    int child[4];
    int n;
};

}  // namespace

struct rs_ctx {
    int device = 0;
    int sm_count = 148;
    std::string error;
    bool has_scene = false;

    // scene (device)
    DevBuf d_nodes, d_xf, d_bvh, d_root_node, d_root_kind, d_root_sphere, d_root_bbox, d_root_gate;
    DevBuf d_mat_hot, d_root_inv_r, d_root_material, d_light_nodes;
    DevBuf d_prog, d_cprog, d_csg, d_root_prog, d_root_list;  // object programs (expression-tree roots)
    DevBuf d_tri_geom, d_tri_normals, d_tri_material, d_lights, d_materials, d_textures, d_texels, d_perm, d_values;
    DScene scene{};
    // scene (host side, kept to re-pad the f32 boxes when the camera moves far away)
    std::vector<HostBvhNode> bvh_host;
    std::vector<HostBvhNode4> bvh4_host;
    std::vector<int> root_kind_host, root_node_host;
    // what the kept tree was built from (bounds + object id of every tree member, before the build re-orders them):
    // a re-upload of unchanged geometry — every pass of the passes loop, every frame over a fixed world — keeps the tree
    std::vector<unsigned char> bvh_key;
    int bvh_key_width = 0, bvh_key_root = 0;
    bool bvh_reused = false;
    // the device copy of the tree (f32 boxes padded for ray origins up to tree_dev_omax, leaf references) is kept with it
    bool tree_dev_valid = false;
    double tree_dev_omax = 0.0;
    // host scratch of rs_upload_scene, kept between calls: a re-upload of a 1 M-triangle world fills ~300 MB of tables,
    // and fresh vectors would page-fault and zero-fill all of it every time
    struct UploadScratch {
        std::vector<int> root_kind, root_material;
        std::vector<double> root_sphere, root_inv_r, tri_geom;
        std::vector<unsigned char> root_gate;
        std::vector<BuildPrim> prims, loose, part[8], part_loose[8];
    } scratch;
    struct ListEntry { double box[6]; int root; bool noclip; };
    std::vector<ListEntry> list_host;  // the world objects with a proper bbox, ascending ids
    bool use_list = false;             // small worlds: the kernels walk the object list instead of the tree (RS_OPT_LIST_MAX)
    bool list_trees_uniform = false;   // the expression trees on the list all have the same program shape (closest_hit's skip-ahead walk)
    int list_max = 12;       // RS_OPT_LIST_MAX; -1 = not set: 12, or 24 for worlds that hold expression trees
    bool list_max_set = false;
    int bvh_root = 0;
    int bvh_width = 2;       // RS_OPT_BVH_WIDTH: 2 = binary 64-byte nodes (default), 4 = QBVH 128-byte nodes
    int extend_mode = 3;     // 0 static spheres only, 1 + triangles, 2 + flat roots (rect / box), 3 everything, 4 = 3 with the skip-ahead list walk
    int shade_generic = 2;  // shade tier: 0 = lean, 1 = + flat roots (no expression trees), 2 = everything
    bool uses_time = false;  // some sphere moves: rays carry their shutter time
    double scene_omax = 0.0, padded_omax = -1.0;
    int n_lights = 0;

    // wave state.  Two complete sets: consecutive waves run on two streams so that the tail of
    // every persistent kernel of one wave (SMs going idle as its last batches finish) is filled by
    // blocks of the other wave's kernel (RS_OPT_WAVE_STREAMS; `ws` is set 0 and the one the probes use)
    struct WaveSet {
        DevBuf d_rec[2], d_time[2], d_hit, d_hit_src, d_cq, d_contrib, d_ctr, d_dbg_cost;
        uint32_t capacity = 0;
    };
    static constexpr int kMaxSets = 4;
    WaveSet wset[kMaxSets];
    DevBuf d_stats;
    DevBuf d_wait_flag;  // one uint32, zeroed at rs_create: raised by a wait kernel of rs_wait_counter that ran out of budget (rs_wait_check)
    uint32_t capacity = 0;
    WaveState ws{}, ws_aux[kMaxSets - 1] = {};
    int wait_budget_ms = 600000;  // RS_OPT_WAIT_BUDGET_MS: how long a wait kernel polls before it gives up (0 = for ever)
    int wave_streams = 4;  // (r02: 2 -> 4 streams and 8 M -> 32 M paths per wave: the tail of every launch — a few very long rays —
                           //  overlaps the next wave's kernels; +33 % on the 1 M-triangle mesh, +6 % on the headline scene)
    cudaStream_t aux_stream[kMaxSets - 1] = {};
    cudaEvent_t ev_fork = nullptr, ev_join[kMaxSets - 1] = {}, ev_acc[kMaxSets] = {};
    int extend_grid = 148, shade_grid = 148, shade_grid_lambert = 148;
    int extend_first_grid = 148;
    bool use_stream = false;  // bounces >= 1 run extend_stream_kernel (triangle-mesh scenes)
    int extend_stream_grid = 148;
    int stage_bounces = 1 << 30, extend_grid_global = 148;  // bounces >= stage_bounces run the unstaged kernel
    int stage_bytes = 0;  // > 0: extend keeps the binary tree + sphere table in shared memory (that many bytes per block)
    int fuse_raygen = 1;  // bounce 0 of a render without pixel mask generates its camera rays inside extend

    // scratch for rs_render / probes
    DevBuf d_accum, d_out, d_mask, d_dbg[11];
    DevBuf d_pass_old, d_pass_comb, d_noise, d_redo_count;  // rs_render_passes
    bool stats_final = false;  // `stats` already holds the device counters (rs_render_passes sums them over its passes)

    // timing
    std::vector<cudaEvent_t> events;
    rs_stats stats{};
};

namespace {

int fail(rs_ctx* ctx, int code, const std::string& msg) {
    if (ctx) ctx->error = msg;
    else g_create_error = msg;
    return code;
}

#define CK(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess) return fail(ctx, RS_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__)); \
    } while (0)

template <class T>
int upload(rs_ctx* ctx, DevBuf& buf, const T* src, size_t n) {
    CK(buf.ensure(std::max<size_t>(n, 1) * sizeof(T)));
    // asynchronous on the legacy stream: a pageable source is staged before the call returns (so `src` may be a
    // temporary), the DMA is not waited for — rs_upload_scene issues some thirty of these and synchronises once at its
    // end, instead of paying thirty round trips (0.4 ms of a 6 ms frame on sdl/example.sdl)
    if (n) CK(cudaMemcpyAsync(buf.p, src, n * sizeof(T), cudaMemcpyHostToDevice, 0));
    return RS_OK;
}

// Host loops over a million world objects (the tables of rs_upload_scene): contiguous chunks on up to 8 threads.
// f(chunk, begin, end); returns the number of chunks used (1 for small n: no thread is started).
constexpr int kMaxChunks = 8;
template <class F>
int parallel_chunks(size_t n, F f) {
    int t = 1;
    if (n >= (size_t)1 << 16) t = (int)std::min<unsigned>(kMaxChunks, std::max(1u, std::thread::hardware_concurrency()));
    if (t == 1) { f(0, (size_t)0, n); return 1; }
    std::vector<std::thread> th;
    for (int k = 1; k < t; ++k) th.emplace_back([&, k] { f(k, n * (size_t)k / (size_t)t, n * (size_t)(k + 1) / (size_t)t); });
    f(0, (size_t)0, n / (size_t)t);
    for (auto& x : th) x.join();
    return t;
}

// ---------------------------------------------------------------------------------------------
// BVH build: binned SAH over the reference bounding boxes of the world objects
// ---------------------------------------------------------------------------------------------
// Conservative bound of the SURFACE of object `ni` (all the points a hit can report), or false
// when there is none (a quadric) or none is known.  It is not the reference's bbox: that one is
// restated by the host layer and is what the BVH leaf gates the object with (bvh.rs:173-176);
// a rotated Intersection, for one, reaches outside it.  The tree may drop a box that lies
// beyond the closest hit so far only if the box holds every hit of its objects, so the tree is
// built over (reference bbox ∪ this bound) and objects without a bound stay out of the tree.
bool geom_bound(const std::vector<rs_node>& nodes, const std::vector<XformPre>& xf, int ni, double* lo, double* hi) {
    const rs_node& n = nodes[ni];
    switch (n.type) {
    case RS_NODE_SPHERE:
        if (n.p[4] != 0.0 || n.p[5] != 0.0 || n.p[6] != 0.0) return false;  // moving: the time range is not known here
        for (int a = 0; a < 3; ++a) { lo[a] = n.p[a] - std::fabs(n.p[3]); hi[a] = n.p[a] + std::fabs(n.p[3]); }
        return true;
    case RS_NODE_BOX:
        for (int a = 0; a < 3; ++a) { lo[a] = std::min(n.p[a], n.p[3 + a]); hi[a] = std::max(n.p[a], n.p[3 + a]); }
        return true;
    case RS_NODE_RECT: {
        // rect.rs: plane `k` on the axis the other two span (RS_RECT_XY: a = x, b = y, k = z, ...)
        const int ax[3][3] = {{0, 1, 2}, {0, 2, 1}, {1, 2, 0}};
        const int* m = ax[n.rect_axis];
        lo[m[0]] = std::min(n.p[1], n.p[2]); hi[m[0]] = std::max(n.p[1], n.p[2]);
        lo[m[1]] = std::min(n.p[3], n.p[4]); hi[m[1]] = std::max(n.p[3], n.p[4]);
        lo[m[2]] = n.p[0]; hi[m[2]] = n.p[0];
        return true;
    }
    case RS_NODE_DIFFERENCE:  // every hit lies on or inside the first operand
    case RS_NODE_MEDIUM:      // inside the boundary
        return geom_bound(nodes, xf, n.child0, lo, hi);
    case RS_NODE_INTERSECTION: {
        double l1[3], h1[3];
        const bool b0 = geom_bound(nodes, xf, n.child0, lo, hi), b1 = geom_bound(nodes, xf, n.child1, l1, h1);
        if (!b0 && !b1) return false;
        if (!b0) { std::memcpy(lo, l1, sizeof(l1)); std::memcpy(hi, h1, sizeof(h1)); return true; }
        if (!b1) return true;
        bool empty = false;
        for (int a = 0; a < 3; ++a) empty |= std::max(lo[a], l1[a]) > std::min(hi[a], h1[a]);
        if (!empty)
            for (int a = 0; a < 3; ++a) { lo[a] = std::max(lo[a], l1[a]); hi[a] = std::min(hi[a], h1[a]); }
        return true;  // (disjoint operands: no hits at all; the first operand's bound will do)
    }
    case RS_NODE_TRANSFORM: {
        double l0[3], h0[3];
        if (!geom_bound(nodes, xf, n.child0, l0, h0)) return false;
        const double* F = xf[n.first_xform].fwd;  // pre-multiplied stack, 3x4 row-major
        for (int a = 0; a < 3; ++a) { lo[a] = INFINITY; hi[a] = -INFINITY; }
        for (int c = 0; c < 8; ++c) {
            const double x = (c & 1) ? h0[0] : l0[0], y = (c & 2) ? h0[1] : l0[1], z = (c & 4) ? h0[2] : l0[2];
            for (int a = 0; a < 3; ++a) {
                const double v = F[4 * a] * x + F[4 * a + 1] * y + F[4 * a + 2] * z + F[4 * a + 3];
                lo[a] = std::min(lo[a], v);
                hi[a] = std::max(hi[a], v);
            }
        }
        return true;
    }
    case RS_NODE_BVH: {  // every hit is a hit of a member
        bool any = false;
        for (int c : {n.child0, n.child1}) {
            if (c < 0) continue;
            double l1[3], h1[3];
            if (!geom_bound(nodes, xf, c, l1, h1)) return false;
            for (int a = 0; a < 3; ++a) { lo[a] = any ? std::min(lo[a], l1[a]) : l1[a]; hi[a] = any ? std::max(hi[a], h1[a]) : h1[a]; }
            any = true;
        }
        if (!any) for (int a = 0; a < 3; ++a) { lo[a] = 0.0; hi[a] = 0.0; }
        return true;
    }
    case RS_NODE_RAYMARCHER:
        // a reported point is within one search step (0.05) of a point that has not escaped |v|^2 <= 8
        // after >= 1 iteration, the first of which is the point itself (raymarching.rs:59-76, 202-241)
        if (n.p[0] < 1.0) return false;
        for (int a = 0; a < 3; ++a) { lo[a] = -3.0; hi[a] = 3.0; }
        return true;
    default:  // RS_NODE_QUADRIC: unbounded
        return false;
    }
}


double half_area(const double* lo, const double* hi) {
    double dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return dx * dy + dy * dz + dz * dx;
}

void box_of(const std::vector<BuildPrim>& prims, size_t lo, size_t hi, double* out) {
    for (int a = 0; a < 3; ++a) { out[a] = DBL_MAX; out[3 + a] = -DBL_MAX; }
    for (size_t i = lo; i < hi; ++i)
        for (int a = 0; a < 3; ++a) {
            out[a] = std::min(out[a], prims[i].lo[a]);
            out[3 + a] = std::max(out[3 + a], prims[i].hi[a]);
        }
}

// returns child reference: >= 0 inner node index, < 0 leaf ~root
int build_bvh(std::vector<BuildPrim>& prims, size_t lo, size_t hi, int depth, std::vector<HostBvhNode>& nodes) {
    if (hi - lo == 1) return ~prims[lo].root;
    size_t mid = lo + (hi - lo) / 2;
    int axis = 0;
    bool split_done = false;
    // small ranges: exact sweep SAH (every split position on every axis, objects sorted by centroid)
    static const int sweep_max = [] { const char* e = std::getenv("RAYSNAIL_B200_SAH_SWEEP"); return e ? std::atoi(e) : 4096; }();
    if (depth < 48 && prims.size() <= (size_t)sweep_max) {  // whole scene small: build time stays negligible
        const size_t n = hi - lo;
        double best_cost = DBL_MAX;
        int best_axis = -1;
        size_t best_k = 0;
        std::vector<double> rarea(n);
        for (int a = 0; a < 3; ++a) {
            std::stable_sort(prims.begin() + lo, prims.begin() + hi, [&](const BuildPrim& x, const BuildPrim& y) { return x.c[a] < y.c[a]; });
            double rl[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, rh[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
            for (size_t k = n; k-- > 1;) {
                for (int d = 0; d < 3; ++d) { rl[d] = std::min(rl[d], prims[lo + k].lo[d]); rh[d] = std::max(rh[d], prims[lo + k].hi[d]); }
                rarea[k] = half_area(rl, rh);
            }
            double ll[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, lh[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
            for (size_t k = 1; k < n; ++k) {
                for (int d = 0; d < 3; ++d) { ll[d] = std::min(ll[d], prims[lo + k - 1].lo[d]); lh[d] = std::max(lh[d], prims[lo + k - 1].hi[d]); }
                const double cost = half_area(ll, lh) * (double)k + rarea[k] * (double)(n - k);
                if (cost < best_cost) { best_cost = cost; best_axis = a; best_k = k; }
            }
        }
        if (best_axis >= 0) {
            axis = best_axis;
            if (axis != 2) std::stable_sort(prims.begin() + lo, prims.begin() + hi, [&](const BuildPrim& x, const BuildPrim& y) { return x.c[axis] < y.c[axis]; });
            mid = lo + best_k;
            split_done = true;
        }
    }
    if (!split_done && depth < 24) {
        double cb[6];
        for (int a = 0; a < 3; ++a) { cb[a] = DBL_MAX; cb[3 + a] = -DBL_MAX; }
        for (size_t i = lo; i < hi; ++i)
            for (int a = 0; a < 3; ++a) { cb[a] = std::min(cb[a], prims[i].c[a]); cb[3 + a] = std::max(cb[3 + a], prims[i].c[a]); }
        constexpr int NB = 16;
        double best_cost = DBL_MAX;
        int best_axis = -1, best_bin = -1;
        for (int a = 0; a < 3; ++a) {
            double ext = cb[3 + a] - cb[a];
            if (!(ext > 0.0)) continue;
            double blo[NB][3], bhi[NB][3];
            int cnt[NB] = {0};
            for (int b = 0; b < NB; ++b)
                for (int k = 0; k < 3; ++k) { blo[b][k] = DBL_MAX; bhi[b][k] = -DBL_MAX; }
            double scale = NB / ext;
            for (size_t i = lo; i < hi; ++i) {
                int b = std::min(NB - 1, std::max(0, (int)((prims[i].c[a] - cb[a]) * scale)));
                cnt[b]++;
                for (int k = 0; k < 3; ++k) { blo[b][k] = std::min(blo[b][k], prims[i].lo[k]); bhi[b][k] = std::max(bhi[b][k], prims[i].hi[k]); }
            }
            double rarea[NB];
            int rcnt[NB];
            double rl[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, rh[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
            int rc = 0;
            for (int b = NB - 1; b > 0; --b) {
                for (int k = 0; k < 3; ++k) { rl[k] = std::min(rl[k], blo[b][k]); rh[k] = std::max(rh[k], bhi[b][k]); }
                rc += cnt[b];
                rarea[b] = rc ? half_area(rl, rh) : 0.0;
                rcnt[b] = rc;
            }
            double ll[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, lh[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
            int lc = 0;
            for (int b = 0; b < NB - 1; ++b) {
                for (int k = 0; k < 3; ++k) { ll[k] = std::min(ll[k], blo[b][k]); lh[k] = std::max(lh[k], bhi[b][k]); }
                lc += cnt[b];
                if (lc == 0 || rcnt[b + 1] == 0) continue;
                double cost = half_area(ll, lh) * lc + rarea[b + 1] * rcnt[b + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = a; best_bin = b; }
            }
        }
        if (best_axis >= 0) {
            axis = best_axis;
            double ext = cb[3 + axis] - cb[axis], scale = NB / ext;
            auto it = std::partition(prims.begin() + lo, prims.begin() + hi, [&](const BuildPrim& p) {
                int b = std::min(NB - 1, std::max(0, (int)((p.c[axis] - cb[axis]) * scale)));
                return b <= best_bin;
            });
            mid = (size_t)(it - prims.begin());
            split_done = mid > lo && mid < hi;
        }
    }
    if (!split_done) {  // median split on the widest centroid axis
        double cb[6];
        for (int a = 0; a < 3; ++a) { cb[a] = DBL_MAX; cb[3 + a] = -DBL_MAX; }
        for (size_t i = lo; i < hi; ++i)
            for (int a = 0; a < 3; ++a) { cb[a] = std::min(cb[a], prims[i].c[a]); cb[3 + a] = std::max(cb[3 + a], prims[i].c[a]); }
        axis = 0;
        for (int a = 1; a < 3; ++a)
            if (cb[3 + a] - cb[a] > cb[3 + axis] - cb[axis]) axis = a;
        mid = lo + (hi - lo) / 2;
        std::nth_element(prims.begin() + lo, prims.begin() + mid, prims.begin() + hi,
                         [&](const BuildPrim& a, const BuildPrim& b) { return a.c[axis] < b.c[axis]; });
    }
    int idx = (int)nodes.size();
    nodes.emplace_back();
    double b0[6], b1[6];
    box_of(prims, lo, mid, b0);
    box_of(prims, mid, hi, b1);
    int c0 = build_bvh(prims, lo, mid, depth + 1, nodes);
    int c1 = build_bvh(prims, mid, hi, depth + 1, nodes);
    HostBvhNode& n = nodes[idx];
    std::memcpy(n.box[0], b0, sizeof(b0));
    std::memcpy(n.box[1], b1, sizeof(b1));
    n.child[0] = c0;
    n.child[1] = c1;
    return idx;
}

float f32_down(double v) {
    float f = (float)v;
    if ((double)f > v) f = std::nextafterf(f, -FLT_MAX);
    return f;
}
float f32_up(double v) {
    float f = (float)v;
    if ((double)f < v) f = std::nextafterf(f, FLT_MAX);
    return f;
}

// f32 node boxes padded outward so that the f32 slab test never rejects a box the f64 ray
// enters: the absolute pad covers the rounding of the ray origin (|o| <= omax) and of the
// bound itself (DESIGN.md "conservative f32 traversal").
// what a leaf holds: the object's id and kind — for a triangle its index in the triangle table instead (the
// record there names the object), so that the test needs no look-up in between
int leaf_ref(const rs_ctx* ctx, int root) {
    const int kind = ctx->root_kind_host[root];
    return ((kind == ROOT_TRIANGLE ? ~ctx->root_node_host[root] : root) << 2) | kind;
}

int upload_bvh(rs_ctx* ctx, double omax) {
    const double k = 1.0 / 2097152.0;  // 2^-21
    // a kept device tree may be padded for a larger origin than this call asks for; padded_omax below reports THAT value,
    // so the object list has to be padded for it too (otherwise a camera between the two values would find the list boxes
    // too tight and prepare_camera, trusting padded_omax, would not re-pad them)
    if (ctx->tree_dev_valid && ctx->tree_dev_omax >= omax) omax = ctx->tree_dev_omax;
    {
        // small worlds: the object list the kernels walk instead of a tree (same outward padding as the node boxes)
        std::vector<float4> list;
        for (const auto& e : ctx->list_host) {
                float lo[3], hi[3];
                for (int a = 0; a < 3; ++a) {
                    lo[a] = f32_down(e.box[a] - k * (omax + std::fabs(e.box[a])) - 1e-7);
                    hi[a] = f32_up(e.box[3 + a] + k * (omax + std::fabs(e.box[3 + a])) + 1e-7);
                }
                const int ref = leaf_ref(ctx, e.root);
                float refbits;
                std::memcpy(&refbits, &ref, 4);
                list.push_back(make_float4(lo[0], lo[1], lo[2], refbits));
                list.push_back(make_float4(hi[0], hi[1], hi[2], e.noclip ? 1.f : 0.f));
        }
        int rcl = upload(ctx, ctx->d_root_list, list.data(), list.size());
        if (rcl != RS_OK) return rcl;
        ctx->scene.root_list = ctx->d_root_list.as<float4>();
        ctx->scene.n_list = (int)(list.size() / 2);
        // (the list is sorted: objects tested regardless of distance, spheres, flat objects, expression trees)
        int lock = 0;
        for (const auto& e : ctx->list_host) lock += ctx->root_kind_host[e.root] != ROOT_GENERIC ? 1 : 0;
        bool sorted = true;
        for (size_t i = 0; i < ctx->list_host.size(); ++i) sorted = sorted && ((ctx->root_kind_host[ctx->list_host[i].root] != ROOT_GENERIC) == ((int)i < lock));
        // lanes may evaluate DIFFERENT trees side by side only when the trees run the same instruction sequence
        // (measured: sdl/quadric.sdl +14 %; with trees of different shapes the lanes diverge inside the evaluator
        // and lockstep is faster: sdl/csg.sdl -15 %, sdl/transforms.sdl -13 %, profiles/r02e_ab_listskip.txt)
        ctx->scene.n_list_lockstep = (sorted && ctx->list_trees_uniform) ? lock : ctx->scene.n_list;
    }
    if (ctx->tree_dev_valid && ctx->tree_dev_omax >= omax) {
        // same tree, same leaf references, boxes already padded for origins at least this far out (a wider pad is
        // still conservative): the nodes in HBM stand
        ctx->padded_omax = ctx->tree_dev_omax;
        return RS_OK;
    }
    ctx->tree_dev_valid = true;
    ctx->tree_dev_omax = omax;
    if (ctx->bvh_width != 4) {  // binary tree, 64-byte nodes
        std::vector<BvhNode> dev2(std::max<size_t>(ctx->bvh_host.size(), 1));
        for (size_t i = 0; i < ctx->bvh_host.size(); ++i) {
            const HostBvhNode& h = ctx->bvh_host[i];
            float lo[2][3], hi[2][3];
            for (int c = 0; c < 2; ++c)
                for (int a = 0; a < 3; ++a) {
                    double l = h.box[c][a], u = h.box[c][3 + a];
                    lo[c][a] = f32_down(l - k * (omax + std::fabs(l)) - 1e-7);
                    hi[c][a] = f32_up(u + k * (omax + std::fabs(u)) + 1e-7);
                }
            BvhNode& d = dev2[i];
            d.bx = make_float4(lo[0][0], hi[0][0], lo[0][1], hi[0][1]);
            d.by = make_float4(lo[1][0], hi[1][0], lo[1][1], hi[1][1]);
            d.bz = make_float4(lo[0][2], hi[0][2], lo[1][2], hi[1][2]);
            // leaf references carry the root kind in their low bits: ~(root << 2 | kind)
            auto enc = [&](int c) { return c >= 0 ? c : ~leaf_ref(ctx, ~c); };
            d.ch = make_int4(enc(h.child[0]), enc(h.child[1]), 0, 0);
        }
        int rc2 = upload(ctx, ctx->d_bvh, dev2.data(), dev2.size());
        if (rc2 != RS_OK) return rc2;
        ctx->scene.bvh2 = ctx->d_bvh.as<BvhNode>();
        ctx->scene.bvh = nullptr;
        ctx->padded_omax = omax;
        return RS_OK;
    }
    std::vector<BvhNode4> dev(std::max<size_t>(ctx->bvh4_host.size(), 1));
    for (size_t i = 0; i < ctx->bvh4_host.size(); ++i) {
        const HostBvhNode4& h = ctx->bvh4_host[i];
        float lo[4][3], hi[4][3];
        int child[4];
        for (int c = 0; c < 4; ++c) {
            if (c >= h.n) {  // unused slot: a degenerate box at +FLT_MAX on every axis: t_near > t_far for any ray
                for (int a = 0; a < 3; ++a) { lo[c][a] = FLT_MAX; hi[c][a] = FLT_MAX; }
                // should a pathological axis-parallel ray ever "hit" it, re-visiting slot 0 is harmless
                child[c] = child[0];
                continue;
            }
            for (int a = 0; a < 3; ++a) {
                double l = h.box[c][a], u = h.box[c][3 + a];
                lo[c][a] = f32_down(l - k * (omax + std::fabs(l)) - 1e-7);
                hi[c][a] = f32_up(u + k * (omax + std::fabs(u)) + 1e-7);
            }
            // leaf references carry the root kind in their low bits: ~(root << 2 | kind)
            const int r = h.child[c];
            child[c] = r >= 0 ? r : ~leaf_ref(ctx, ~r);
        }
        BvhNode4& d = dev[i];
        d.lox = make_float4(lo[0][0], lo[1][0], lo[2][0], lo[3][0]);
        d.hix = make_float4(hi[0][0], hi[1][0], hi[2][0], hi[3][0]);
        d.loy = make_float4(lo[0][1], lo[1][1], lo[2][1], lo[3][1]);
        d.hiy = make_float4(hi[0][1], hi[1][1], hi[2][1], hi[3][1]);
        d.loz = make_float4(lo[0][2], lo[1][2], lo[2][2], lo[3][2]);
        d.hiz = make_float4(hi[0][2], hi[1][2], hi[2][2], hi[3][2]);
        d.child = make_int4(child[0], child[1], child[2], child[3]);
        d.pad = make_int4(0, 0, 0, 0);
    }
    int rc = upload(ctx, ctx->d_bvh, dev.data(), dev.size());
    if (rc != RS_OK) return rc;
    ctx->scene.bvh = ctx->d_bvh.as<BvhNode4>();
    ctx->scene.bvh2 = nullptr;
    ctx->padded_omax = omax;
    return RS_OK;
}

// Collapse the binary SAH tree into the 4-wide QBVH: a node adopts its grandchildren, always
// opening the inner child with the largest surface area first.  Returns the wide node index
// and the stack entries a traversal may need below it (3 pushes per level at most).
int collapse_bvh4(const std::vector<HostBvhNode>& bin, int bi, std::vector<HostBvhNode4>& out, int& need) {
    struct Entry { double box[6]; int ref; };
    Entry e[4];
    int n = 2;
    for (int c = 0; c < 2; ++c) { std::memcpy(e[c].box, bin[bi].box[c], sizeof(e[c].box)); e[c].ref = bin[bi].child[c]; }
    while (n < 4) {
        int best = -1;
        double best_area = -1.0;
        for (int c = 0; c < n; ++c)
            if (e[c].ref >= 0) {
                double a = half_area(e[c].box, e[c].box + 3);
                if (a > best_area) { best_area = a; best = c; }
            }
        if (best < 0) break;
        const HostBvhNode& b = bin[e[best].ref];
        std::memcpy(e[best].box, b.box[0], sizeof(e[best].box));
        e[best].ref = b.child[0];
        std::memcpy(e[n].box, b.box[1], sizeof(e[n].box));
        e[n].ref = b.child[1];
        ++n;
    }
    const int idx = (int)out.size();
    out.emplace_back();
    int child[4], deepest = 0;
    for (int c = 0; c < n; ++c) {
        if (e[c].ref >= 0) {
            int sub = 0;
            child[c] = collapse_bvh4(bin, e[c].ref, out, sub);
            deepest = std::max(deepest, sub);
        } else {
            child[c] = e[c].ref;
        }
    }
    HostBvhNode4& w = out[idx];
    w.n = n;
    for (int c = 0; c < n; ++c) { std::memcpy(w.box[c], e[c].box, sizeof(e[c].box)); w.child[c] = child[c]; }
    need = 3 + deepest;
    return idx;
}

int tree_depth(const rs_scene_desc* s, int ni) {
    const rs_node& n = s->nodes[ni];
    switch (n.type) {
        case RS_NODE_DIFFERENCE:
        case RS_NODE_INTERSECTION: return 1 + std::max(tree_depth(s, n.child0), tree_depth(s, n.child1));
        case RS_NODE_TRANSFORM:
        case RS_NODE_MEDIUM: return 1 + tree_depth(s, n.child0);
        case RS_NODE_BVH: {  // one level for the whole BVH object, however deep its own tree is (ProgramBuilder::emit keeps one running result)
            int d = 0;
            std::vector<int> st{ni};
            while (!st.empty()) {
                const rs_node& b = s->nodes[st.back()];
                st.pop_back();
                for (int c : {b.child0, b.child1}) {
                    if (c < 0) continue;
                    if (s->nodes[c].type == RS_NODE_BVH) st.push_back(c);
                    else d = std::max(d, tree_depth(s, c));
                }
            }
            return 1 + d;
        }
        default: return 1;
    }
}

void mat3x4_mul(const long double* a, const double* b16, long double* out) {  // out = b(4x4 affine) * a(3x4 affine)
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 4; ++c) {
            long double s = 0;
            for (int k = 0; k < 3; ++k) s += (long double)b16[4 * r + k] * a[4 * k + c];
            if (c == 3) s += (long double)b16[4 * r + 3];
            out[4 * r + c] = s;
        }
}

// ---------------------------------------------------------------------------------------------
// Object programs (rt_scene.cuh "Object programs"): post-order flattening of a world object's expression tree
// ---------------------------------------------------------------------------------------------
struct ProgramBuilder {
    const std::vector<rs_node>& nodes;
    std::vector<XformPre>& xf;  // per-TfFacade pre-multiplied stacks, composed world -> leaf chains are appended
    std::vector<ObjInstr> prog;
    std::vector<int2> cprog;
    std::vector<CsgInfo> csg;
    int max_stack = 0;
    bool nested_medium = false;
    bool bvh_in_csg = false;  // BVH::contains is unimplemented!() in the reference (bvh.rs:194-200)

    // (parent chain `outer`, TfFacade `inner` below it): inv = inner.inv o outer.inv, fwd = outer.fwd o inner.fwd
    int compose(int outer, int inner) {
        if (outer < 0) return inner;
        XformPre r;
        auto mul = [](const double* a, const double* b, double* out) {  // out = a o b (apply b first)
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 4; ++j) {
                    long double acc = j == 3 ? (long double)a[4 * i + 3] : 0.0L;
                    for (int k = 0; k < 3; ++k) acc += (long double)a[4 * i + k] * (long double)b[4 * k + j];
                    out[4 * i + j] = (double)acc;
                }
        };
        mul(xf[inner].inv, xf[outer].inv, r.inv);
        mul(xf[outer].fwd, xf[inner].fwd, r.fwd);
        xf.push_back(r);
        return (int)xf.size() - 1;
    }
    static bool primitive(int type) {
        return type == RS_NODE_SPHERE || type == RS_NODE_BOX || type == RS_NODE_RECT || type == RS_NODE_QUADRIC || type == RS_NODE_RAYMARCHER;
    }
    void cemit(int ni, int cur) {  // `Hittable::contains`
        const rs_node& n = nodes[ni];
        if (n.type == RS_NODE_TRANSFORM) { cemit(n.child0, compose(cur, n.first_xform)); return; }
        if (n.type == RS_NODE_DIFFERENCE || n.type == RS_NODE_INTERSECTION) {
            cemit(n.child0, cur);
            cemit(n.child1, cur);
            cprog.push_back(make_int2((int)(((uint32_t)(n.type == RS_NODE_INTERSECTION ? RS_C_AND : RS_C_ANDNOT)) << 28), -1));
            return;
        }
        if (n.type == RS_NODE_BVH) bvh_in_csg = true;
        if (primitive(n.type) && n.type != RS_NODE_RECT) cprog.push_back(make_int2(ni | (RS_C_LEAF << 28), cur));
        else cprog.push_back(make_int2((int)((uint32_t)RS_C_FALSE << 28), -1));  // rect, medium (hit.rs default / unimplemented)
    }
    // returns the operand stack depth the subtree needs
    int emit(int ni, int cur, bool in_medium) {
        const rs_node& n = nodes[ni];
        switch (n.type) {
            case RS_NODE_TRANSFORM: return emit(n.child0, compose(cur, n.first_xform), in_medium);
            case RS_NODE_DIFFERENCE:
            case RS_NODE_INTERSECTION: {
                const int d0 = emit(n.child0, cur, in_medium);
                const size_t jm = prog.size();
                prog.push_back(ObjInstr{OP_JMISS, ni, -1, 0});
                const int d1 = emit(n.child1, cur, in_medium);
                CsgInfo ci;
                ci.c0_off = (int)cprog.size();
                cemit(n.child0, cur);
                ci.c0_len = (int)cprog.size() - ci.c0_off;
                ci.c1_off = (int)cprog.size();
                cemit(n.child1, cur);
                ci.c1_len = (int)cprog.size() - ci.c1_off;
                csg.push_back(ci);
                prog.push_back(ObjInstr{n.type == RS_NODE_DIFFERENCE ? OP_DIFF : OP_INTER, ni, cur, (int)csg.size() - 1});
                prog[jm].aux = (int)(prog.size() - 1 - jm);  // pc += aux, then ++pc: the instruction after the combiner
                return std::max(d0, 1 + d1);
            }
            case RS_NODE_MEDIUM: {
                if (in_medium) nested_medium = true;
                prog.push_back(ObjInstr{OP_MED_BEGIN, ni, cur, 0});
                const int d0 = emit(n.child0, cur, true);
                const size_t mid = prog.size();
                prog.push_back(ObjInstr{OP_MED_MID, ni, cur, 0});
                emit(n.child0, cur, true);
                prog.push_back(ObjInstr{OP_MED_END, ni, cur, 0});
                prog[mid].aux = (int)(prog.size() - 1 - mid);
                return d0;
            }
            case RS_NODE_BVH: {  // bvh.rs:173-192
                // "The closer of my two children" is associative, and ties go to the left child, so a BVH object returns the
                // closest of its members with ties to the first one in left-to-right order — whatever the shape of the tree.
                // The program therefore keeps ONE running result for the whole BVH object instead of one operand per level
                // (a balanced tree over the 400 boxes of the reference's all_feature scene, examples/common/scene.rs:361,
                // would need 10 operands that way):
                //     BVH_BEGIN top        bbox gate of the object; on a miss: push a miss, jump to the end
                //     MISS                 the running result
                //     BVH_GATE inner ...   bbox gate of an inner node: on a miss skip its subtree (nothing is pushed)
                //     <member> BVH_END     the member's own program, then: keep the closer of (running result, member)
                // Every gate the reference applies is still applied, in the reference's order.
                const size_t begin = prog.size();
                prog.push_back(ObjInstr{OP_BVH_BEGIN, ni, cur, 0});
                prog.push_back(ObjInstr{OP_MISS, ni, cur, 0});
                int depth = 0;
                struct Frame { int node; size_t gate; int next; };  // explicit stack: a degenerate tree may be as deep as it has nodes
                std::vector<Frame> st;
                st.push_back(Frame{ni, (size_t)-1, 0});
                while (!st.empty()) {
                    Frame& f = st.back();
                    if (f.next == 2) {
                        if (f.gate != (size_t)-1) prog[f.gate].aux = (int)(prog.size() - 1 - f.gate);  // pc += aux, then ++pc
                        st.pop_back();
                        continue;
                    }
                    const rs_node& b = nodes[f.node];
                    const int c = f.next++ == 0 ? b.child0 : b.child1;
                    if (c < 0) continue;  // a child that is None contributes nothing
                    if (nodes[c].type == RS_NODE_BVH) {
                        prog.push_back(ObjInstr{OP_BVH_GATE, c, cur, 0});
                        st.push_back(Frame{c, prog.size() - 1, 0});  // (invalidates f)
                    } else {
                        depth = std::max(depth, emit(c, cur, in_medium));
                        prog.push_back(ObjInstr{OP_BVH_END, ni, cur, 0});
                    }
                }
                prog[begin].aux = (int)(prog.size() - 1 - begin);  // on a miss of the top gate: past the last instruction
                return 1 + std::max(depth, 1);
            }
            default:
                prog.push_back(ObjInstr{OP_LEAF, ni, cur, 0});
                return 1;
        }
    }
};

int ensure_wave_set(rs_ctx* ctx, int k, uint32_t cap) {
    rs_ctx::WaveSet& w = ctx->wset[k];
    WaveState& ws = k == 0 ? ctx->ws : ctx->ws_aux[k - 1];
    CK(ctx->d_stats.ensure(sizeof(DeviceStats)));
    ws.stats = ctx->d_stats.as<DeviceStats>();  // shared by both sets (atomic counters)
    ws.uses_time = ctx->uses_time ? 1 : 0;
    if (cap <= w.capacity) return RS_OK;
    for (int b = 0; b < 2; ++b) {
        CK(w.d_rec[b].ensure((size_t)cap * 4 * sizeof(uint4)));
        CK(w.d_time[b].ensure((size_t)cap * sizeof(double)));
        ws.buf[b].rec = w.d_rec[b].as<uint4>();
        ws.buf[b].time = w.d_time[b].as<double>();
    }
    CK(w.d_hit.ensure((size_t)cap * sizeof(HitA)));
    CK(w.d_hit_src.ensure((size_t)cap * sizeof(int2)));
    CK(w.d_cq.ensure((size_t)cap * RS_NUM_CLASSES * sizeof(uint2)));
    CK(w.d_contrib.ensure((size_t)cap * sizeof(float4)));
    CK(w.d_ctr.ensure(sizeof(WaveCounters)));
    ws.hit = w.d_hit.as<HitA>();
    ws.hit_src = w.d_hit_src.as<int2>();
    ws.class_queue = w.d_cq.as<uint2>();
    ws.contrib = w.d_contrib.as<float4>();
    ws.ctr = w.d_ctr.as<WaveCounters>();
#if defined(RS_TUNE_COST_DUMP) && RS_TUNE_COST_DUMP
    CK(w.d_dbg_cost.ensure((size_t)cap * sizeof(int2)));
    CK(cudaMemset(w.d_dbg_cost.p, 0, (size_t)cap * sizeof(int2)));
    ws.dbg_cost = w.d_dbg_cost.as<int2>();
#endif
    ws.capacity = cap;
    w.capacity = cap;
    return RS_OK;
}

int ensure_wave(rs_ctx* ctx, uint32_t cap, int n_sets = 1) {
    for (int k = 0; k < n_sets; ++k) {
        int rc = ensure_wave_set(ctx, k, cap);
        if (rc != RS_OK) return rc;
    }
    ctx->capacity = ctx->wset[0].capacity;
    return RS_OK;
}

int check_render_args(rs_ctx* ctx, const rs_camera* cam, const rs_render_params* p) {
    if (!ctx) return RS_ERR_INVALID_ARGUMENT;
    if (!ctx->has_scene) return fail(ctx, RS_ERR_NO_SCENE, "no scene uploaded");
    if (!cam || !p) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "null camera or params");
    if (cam->width <= 0 || cam->height <= 0 || (int64_t)cam->width * cam->height > (int64_t)1 << 30)
        return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad image size");
    if (p->sqrt_spp <= 0 || p->sqrt_spp > 46340) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "sqrt_spp must be in 1..46340");
    if (p->depth < 0 || p->depth > RS_MAX_DEPTH) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "depth must be in 0..64");
    if (p->stratum_row_begin < 0 || p->stratum_row_end > p->sqrt_spp || p->stratum_row_begin > p->stratum_row_end)
        return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad stratum row range");
    if (ctx->n_lights == 0)
        return fail(ctx, RS_ERR_INVALID_ARGUMENT, "world.lights is empty (the reference panics in HittableList::random)");
    return RS_OK;
}

FrameParams make_frame(const rs_camera* cam, const rs_render_params* p) {
    FrameParams fp;
    std::memset(&fp, 0, sizeof(fp));
    fp.cam = *cam;
    fp.sqrt_spp = p->sqrt_spp;
    fp.magic_q = div_magic((uint32_t)p->sqrt_spp);
    fp.depth = p->depth;
    fp.stratum_row_begin = p->stratum_row_begin;
    fp.flags = p->flags;
    fp.key = make_uint2((uint32_t)p->seed, (uint32_t)(p->seed >> 32));
    for (uint32_t r = 0; r < 10; ++r) fp.rk.rk[r] = make_uint2(fp.key.x + r * 0x9E3779B9u, fp.key.y + r * 0xBB67AE85u);
    return fp;
}

int prepare_camera(rs_ctx* ctx, const rs_camera* cam) {
    double om = ctx->scene_omax;
    for (int a = 0; a < 3; ++a) om = std::max(om, std::fabs(cam->origin[a]) + std::fabs(cam->aperture) + 1e-3);
    if (om > ctx->padded_omax) {
        const int rc = upload_bvh(ctx, om * 1.25);
        if (rc != RS_OK) return rc;
        CK(cudaStreamSynchronize(0));  // (the uploads are asynchronous on the legacy stream; the render may use another one)
    }
    return RS_OK;
}

struct Timing {
    rs_ctx* ctx;
    cudaStream_t stream;
    bool on;
    size_t used = 0;
    std::vector<int> tags;  // category of the interval that starts at event i
    void mark(int tag) {
        if (!on) return;
        if (used == ctx->events.size()) {
            cudaEvent_t e;
            cudaEventCreate(&e);
            ctx->events.push_back(e);
        }
        cudaEventRecord(ctx->events[used++], stream);
        tags.push_back(tag);
    }
};

// the wavefront loop
int render_into(rs_ctx* ctx, const rs_camera* cam, const rs_render_params* p, const unsigned char* d_mask, float4* d_accum,
                cudaStream_t stream, bool timing) {
    const uint32_t np = (uint32_t)cam->width * (uint32_t)cam->height;
    const uint32_t s_total = (uint32_t)(p->stratum_row_end - p->stratum_row_begin) * (uint32_t)p->sqrt_spp;
    ctx->stats = rs_stats{};
    ctx->stats_final = false;
    if (s_total == 0 || np == 0) {  // nothing to render (a device beyond the strata rows): the zeroed counters above ARE the result;
        ctx->stats_final = true;    // the device-side ones hold an earlier frame's values, or do not exist yet on a fresh context
        return RS_OK;
    }
    // default: 64 Mi paths per wave (272 B of wave state per path: 18 GB per wave set, 73 of the 180 GB for the four sets).
    // 32 -> 64 Mi: +0.9 % on the headline scene, +4.6 % on the 1 M-triangle mesh; 128 Mi another +0.4 / +1.5 % for twice
    // the memory (profiles/r02h_ab_wave.txt, r02h_ab_wave2.txt)
    uint32_t cap = p->max_wave_paths > 0 ? (uint32_t)p->max_wave_paths : (1u << 26);
    cap = std::max(cap, 1024u);
    uint32_t P, K, n_waves;
    int n_sets;
    auto partition = [&] {
        P = std::min(np, cap);
        K = std::max(1u, std::min(s_total, cap / P));
        n_waves = ((np + P - 1) / P) * ((s_total + K - 1) / K);
        // kernel timing needs the kernels back to back on one stream
        n_sets = timing ? 1 : (int)std::min<uint32_t>((uint32_t)ctx->wave_streams, n_waves);
    };
    partition();
    if (p->max_wave_paths <= 0) {
        // The default is sized for an otherwise empty 180 GB card.  Where less is free — a smaller device, several contexts
        // or ranks on one GPU, a host application with its own buffers — halve the wave until the sets that still have to
        // grow fit into 80 % of what is free now (the image does not depend on the wave size beyond the rounding of the
        // per-wave sums); 1 Mi paths (1.1 GB for four sets) is the floor, below it the allocation may simply fail.
        constexpr size_t kBytesPerPath = 272;
        auto missing = [&] {
            size_t need = 0;
            for (int k = 0; k < n_sets; ++k)
                if (ctx->wset[k].capacity < P * K) need += (size_t)P * K * kBytesPerPath;
            return need;
        };
        size_t free_b = 0, total_b = 0;
        // (the query is not free — tens of microseconds to a millisecond — so frames whose buffers exist never make it)
        if (missing() > 0) {
            if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) {
                while (cap > (1u << 20) && missing() > free_b / 5 * 4) {
                    cap >>= 1;
                    partition();
                }
            } else {
                cudaGetLastError();
            }
        }
    }
    int rc = ensure_wave(ctx, P * K, n_sets);
    if (rc != RS_OK) {
        if (ctx->error.find("out of memory") != std::string::npos)
            return fail(ctx, RS_ERR_OUT_OF_MEMORY, "wave buffers (" + std::to_string(n_sets) + " sets of " + std::to_string((uint64_t)P * K) +
                                                       " paths, 272 B each) do not fit: lower rs_render_params.max_wave_paths or RS_OPT_WAVE_STREAMS");
        return rc;
    }
    rc = prepare_camera(ctx, cam);
    if (rc != RS_OK) return rc;
    FrameParams fp = make_frame(cam, p);
    fp.mask = d_mask;
#if RS_CHECKED
    // negative control of the checked build (tests/tools/checked_run.sh): the first 32 paths of every wave also try to
    // store their record one whole buffer too far; RS_IN_BOUNDS must catch (and skip) the stores and the call must fail
    static const bool selftest = [] { const char* e = std::getenv("RAYSNAIL_B200_CHECK_SELFTEST"); return e && std::atoi(e) != 0; }();
    if (selftest) fp.flags |= 0x80000000u;
#endif
    CK(cudaMemsetAsync(ctx->ws.stats, 0, sizeof(DeviceStats), stream));
    if (n_sets > 1) {
        if (!ctx->ev_fork) {
            CK(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
            for (int k = 0; k < rs_ctx::kMaxSets; ++k) CK(cudaEventCreateWithFlags(&ctx->ev_acc[k], cudaEventDisableTiming));
            for (int k = 0; k < rs_ctx::kMaxSets - 1; ++k) {
                CK(cudaStreamCreateWithFlags(&ctx->aux_stream[k], cudaStreamNonBlocking));
                CK(cudaEventCreateWithFlags(&ctx->ev_join[k], cudaEventDisableTiming));
            }
        }
        // fork: everything already queued on the caller's stream (e.g. the clear of d_accum) comes first
        CK(cudaEventRecord(ctx->ev_fork, stream));
        for (int k = 1; k < n_sets; ++k) CK(cudaStreamWaitEvent(ctx->aux_stream[k - 1], ctx->ev_fork, 0));
    }
    Timing tm{ctx, stream, timing};
    uint64_t launches = 0;
    uint32_t wave = 0;
    for (uint32_t pixel0 = 0; pixel0 < np; pixel0 += P) {
        fp.pixel0 = pixel0;
        fp.n_pixels = std::min(P, np - pixel0);
        fp.magic_pixels = div_magic(fp.n_pixels);
        for (uint32_t sample0 = 0; sample0 < s_total; sample0 += K, ++wave) {
            fp.sample0 = sample0;
            fp.n_samples = std::min(K, s_total - sample0);
            const int set = (int)(wave % (uint32_t)n_sets);
            const WaveState& ws = set ? ctx->ws_aux[set - 1] : ctx->ws;
            cudaStream_t st = set ? ctx->aux_stream[set - 1] : stream;
            CK(cudaMemsetAsync(ws.ctr, 0, sizeof(WaveCounters), st));
            const bool fused_first = ctx->fuse_raygen && !fp.mask && ctx->bvh_width == 2 && p->depth > 0;
            tm.mark(0);
            if (!fused_first) {
                launch_raygen(fp, ws, st);
                ++launches;
            }
            for (int b = 0; b < p->depth; ++b) {
                tm.mark(1);
                if (b == 0 && fused_first)
                    launch_extend_first(ctx->scene, fp, ws, ctx->extend_mode, ctx->extend_first_grid, ctx->stage_bytes, st);
                else
                    if (ctx->use_stream)
                        launch_extend_stream(ctx->scene, fp, ws, b, ctx->extend_mode, 1, ctx->extend_stream_grid, st);
                    else
                        launch_extend(ctx->scene, fp, ws, b, ctx->extend_mode, ctx->bvh_width == 4, 1, b < ctx->stage_bounces ? ctx->extend_grid : ctx->extend_grid_global,
                                      b < ctx->stage_bounces ? ctx->stage_bytes : 0, st);
#if defined(RS_TUNE_COST_DUMP) && RS_TUNE_COST_DUMP
                // profiling builds: RAYSNAIL_B200_COST_DUMP=<file>:<bounce> writes the rays of that bounce of wave 0
                // and their traversal costs: header (cap, nf, nb), cap x 64-byte records, cap x int2
                if (const char* spec = std::getenv("RAYSNAIL_B200_COST_DUMP")) {
                    std::string sp(spec);
                    const size_t colon = sp.rfind(':');
                    const int want = colon == std::string::npos ? 1 : std::atoi(sp.c_str() + colon + 1);
                    if (wave == 0 && b == want && ws.dbg_cost) {
                        CK(cudaStreamSynchronize(st));
                        WaveCounters hc;
                        CK(cudaMemcpy(&hc, ws.ctr, sizeof(hc), cudaMemcpyDeviceToHost));
                        const uint32_t capn = ws.capacity, nf = (b == 0 && fused_first) ? fp.n_pixels * fp.n_samples : hc.n_rays[b], nb = (b == 0 && fused_first) ? 0u : hc.n_back[b];
                        std::vector<uint4> rec((size_t)capn * 4);
                        std::vector<int2> cost(capn);
                        CK(cudaMemcpy(rec.data(), ws.buf[b & 1].rec, rec.size() * sizeof(uint4), cudaMemcpyDeviceToHost));
                        CK(cudaMemcpy(cost.data(), ws.dbg_cost, cost.size() * sizeof(int2), cudaMemcpyDeviceToHost));
                        if (FILE* f = std::fopen(sp.substr(0, colon).c_str(), "wb")) {
                            const uint32_t hdr[4] = {capn, nf, nb, (uint32_t)b};
                            std::fwrite(hdr, sizeof(hdr), 1, f);
                            std::fwrite(rec.data(), sizeof(uint4), rec.size(), f);
                            std::fwrite(cost.data(), sizeof(int2), cost.size(), f);
                            std::fclose(f);
                        }
                    }
                }
#endif
                tm.mark(2);
                if (ctx->shade_generic == 2 && !shade_split_trees()) {
                    launch_shade(ctx->scene, fp, ws, b, 2, 0, ctx->shade_grid, nullptr, st);
                } else {  // lean / flat scenes: the Lambertian hits, then the other classes (see shade_kernel)
                    launch_shade(ctx->scene, fp, ws, b, ctx->shade_generic, 1, ctx->shade_grid_lambert, nullptr, st);
                    launch_shade(ctx->scene, fp, ws, b, ctx->shade_generic, 2, ctx->shade_grid, nullptr, st);
                    ++launches;
                }
                launches += 2;
            }
            tm.mark(3);
            // the accumulators are read-modify-written without atomics: one accumulate at a time, in wave order
            if (n_sets > 1 && wave > 0) CK(cudaStreamWaitEvent(st, ctx->ev_acc[(set + n_sets - 1) % n_sets], 0));
            launch_accumulate(fp, ws, d_accum, st);
            if (n_sets > 1) CK(cudaEventRecord(ctx->ev_acc[set], st));
            ++launches;
        }
    }
    for (int k = 1; k < n_sets; ++k) {  // join: the caller's stream continues after all streams are done
        CK(cudaEventRecord(ctx->ev_join[k - 1], ctx->aux_stream[k - 1]));
        CK(cudaStreamWaitEvent(stream, ctx->ev_join[k - 1], 0));
    }
    tm.mark(4);
    CK(cudaGetLastError());
    ctx->stats.n_launches = launches;
    if (timing) {
        CK(cudaStreamSynchronize(stream));
        double ms[5] = {0, 0, 0, 0, 0};
        for (size_t i = 0; i + 1 < tm.used; ++i) {
            float e = 0.f;
            cudaEventElapsedTime(&e, ctx->events[i], ctx->events[i + 1]);
            ms[tm.tags[i]] += e;
        }
        float total = 0.f;
        if (tm.used >= 2) cudaEventElapsedTime(&total, ctx->events[0], ctx->events[tm.used - 1]);
        ctx->stats.ms_raygen = ms[0];
        ctx->stats.ms_extend = ms[1];
        ctx->stats.ms_shade = ms[2];
        ctx->stats.ms_accumulate = ms[3];
        ctx->stats.ms_total = total;
    }
    return RS_OK;
}

int fetch_device_stats(rs_ctx* ctx, cudaStream_t stream) {
    if (ctx->stats_final) return RS_OK;  // an empty render: nothing was counted on the device
    DeviceStats ds;
    CK(cudaMemcpyAsync(&ds, ctx->ws.stats, sizeof(ds), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    ctx->stats.n_samples = ds.n_samples;
    ctx->stats.n_segments = ds.n_segments;
    if (RS_CHECKED) {  // the library's own memcheck (rt_math.cuh): a render with a failed bounds check is an error
        CK(cudaDeviceSynchronize());
        const unsigned long long v = read_violations();
        if (v) return fail(ctx, RS_ERR_CUDA, "RS_CHECKED: " + std::to_string(v) + " out-of-bounds accesses caught on the device");
    }
    return RS_OK;
}

}  // namespace

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" {

int rs_abi_version(void) { return RS_ABI_VERSION; }

int rs_abi_struct_size(int which) {
    switch (which) {
        case 0: return (int)sizeof(rs_node);
        case 1: return (int)sizeof(rs_xform);
        case 2: return (int)sizeof(rs_material);
        case 3: return (int)sizeof(rs_texture);
        case 4: return (int)sizeof(rs_scene_desc);
        case 5: return (int)sizeof(rs_camera);
        case 6: return (int)sizeof(rs_render_params);
        case 7: return (int)sizeof(rs_stats);
        default: return -1;
    }
}

int rs_create(rs_ctx** out, int device) {
    if (!out) return RS_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(nullptr, RS_ERR_NO_DEVICE, std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                                                   " (this backend has no CPU fallback)");
    if (device < 0 || device >= count) return fail(nullptr, RS_ERR_INVALID_ARGUMENT, "device index out of range");
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(nullptr, RS_ERR_CUDA, cudaGetErrorString(e));
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) return fail(nullptr, RS_ERR_CUDA, cudaGetErrorString(e));
    if (prop.major < 10)
        return fail(nullptr, RS_ERR_NO_DEVICE, std::string("device '") + prop.name + "' is not sm_100-class; this library is built for sm_100a only");
    rs_ctx* ctx = new rs_ctx();
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    ctx->extend_grid = ctx->sm_count * extend_blocks_per_sm(3, 0, 0);
    ctx->shade_grid = ctx->sm_count * shade_blocks_per_sm(2, 0);
    // allocated here, not at the first rs_wait_counter: by then peers may already be spinning, and cudaMalloc can wait for them
    if (ctx->d_wait_flag.ensure(256) != cudaSuccess || cudaMemset(ctx->d_wait_flag.p, 0, 256) != cudaSuccess) {
        const std::string why = std::string("rs_create: ") + cudaGetErrorString(cudaGetLastError());
        delete ctx;
        return fail(nullptr, RS_ERR_CUDA, why);
    }
    *out = ctx;
    return RS_OK;
}

void rs_destroy(rs_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    for (cudaEvent_t e : ctx->events) cudaEventDestroy(e);
    if (ctx->ev_fork) {
        cudaEventDestroy(ctx->ev_fork);
        for (cudaEvent_t e : ctx->ev_acc) if (e) cudaEventDestroy(e);
        for (cudaEvent_t e : ctx->ev_join) if (e) cudaEventDestroy(e);
        for (cudaStream_t q : ctx->aux_stream) if (q) cudaStreamDestroy(q);
    }
    delete ctx;
}

const char* rs_last_error(const rs_ctx* ctx) { return ctx ? ctx->error.c_str() : g_create_error.c_str(); }

int rs_set_option(rs_ctx* ctx, int option, int value) {
    if (!ctx) return RS_ERR_INVALID_ARGUMENT;
    if (option == RS_OPT_BVH_WIDTH && (value == 2 || value == 4)) {
        if (ctx->bvh_width != value) ctx->has_scene = false;  // the tree must be rebuilt by the next upload
        ctx->bvh_width = value;
        return RS_OK;
    }
    if (option == RS_OPT_LIST_MAX && value >= 0 && value <= 64) {
        if (ctx->list_max != value) ctx->has_scene = false;
        ctx->list_max = value;
        ctx->list_max_set = true;
        return RS_OK;
    }
    if (option == RS_OPT_WAVE_STREAMS && value >= 1 && value <= rs_ctx::kMaxSets) {
        ctx->wave_streams = value;
        return RS_OK;
    }
    if (option == RS_OPT_WAIT_BUDGET_MS && value >= 0) {
        ctx->wait_budget_ms = value;
        return RS_OK;
    }
    return fail(ctx, RS_ERR_INVALID_ARGUMENT, "unknown option or value");
}

int rs_upload_scene(rs_ctx* ctx, const rs_scene_desc* s) {
    if (!ctx || !s) return RS_ERR_INVALID_ARGUMENT;
    CK(cudaSetDevice(ctx->device));
    // RAYSNAIL_B200_UPLOAD_TRACE=1: wall time of every stage of the upload on stderr
    static const bool trace = [] { const char* e = std::getenv("RAYSNAIL_B200_UPLOAD_TRACE"); return e && std::atoi(e) != 0; }();
    auto t_last = std::chrono::steady_clock::now();
    auto stage = [&](const char* what) {
        if (!trace) return;
        const auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[rs_upload_scene] %-28s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };
    ctx->has_scene = false;
    // ---- validation -------------------------------------------------------------------------
    if (s->n_nodes < 0 || s->n_roots < 0 || s->n_tris < 0 || s->n_materials <= 0 || s->n_textures < 0)
        return fail(ctx, RS_ERR_INVALID_ARGUMENT, "negative counts / no materials");
    if (s->n_materials > 65535) return fail(ctx, RS_ERR_UNSUPPORTED, "more than 65535 materials");
    if (s->default_material < 0 || s->default_material >= s->n_materials) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad default material");
    for (int i = 0; i < s->n_textures; ++i) {
        const rs_texture& t = s->textures[i];
        if (t.type < RS_TEX_SOLID || t.type > RS_TEX_PERLIN) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad texture type");
        if (t.type == RS_TEX_CHECKER && (t.odd < 0 || t.odd >= i || t.even < 0 || t.even >= i)) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "checker children must precede it");
        if (t.type == RS_TEX_IMAGE && (t.width <= 0 || t.height <= 0 || t.texel_offset < 0 || (int64_t)t.texel_offset + (int64_t)t.width * t.height * 4 > s->n_texel_bytes))
            return fail(ctx, RS_ERR_INVALID_ARGUMENT, "image texture out of range");
        if (t.type == RS_TEX_PERLIN) {
            int pc = t.point_count;
            if (pc <= 0 || (pc & (pc - 1)) || t.perm_offset < 0 || t.perm_offset + 3 * pc > s->n_perlin_perm || t.value_offset < 0 ||
                t.value_offset + pc * (t.perlin_vector ? 3 : 1) > s->n_perlin_values)
                return fail(ctx, RS_ERR_INVALID_ARGUMENT, "perlin tables out of range");
        }
    }
    std::vector<int> mixed_depth(s->n_materials, 0);
    for (int i = 0; i < s->n_materials; ++i) {
        const rs_material& m = s->materials[i];
        if (m.type < RS_MAT_LAMBERTIAN || m.type > RS_MAT_MIXED) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad material type");
        bool needs_tex = m.type == RS_MAT_LAMBERTIAN || m.type == RS_MAT_METAL || m.type == RS_MAT_DIFFUSE_METAL || m.type == RS_MAT_DIFFUSE_LIGHT || m.type == RS_MAT_BLINN_PHONG;
        if (needs_tex && (m.texture < 0 || m.texture >= s->n_textures)) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "material texture out of range");
        if (m.type == RS_MAT_MIXED && (m.mat1 < 0 || m.mat1 >= i || m.mat2 < 0 || m.mat2 >= i)) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "mixed material children must precede it");
        // the kernels resolve a MixedMaterial chain (mixed_material.rs:43-51) in at most RS_MAX_MIXED_DEPTH draws
        mixed_depth[i] = m.type == RS_MAT_MIXED ? 1 + std::max(mixed_depth[m.mat1], mixed_depth[m.mat2]) : 0;
        if (mixed_depth[i] > RS_MAX_MIXED_DEPTH) return fail(ctx, RS_ERR_UNSUPPORTED, "MixedMaterial nested deeper than RS_MAX_MIXED_DEPTH (8)");
    }
    std::vector<rs_node> nodes(s->nodes, s->nodes + s->n_nodes);
    std::vector<XformPre> xf;
    for (int i = 0; i < s->n_nodes; ++i) {
        rs_node& n = nodes[i];
        if (n.type < RS_NODE_SPHERE || n.type > RS_NODE_BVH) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad node type");
        if (n.type == RS_NODE_BVH && (n.child0 >= i || n.child1 >= i || n.child0 < -1 || n.child1 < -1))
            return fail(ctx, RS_ERR_INVALID_ARGUMENT, "the children of a BVH node must precede it (or be -1)");
        if (n.material >= s->n_materials) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "node material out of range");
        bool two = n.type == RS_NODE_DIFFERENCE || n.type == RS_NODE_INTERSECTION, one = n.type == RS_NODE_TRANSFORM || n.type == RS_NODE_MEDIUM;
        if ((two || one) && (n.child0 < 0 || n.child0 >= i)) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "child0 must precede its parent");
        if (two && (n.child1 < 0 || n.child1 >= i)) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "child1 must precede its parent");
        if (n.type == RS_NODE_MEDIUM && n.material < 0) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "medium needs its isotropic material");
        if (n.type == RS_NODE_RECT && (n.rect_axis < 0 || n.rect_axis > 2)) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad rect axis");
        // device copy only: 1/r for `(p - c) / r` = `(p - c) * (1.0 / r)` (sphere.rs:56-58, vec3.rs:486-500);
        // the same IEEE quotient the device would compute per hit
        if (n.type == RS_NODE_SPHERE) n.p[7] = 1.0 / n.p[3];
        if (n.type == RS_NODE_TRANSFORM) {
            if (n.first_xform < 0 || n.n_xform < 0 || n.first_xform + n.n_xform > s->n_xforms) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "transform stack out of range");
            // pre-multiply the stack (transform.rs:133-157): forward applies entry 0 first,
            // inverse applies the inverses from the last entry to the first
            long double F[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0}, J[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0}, T[12];
            for (int k = 0; k < n.n_xform; ++k) { mat3x4_mul(F, s->xforms[n.first_xform + k].m, T); std::memcpy(F, T, sizeof(T)); }
            // the reference applies inv_{n-1} first: composite = inv_0 * inv_1 * ... * inv_{n-1}
            for (int k = n.n_xform - 1; k >= 0; --k) { mat3x4_mul(J, s->xforms[n.first_xform + k].inv, T); std::memcpy(J, T, sizeof(T)); }
            XformPre x;
            for (int k = 0; k < 12; ++k) { x.fwd[k] = (double)F[k]; x.inv[k] = (double)J[k]; }
            n.first_xform = (int)xf.size();
            n.n_xform = 1;
            xf.push_back(x);
        }
    }
    for (int i = 0; i < s->n_roots; ++i) {
        int r = s->roots[i];
        if (r >= 0 ? r >= s->n_nodes : ~r >= s->n_tris) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "root out of range");
        if (r >= 0 && tree_depth(s, r) > RS_MAX_TREE_DEPTH) return fail(ctx, RS_ERR_UNSUPPORTED, "object tree deeper than RS_MAX_TREE_DEPTH (8)");
    }
    for (int i = 0; i < s->n_lights; ++i) {
        int r = s->lights[i];
        if (r < 0 || r >= s->n_nodes) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "light out of range");
        if (tree_depth(s, r) > RS_MAX_TREE_DEPTH) return fail(ctx, RS_ERR_UNSUPPORTED, "light object tree too deep");
    }
    for (int i = 0; i < s->n_tris; ++i)
        if (s->tri_material[i] >= s->n_materials) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "triangle material out of range");

    stage("validation");
    // ---- root tables + BVH ------------------------------------------------------------------
    rs_ctx::UploadScratch& sc = ctx->scratch;
    std::vector<int>& root_kind = sc.root_kind;
    std::vector<double>& root_sphere = sc.root_sphere;
    std::vector<unsigned char>& root_gate = sc.root_gate;
    std::vector<BuildPrim>&prims = sc.prims, &loose = sc.loose;  // loose: no bound on the surface; searched as a list, never dropped by distance
    root_kind.resize(s->n_roots);
    root_sphere.resize((size_t)s->n_roots * 4);
    root_gate.resize(s->n_roots);
    loose.clear();  // (`prims` is overwritten whole below: clearing it first would make the resize zero-fill 80 MB per million objects)
    bool any_quadric_root = false;
    double omax = 1.0;
    {
        bool anyq[kMaxChunks] = {};
        double om[kMaxChunks];
        for (double& v : om) v = 1.0;
        // every chunk appends to its own lists; they are concatenated in chunk order, i.e. in ascending object id
        const int chunks = parallel_chunks((size_t)s->n_roots, [&](int chunk, size_t begin, size_t end) {
            std::vector<BuildPrim>&my = sc.part[chunk], &my_loose = sc.part_loose[chunk];
            my.clear();
            my_loose.clear();
            bool my_anyq = false;  // (chunk-local, stored once at the end: the shared arrays would bounce one cache line between the threads)
            double my_om = 1.0;
            for (size_t ii = begin; ii < end; ++ii) {
                const int i = (int)ii;
                int r = s->roots[i];
                const double* bb = s->root_bbox + 6 * (size_t)i;
                root_gate[i] = 0;
                for (int k = 0; k < 4; ++k) root_sphere[4 * (size_t)i + k] = 0.0;
                if (r < 0) {
                    root_kind[i] = ROOT_TRIANGLE;
                } else {
                    const rs_node& n = nodes[r];
                    bool static_sphere = n.type == RS_NODE_SPHERE && n.p[4] == 0.0 && n.p[5] == 0.0 && n.p[6] == 0.0;
                    if (static_sphere) {
                        root_kind[i] = ROOT_SPHERE;
                        for (int k = 0; k < 4; ++k) root_sphere[4 * (size_t)i + k] = n.p[k];
                    } else {
                        // rect / box, bare or under one TfFacade: inline distance tests in extend
                        const rs_node& inner = n.type == RS_NODE_TRANSFORM ? nodes[n.child0] : n;
                        // (a quadric root takes the same inline route, but only the generic-tier kernels hold its code)
                        root_kind[i] = (inner.type == RS_NODE_BOX || inner.type == RS_NODE_RECT || inner.type == RS_NODE_QUADRIC) ? ROOT_FLAT : ROOT_GENERIC;
                        my_anyq |= inner.type == RS_NODE_QUADRIC;
                        // spheres / boxes / rects always lie inside their reference bbox; everything
                        // else is clipped by it the way the reference's BVH leaf clips it
                        root_gate[i] = !(n.type == RS_NODE_SPHERE || n.type == RS_NODE_BOX || n.type == RS_NODE_RECT);
                    }
                }
                BuildPrim p;
                p.pad = 0;
                bool degenerate = false;
                for (int a = 0; a < 3; ++a) {
                    double lo = bb[a], hi = bb[3 + a];
                    if (!(lo == lo) || !(hi == hi)) degenerate = true;
                    lo = std::max(lo, -1e30);
                    hi = std::min(hi, 1e30);
                    // AABB::hit rejects t_max <= t_min (aabb.rs:33): a box with no thickness on an axis
                    // (e.g. an axis-aligned triangle, triangle_mesh.rs:43-45) is never hit
                    if (!(hi > lo)) degenerate = true;
                    p.lo[a] = lo;
                    p.hi[a] = hi;
                    p.c[a] = 0.5 * (lo + hi);
                }
                if (degenerate) continue;
                p.root = i;
                // objects the bbox gates (not a bare sphere / box / rect): their surface may reach outside it (geom_bound)
                bool bounded = true;
                if (r >= 0 && root_gate[i]) {
                    double glo[3], ghi[3];
                    bounded = geom_bound(nodes, xf, r, glo, ghi);
                    if (bounded)
                        for (int a = 0; a < 3; ++a) {
                            p.lo[a] = std::min(p.lo[a], std::max(glo[a], -1e30));
                            p.hi[a] = std::max(p.hi[a], std::min(ghi[a], 1e30));
                            p.c[a] = 0.5 * (p.lo[a] + p.hi[a]);
                        }
                }
                for (int a = 0; a < 3; ++a) my_om = std::max(my_om, std::max(std::fabs(p.lo[a]), std::fabs(p.hi[a])));
                (bounded ? my : my_loose).push_back(p);
            }
            anyq[chunk] = my_anyq;
            om[chunk] = my_om;
        });
        size_t off[kMaxChunks + 1] = {0};
        for (int c = 0; c < chunks; ++c) {
            off[c + 1] = off[c] + sc.part[c].size();
            loose.insert(loose.end(), sc.part_loose[c].begin(), sc.part_loose[c].end());
            any_quadric_root |= anyq[c];
            omax = std::max(omax, om[c]);
        }
        prims.resize(off[chunks]);
        parallel_chunks((size_t)chunks << 16, [&](int, size_t begin, size_t end) {  // (one chunk of this loop = one part)
            for (size_t c = begin >> 16; c < (end >> 16); ++c)
                if (!sc.part[c].empty()) std::memcpy(prims.data() + off[c], sc.part[c].data(), sc.part[c].size() * sizeof(BuildPrim));
        });
    }
    stage("root tables");
    // small worlds are searched as one list (RS_OPT_LIST_MAX); otherwise only the loose objects are
    if (const char* e = std::getenv("RAYSNAIL_B200_LIST_MAX")) { ctx->list_max = std::min(std::max(std::atoi(e), 0), 64); ctx->list_max_set = true; }  // same range as RS_OPT_LIST_MAX
    bool flat_or_tree = !loose.empty(), any_generic = false;
    for (const BuildPrim& p : prims) {
        flat_or_tree |= root_kind[p.root] == ROOT_FLAT || root_kind[p.root] == ROOT_GENERIC;
        any_generic |= root_kind[p.root] == ROOT_GENERIC;
    }
    // default: 12 objects; 24 when the world holds expression trees (a leaf that every lane reaches at another time costs
    // a tree evaluation per lane; sdl/declares.sdl, 17 objects: 423 -> 471 M samples/s as a list, profiles/r02e_ab_listskip2.txt)
    const int list_limit = ctx->list_max_set ? ctx->list_max : (any_generic ? 24 : 12);
    ctx->use_list = flat_or_tree && ctx->bvh_width == 2 && (int)(prims.size() + loose.size()) <= list_limit;
    ctx->list_host.clear();
    auto list_add = [&](const BuildPrim& p, bool noclip) {
        rs_ctx::ListEntry e;
        for (int a = 0; a < 3; ++a) { e.box[a] = p.lo[a]; e.box[3 + a] = p.hi[a]; }
        e.root = p.root;
        e.noclip = noclip;
        ctx->list_host.push_back(e);
    };
    for (const BuildPrim& p : loose) list_add(p, true);
    if (ctx->use_list) {
        for (const BuildPrim& p : prims) list_add(p, false);
        prims.clear();  // no tree
    }
    // any fixed order gives the same hit (the tie rule does not depend on it).  Objects that are tested whatever
    // the closest hit is come first, then the others by the cost of their test, so that the cheap ones'
    // hits spare the dear ones: sphere, rect / box, expression tree
    auto list_rank = [&](const rs_ctx::ListEntry& e) { return e.noclip ? -1 : root_kind[e.root] == ROOT_SPHERE ? 0 : root_kind[e.root] == ROOT_GENERIC ? 2 : 1; };
    std::sort(ctx->list_host.begin(), ctx->list_host.end(), [&](const rs_ctx::ListEntry& a, const rs_ctx::ListEntry& b) {
        const int ra = list_rank(a), rb = list_rank(b);
        return ra != rb ? ra < rb : a.root < b.root;
    });
    const bool any_object = !prims.empty() || !ctx->list_host.empty();
    // (leaf references are made of the kinds and the node / triangle numbers of the objects)
    const bool same_refs = ctx->root_kind_host == root_kind && ctx->root_node_host.size() == (size_t)s->n_roots &&
                           (s->n_roots == 0 || std::memcmp(ctx->root_node_host.data(), s->roots, (size_t)s->n_roots * sizeof(int)) == 0);
    ctx->root_kind_host = root_kind;
    ctx->root_node_host.assign(s->roots, s->roots + s->n_roots);
    int bvh_root = (int)0x80000000;  // RS_BVH_SENTINEL: no tree
    // The tree holds object ids and f64 bounds only (kinds and leaf references are filled in by upload_bvh), so it
    // is a function of `prims` and the width alone: the SAH build (0.9 s for a 1 M-triangle mesh, the whole
    // cost of a re-upload) is skipped when neither changed.  RAYSNAIL_B200_BVH_CACHE=0 always rebuilds.
    static const bool cache_on = [] { const char* e = std::getenv("RAYSNAIL_B200_BVH_CACHE"); return !e || std::atoi(e) != 0; }();
    const size_t key_bytes = prims.size() * sizeof(BuildPrim);
    ctx->bvh_reused = cache_on && prims.size() > 1 && ctx->bvh_key_width == ctx->bvh_width && ctx->bvh_key.size() == key_bytes;
    if (ctx->bvh_reused) {
        bool differ[kMaxChunks] = {};
        const int chunks = parallel_chunks(key_bytes, [&](int chunk, size_t begin, size_t end) {
            differ[chunk] = std::memcmp(ctx->bvh_key.data() + begin, reinterpret_cast<const unsigned char*>(prims.data()) + begin, end - begin) != 0;
        });
        for (int c = 0; c < chunks; ++c) ctx->bvh_reused = ctx->bvh_reused && !differ[c];
    }
    ctx->tree_dev_valid = ctx->tree_dev_valid && ctx->bvh_reused && same_refs;
    if (ctx->bvh_reused) {
        bvh_root = ctx->bvh_key_root;
    } else {
        ctx->bvh_key_width = 0;  // (nothing valid is kept while the build below can still fail)
        ctx->bvh_host.clear();
        ctx->bvh4_host.clear();
        if (cache_on && prims.size() > 1) ctx->bvh_key.assign(reinterpret_cast<const unsigned char*>(prims.data()), reinterpret_cast<const unsigned char*>(prims.data()) + key_bytes);
        else std::vector<unsigned char>().swap(ctx->bvh_key);
        if (!prims.empty()) bvh_root = build_bvh(prims, 0, prims.size(), 0, ctx->bvh_host);
        if (!prims.empty() && bvh_root < 0) {
            bvh_root = ~leaf_ref(ctx, ~bvh_root);  // single object
        } else if (!prims.empty() && ctx->bvh_width == 4) {
            int need = 0;
            bvh_root = collapse_bvh4(ctx->bvh_host, bvh_root, ctx->bvh4_host, need);
            if (need + 1 > RS_TRAVERSAL_STACK) return fail(ctx, RS_ERR_UNSUPPORTED, "world BVH too deep for the traversal stack");
            std::vector<HostBvhNode>().swap(ctx->bvh_host);  // only the collapsed tree is kept
        }
        if (!ctx->bvh_key.empty()) { ctx->bvh_key_width = ctx->bvh_width; ctx->bvh_key_root = bvh_root; }
    }
    stage(ctx->bvh_reused ? "tree (kept)" : "tree (built)");
    // (binary tree: depth <= 24 SAH levels + log2(n) median levels, one push per level: fits)
    ctx->scene_omax = omax;
    // object programs of the expression-tree roots
    ProgramBuilder pb{nodes, xf};
    std::vector<int2> root_prog(std::max(s->n_roots, 1), make_int2(0, 0));
    for (int i = 0; i < s->n_roots; ++i) {
        if (root_kind[i] != ROOT_GENERIC) continue;
        const int first = (int)pb.prog.size();
        const int depth = pb.emit(s->roots[i], -1, false);
        if (depth > RS_EVAL_STACK) return fail(ctx, RS_ERR_UNSUPPORTED, "object tree needs more than RS_EVAL_STACK operands");
        root_prog[i] = make_int2(first, (int)pb.prog.size() - first);
    }
    if (pb.nested_medium) return fail(ctx, RS_ERR_UNSUPPORTED, "a ConstantMedium inside the boundary of another ConstantMedium is not supported");
    if (pb.bvh_in_csg) return fail(ctx, RS_ERR_UNSUPPORTED, "a BVH cannot be a CSG operand: BVH::contains is unimplemented!() in the reference (bvh.rs:194-200)");
    {
        // do the expression trees on the object list share one program shape (same operations, same primitive kinds,
        // same skip distances)?  Then closest_hit lets every lane walk to its own next tree (upload_bvh).
        int first_root = -1, n_trees = 0;
        bool uniform = true;
        for (const auto& e : ctx->list_host) {
            if (root_kind[e.root] != ROOT_GENERIC) continue;
            ++n_trees;
            if (first_root < 0) { first_root = e.root; continue; }
            const int2 a = root_prog[first_root], b = root_prog[e.root];
            if (a.y != b.y) { uniform = false; break; }
            for (int k = 0; k < a.y && uniform; ++k) {
                const ObjInstr &ia = pb.prog[a.x + k], &ib = pb.prog[b.x + k];
                uniform = ia.op == ib.op && (ia.xf >= 0) == (ib.xf >= 0);
                if (uniform && ia.op == OP_LEAF) uniform = nodes[ia.node].type == nodes[ib.node].type;
                if (uniform && (ia.op == OP_JMISS || ia.op == OP_MED_MID)) uniform = ia.aux == ib.aux;
            }
            if (!uniform) break;
        }
        ctx->list_trees_uniform = uniform && n_trees >= 2;
        if (const char* e = std::getenv("RAYSNAIL_B200_LIST_SKIP")) ctx->list_trees_uniform = std::atoi(e) != 0;
    }
    // which kernel specialisations this scene needs
    bool any_tri = false, any_flat = false, any_tree = RS_TUNE_QUADRIC_FLAT ? false : any_quadric_root;
    auto needs = [&](int root) {
        any_tri |= root_kind[root] == ROOT_TRIANGLE;
        any_flat |= root_kind[root] == ROOT_FLAT;
        any_tree |= root_kind[root] == ROOT_GENERIC;
    };
    for (const BuildPrim& p : prims) needs(p.root);
    for (const auto& e : ctx->list_host) needs(e.root);
    ctx->extend_mode = any_tree ? 3 : (any_flat ? 2 : (any_tri ? 1 : 0));
    // 4 = 3 + the skip-ahead walk over uniformly shaped trees on the object list (its own kernel: compiled into the
    // generic one the extra code costs the other scenes 4-8 % in spills, profiles/r02e_ab_listskip2.txt)
    if (ctx->extend_mode == 3 && ctx->list_trees_uniform && ctx->bvh_width == 2) ctx->extend_mode = 4;
    ctx->uses_time = false;
    for (const rs_node& n : nodes)
        if (n.type == RS_NODE_SPHERE && (n.p[4] != 0.0 || n.p[5] != 0.0 || n.p[6] != 0.0)) ctx->uses_time = true;
    bool lights_simple = true;  // lights the lean shade kernel samples inline: plain spheres and rects
    for (int i = 0; i < s->n_lights; ++i) {
        const rs_node& n = nodes[s->lights[i]];
        lights_simple &= (n.type == RS_NODE_SPHERE || n.type == RS_NODE_RECT);
    }
    ctx->shade_generic = (any_tree || !lights_simple) ? 2 : (any_flat ? 1 : 0);
    // lean scenes whose binary tree and sphere table fit: the staged extend kernels (tree in shared memory)
    ctx->stage_bytes = 0;
    {
        const size_t tree_bytes = ctx->bvh_host.size() * sizeof(BvhNode), sphere_bytes = (size_t)s->n_roots * sizeof(double4);
        int want = 1;
        if (const char* e = std::getenv("RAYSNAIL_B200_STAGE_TREE")) want = std::atoi(e);
        if (want && ctx->extend_mode <= 1 && ctx->bvh_width == 2 && !ctx->bvh_host.empty() && tree_bytes + sphere_bytes <= (size_t)RS_STAGE_MAX_BYTES) {
            const int bytes = (int)(tree_bytes + sphere_bytes);
            const int b0 = extend_blocks_per_sm(ctx->extend_mode, 0, bytes), b1 = extend_first_blocks_per_sm(ctx->extend_mode, bytes);
            if (b0 > 0 && b1 > 0) ctx->stage_bytes = bytes;
        }
    }
    ctx->extend_grid = ctx->sm_count * extend_blocks_per_sm(ctx->extend_mode, ctx->bvh_width == 4, ctx->stage_bytes);
    {
        int want = ctx->extend_mode == 1 ? 1 : 0;  // meshes: per-ray cost varies by orders of magnitude
        if (const char* e = std::getenv("RAYSNAIL_B200_STREAM")) want = std::atoi(e);
        ctx->use_stream = want && ctx->extend_mode <= 1 && ctx->bvh_width == 2 && !ctx->bvh_host.empty() && bvh_root >= 0;
        if (ctx->use_stream) ctx->extend_stream_grid = ctx->sm_count * extend_stream_blocks_per_sm(ctx->extend_mode);
    }
    ctx->extend_grid_global = ctx->sm_count * extend_blocks_per_sm(ctx->extend_mode, ctx->bvh_width == 4, 0);
    if (const char* e = std::getenv("RAYSNAIL_B200_STAGE_BOUNCES")) ctx->stage_bounces = std::atoi(e);
    const bool one_shade = ctx->shade_generic == 2 && !shade_split_trees();
    ctx->shade_grid = ctx->sm_count * shade_blocks_per_sm(ctx->shade_generic, one_shade ? 0 : 2);
    ctx->shade_grid_lambert = ctx->sm_count * shade_blocks_per_sm(ctx->shade_generic, one_shade ? 0 : 1);
    ctx->extend_first_grid = ctx->sm_count * extend_first_blocks_per_sm(ctx->extend_mode, ctx->stage_bytes);
    if (const char* e = std::getenv("RAYSNAIL_B200_FUSE_RAYGEN")) ctx->fuse_raygen = std::atoi(e);
    // tuning hooks: fewer resident blocks per SM than the occupancy limit (profiling experiments)
    if (const char* e = std::getenv("RAYSNAIL_B200_EXTEND_BLOCKS_PER_SM")) { int v = std::atoi(e); if (v > 0) ctx->extend_grid = std::min(ctx->extend_grid, ctx->sm_count * v); }
    if (const char* e = std::getenv("RAYSNAIL_B200_SHADE_BLOCKS_PER_SM")) { int v = std::atoi(e); if (v > 0) ctx->shade_grid = std::min(ctx->shade_grid, ctx->sm_count * v); }
    if (const char* e = std::getenv("RAYSNAIL_B200_SHADE_LAMBERT_BLOCKS_PER_SM")) { int v = std::atoi(e); if (v > 0) ctx->shade_grid_lambert = std::min(ctx->shade_grid_lambert, ctx->sm_count * v); }


    // triangles: p0, (a,b,c) = p0 - p1, (d,e,f) = p0 - p2 (triangle_mesh.rs:51-56)
    // RS_TRI_DOUBLES per record (load_triangle); word 9 names the world object the triangle is
    constexpr size_t TD = RS_TRI_DOUBLES;
    std::vector<double>& tri_geom = sc.tri_geom;
    tri_geom.resize((size_t)s->n_tris * TD);
    parallel_chunks((size_t)s->n_tris, [&](int, size_t begin, size_t end) {
        for (size_t i = begin; i < end; ++i) {
            const double* v = s->tri_verts + 9 * i;
            double* g = &tri_geom[TD * i];
            for (size_t k = 10; k < TD; ++k) g[k] = 0.0;
            g[0] = v[0]; g[1] = v[1]; g[2] = v[2];
            g[3] = v[0] - v[3]; g[4] = v[1] - v[4]; g[5] = v[2] - v[5];
            g[6] = v[0] - v[6]; g[7] = v[1] - v[7]; g[8] = v[2] - v[8];
            const int64_t none = -1;
            std::memcpy(&g[9], &none, 8);
        }
    });
    for (int i = 0; i < s->n_roots; ++i)
        if (s->roots[i] < 0) {
            double& slot = tri_geom[TD * (size_t)(~s->roots[i]) + 9];
            int64_t root;
            std::memcpy(&root, &slot, 8);
            if (root != -1) return fail(ctx, RS_ERR_UNSUPPORTED, "a triangle is listed as two world objects");
            root = i;
            std::memcpy(&slot, &root, 8);
        }

    stage("programs + triangle records");
    // ---- upload -------------------------------------------------------------------------------
    int rc;
    if ((rc = upload(ctx, ctx->d_nodes, nodes.data(), nodes.size())) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_xf, xf.data(), xf.size())) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_prog, pb.prog.data(), pb.prog.size())) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_cprog, pb.cprog.data(), pb.cprog.size())) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_csg, pb.csg.data(), pb.csg.size())) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_root_prog, root_prog.data(), root_prog.size())) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_root_node, s->roots, (size_t)s->n_roots)) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_root_kind, root_kind.data(), root_kind.size())) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_root_sphere, root_sphere.data(), root_sphere.size())) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_root_bbox, s->root_bbox, (size_t)s->n_roots * 6)) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_root_gate, root_gate.data(), root_gate.size())) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_tri_geom, tri_geom.data(), tri_geom.size())) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_tri_normals, s->tri_normals, (size_t)s->n_tris * 9)) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_tri_material, s->tri_material, (size_t)s->n_tris)) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_lights, s->lights, (size_t)s->n_lights)) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_materials, s->materials, (size_t)s->n_materials)) != RS_OK) return rc;
    {
        // flattened per-root / per-material / per-light tables (see MatHot in rt_scene.cuh)
        std::vector<double>& root_inv_r = sc.root_inv_r;
        std::vector<int>& root_material = sc.root_material;
        root_inv_r.resize(s->n_roots);
        root_material.resize(s->n_roots);
        bool all_spheres = s->n_roots > 0;
        for (int i = 0; i < s->n_roots; ++i) {
            const int r = s->roots[i];
            root_inv_r[i] = 0.0;
            root_material[i] = -1;
            if (root_kind[i] == ROOT_SPHERE) { root_inv_r[i] = nodes[r].p[7]; root_material[i] = nodes[r].material; }
            else if (root_kind[i] == ROOT_TRIANGLE) root_material[i] = s->tri_material[~r];
            else if (root_kind[i] == ROOT_FLAT) root_material[i] = (nodes[r].type == RS_NODE_TRANSFORM ? nodes[nodes[r].child0] : nodes[r]).material;
            all_spheres &= root_kind[i] == ROOT_SPHERE;
        }
        std::vector<MatHot> hot(s->n_materials);
        for (int i = 0; i < s->n_materials; ++i) {
            const rs_material& m = s->materials[i];
            MatHot& h = hot[i];
            std::memset(&h, 0, sizeof(h));
            h.a = m.a; h.b = m.b; h.phong_factor = m.phong_factor; h.phong_exponent = m.phong_exponent; h.texture = m.texture;
            int mode = 2, needs_uv = 0;
            if (m.type == RS_MAT_DIELECTRIC || m.type == RS_MAT_ISOTROPIC || m.type == RS_MAT_MIXED) {
                mode = 0;  // constant colour (dielectric.rs:76, isotropic.rs:30); Mixed is resolved before shading
                h.c0 = make_float4(m.color[0], m.color[1], m.color[2], m.color[3]);
            } else if (m.texture >= 0) {
                const rs_texture& t = s->textures[m.texture];
                if (t.type == RS_TEX_SOLID) {
                    mode = 0;
                    h.c0 = make_float4(t.color[0], t.color[1], t.color[2], t.color[3]);
                } else if (t.type == RS_TEX_CHECKER && s->textures[t.odd].type == RS_TEX_SOLID && s->textures[t.even].type == RS_TEX_SOLID) {
                    mode = 1;
                    const rs_texture &e = s->textures[t.even], &o = s->textures[t.odd];
                    h.c0 = make_float4(e.color[0], e.color[1], e.color[2], e.color[3]);
                    h.c1 = make_float4(o.color[0], o.color[1], o.color[2], o.color[3]);
                    h.scale = t.scale;
                }
            }
            if (mode == 2 && m.texture >= 0) {
                // does a lookup read (u, v)?  an image, or a checker that (transitively) nests one
                std::vector<int> todo{m.texture};
                for (size_t k = 0; k < todo.size() && k < 64; ++k) {
                    const rs_texture& t = s->textures[todo[k]];
                    if (t.type == RS_TEX_IMAGE) needs_uv = 1;
                    if (t.type == RS_TEX_CHECKER) { todo.push_back(t.odd); todo.push_back(t.even); }
                }
            }
            h.type_flags = m.type | (mode << 8) | ((m.has_curve ? 1 : 0) << 16) | (needs_uv << 17);
        }
        std::vector<rs_node> light_nodes(s->n_lights);
        for (int i = 0; i < s->n_lights; ++i) light_nodes[i] = nodes[s->lights[i]];
        if ((rc = upload(ctx, ctx->d_root_inv_r, root_inv_r.data(), root_inv_r.size())) != RS_OK) return rc;
        if ((rc = upload(ctx, ctx->d_root_material, root_material.data(), root_material.size())) != RS_OK) return rc;
        if ((rc = upload(ctx, ctx->d_mat_hot, hot.data(), hot.size())) != RS_OK) return rc;
        if ((rc = upload(ctx, ctx->d_light_nodes, light_nodes.data(), light_nodes.size())) != RS_OK) return rc;
        ctx->scene.all_spheres = all_spheres ? 1 : 0;
    }
    {
        // device copy only: `reserved` = 1 when a lookup in this texture reads (u, v) — an image,
        // or a checker that (transitively) nests one (checker.rs is generic over its children)
        std::vector<rs_texture> tex(s->textures, s->textures + s->n_textures);
        for (auto& t : tex) t.reserved = t.type == RS_TEX_IMAGE ? 1 : 0;
        for (int pass = 0; pass < 16; ++pass) {
            bool changed = false;
            for (auto& t : tex)
                if (t.type == RS_TEX_CHECKER && !t.reserved && (tex[t.odd].reserved || tex[t.even].reserved)) { t.reserved = 1; changed = true; }
            if (!changed) break;
        }
        if ((rc = upload(ctx, ctx->d_textures, tex.data(), tex.size())) != RS_OK) return rc;
    }
    if ((rc = upload(ctx, ctx->d_texels, s->texels, (size_t)s->n_texel_bytes)) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_perm, s->perlin_perm, (size_t)s->n_perlin_perm)) != RS_OK) return rc;
    if ((rc = upload(ctx, ctx->d_values, s->perlin_values, (size_t)s->n_perlin_values)) != RS_OK) return rc;

    stage("uploads");
    DScene& S = ctx->scene;
    S.nodes = ctx->d_nodes.as<rs_node>();
    S.xf = ctx->d_xf.as<XformPre>();
    S.bvh_root = bvh_root;
    S.any_gate = 0;
    for (unsigned char g : root_gate) S.any_gate |= g;
    S.bvh2_bytes = ctx->stage_bytes ? (uint32_t)(ctx->bvh_host.size() * sizeof(BvhNode)) : 0u;
    S.sphere_bytes = ctx->stage_bytes ? (uint32_t)((size_t)s->n_roots * sizeof(double4)) : 0u;
    S.n_roots = any_object ? s->n_roots : 0;
    S.root_node = ctx->d_root_node.as<int>();
    S.root_kind = ctx->d_root_kind.as<int>();
    S.root_sphere = ctx->d_root_sphere.as<double4>();
    S.root_bbox = ctx->d_root_bbox.as<double>();
    S.root_gate = ctx->d_root_gate.as<unsigned char>();
    S.obj.prog = ctx->d_prog.as<ObjInstr>();
    S.obj.cprog = ctx->d_cprog.as<int2>();
    S.obj.csg = ctx->d_csg.as<CsgInfo>();
    S.obj.root_prog = ctx->d_root_prog.as<int2>();
    S.tri_geom = ctx->d_tri_geom.as<double2>();
    S.tri_normals = ctx->d_tri_normals.as<double>();
    S.tri_material = ctx->d_tri_material.as<int>();
    S.lights = ctx->d_lights.as<int>();
    S.n_lights = s->n_lights;
    S.materials = ctx->d_materials.as<rs_material>();
    S.mat_hot = ctx->d_mat_hot.as<MatHot>();
    S.root_inv_r = ctx->d_root_inv_r.as<double>();
    S.root_material = ctx->d_root_material.as<int>();
    S.light_nodes = ctx->d_light_nodes.as<rs_node>();
    S.textures = ctx->d_textures.as<rs_texture>();
    S.texels = ctx->d_texels.as<unsigned char>();
    S.perlin_perm = ctx->d_perm.as<int>();
    S.perlin_values = ctx->d_values.as<double>();
    S.default_material = s->default_material;
    S.bg0 = make_float4(s->bg0[0], s->bg0[1], s->bg0[2], s->bg0[3]);
    S.bg1 = make_float4(s->bg1[0], s->bg1[1], s->bg1[2], s->bg1[3]);
    ctx->n_lights = s->n_lights;
    ctx->padded_omax = -1.0;
    if ((rc = upload_bvh(ctx, omax * 1.25)) != RS_OK) return rc;
    CK(cudaStreamSynchronize(0));  // every table is in HBM before any stream may render from it
    stage("tree conversion + upload");
    ctx->has_scene = true;
    return RS_OK;
}

int rs_render_accumulate(rs_ctx* ctx, const rs_camera* cam, const rs_render_params* p, const uint8_t* d_mask, float* d_accum, void* stream) {
    int rc = check_render_args(ctx, cam, p);
    if (rc != RS_OK) return rc;
    if (!d_accum) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "null accumulation buffer");
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    rc = render_into(ctx, cam, p, d_mask, reinterpret_cast<float4*>(d_accum), st, (p->flags & RS_FLAG_KERNEL_TIMING) != 0);
    if (rc != RS_OK) return rc;
    return RS_OK;
}

int rs_finalize(rs_ctx* ctx, const float* d_accum, const uint8_t* d_mask, float* d_out, int32_t width, int32_t height, int32_t n_samples,
                int32_t gamma, void* stream) {
    if (!ctx || !d_accum || !d_out || width <= 0 || height <= 0 || n_samples <= 0) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad finalize arguments");
    CK(cudaSetDevice(ctx->device));
    launch_finalize(reinterpret_cast<const float4*>(d_accum), d_mask, reinterpret_cast<float4*>(d_out), (int64_t)width * height, n_samples, gamma,
                    (cudaStream_t)stream);
    CK(cudaGetLastError());
    return RS_OK;
}

int rs_device_alloc(rs_ctx* ctx, int64_t bytes, void** d_ptr) {
    if (!ctx || !d_ptr || bytes <= 0) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad rs_device_alloc arguments");
    CK(cudaSetDevice(ctx->device));
    void* p = nullptr;
    CK(cudaMalloc(&p, (size_t)bytes));
    CK(cudaMemset(p, 0, (size_t)bytes));
    *d_ptr = p;
    return RS_OK;
}

int rs_device_free(rs_ctx* ctx, void* d_ptr) {
    if (!ctx) return RS_ERR_INVALID_ARGUMENT;
    CK(cudaSetDevice(ctx->device));
    CK(cudaFree(d_ptr));
    return RS_OK;
}

int rs_ipc_export(rs_ctx* ctx, const void* d_ptr, uint8_t handle[RS_IPC_HANDLE_BYTES]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == RS_IPC_HANDLE_BYTES, "cudaIpcMemHandle_t is 64 bytes");
    if (!ctx || !d_ptr || !handle) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad rs_ipc_export arguments");
    CK(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, const_cast<void*>(d_ptr)));
    std::memcpy(handle, &h, sizeof(h));
    return RS_OK;
}

int rs_ipc_open(rs_ctx* ctx, const uint8_t handle[RS_IPC_HANDLE_BYTES], void** d_ptr) {
    if (!ctx || !d_ptr || !handle) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad rs_ipc_open arguments");
    CK(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    CK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *d_ptr = p;
    return RS_OK;
}

int rs_ipc_close(rs_ctx* ctx, void* d_ptr) {
    if (!ctx) return RS_ERR_INVALID_ARGUMENT;
    CK(cudaSetDevice(ctx->device));
    CK(cudaIpcCloseMemHandle(d_ptr));
    return RS_OK;
}

int rs_finalize_peers(rs_ctx* ctx, const float* const* d_accums, int32_t n_peers, const uint8_t* d_mask, float* d_out, int64_t pixel_begin,
                      int64_t pixel_end, int32_t n_samples, int32_t gamma, void* stream) {
    if (!ctx || !d_accums || !d_out || n_peers <= 0 || n_peers > RS_MAX_PEERS || n_samples <= 0 || pixel_begin < 0 || pixel_end < pixel_begin)
        return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad rs_finalize_peers arguments");
    CK(cudaSetDevice(ctx->device));
    PeerAccums pa;
    std::memset(&pa, 0, sizeof(pa));
    pa.n = n_peers;
    for (int k = 0; k < n_peers; ++k) {
        if (!d_accums[k]) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "null peer accumulator");
        pa.p[k] = reinterpret_cast<const float4*>(d_accums[k]);
    }
    launch_finalize_peers(pa, d_mask, reinterpret_cast<float4*>(d_out), pixel_begin, pixel_end, n_samples, gamma, (cudaStream_t)stream);
    CK(cudaGetLastError());
    return RS_OK;
}

int rs_render(rs_ctx* ctx, const rs_camera* cam, const rs_render_params* p, const uint8_t* pixel_mask, float* out_rgba) {
    int rc = check_render_args(ctx, cam, p);
    if (rc != RS_OK) return rc;
    if (!out_rgba) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "null output image");
    CK(cudaSetDevice(ctx->device));
    const size_t np = (size_t)cam->width * cam->height;
    CK(ctx->d_accum.ensure(np * sizeof(float4)));
    CK(ctx->d_out.ensure(np * sizeof(float4)));
    const unsigned char* d_mask = nullptr;
    // a mask that selects every pixel is no mask: the frame then takes the path that generates the camera rays
    // inside extend instead of compacting the selected pixels first
    if (pixel_mask && std::memchr(pixel_mask, 0, np) == nullptr) pixel_mask = nullptr;
    if (pixel_mask) {
        CK(ctx->d_mask.ensure(np));
        CK(cudaMemcpyAsync(ctx->d_mask.p, pixel_mask, np, cudaMemcpyHostToDevice, 0));
        d_mask = ctx->d_mask.as<unsigned char>();
    }
    CK(cudaMemsetAsync(ctx->d_accum.p, 0, np * sizeof(float4), 0));
    rc = render_into(ctx, cam, p, d_mask, ctx->d_accum.as<float4>(), 0, (p->flags & RS_FLAG_KERNEL_TIMING) != 0);
    if (rc != RS_OK) return rc;
    const int n = (p->stratum_row_end - p->stratum_row_begin) * p->sqrt_spp;
    launch_finalize(ctx->d_accum.as<float4>(), d_mask, ctx->d_out.as<float4>(), (int64_t)np, std::max(n, 1), p->gamma, 0);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out_rgba, ctx->d_out.p, np * sizeof(float4), cudaMemcpyDeviceToHost, 0));
    CK(cudaStreamSynchronize(0));
    return fetch_device_stats(ctx, 0);
}

int rs_combine_passes(rs_ctx* ctx, const float* d_old, const float* d_new, float* d_out, int64_t n, float pass, void* stream) {
    if (!ctx || !d_old || !d_new || !d_out || n < 0) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad combine arguments");
    CK(cudaSetDevice(ctx->device));
    launch_combine(reinterpret_cast<const float4*>(d_old), reinterpret_cast<const float4*>(d_new), reinterpret_cast<float4*>(d_out), n, pass, (cudaStream_t)stream);
    CK(cudaGetLastError());
    return RS_OK;
}

int rs_noise_map(rs_ctx* ctx, const float* d_pixels, float* d_noise, int32_t width, int32_t height, int32_t quirk, void* stream) {
    if (!ctx || !d_pixels || !d_noise || width <= 0 || height <= 0) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad noise arguments");
    CK(cudaSetDevice(ctx->device));
    launch_noise(reinterpret_cast<const float4*>(d_pixels), d_noise, width, height, quirk, (cudaStream_t)stream);
    CK(cudaGetLastError());
    return RS_OK;
}

int rs_quantize_rgb8(rs_ctx* ctx, const float* d_pixels, uint8_t* d_rgb, int64_t n, void* stream) {
    if (!ctx || !d_pixels || !d_rgb || n < 0) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad quantize arguments");
    CK(cudaSetDevice(ctx->device));
    launch_quantize(reinterpret_cast<const float4*>(d_pixels), d_rgb, n, (cudaStream_t)stream);
    CK(cudaGetLastError());
    return RS_OK;
}

uint64_t rs_pass_seed(uint64_t seed, int32_t pass) { return seed + (uint64_t)(int64_t)pass * 0x9E3779B97F4A7C15ull; }

// The passes loop of the CLI (raysnail.rs:379-427), device-resident: render -> combine -> noise -> redo map.
int rs_render_passes(rs_ctx* ctx, const rs_camera* cam, const rs_render_params* p, int32_t n_passes, int32_t redo_mode, float noise_threshold,
                     float* out_rgba, uint8_t* out_rgb8, int64_t* redo_counts) {
    int rc = check_render_args(ctx, cam, p);
    if (rc != RS_OK) return rc;
    if (!out_rgba || n_passes < 0 || redo_mode < RS_REDO_REFERENCE || redo_mode > RS_REDO_NOISE_FIXED)
        return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad rs_render_passes arguments");
    CK(cudaSetDevice(ctx->device));
    const size_t np = (size_t)cam->width * cam->height;
    CK(ctx->d_accum.ensure(np * sizeof(float4)));
    CK(ctx->d_out.ensure(np * sizeof(float4)));
    CK(ctx->d_pass_old.ensure(np * sizeof(float4)));
    CK(ctx->d_pass_comb.ensure(np * sizeof(float4)));
    CK(ctx->d_noise.ensure(np * sizeof(float)));
    CK(ctx->d_mask.ensure(np));
    CK(ctx->d_redo_count.ensure((size_t)std::max(n_passes, 1) * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(ctx->d_redo_count.p, 0, (size_t)std::max(n_passes, 1) * sizeof(unsigned long long), 0));
    float4* d_old = ctx->d_pass_old.as<float4>();
    float4* d_comb = ctx->d_pass_comb.as<float4>();
    // "in the first pass all pixels must be calculated": old = [0,0,0,1], redo = 1 (raysnail.rs:318-326)
    launch_fill(d_old, (int64_t)np, make_float4(0.f, 0.f, 0.f, 1.f), 0);
    const int n = (p->stratum_row_end - p->stratum_row_begin) * p->sqrt_spp;
    rs_stats total{};
    for (int pass = 0; pass < n_passes; ++pass) {
        rs_render_params pp = *p;
        pp.seed = rs_pass_seed(p->seed, pass);
        pp.flags &= ~RS_FLAG_KERNEL_TIMING;
        // RS_REDO_REFERENCE: the controller's map is the stale all-ones clone => every pixel, every pass
        const unsigned char* d_mask = (redo_mode != RS_REDO_REFERENCE && pass > 0) ? ctx->d_mask.as<unsigned char>() : nullptr;
        CK(cudaMemsetAsync(ctx->d_accum.p, 0, np * sizeof(float4), 0));
        rc = render_into(ctx, cam, &pp, d_mask, ctx->d_accum.as<float4>(), 0, false);
        if (rc != RS_OK) return rc;
        total.n_launches += ctx->stats.n_launches + 5;
        launch_finalize(ctx->d_accum.as<float4>(), d_mask, ctx->d_out.as<float4>(), (int64_t)np, std::max(n, 1), p->gamma, 0);
        launch_combine(d_old, ctx->d_out.as<float4>(), d_comb, (int64_t)np, (float)pass, 0);
        launch_noise(d_comb, ctx->d_noise.as<float>(), cam->width, cam->height, redo_mode == RS_REDO_NOISE_FIXED ? 0 : 1, 0);
        launch_redo_mask(ctx->d_noise.as<float>(), ctx->d_mask.as<unsigned char>(), (int64_t)np, noise_threshold,
                         ctx->d_redo_count.as<unsigned long long>() + pass, 0);
        std::swap(d_old, d_comb);
        CK(cudaGetLastError());
        DeviceStats ds;  // the per-pass path counters (also orders the host loop behind the device: one pass in flight)
        CK(cudaMemcpyAsync(&ds, ctx->ws.stats, sizeof(ds), cudaMemcpyDeviceToHost, 0));
        CK(cudaStreamSynchronize(0));
        total.n_samples += ds.n_samples;
        total.n_segments += ds.n_segments;
    }
    CK(cudaMemcpyAsync(out_rgba, d_old, np * sizeof(float4), cudaMemcpyDeviceToHost, 0));
    if (out_rgb8) {
        unsigned char* d_rgb = reinterpret_cast<unsigned char*>(ctx->d_out.p);  // W*H*3 <= W*H*16 bytes: reuse
        launch_quantize(d_old, d_rgb, (int64_t)np, 0);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(out_rgb8, d_rgb, np * 3, cudaMemcpyDeviceToHost, 0));
    }
    if (redo_counts && n_passes > 0) {
        std::vector<unsigned long long> c((size_t)n_passes);
        CK(cudaMemcpyAsync(c.data(), ctx->d_redo_count.p, (size_t)n_passes * sizeof(unsigned long long), cudaMemcpyDeviceToHost, 0));
        CK(cudaStreamSynchronize(0));
        for (int k = 0; k < n_passes; ++k) redo_counts[k] = (int64_t)c[(size_t)k];
    }
    CK(cudaStreamSynchronize(0));
    ctx->stats = total;
    ctx->stats_final = true;
    return RS_OK;
}

// Philox4x32-10 on the host: the same definition as rt_math.cuh's philox4x32_10 (integer arithmetic only)
int rs_philox4x32_10(const uint32_t counter[4], const uint32_t key[2], uint32_t out[4]) {
    if (!counter || !key || !out) return RS_ERR_INVALID_ARGUMENT;
    uint32_t c0 = counter[0], c1 = counter[1], c2 = counter[2], c3 = counter[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
    return RS_OK;
}

int rs_philox_uniform(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t depth, uint32_t purpose, uint32_t call, double out4[4]) {
    if (!out4 || purpose > 0xffu || call > 0xffffffu) return RS_ERR_INVALID_ARGUMENT;
    const uint32_t ctr[4] = {pixel, sample, depth, purpose | (call << 8)}, key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    uint32_t w[4];
    rs_philox4x32_10(ctr, key, w);
    for (int k = 0; k < 4; ++k) out4[k] = (double)(w[k] >> 8) * (1.0 / 16777216.0);
    return RS_OK;
}

static int run_probe(rs_ctx* ctx, const rs_camera* cam, const rs_render_params* p, int32_t sample_index, bool shade, void* host_out[11],
                     const size_t elem_bytes[11]) {
    int rc = check_render_args(ctx, cam, p);
    if (rc != RS_OK) return rc;
    if (sample_index < 0 || sample_index >= p->sqrt_spp * p->sqrt_spp) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "sample index out of range");
    if (p->depth < 1) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "probes need depth >= 1");
    CK(cudaSetDevice(ctx->device));
    const uint32_t np = (uint32_t)cam->width * (uint32_t)cam->height;
    rc = ensure_wave(ctx, np);
    if (rc != RS_OK) return rc;
    rc = prepare_camera(ctx, cam);
    if (rc != RS_OK) return rc;
    FrameParams fp = make_frame(cam, p);
    fp.mask = nullptr;
    fp.pixel0 = 0;
    fp.n_pixels = np;
    fp.magic_pixels = div_magic(fp.n_pixels);
    // local sample index such that s_j * sqrt_spp + s_i == sample_index
    fp.stratum_row_begin = 0;
    fp.sample0 = (uint32_t)sample_index;
    fp.n_samples = 1;
    void* dev[11];
    for (int i = 0; i < 11; ++i) {
        dev[i] = nullptr;
        if (host_out[i]) {
            CK(ctx->d_dbg[i].ensure((size_t)np * elem_bytes[i]));
            CK(cudaMemsetAsync(ctx->d_dbg[i].p, 0, (size_t)np * elem_bytes[i], 0));
            dev[i] = ctx->d_dbg[i].p;
        }
    }
    DebugOut dbg;
    dbg.obj_id = (int*)dev[0]; dbg.t = (double*)dev[1]; dbg.normal3 = (double*)dev[2]; dbg.point3 = (double*)dev[3];
    dbg.front = (int*)dev[4]; dbg.material = (int*)dev[5]; dbg.ray_od6 = (double*)dev[6];
    dbg.kind = (int*)dev[7]; dbg.next_od6 = (double*)dev[8]; dbg.weight3 = (double*)dev[9]; dbg.emitted3 = (double*)dev[10];
    CK(cudaMemsetAsync(ctx->ws.ctr, 0, sizeof(WaveCounters), 0));
    CK(cudaMemsetAsync(ctx->ws.stats, 0, sizeof(DeviceStats), 0));
    launch_raygen(fp, ctx->ws, 0);
    launch_extend(ctx->scene, fp, ctx->ws, 0, ctx->extend_mode, ctx->bvh_width == 4, 0, ctx->extend_grid, ctx->stage_bytes, 0);  // probes: misses stay visible to shade
    const bool one_shade = ctx->shade_generic == 2 && !shade_split_trees();
    if (shade && one_shade) {
        launch_shade(ctx->scene, fp, ctx->ws, 0, 2, 0, ctx->shade_grid, &dbg, 0);
    } else if (shade) {
        launch_shade(ctx->scene, fp, ctx->ws, 0, ctx->shade_generic, 1, ctx->shade_grid_lambert, &dbg, 0);
        launch_shade(ctx->scene, fp, ctx->ws, 0, ctx->shade_generic, 2, ctx->shade_grid, &dbg, 0);
    } else {
        launch_probe_primary(ctx->scene, fp, ctx->ws, dbg, 0);
    }
    CK(cudaGetLastError());
    for (int i = 0; i < 11; ++i)
        if (host_out[i]) CK(cudaMemcpyAsync(host_out[i], dev[i], (size_t)np * elem_bytes[i], cudaMemcpyDeviceToHost, 0));
    CK(cudaStreamSynchronize(0));
    return RS_OK;
}

int rs_debug_primary(rs_ctx* ctx, const rs_camera* cam, const rs_render_params* p, int32_t sample_index, int32_t* obj_id, double* t,
                     double* normal3, double* point3, int32_t* front, int32_t* material, double* ray_od6) {
    void* host[11] = {obj_id, t, normal3, point3, front, material, ray_od6, nullptr, nullptr, nullptr, nullptr};
    const size_t eb[11] = {4, 8, 24, 24, 4, 4, 48, 4, 48, 24, 24};
    return run_probe(ctx, cam, p, sample_index, false, host, eb);
}

int rs_debug_scatter(rs_ctx* ctx, const rs_camera* cam, const rs_render_params* p, int32_t sample_index, int32_t* kind, double* next_od6,
                     double* weight3, double* emitted3) {
    void* host[11] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, kind, next_od6, weight3, emitted3};
    const size_t eb[11] = {4, 8, 24, 24, 4, 4, 48, 4, 48, 24, 24};
    return run_probe(ctx, cam, p, sample_index, true, host, eb);
}

// ---------------------------------------------------------------------------------------------
// several devices behind one call: the scene is replicated, the strata rows of
// every pixel are split across the devices, every device sum-reduces + finalizes its slice of
// the pixels out of all accumulators over NVLink peer memory into device 0's image.  One host
// thread per device enqueues; the devices order themselves with arrival counters in device 0's
// memory (signal_kernel / wait_counter_kernel) — no host barrier between render and combine.
// ---------------------------------------------------------------------------------------------
}  // extern "C"

struct rs_multi {
    std::vector<rs_ctx*> ctx;
    std::vector<int> device;
    std::vector<cudaStream_t> stream;
    std::vector<DevBuf> accum, mask;
    DevBuf out0, counters;  // device 0: the image; [0] = accumulators complete, [1] = slices written, [2] = a wait kernel gave up
    uint32_t epoch = 0;     // frames completed since the counters were last zeroed: the counters stand at n * epoch between calls
    int inject_fault = 0;   // RS_OPT_INJECT_FAULT (tests): 1 = the last device fails in phase 1, 2 = between its two signals; one shot
    std::string error;
};

namespace {
int multi_fail(rs_multi* m, int code, const std::string& msg) {
    if (m) m->error = msg;
    else g_create_error = msg;
    return code;
}
// run f(k) for every device on its own host thread; first non-zero return wins
template <class F>
int for_each_device(rs_multi* m, F f) {
    const int n = (int)m->ctx.size();
    std::vector<int> rc(n, RS_OK);
    std::vector<std::thread> th;
    for (int k = 1; k < n; ++k) th.emplace_back([&, k] { rc[k] = f(k); });
    rc[0] = f(0);
    for (auto& t : th) t.join();
    for (int k = 0; k < n; ++k)
        if (rc[k] != RS_OK) return multi_fail(m, rc[k], "device " + std::to_string(m->device[k]) + ": " + rs_last_error(m->ctx[k]));
    return RS_OK;
}
// After a frame that failed part-way: some devices may have signalled and others not, so the counters no longer
// stand at n * epoch and the next frame's waits would never be satisfied.  Let every stream drain (wait kernels give
// up after the budget), then start counting from zero again.  Keeps m->error (the reason of the failure).
void multi_rebase(rs_multi* m) {
    for (size_t k = 0; k < m->ctx.size(); ++k) {
        cudaSetDevice(m->device[k]);
        if (m->stream[k]) cudaStreamSynchronize(m->stream[k]);
        cudaGetLastError();
    }
    cudaSetDevice(m->device[0]);
    if (m->counters.p) cudaMemset(m->counters.p, 0, 4 * sizeof(uint32_t));
    m->epoch = 0;
}
}  // namespace

extern "C" {

int rs_create_multi(rs_multi** out, const int32_t* device_ids, int32_t n) {
    if (!out || !device_ids || n <= 0 || n > RS_MAX_PEERS) return multi_fail(nullptr, RS_ERR_INVALID_ARGUMENT, "rs_create_multi: 1..16 devices");
    *out = nullptr;
    rs_multi* m = new rs_multi();
    m->device.assign(device_ids, device_ids + n);
    m->ctx.assign(n, nullptr);
    m->stream.assign(n, nullptr);
    m->accum.resize(n);
    m->mask.resize(n);
    for (int k = 0; k < n; ++k) {
        int rc = rs_create(&m->ctx[k], m->device[k]);
        if (rc != RS_OK) {
            const std::string why = g_create_error;
            rs_destroy_multi(m);
            return multi_fail(nullptr, rc, why);
        }
    }
    // every device reads every accumulator and writes device 0's image and counters
    for (int a = 0; a < n; ++a)
        for (int b = 0; b < n; ++b) {
            if (m->device[a] == m->device[b]) continue;  // (the same device listed twice: separate contexts, no peer mapping needed)
            int can = 0;
            cudaDeviceCanAccessPeer(&can, m->device[a], m->device[b]);
            if (!can) {
                rs_destroy_multi(m);
                return multi_fail(nullptr, RS_ERR_UNSUPPORTED, "devices " + std::to_string(device_ids[a]) + " and " + std::to_string(device_ids[b]) + " are not peers");
            }
            cudaSetDevice(m->device[a]);
            cudaError_t e = cudaDeviceEnablePeerAccess(m->device[b], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                rs_destroy_multi(m);
                return multi_fail(nullptr, RS_ERR_CUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e));
            }
            cudaGetLastError();
        }
    // A device listed more than once (a test set-up): its contexts share the device's hardware queues, and a
    // wait kernel that spins at the head of a queue another context's render also maps to would never see
    // that render finish.  One stream per such context keeps every stream on a queue of its own
    // (8 per device by default).
    for (int k = 0; k < n; ++k) {
        int same = 0;
        for (int j = 0; j < n; ++j) same += m->device[j] == m->device[k];
        if (same > 4) {
            rs_destroy_multi(m);
            return multi_fail(nullptr, RS_ERR_INVALID_ARGUMENT, "rs_create_multi: a device may be listed at most 4 times");
        }
        if (same > 1) rs_set_option(m->ctx[k], RS_OPT_WAVE_STREAMS, 1);
    }
    for (int k = 0; k < n; ++k) {
        cudaSetDevice(m->device[k]);
        if (cudaStreamCreateWithFlags(&m->stream[k], cudaStreamNonBlocking) != cudaSuccess) {
            rs_destroy_multi(m);
            return multi_fail(nullptr, RS_ERR_CUDA, "cudaStreamCreate failed");
        }
    }
    cudaSetDevice(m->device[0]);
    if (m->counters.ensure(4 * sizeof(uint32_t)) != cudaSuccess || cudaMemset(m->counters.p, 0, 4 * sizeof(uint32_t)) != cudaSuccess) {
        rs_destroy_multi(m);
        return multi_fail(nullptr, RS_ERR_CUDA, "counter allocation failed");
    }
    *out = m;
    return RS_OK;
}

void rs_destroy_multi(rs_multi* m) {
    if (!m) return;
    for (size_t k = 0; k < m->ctx.size(); ++k) {
        cudaSetDevice(m->device[k]);
        if (m->stream[k]) { cudaStreamSynchronize(m->stream[k]); cudaStreamDestroy(m->stream[k]); }
        m->accum[k].release();
        m->mask[k].release();
    }
    if (!m->device.empty()) {
        cudaSetDevice(m->device[0]);
        m->out0.release();
        m->counters.release();
    }
    for (rs_ctx* c : m->ctx) rs_destroy(c);
    delete m;
}

const char* rs_multi_last_error(const rs_multi* m) { return m ? m->error.c_str() : g_create_error.c_str(); }
int rs_multi_device_count(const rs_multi* m) { return m ? (int)m->ctx.size() : 0; }

int rs_multi_set_option(rs_multi* m, int option, int value) {
    if (!m) return RS_ERR_INVALID_ARGUMENT;
    if (option == RS_OPT_INJECT_FAULT) {
        if (value < 0 || value > 2) return multi_fail(m, RS_ERR_INVALID_ARGUMENT, "RS_OPT_INJECT_FAULT: 0, 1 or 2");
        m->inject_fault = value;
        return RS_OK;
    }
    for (size_t k = 0; k < m->ctx.size(); ++k) {
        if (option == RS_OPT_WAVE_STREAMS) {  // contexts that share a device keep their single stream (rs_create_multi)
            int same = 0;
            for (size_t j = 0; j < m->ctx.size(); ++j) same += m->device[j] == m->device[k];
            if (same > 1) continue;
        }
        int rc = rs_set_option(m->ctx[k], option, value);
        if (rc != RS_OK) return multi_fail(m, rc, rs_last_error(m->ctx[k]));
    }
    return RS_OK;
}

int rs_multi_upload_scene(rs_multi* m, const rs_scene_desc* s) {
    if (!m || !s) return multi_fail(m, RS_ERR_INVALID_ARGUMENT, "null scene");
    return for_each_device(m, [&](int k) { return rs_upload_scene(m->ctx[k], s); });
}

int rs_multi_render(rs_multi* m, const rs_camera* cam, const rs_render_params* p, const uint8_t* pixel_mask, float* out_rgba, rs_stats* stats_sum) {
    if (!m || !cam || !p || !out_rgba) return multi_fail(m, RS_ERR_INVALID_ARGUMENT, "bad rs_multi_render arguments");
    const int n = (int)m->ctx.size();
    for (int k = 0; k < n; ++k) {
        int rc = check_render_args(m->ctx[k], cam, p);
        if (rc != RS_OK) return multi_fail(m, rc, rs_last_error(m->ctx[k]));
    }
    const size_t np = (size_t)cam->width * cam->height;
    if (pixel_mask && std::memchr(pixel_mask, 0, np) == nullptr) pixel_mask = nullptr;  // (as rs_render)
    const int rows0 = p->stratum_row_begin, rows = p->stratum_row_end - p->stratum_row_begin;
    const int n_total = std::max(rows * p->sqrt_spp, 1);
    // buffers first (allocation synchronises), so that nothing but launches follows
    for (int k = 0; k < n; ++k) {
        cudaSetDevice(m->device[k]);
        if (m->accum[k].ensure(np * sizeof(float4)) != cudaSuccess || (pixel_mask && m->mask[k].ensure(np) != cudaSuccess))
            return multi_fail(m, RS_ERR_OUT_OF_MEMORY, "accumulator allocation failed");
    }
    cudaSetDevice(m->device[0]);
    if (m->out0.ensure(np * sizeof(float4)) != cudaSuccess) return multi_fail(m, RS_ERR_OUT_OF_MEMORY, "image allocation failed");
    // the counters stand at n * epoch; this frame brings them to n * (epoch + 1).  epoch itself moves only once the frame
    // is through: a frame that fails part-way re-bases the counters (multi_rebase) instead of leaving them short
    const uint32_t expect = (uint32_t)n * (m->epoch + 1);
    uint32_t* const arrived = m->counters.as<uint32_t>();
    uint32_t* const written = arrived + 1;
    uint32_t* const gave_up = arrived + 2;
    const unsigned long long budget_ns = (unsigned long long)m->ctx[0]->wait_budget_ms * 1000000ull;
    const int fault = m->inject_fault;
    m->inject_fault = 0;
    PeerAccums pa;
    std::memset(&pa, 0, sizeof(pa));
    pa.n = n;
    for (int k = 0; k < n; ++k) pa.p[k] = m->accum[k].as<float4>();
    // Phase 1, every device: enqueue the render.  Whatever allocates happens here (the wave buffers on a first
    // call): with peer access enabled cudaMalloc synchronises with the peer devices, and it would wait forever for
    // a peer that already spins in phase 2's wait kernel.  The host threads join before phase 2 — they wait for
    // each other's ENQUEUE only, the devices keep running.
    int rc = for_each_device(m, [&](int k) -> int {
        rs_ctx* ctx = m->ctx[k];
        cudaStream_t st = m->stream[k];
        CK(cudaSetDevice(m->device[k]));
        const unsigned char* d_mask = nullptr;
        if (pixel_mask) {
            CK(cudaMemcpyAsync(m->mask[k].p, pixel_mask, np, cudaMemcpyHostToDevice, st));
            d_mask = m->mask[k].as<unsigned char>();
        }
        CK(cudaMemsetAsync(m->accum[k].p, 0, np * sizeof(float4), st));
        rs_render_params pk = *p;  // this device's strata rows: contiguous, balanced (the split of distributed.py)
        pk.stratum_row_begin = rows0 + (int)(((int64_t)rows * k) / n);
        pk.stratum_row_end = rows0 + (int)(((int64_t)rows * (k + 1)) / n);
        pk.flags &= ~RS_FLAG_KERNEL_TIMING;
        if (fault == 1 && k == n - 1) return fail(ctx, RS_ERR_CUDA, "injected fault in phase 1 (RS_OPT_INJECT_FAULT)");
        return render_into(ctx, cam, &pk, d_mask, m->accum[k].as<float4>(), st, false);
    });
    if (rc != RS_OK) {  // nobody has signalled yet: drain the renders that were queued and leave the counters where they were
        multi_rebase(m);
        return rc;
    }
    // Phase 2, every device: arrive, wait for all accumulators, reduce + finalize this device's slice of the pixels
    // into device 0's image, arrive again; device 0 waits for all slices and copies the image out.
    rc = for_each_device(m, [&](int k) -> int {
        rs_ctx* ctx = m->ctx[k];
        cudaStream_t st = m->stream[k];
        CK(cudaSetDevice(m->device[k]));
        const unsigned char* d_mask = pixel_mask ? m->mask[k].as<unsigned char>() : nullptr;
        launch_signal(arrived, st);
        launch_wait_counter(arrived, expect, st, budget_ns, gave_up);
        if (fault == 2 && k == n - 1) return fail(ctx, RS_ERR_CUDA, "injected fault between the two signals (RS_OPT_INJECT_FAULT)");
        const int64_t per = ((((int64_t)np + n - 1) / n) + 15) / 16 * 16;
        const int64_t b = std::min<int64_t>((int64_t)np, k * per), e = std::min<int64_t>((int64_t)np, (k + 1) * per);
        launch_finalize_peers(pa, d_mask, m->out0.as<float4>(), b, e, n_total, p->gamma, st);
        launch_signal(written, st);
        CK(cudaGetLastError());
        if (k == 0) {
            launch_wait_counter(written, expect, st, budget_ns, gave_up);
            CK(cudaMemcpyAsync(out_rgba, m->out0.p, np * sizeof(float4), cudaMemcpyDeviceToHost, st));
        }
        CK(cudaStreamSynchronize(st));
        return fetch_device_stats(ctx, st);
    });
    if (rc == RS_OK) {  // every stream has drained: did a wait kernel run out of patience?
        uint32_t flag = 0;
        cudaSetDevice(m->device[0]);
        if (cudaMemcpy(&flag, gave_up, sizeof(flag), cudaMemcpyDeviceToHost) != cudaSuccess) rc = multi_fail(m, RS_ERR_CUDA, "reading the wait flag failed");
        else if (flag) rc = multi_fail(m, RS_ERR_CUDA, "a device waited longer than RS_OPT_WAIT_BUDGET_MS for its peers");
    }
    if (rc != RS_OK) {
        multi_rebase(m);
        return rc;
    }
    ++m->epoch;
    if (stats_sum) {
        *stats_sum = rs_stats{};
        for (int k = 0; k < n; ++k) {
            rs_stats sk;
            rs_get_stats(m->ctx[k], &sk);
            stats_sum->n_samples += sk.n_samples;
            stats_sum->n_segments += sk.n_segments;
            stats_sum->n_launches += sk.n_launches + 4;
        }
    }
    return RS_OK;
}

// the same counters for callers that run one PROCESS per device (torchrun): the counter lives in memory
// every rank mapped (rs_ipc_open)
int rs_signal(rs_ctx* ctx, uint32_t* d_counter, void* stream) {
    if (!ctx || !d_counter) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "null counter");
    CK(cudaSetDevice(ctx->device));
    launch_signal(d_counter, (cudaStream_t)stream);
    CK(cudaGetLastError());
    return RS_OK;
}
int rs_wait_counter(rs_ctx* ctx, const uint32_t* d_counter, uint32_t expect, void* stream) {
    if (!ctx || !d_counter) return fail(ctx, RS_ERR_INVALID_ARGUMENT, "null counter");
    CK(cudaSetDevice(ctx->device));
    launch_wait_counter(d_counter, expect, (cudaStream_t)stream, (unsigned long long)ctx->wait_budget_ms * 1000000ull, ctx->d_wait_flag.as<uint32_t>());
    CK(cudaGetLastError());
    return RS_OK;
}
int rs_wait_check(rs_ctx* ctx, void* stream) {
    if (!ctx) return RS_ERR_INVALID_ARGUMENT;
    CK(cudaSetDevice(ctx->device));
    uint32_t flag = 0;
    CK(cudaMemcpyAsync(&flag, ctx->d_wait_flag.p, sizeof(flag), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    CK(cudaStreamSynchronize((cudaStream_t)stream));
    if (!flag) return RS_OK;
    CK(cudaMemsetAsync(ctx->d_wait_flag.p, 0, sizeof(flag), (cudaStream_t)stream));
    return fail(ctx, RS_ERR_CUDA, "a wait kernel gave up after RS_OPT_WAIT_BUDGET_MS = " + std::to_string(ctx->wait_budget_ms) +
                                      " ms: a peer never signalled; what was queued behind that wait ran on incomplete data");
}

int rs_get_stats(const rs_ctx* ctx, rs_stats* out) {
    if (!ctx || !out) return RS_ERR_INVALID_ARGUMENT;
    rs_ctx* c = const_cast<rs_ctx*>(ctx);
    if (c->capacity && !c->stats_final) {
        cudaSetDevice(c->device);
        DeviceStats ds;
        if (cudaMemcpy(&ds, c->ws.stats, sizeof(ds), cudaMemcpyDeviceToHost) == cudaSuccess) {
            c->stats.n_samples = ds.n_samples;
            c->stats.n_segments = ds.n_segments;
        }
    }
    *out = c->stats;
    return RS_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------------
// rs_redo_map_static — the redo-map step of the passes loop for callers that must not allocate.
//
// Who that is: a host that has handed the whole card to its own memory pool and leaves the CUDA heap nothing to
// give; a per-frame loop in which a cudaMalloc would stall every stream of the process (cudaMalloc synchronises
// the device); and, in this very library, code that runs between rs_signal and rs_wait_counter on several GPUs —
// with peer access on, cudaMalloc waits for the peer devices, and a peer that is already spinning in a wait
// kernel never answers (see the note in rs_multi_render).  rs_render_passes keeps the intermediate noise map in a
// cudaMalloc'ed context buffer; here it lives in the fixed array below, which belongs to the loaded library, not
// to a context: every GPU that loads the library carries these 16 MiB until the process exits.
//
// `rs_static_noise` as written in HOST code is only a handle for the runtime, not an address a kernel can use;
// the address on a given GPU comes from cudaGetSymbolAddress, asked once per device with that device current.
// ------------------------------------------------------------------------------------------------------------
static_assert(RS_STATIC_MAX_PIXELS == (1 << 22), "the header states 4 Mi pixels");
static __device__ __align__(256) float rs_static_noise[RS_STATIC_MAX_PIXELS];

namespace {
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mutex;
float* g_static_noise_ptr[kStaticMaxDevices] = {};  // device address of rs_static_noise on device k; nullptr = not asked yet

// set-up, once per device (the caller has made `device` current).  The first use of the array on a device may load the
// library's code onto it, which is why that first call belongs before any stream capture.
int static_noise_address(rs_ctx* ctx, int device, float** out) {
    if (device < 0 || device >= kStaticMaxDevices) return fail(ctx, RS_ERR_UNSUPPORTED, "rs_redo_map_static: device ordinal beyond 63");
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_noise_ptr[device]) {
        float* p = nullptr;
        CK(cudaGetSymbolAddress((void**)&p, rs_static_noise));
        if (!p) return fail(ctx, RS_ERR_CUDA, "cudaGetSymbolAddress(rs_static_noise) returned a null address");
        g_static_noise_ptr[device] = p;
    }
    *out = g_static_noise_ptr[device];
    return RS_OK;
}
}  // namespace

extern "C" int rs_redo_map_static(rs_ctx* ctx, const float* d_pixels, uint8_t* d_redo_mask, uint64_t* d_redo_count, int32_t width,
                                  int32_t height, int32_t quirk, float noise_threshold, void* stream) {
    if (!ctx || !d_pixels || !d_redo_mask || !d_redo_count || width <= 0 || height <= 0)
        return fail(ctx, RS_ERR_INVALID_ARGUMENT, "bad rs_redo_map_static arguments");
    const int64_t np = (int64_t)width * height;
    if (np > (int64_t)RS_STATIC_MAX_PIXELS)
        return fail(ctx, RS_ERR_UNSUPPORTED, "rs_redo_map_static: " + std::to_string(width) + " x " + std::to_string(height) + " = " +
                                                 std::to_string(np) + " pixels exceed the fixed scratch of RS_STATIC_MAX_PIXELS = " +
                                                 std::to_string((long long)RS_STATIC_MAX_PIXELS) +
                                                 " pixels; use rs_noise_map with a buffer of your own or rs_render_passes");
    CK(cudaSetDevice(ctx->device));
    float* d_noise = nullptr;
    const int rc = static_noise_address(ctx, ctx->device, &d_noise);
    if (rc != RS_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // the count is summed with atomics, so it has to start from zero; the noise map needs no clearing: its first
    // width*height entries are all rewritten by the first kernel before the second one reads them
    CK(cudaMemsetAsync(d_redo_count, 0, sizeof(uint64_t), st));
    launch_noise(reinterpret_cast<const float4*>(d_pixels), d_noise, width, height, quirk, st);
    launch_redo_mask(d_noise, d_redo_mask, np, noise_threshold, reinterpret_cast<unsigned long long*>(d_redo_count), st);
    CK(cudaGetLastError());
    return RS_OK;
}
