// grt_internal.h -- device data layout shared by the build and trace translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/grt.h"

namespace grt {

constexpr int WIDE = 8;            // children per wide node
#ifndef GRT_LEAF_MAX
#define GRT_LEAF_MAX 1
#endif
constexpr int LEAF_MAX = GRT_LEAF_MAX;  // proxies (or triangles) per leaf child: runs of 1..LEAF_MAX Morton-consecutive records (<= 8)
constexpr uint32_t CHILD_EMPTY = 0xFFFFFFFFu;
constexpr uint32_t CHILD_LEAF = 0x80000000u;  // leaf: CHILD_LEAF | (count-1) << 28 | start (28 bits)
constexpr uint32_t INVALID_ID = 0xFFFFFFFFu;

// Packed per-particle geometry record, 64 B, in Morton (BVH leaf) order.  Four 128-bit loads.
//   q0 = A00 A01 A02 A10   q1 = A11 A12 A20 A21   q2 = A22 mu.x mu.y mu.z   q3 = s opacity id(bits) 0
// A = diag(1/scale) * R^T  (shaders/tracer.cuh:194-201), s = proxy in-sphere radius in Gaussian space
// (src/GaussianTracer.cpp:306).
struct __align__(16) PRec {
    float4 q0, q1, q2, q3;
};
static_assert(sizeof(PRec) == 64, "PRec");

// SH coefficients in the same order: 16 x float3 = 12 float4 per particle (shaders/tracer.cuh:218-221).
constexpr int SH_F4 = 12;

// 8-wide BVH node, SoA so that one warp-uniform 128-bit load fetches 4 children's plane.
struct __align__(16) WideNode {
    float lox[WIDE], loy[WIDE], loz[WIDE], hix[WIDE], hiy[WIDE], hiz[WIDE];
    uint32_t child[WIDE];
    uint32_t pad[WIDE];
};
static_assert(sizeof(WideNode) == 256, "WideNode");

// Triangle record for the reflection primitives, 48 B: v0|mesh, e1|face, e2|0  (e = v1-v0, v2-v0)
struct __align__(16) TriRec {
    float4 a, b, c;
};

// What Params.handle / Params.mesh_handle point at (device memory; the value is validated against the
// context's own handles on the host before launch).
struct AccelHeader {
    uint64_t magic;
    const WideNode* nodes;
    const PRec* recs;     // particles
    const float4* sh;     // particles
    const TriRec* tris;   // meshes
    uint32_t n_prims;
    uint32_t n_nodes;
};
constexpr uint64_t ACCEL_MAGIC_PARTICLES = 0x4752545041525431ull;  // "GRTPART1"
constexpr uint64_t ACCEL_MAGIC_MESH = 0x4752544d45534831ull;       // "GRTMESH1"

// reference's Mesh (src/Parameters.h:36-40)
struct MeshDev {
    uint3* faces;
    float3* vertex_normals;
};

struct DeviceCounters {
    unsigned long long rays, segments, rounds, hits_gathered, hits_shaded, warp_nodes, warp_leaf_prims, warp_rounds,
        stack_overflow, warp_prims_hit, warp_consume_iters, lanes_repass, order_violations, warp_dfs_fallbacks, warp_candidate_iters;
    unsigned long long dbg[8];  // free-form experiment counters (tools/prof_trace.py prints them)
    unsigned long long warp_per_ray_passes, warp_composite_rounds, warp_loop_iters, warp_forced_flushes, warp_dfs_iters, warp_list_full;
};

// Extra kernel arguments next to the 144-byte Params.
struct LaunchExtras {
    const WideNode* p_nodes;
    const PRec* p_recs;
    const float4* p_sh;
    uint32_t p_count;
    const WideNode* m_nodes;
    const TriRec* m_tris;
    uint32_t m_count;
    uint32_t x0, y0, w, h;  // pixel rectangle
    float* rgb32f;
    uint32_t* trace_ids;
    uint32_t* trace_counts;
    uint32_t tx0, ty0, tw, th, trace_max;
    DeviceCounters* counters;
    uint32_t* tile_counter;  // persistent-warp scheduler: next 8x4 tile to hand out (zeroed before the launch)
    // row-band partition of one frame over several GPUs: bands of band_groups 8-row groups; this launch owns the bands b with
    // b % band_stride == band_offset (band_stride <= 1: the whole rectangle)
    uint32_t band_groups, band_stride, band_offset;
    // wavefront mode (scenes with reflection primitives, trace.cu): payload array, tile lists of this / the next bounce level
    uint32_t* wf_state;
    const uint32_t* wf_list_in;
    uint32_t* wf_list_out;
    const uint32_t* wf_count_in;
    uint32_t* wf_count_out;
    uint32_t wf_level;
    uint32_t* wf_list_in_w;   // phase 1 with wf_classify: the list it sorts the tiles into (classes 0 / 1), wf_list_out takes class 2
    uint32_t wf_classify;
};
// Device buffers of the wavefront mode, owned by the context and sized for the largest launch rectangle so far.
constexpr uint32_t WF_MAX_LEVELS = 1056;                 // > TIMEOUT_ITERATIONS + MAX_BOUNCES segments of one ray (tracer.cu:100-104)
constexpr uint32_t WF_CTRL_WORDS = 3u * WF_MAX_LEVELS + 8u;
struct WavefrontBuffers {
    uint32_t* state = nullptr;      // tiles x WF_FIELDS x 32 words
    uint32_t* list[2] = {nullptr, nullptr};
    uint32_t* ctrl = nullptr;       // WF_CTRL_WORDS
    size_t state_bytes = 0, list_bytes = 0;
};
size_t wavefront_state_bytes(const LaunchExtras& ex);
size_t wavefront_list_bytes(const LaunchExtras& ex);

// ---- generic LBVH builder (build.cu) ----
// Scratch arena of a context: the builder's ~20 temporaries are bump-allocated from one device block that is kept between
// builds (cudaMalloc/cudaFree cost more than the build kernels: 13 ms -> ~3 ms per 1M-particle rebuild).  A build that
// needs more than the block holds falls back to cudaMalloc for the overflow and the block grows before the next build.
struct BuildScratch {
    char* base = nullptr;
    size_t cap = 0, off = 0, want = 0;
    void begin();                 // start of a build: grow to what the last build asked for, rewind
    void* get(size_t bytes);      // nullptr when the block is exhausted (caller cudaMallocs)
    void release();               // free the block
};
struct LbvhResult {
    WideNode* nodes = nullptr;  // cudaMalloc'd (nodes_cap entries), or reuse_nodes when the tree fits into it
    uint32_t n_nodes = 0;
    uint32_t nodes_cap = 0;     // capacity of a freshly cudaMalloc'd `nodes` (0 when reuse_nodes was kept)
    uint32_t* perm = nullptr;   // sorted position -> input index, n_live entries valid (n entries; scratch unless perm_owned)
    bool perm_owned = true;     // perm was cudaMalloc'd: the caller frees it
    WideNode* reuse_nodes = nullptr;  // in: a node array of an earlier build the caller is willing to overwrite
    uint32_t reuse_cap = 0;           // in: its capacity in nodes
    uint32_t n_live = 0;
    uint32_t levels = 0;
    float lo[3], hi[3];
};
// d_lo/d_hi: n boxes (float4, w ignored); dead boxes must have lo.x > hi.x (they sort last and are dropped).
// Returns cudaError_t as int; launches counted into *launches.
int lbvh_build(const float4* d_lo, const float4* d_hi, uint32_t n, cudaStream_t stream, LbvhResult* out,
               unsigned long long* launches, BuildScratch* scratch = nullptr);
int radix_sort_pairs(uint64_t* d_keys, uint64_t* d_keys_tmp, uint32_t* d_vals, uint32_t* d_vals_tmp, uint32_t n,
                     cudaStream_t stream, unsigned long long* launches, bool* result_in_tmp, BuildScratch* scratch = nullptr, int passes = 8);

// The same sort without any allocation: temporaries in a file-scope device array of build.cu (at most GRT_NOALLOC_SORT_MAX_N pairs;
// in place; calls that overlap on one GPU race on the array).
#ifndef GRT_NOALLOC_SORT_MAX_N
#define GRT_NOALLOC_SORT_MAX_N (1u << 20)
#endif
int radix_sort_pairs_noalloc(int device, uint64_t* d_keys, uint32_t* d_vals, uint32_t n, cudaStream_t stream, unsigned long long* launches);
uint32_t radix_sort_pairs_noalloc_max();

// particle pack kernels (build.cu)
int particles_compute_boxes(const void* d_particles236, const float* d_s, uint32_t n, float4* d_lo, float4* d_hi,
                            cudaStream_t stream, unsigned long long* launches);
int particles_gather(const void* d_particles236, const float* d_s, const uint32_t* d_perm, uint32_t n_live, PRec* d_recs,
                     float4* d_sh, cudaStream_t stream, unsigned long long* launches);
// raw PLY vertex rows -> records + proxy radii (ingest.cu).  off[k]: byte offset inside a row of the k-th of the 59 properties
// x y z scale_0-2 rot_0-3 opacity f_dc_0-2 f_rest_0-44; aligned: stride and every offset are multiples of 4.
struct PlyLayout {
    uint32_t stride, aligned;
    uint32_t off[59];
};
int particles_activate(const unsigned char* d_rows, const PlyLayout& L, uint32_t n, float alpha_min, float* d_out59, float* d_s,
                       cudaStream_t stream, unsigned long long* launches);
// object-space primitives -> world-space triangle soup + transformed normals on the device (build.cu)
struct MeshPrep {
    float T[16];            // glm column-major 4x4
    uint32_t v_off, i_off;  // first vertex / first index of this mesh in the concatenated arrays
    uint32_t tri_off, pad;  // first triangle
};
int mesh_prepare(const float* d_ov, const float* d_on, const uint32_t* d_idx, const MeshPrep* d_prep, uint32_t n_meshes, uint32_t n_verts,
                 uint32_t n_tris, float* d_vw, float* d_nw, float* d_tri9, uint2* d_meta, cudaStream_t stream, unsigned long long* launches);
// triangle soup: d_tri9 = n x 9 floats (world-space v0,v1,v2), d_meta = n x (mesh, face)
int mesh_compute_boxes(const float* d_tri9, uint32_t n_tris, float4* d_lo, float4* d_hi, cudaStream_t stream,
                       unsigned long long* launches);
int mesh_gather(const float* d_tri9, const uint2* d_meta, const uint32_t* d_perm, uint32_t n, TriRec* out,
                cudaStream_t stream, unsigned long long* launches);

// trace kernel (trace.cu)
// wf == nullptr (or no primitives): one monolithic launch.  *launches is advanced by the kernels launched.
int launch_trace(const grtParams& p, const LaunchExtras& ex, bool counters, cudaStream_t stream, int sm_count, bool* smem_attr_set,
                 const WavefrontBuffers* wf, unsigned long long* launches);

}  // namespace grt
