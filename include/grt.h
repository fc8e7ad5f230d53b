/*
 * grt.h -- C ABI of libgrt.so, the B200 (sm_100a) software Gaussian ray tracer.
 *
 * This is the drop-in boundary for the one hot path of Ray-Studio2/gaussian-ray-tracing: everything the
 * reference host does through OptiX for that path.  Each entry point names the reference call it replaces
 * (file:line relative to the reference repository).  Plain pointers and sizes only; no C++/torch types.
 *
 * Conventions: every function returns 0 on success, non-zero on failure (grtGetLastError() gives the
 * text; the C++ facade turns it into std::runtime_error like the reference's CUDA_CHECK/OPTIX_CHECK,
 * src/Exception.h:31-80).  `stream` arguments are cudaStream_t passed as void* (NULL = default stream).
 * A context is bound to one GPU and is not thread-safe.  There is NO CPU fallback: without a CUDA
 * device grtCreate fails.
 */
#ifndef GRT_H
#define GRT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GRT_VERSION 100

/* Byte-identical to the reference's `struct Params` (src/Parameters.h:42-74): 144 bytes, offsets
 * out@0 width@8 height@12 sh_degree_max@16 eye@20 U@32 V@44 W@56 t_min@68 t_max@72 minTransmittance@76
 * alpha_min@80 handle@88 d_particles@96 mesh_handle@104 mode_fisheye@112 d_meshes@120 type@128 traceState@136. */
typedef struct grtParams {
    void* output_buffer;      /* device uchar3*, width*height, index y*width+x */
    uint32_t width;
    uint32_t height;
    uint32_t sh_degree_max;   /* 0..3 */
    float eye[3];
    float U[3];
    float V[3];
    float W[3];
    float t_min;
    float t_max;
    float minTransmittance;
    float alpha_min;
    uint64_t handle;          /* opaque: from grtBuildParticleAccel (replaces OptixTraversableHandle) */
    void* d_particles;        /* device copy of the 236-byte GaussianParticle records */
    uint64_t mesh_handle;     /* opaque: from grtBuildMeshAccel; 0 = no reflection primitives */
    uint8_t mode_fisheye;     /* C++ bool */
    void* d_meshes;           /* device Mesh{uint3* faces; float3* vertex_normals}[] (src/Parameters.h:36-40) */
    int32_t type;             /* MeshType: 0 MIRROR, 1 NORMAL, 2 GLASS */
    uint32_t* traceState;     /* kept for layout compatibility; unused (state lives in registers) */
} grtParams;

/* One reflection primitive as the reference's `Primitive` holds it (src/geometry/Primitives.h:14-29):
 * object-space vertices/normals, triangle indices and a glm column-major 4x4 transform. */
typedef struct grtMeshDesc {
    const float* vertices;    /* vertex_count x 3 */
    const float* normals;     /* vertex_count x 3 */
    const uint32_t* indices;  /* index_count (3 per triangle) */
    uint32_t vertex_count;
    uint32_t index_count;
    float transform[16];      /* column-major (glm::mat4) */
} grtMeshDesc;

typedef struct grtBuildInfo {
    uint64_t n_particles;     /* records uploaded */
    uint64_t n_live;          /* opacity > alpha_min (the rest can never be hit: SURVEY app. A item 4) */
    uint32_t n_wide_nodes;
    uint32_t wide_depth_levels;
    uint32_t n_mesh_triangles;
    uint32_t n_mesh_wide_nodes;
    float build_ms;           /* device time of the last particle-accel build */
    float mesh_build_ms;
    float scene_lo[3], scene_hi[3];
} grtBuildInfo;

typedef struct grtCounters {
    uint64_t rays;
    uint64_t segments;        /* traceGaussians-equivalents */
    uint64_t rounds;          /* k-buffer gathers (optixTrace-equivalents), per ray */
    uint64_t hits_gathered;   /* h_g */
    uint64_t hits_shaded;     /* h_s */
    uint64_t warp_nodes;      /* wide nodes expanded, per warp */
    uint64_t warp_leaf_prims; /* proxy records tested, per warp (x32 lanes of arithmetic) */
    uint64_t warp_rounds;
    uint64_t stack_overflow;  /* must stay 0 */
    uint64_t kernel_launches; /* launches issued by this context since creation (all kernels) */
    uint64_t warp_prims_hit;  /* proxy records that at least one lane of the packet hit */
    uint64_t warp_consume_iters; /* warp iterations of the shading loop (SIMD efficiency = h_g / (32 x this)) */
    uint64_t lanes_repass;    /* rays that needed another traversal pass because their pending buffer overflowed */
    uint64_t order_violations;/* face hits found behind the already composited front; must stay 0 */
    uint64_t warp_dfs_fallbacks; /* packets whose best-first queue filled up and that redid the pass depth-first */
    uint64_t warp_candidate_iters; /* warp iterations of the per-ray exact proxy test loop */
    uint64_t dbg[8];          /* free-form profiling counters of the counting kernel variant (meaning: DESIGN.md) */
    uint64_t warp_per_ray_passes; /* traversal passes that ran ray by ray (divergent secondary packets, trace_rays_dfs) */
    uint64_t warp_composite_rounds; /* compositing rounds (work list + shade + accumulate), per warp */
    uint64_t warp_loop_iters;  /* iterations of the traversal loop, per warp */
    uint64_t warp_forced_flushes; /* flushes that intersected waiting slots behind the frontier because the slots ran out */
    uint64_t warp_dfs_iters;   /* iterations of the per-ray walk's loop (trace_rays_dfs), per warp */
    uint64_t warp_list_full;   /* record pops that left records in front of the nearest node behind (new-record list full) */
} grtCounters;

typedef struct grtContext grtContext;

/* -- lifetime (replaces optixInit/optixDeviceContextCreate + module/pipeline/SBT creation,
 *    src/GaussianTracer.cpp:72-295: there is nothing to JIT, the kernels are linked in) -- */
int grtCreate(int device, grtContext** ctx_out);
int grtDestroy(grtContext* ctx);
const char* grtGetLastError(const grtContext* ctx); /* ctx may be NULL: error of a failed grtCreate */

/* -- scene (replaces the particle upload of initializeParams, src/GaussianTracer.cpp:491-502) --
 * host_particles: n records of 236 bytes (GaussianParticle, src/GaussianData.h:12-20).
 * alpha_min sizes the proxies: s = sqrt(2 ln(opacity/alpha_min)) (src/GaussianTracer.cpp:306). */
int grtSetParticles(grtContext* ctx, const void* host_particles, uint64_t n, float alpha_min);

/* Device-side ingest (replaces the activation loop of GaussianData::parse, src/GaussianData.cpp:97-131, and the proxy-size
 * loop of src/GaussianTracer.cpp:304-306): host_rows = n raw PLY vertex rows of row_stride bytes as they lie in the file,
 * prop_offsets59[k] = byte offset inside a row of the k-th of the float properties x y z scale_0-2 rot_0-3 opacity
 * f_dc_0-2 f_rest_0-44.  exp / sigmoid / quaternion normalisation / SH re-ordering run in one CUDA kernel; the resident
 * result is the same 236-byte record array grtSetParticles would have uploaded. */
int grtSetParticlesRaw(grtContext* ctx, const void* host_rows, uint64_t n, uint32_t row_stride, const uint32_t* prop_offsets59, float alpha_min);
/* GaussianData::GaussianData(path) + parse (src/GaussianData.cpp:8-132) in one call: own binary-LE reader (mmap, no
 * column copies), then grtSetParticlesRaw.  center3_out = GaussianData::getCenter (:140-152).  Errors like the
 * reference: unreadable file, missing property (happly.h:857). */
int grtLoadPly(grtContext* ctx, const char* path, float alpha_min, uint64_t* n_out, float* center3_out);
/* GaussianData::getParticles equivalent: downloads the first n resident 236-byte records. */
int grtGetParticles(grtContext* ctx, void* host_out236, uint64_t n);

/* replaces createGaussianParticlesBVH: createGAS + N x createIAS + optixAccelBuild
 * (src/GaussianTracer.cpp:297-317, 379, 401-473).  LBVH on the device.  handle_out -> Params.handle,
 * d_particles_out -> Params.d_particles. */
int grtBuildParticleAccel(grtContext* ctx, void* stream, uint64_t* handle_out, void** d_particles_out);

/* replaces createGAS/createIAS/buildAccelationStructure for meshes + sendGeometryAttributesToDevice
 * (src/GaussianTracer.cpp:578-709, 711-794).  n == 0 clears them (removePrimitive, :553-566):
 * mesh_handle_out = 0, d_meshes_out = NULL. */
int grtSetMeshes(grtContext* ctx, const grtMeshDesc* meshes, uint32_t n);
int grtBuildMeshAccel(grtContext* ctx, void* stream, uint64_t* mesh_handle_out, void** d_meshes_out);

/* -- render (replaces cudaMemcpyAsync(d_params) + optixLaunch(w,h,1), src/GaussianTracer.cpp:517-534) --
 * Asynchronous on `stream`.  host_params is read before returning. */
int grtLaunch(grtContext* ctx, void* stream, const grtParams* host_params);
/* Only the pixel rectangle [x0,x0+w) x [y0,y0+h) of the frame is traced (multi-GPU tile partition). */
int grtLaunchTile(grtContext* ctx, void* stream, const grtParams* host_params, uint32_t x0, uint32_t y0, uint32_t w, uint32_t h);
/* One frame split over several GPUs in ONE launch per GPU: the frame's rows are cut into bands of band_rows rows (a multiple of
 * 8, the packet grid) and this launch traces the bands b with b % band_stride == band_offset (rank r of G: stride G, offset r).
 * The union over the offsets 0..band_stride-1 is bit-identical to grtLaunch's frame.  (SURVEY 8e "Partitioning"; the reference is
 * single-GPU, src/GaussianTracer.h:50.) */
int grtLaunchRows(grtContext* ctx, void* stream, const grtParams* host_params, uint32_t band_rows, uint32_t band_stride, uint32_t band_offset);
/* n cameras, one full frame each (output_buffer per entry). */
int grtLaunchBatch(grtContext* ctx, void* stream, const grtParams* host_params_array, uint32_t n);
/* End-to-end convenience with HOST output: traces rows [y0,y0+h) into a context-owned device frame,
 * copies those rows to host_rgb8 + 3*width*y0 (pinned or pageable) and synchronises.
 * host_params->output_buffer is ignored. h == 0 means the full frame. */
int grtRenderHost(grtContext* ctx, const grtParams* host_params, uint8_t* host_rgb8, uint32_t y0, uint32_t h);

/* -- multi-GPU (the reference is single-GPU, src/GaussianTracer.h:50): the scene and its accel are replicated per GPU,
 *    rays are partitioned (cameras k mod G, or row bands of one frame: grtLaunchRows) and the ONLY exchange is the gather of RGB8
 *    frame data to a root rank -- grouped ncclSend/ncclRecv over NVLink, asynchronous on the caller's stream.  NCCL is bound at
 *    run time (dlopen of libnccl.so.2); one context = one rank.  The 128-byte id comes from grtCommGetUniqueId on one rank and
 *    reaches the others by any host channel (file, socket, MPI, torch.distributed broadcast). -- */
int grtCommGetUniqueId(uint8_t out128[128]);
int grtCommInit(grtContext* ctx, int nranks, int rank, const uint8_t id128[128]);
int grtCommDestroy(grtContext* ctx);
/* every rank: `bytes` from d_send; root: rank r's block lands at d_recv_all + r*bytes (non-roots pass NULL) */
int grtGatherFrames(grtContext* ctx, void* stream, const void* d_send, size_t bytes, void* d_recv_all, int root);
/* row bands of one width x height RGB8 frame: band b (band_rows rows, a multiple of 8) was traced by rank b % nranks into ITS d_frame;
 * afterwards the root's d_frame holds the whole frame */
int grtGatherBands(grtContext* ctx, void* stream, void* d_frame, uint32_t width, uint32_t height, uint32_t band_rows, int root);

/* -- parity / profiling side channels (not in the reference; Params layout untouched) --
 * d_rgb32f: width*height*3 floats, accumColor before clamp+quantise (NULL disables).
 * Hit-ID traces for the pixel rectangle (tx0,ty0,tw,th): the first trace_max consumed particle IDs per
 * pixel go to d_trace_ids[((y-ty0)*tw+(x-tx0))*trace_max + k], the total count to d_trace_counts[...]. */
int grtSetDebugOutputs(grtContext* ctx, float* d_rgb32f, uint32_t* d_trace_ids, uint32_t* d_trace_counts,
                       uint32_t tx0, uint32_t ty0, uint32_t tw, uint32_t th, uint32_t trace_max);
/* sizeof / offsetof of grtParams as compiled into this library, in the order sizeof, output_buffer, width, height, sh_degree_max, eye,
 * U, V, W, t_min, t_max, minTransmittance, alpha_min, handle, d_particles, mesh_handle, mode_fisheye, d_meshes, type, traceState,
 * then the particle record size (236) and sizeof(Mesh).  tests/test_ref_pins.py compares them with the numbers the REFERENCE's own
 * src/Parameters.h gives (oracle/_ref).  Needs no GPU. */
int grtGetAbiLayout(uint64_t out32[32]);
int grtEnableCounters(grtContext* ctx, int enable); /* selects the counting kernel variant */
int grtGetCounters(grtContext* ctx, grtCounters* out, int reset);
/* Scenes with reflection primitives are traced level by level (wavefront: per bounce one launch for traceMesh + closest hit,
 * shaders/tracer.cu:155-187, and one for the Gaussian segment, shaders/tracer.cuh:328-394) instead of by one launch that runs
 * the whole bounce loop of shaders/tracer.cu:58-106, for the first bounce level(s); what is still alive after them finishes in
 * one more launch.  Same pixels either way; enable = 0 selects the single launch.  Asynchronous like grtLaunch.  The per-pixel
 * payload (30 words per pixel of the launch rectangle) lives in ONE context-owned buffer: launches of a scene with primitives
 * must not overlap on the GPU (same stream, or ordered by events); scenes without primitives are not affected. */
int grtSetWavefront(grtContext* ctx, int enable);
int grtGetBuildInfo(grtContext* ctx, grtBuildInfo* out);
/* Walks the built wide BVH on the host: every live record referenced exactly once, every leaf box inside
 * its ancestors' boxes.  Returns 0 if valid. which: 0 particles, 1 meshes. */
int grtValidateAccel(grtContext* ctx, int which);
/* Last kernel's device time (cudaEvent pair recorded around the most recent grtLaunch*), ms. */
int grtLastLaunchMs(grtContext* ctx, float* ms_out);

/* -- small device-memory helpers so that a C/ctypes caller needs no other CUDA binding -- */
int grtMalloc(grtContext* ctx, size_t bytes, void** d_ptr_out);
int grtFree(grtContext* ctx, void* d_ptr);
int grtMallocHost(grtContext* ctx, size_t bytes, void** h_ptr_out); /* pinned */
int grtFreeHost(grtContext* ctx, void* h_ptr);
int grtMemcpyDtoH(grtContext* ctx, void* dst_host, const void* src_dev, size_t bytes, void* stream);
int grtMemcpyHtoD(grtContext* ctx, void* dst_dev, const void* src_host, size_t bytes, void* stream);
int grtMemset(grtContext* ctx, void* d_ptr, int value, size_t bytes, void* stream);
int grtStreamCreate(grtContext* ctx, void** stream_out);
int grtStreamDestroy(grtContext* ctx, void* stream);
int grtStreamSynchronize(grtContext* ctx, void* stream);
/* Sorts n (key,value) pairs with the hand-written device radix sort used by the LBVH build (test hook). */
int grtDebugRadixSort(grtContext* ctx, uint64_t* host_keys, uint32_t* host_vals, uint64_t n);

/* Allocation-free variant of the (key, value) radix sort above.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops).
 * Sorts n pairs IN PLACE in the caller's device buffers d_keys[n] (uint64) / d_vals[n] (uint32), stable, asynchronously on
 * `stream`.  The sort's temporaries (ping-pong buffers, histograms) live in a file-scope device array of the library
 * (13.1 MB, held by every GPU that loads the library until exit), so n is limited to GRT_NOALLOC_SORT_MAX_N = 1 048 576
 * pairs: a larger n is rejected with an error.  Calls no cudaMalloc / cudaFree.  All calls on one GPU share the array:
 * calls that overlap on the GPU (different streams) race; back-to-back calls on one stream are fine. */
#define GRT_NOALLOC_SORT_MAX_N (1u << 20)
int grtSortPairsNoAlloc(grtContext* ctx, uint64_t* d_keys, uint32_t* d_vals, uint64_t n, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GRT_H */
