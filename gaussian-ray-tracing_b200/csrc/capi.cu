// capi.cu -- implementation of the C ABI declared in include/grt.h (the drop-in boundary).
// Host-side only: owns device memory, validates handles, launches the kernels of build.cu / trace.cu.
// There is no CPU fallback anywhere in this library.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <dlfcn.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "grt_internal.h"

using namespace grt;

static_assert(sizeof(grtParams) == 144, "grtParams must match the reference Params (144 bytes)");
static_assert(offsetof(grtParams, width) == 8 && offsetof(grtParams, eye) == 20 && offsetof(grtParams, U) == 32 &&
                  offsetof(grtParams, V) == 44 && offsetof(grtParams, W) == 56 && offsetof(grtParams, t_min) == 68 &&
                  offsetof(grtParams, alpha_min) == 80 && offsetof(grtParams, handle) == 88 && offsetof(grtParams, d_particles) == 96 &&
                  offsetof(grtParams, mesh_handle) == 104 && offsetof(grtParams, mode_fisheye) == 112 &&
                  offsetof(grtParams, d_meshes) == 120 && offsetof(grtParams, type) == 128 && offsetof(grtParams, traceState) == 136,
              "grtParams offsets must match src/Parameters.h:42-74");

struct grtContext {
    int device = 0;
    std::string last_error;
    unsigned long long launches = 0;
    // particles
    uint64_t n_particles = 0;
    float alpha_min_build = 0.01f;
    void* d_particles = nullptr;  // 236-byte records
    float* d_s = nullptr;         // proxy in-sphere radius per particle (0 = culled)
    WideNode* p_nodes = nullptr;
    PRec* p_recs = nullptr;
    float4* p_sh = nullptr;
    AccelHeader* d_p_header = nullptr;
    uint32_t p_live = 0, p_nnodes = 0, p_levels = 0;
    uint32_t p_nodes_cap = 0, p_recs_cap = 0;  // capacities of p_nodes / p_recs+p_sh: rebuilds reuse the arrays
    uint64_t particles_cap = 0;                // capacity (records) of d_particles / d_s
    bool p_valid = false;                      // the particle accel matches the resident particle set
    // meshes
    std::vector<float> h_ov, h_on;              // object-space vertices / normals of all meshes, concatenated (as grtSetMeshes got them)
    std::vector<uint32_t> h_idx;                // triangle indices of all meshes, concatenated (= the reference's Mesh.faces arrays)
    std::vector<MeshPrep> h_prep;               // per mesh: transform and offsets; world-space data is produced on the device
    std::vector<void*> mesh_allocs;             // mesh device arrays that did not fit into mesh_pool (freed on the next build)
    BuildScratch mesh_pool;                     // faces / normals / Mesh table / triangle records / header of the current mesh accel
    uint32_t m_nodes_cap = 0;                   // capacity of m_nodes (kept across rebuilds)
    MeshDev* d_meshes = nullptr;
    WideNode* m_nodes = nullptr;
    TriRec* m_tris = nullptr;
    AccelHeader* d_m_header = nullptr;
    uint32_t m_count = 0, m_nnodes = 0;
    // debug / counters
    float* dbg_rgb32f = nullptr;
    uint32_t* dbg_trace_ids = nullptr;
    uint32_t* dbg_trace_counts = nullptr;
    uint32_t tx0 = 0, ty0 = 0, tw = 0, th = 0, trace_max = 0;
    bool counters_on = false;
    DeviceCounters* d_counters = nullptr;
    uint32_t* d_tile_counter = nullptr;  // TILE_RING x TILE_SLOT_WORDS: one scheduler counter block per launch in flight
    uint32_t tile_seq = 0;
    WavefrontBuffers wf;                 // wavefront mode of scenes with primitives (trace.cu), grown on demand
    bool wavefront = true;               // grtSetWavefront; GRT_WAVEFRONT=0 in the environment turns it off at grtCreate
    bool smem_attr_set = false;          // cudaFuncAttributeMaxDynamicSharedMemorySize is per DEVICE: kept per context, not per process
    BuildScratch scratch;  // builder temporaries, kept between builds
    int sm_count = 0;
    // e2e frame
    uint8_t* d_frame = nullptr;
    size_t frame_bytes = 0;
    cudaStream_t own_stream = nullptr;
    cudaEvent_t ev_b0 = nullptr, ev_b1 = nullptr;  // build timing
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool ev_valid = false;
    grtBuildInfo info{};
    // multi-GPU exchange (framebuffer gather): an NCCL communicator of one rank per GPU, bound through dlopen
    void* comm = nullptr;
    int comm_rank = 0, comm_nranks = 1;
};

// ---- NCCL, bound at run time (libnccl.so.2: the copy already loaded into the process -- torch's -- or the system one), so that
// libgrt.so loads on machines without NCCL and single-GPU users pay nothing.  Only the calls the framebuffer gather needs.
namespace {
struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(void*) = nullptr;
    void* CommInitRank = nullptr;   // int ncclCommInitRank(ncclComm_t*, int, ncclUniqueId /* 128-byte struct BY VALUE */, int): cast at the call
    int (*CommDestroy)(void*) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*Send)(const void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
};
struct NcclId { char b[128]; };
NcclApi g_nccl;
const char* nccl_bind() {
    if (g_nccl.lib) return nullptr;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return "libnccl.so.2 not found (multi-GPU gather needs NCCL)";
    auto sym = [&](const char* n) { return dlsym(h, n); };
    g_nccl.GetUniqueId = (int (*)(void*))sym("ncclGetUniqueId");
    g_nccl.CommInitRank = sym("ncclCommInitRank");
    g_nccl.CommDestroy = (int (*)(void*))sym("ncclCommDestroy");
    g_nccl.GroupStart = (int (*)())sym("ncclGroupStart");
    g_nccl.GroupEnd = (int (*)())sym("ncclGroupEnd");
    g_nccl.Send = (int (*)(const void*, size_t, int, int, void*, cudaStream_t))sym("ncclSend");
    g_nccl.Recv = (int (*)(void*, size_t, int, int, void*, cudaStream_t))sym("ncclRecv");
    g_nccl.GetErrorString = (const char* (*)(int))sym("ncclGetErrorString");
    g_nccl.GetVersion = (int (*)(int*))sym("ncclGetVersion");
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.CommDestroy || !g_nccl.GroupStart || !g_nccl.GroupEnd || !g_nccl.Send || !g_nccl.Recv)
        return "libnccl.so.2 lacks a required symbol";
    g_nccl.lib = h;
    return nullptr;
}
}  // namespace
#define NC(ctx, call)                                                                                                          \
    do {                                                                                                                       \
        int r_ = (call);                                                                                                       \
        if (r_) return fail(ctx, "%s failed: %s (%s:%d)", #call, g_nccl.GetErrorString ? g_nccl.GetErrorString(r_) : "nccl error", __FILE__, __LINE__); \
    } while (0)

static std::string g_create_error;

static int fail(grtContext* ctx, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (ctx) ctx->last_error = buf; else g_create_error = buf;
    return 1;
}
#define CU(ctx, call)                                                                              \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(ctx, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)
#define DEV(ctx) CU(ctx, cudaSetDevice((ctx)->device))  /* a context is bound to one GPU; the caller's current device may be another */
constexpr uint32_t TILE_RING = 64, TILE_SLOT_WORDS = 256;
#ifdef GRT_SM_REGIONS
#define GRT_SM_REGIONS_HOST GRT_SM_REGIONS
#else
#define GRT_SM_REGIONS_HOST 0
#endif
#define RC(ctx, call)                                                                                            \
    do {                                                                                                         \
        int r_ = (call);                                                                                         \
        if (r_) return fail(ctx, "%s failed: %s (%s:%d)", #call, cudaGetErrorString((cudaError_t)r_), __FILE__, __LINE__); \
    } while (0)

static void free_particle_accel(grtContext* c) {
    cudaFree(c->p_nodes); cudaFree(c->p_recs); cudaFree(c->p_sh); cudaFree(c->d_p_header);
    c->p_nodes = nullptr; c->p_recs = nullptr; c->p_sh = nullptr; c->d_p_header = nullptr;
    c->p_live = c->p_nnodes = c->p_levels = 0;
    c->p_nodes_cap = c->p_recs_cap = 0;
    c->p_valid = false;
}
// d_particles / d_s hold at least n records (kept when large enough: per-frame particle updates do not reallocate)
static int reserve_particles(grtContext* c, uint64_t n) {
    if (n <= c->particles_cap && c->d_particles && c->d_s) return 0;
    cudaFree(c->d_particles); cudaFree(c->d_s);
    c->d_particles = nullptr; c->d_s = nullptr; c->particles_cap = 0;
    cudaError_t e = cudaMalloc(&c->d_particles, n * 236ull);
    if (e == cudaSuccess) e = cudaMalloc(&c->d_s, n * sizeof(float));
    if (e != cudaSuccess) return (int)e;
    c->particles_cap = n;
    return 0;
}
static void free_mesh_accel(grtContext* c) {
    // the arrays live in mesh_pool (or in mesh_allocs when the pool was too small); the node array is kept for the next build
    for (void* p : c->mesh_allocs) cudaFree(p);
    c->mesh_allocs.clear();
    c->m_tris = nullptr; c->d_m_header = nullptr; c->d_meshes = nullptr;
    c->m_count = c->m_nnodes = 0;
}

extern "C" {

int grtCreate(int device, grtContext** ctx_out) {
    if (!ctx_out) return fail(nullptr, "grtCreate: ctx_out is NULL");
    *ctx_out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(nullptr, "grtCreate: no CUDA device (%s); libgrt has no CPU fallback", cudaGetErrorString(e));
    if (device < 0 || device >= count) return fail(nullptr, "grtCreate: device %d out of range (0..%d)", device, count - 1);
    CU(nullptr, cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(nullptr, cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(nullptr, "grtCreate: device %d is sm_%d%d; libgrt is built for sm_100a only", device, prop.major, prop.minor);
    grtContext* c = new grtContext();
    c->device = device;
    { const char* e = getenv("GRT_WAVEFRONT"); if (e && e[0] == '0') c->wavefront = false; }
    c->sm_count = prop.multiProcessorCount;
    if (cudaMalloc(&c->d_tile_counter, TILE_RING * TILE_SLOT_WORDS * sizeof(uint32_t)) != cudaSuccess || cudaMalloc(&c->d_counters, sizeof(DeviceCounters)) != cudaSuccess || cudaMemset(c->d_counters, 0, sizeof(DeviceCounters)) != cudaSuccess ||
        cudaStreamCreate(&c->own_stream) != cudaSuccess || cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess) {
        delete c;
        return fail(nullptr, "grtCreate: device allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
    }
    *ctx_out = c;
    return 0;
}

int grtDestroy(grtContext* c) {
    if (!c) return 0;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    free_particle_accel(c);
    free_mesh_accel(c);
    cudaFree(c->d_particles); cudaFree(c->d_s); cudaFree(c->d_counters); cudaFree(c->d_tile_counter); cudaFree(c->d_frame);
    cudaFree(c->wf.state); cudaFree(c->wf.list[0]); cudaFree(c->wf.list[1]); cudaFree(c->wf.ctrl);
    c->scratch.release();
    c->mesh_pool.release();
    cudaFree(c->m_nodes);
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    if (c->ev_b0) cudaEventDestroy(c->ev_b0);
    if (c->ev_b1) cudaEventDestroy(c->ev_b1);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    delete c;
    return 0;
}

const char* grtGetLastError(const grtContext* ctx) { return ctx ? ctx->last_error.c_str() : g_create_error.c_str(); }

// ---------------------------------------------------------------- memory helpers
int grtMalloc(grtContext* c, size_t bytes, void** out) {
    if (!c || !out) return fail(c, "grtMalloc: bad argument");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMalloc(out, bytes ? bytes : 1));
    return 0;
}
int grtFree(grtContext* c, void* p) { if (!c) return 1; DEV(c); CU(c, cudaFree(p)); return 0; }
int grtMallocHost(grtContext* c, size_t bytes, void** out) {
    if (!c || !out) return fail(c, "grtMallocHost: bad argument");
    DEV(c);
    CU(c, cudaMallocHost(out, bytes ? bytes : 1));
    return 0;
}
int grtFreeHost(grtContext* c, void* p) { if (!c) return 1; DEV(c); CU(c, cudaFreeHost(p)); return 0; }
int grtMemcpyDtoH(grtContext* c, void* dst, const void* src, size_t bytes, void* stream) {
    if (!c) return 1;
    DEV(c);
    CU(c, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    CU(c, cudaStreamSynchronize((cudaStream_t)stream));
    return 0;
}
int grtMemcpyHtoD(grtContext* c, void* dst, const void* src, size_t bytes, void* stream) {
    if (!c) return 1;
    DEV(c);
    CU(c, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    CU(c, cudaStreamSynchronize((cudaStream_t)stream));
    return 0;
}
int grtMemset(grtContext* c, void* p, int value, size_t bytes, void* stream) {
    if (!c) return 1;
    DEV(c);
    CU(c, cudaMemsetAsync(p, value, bytes, (cudaStream_t)stream));
    return 0;
}
int grtStreamCreate(grtContext* c, void** out) {
    if (!c || !out) return 1;
    cudaStream_t s;
    DEV(c);
    CU(c, cudaStreamCreate(&s));
    *out = (void*)s;
    return 0;
}
int grtStreamDestroy(grtContext* c, void* s) { if (!c) return 1; DEV(c); CU(c, cudaStreamDestroy((cudaStream_t)s)); return 0; }
int grtStreamSynchronize(grtContext* c, void* s) { if (!c) return 1; DEV(c); CU(c, cudaStreamSynchronize((cudaStream_t)s)); return 0; }

// ---------------------------------------------------------------- particles
int grtSetParticles(grtContext* c, const void* host_particles, uint64_t n, float alpha_min) {
    if (!c) return 1;
    if (n && !host_particles) return fail(c, "grtSetParticles: host_particles is NULL");
    if (n >= (1ull << 28)) return fail(c, "grtSetParticles: %llu particles exceed the 2^28 limit", (unsigned long long)n);
    CU(c, cudaSetDevice(c->device));
    c->p_valid = false;  // the accel (if any) belongs to the previous particle set; its arrays are reused by the next build
    c->n_particles = n;
    c->alpha_min_build = alpha_min;
    if (n == 0) return 0;
    // Proxy size, exactly as the reference host computes it (src/GaussianTracer.cpp:304-306):
    //   s = sqrtf(2 ln(opacity / alpha_min)); opacity <= alpha_min gives 0/NaN -> the instance can never be hit
    //    -> s = 0 marks the particle culled.
    std::vector<float> s(n);
    const float* raw = reinterpret_cast<const float*>(host_particles);
    for (uint64_t i = 0; i < n; ++i) {
        const float opacity = raw[i * 59 + 10];
        float v = 0.0f;
        if (opacity > alpha_min) {
            v = sqrtf(2.0f * logf(opacity / alpha_min));
            if (!(v > 0.0f) || !std::isfinite(v)) v = 0.0f;
        }
        s[i] = v;
    }
    RC(c, reserve_particles(c, n));
    CU(c, cudaMemcpy(c->d_particles, host_particles, n * 236ull, cudaMemcpyHostToDevice));
    CU(c, cudaMemcpy(c->d_s, s.data(), n * sizeof(float), cudaMemcpyHostToDevice));
    return 0;
}

// ---------------------------------------------------------------- device-side PLY ingest
static const char* const kPlyProps[59] = {
    "x", "y", "z", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3", "opacity", "f_dc_0", "f_dc_1", "f_dc_2",
    "f_rest_0", "f_rest_1", "f_rest_2", "f_rest_3", "f_rest_4", "f_rest_5", "f_rest_6", "f_rest_7", "f_rest_8", "f_rest_9",
    "f_rest_10", "f_rest_11", "f_rest_12", "f_rest_13", "f_rest_14", "f_rest_15", "f_rest_16", "f_rest_17", "f_rest_18", "f_rest_19",
    "f_rest_20", "f_rest_21", "f_rest_22", "f_rest_23", "f_rest_24", "f_rest_25", "f_rest_26", "f_rest_27", "f_rest_28", "f_rest_29",
    "f_rest_30", "f_rest_31", "f_rest_32", "f_rest_33", "f_rest_34", "f_rest_35", "f_rest_36", "f_rest_37", "f_rest_38", "f_rest_39",
    "f_rest_40", "f_rest_41", "f_rest_42", "f_rest_43", "f_rest_44"};

int grtSetParticlesRaw(grtContext* c, const void* host_rows, uint64_t n, uint32_t row_stride, const uint32_t* off59, float alpha_min) {
    if (!c) return 1;
    if (n && (!host_rows || !off59)) return fail(c, "grtSetParticlesRaw: NULL argument");
    if (n >= (1ull << 28)) return fail(c, "grtSetParticlesRaw: %llu particles exceed the 2^28 limit", (unsigned long long)n);
    if (n && (row_stride < 59 * 4 || row_stride > 1024)) return fail(c, "grtSetParticlesRaw: row stride %u outside [236, 1024]", row_stride);
    PlyLayout L{};
    L.stride = row_stride;
    L.aligned = (row_stride & 3u) == 0;
    for (int k = 0; k < 59 && n; ++k) {
        if (off59[k] + 4u > row_stride) return fail(c, "grtSetParticlesRaw: property %s at byte %u does not fit a %u-byte row", kPlyProps[k], off59[k], row_stride);
        L.off[k] = off59[k];
        if (off59[k] & 3u) L.aligned = 0;
    }
    CU(c, cudaSetDevice(c->device));
    c->p_valid = false;
    c->n_particles = n;
    c->alpha_min_build = alpha_min;
    if (n == 0) return 0;
    unsigned char* d_rows = nullptr;
    CU(c, cudaMalloc(&d_rows, n * (size_t)row_stride));
    cudaError_t e = (cudaError_t)reserve_particles(c, n);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_rows, host_rows, n * (size_t)row_stride, cudaMemcpyHostToDevice, c->own_stream);
    int rc = 0;
    if (e == cudaSuccess) rc = particles_activate(d_rows, L, (uint32_t)n, alpha_min, reinterpret_cast<float*>(c->d_particles), c->d_s, c->own_stream, &c->launches);
    if (e == cudaSuccess && rc == 0) e = cudaStreamSynchronize(c->own_stream);
    cudaFree(d_rows);
    if (e != cudaSuccess || rc != 0) {
        c->n_particles = 0;
        return fail(c, "grtSetParticlesRaw: %s", cudaGetErrorString(e != cudaSuccess ? e : (cudaError_t)rc));
    }
    return 0;
}

static int ply_type_size(const std::string& t) {
    if (t == "char" || t == "uchar" || t == "int8" || t == "uint8") return 1;
    if (t == "short" || t == "ushort" || t == "int16" || t == "uint16") return 2;
    if (t == "int" || t == "uint" || t == "float" || t == "int32" || t == "uint32" || t == "float32") return 4;
    if (t == "double" || t == "float64") return 8;
    return 0;
}

// Own reader for the file layout GaussianData::parse accepts (src/GaussianData.cpp:25-95): binary little-endian, a "vertex"
// element with the 59 named float properties in any order, extra properties ignored, missing ones are an error (happly.h:857).
int grtLoadPly(grtContext* c, const char* path, float alpha_min, uint64_t* n_out, float* center3_out) {
    if (!c) return 1;
    if (!path) return fail(c, "grtLoadPly: path is NULL");
    const int fd = open(path, O_RDONLY);
    if (fd < 0) return fail(c, "grtLoadPly: cannot open %s", path);
    struct stat st;
    if (fstat(fd, &st) != 0 || st.st_size <= 0) { close(fd); return fail(c, "grtLoadPly: cannot stat %s", path); }
    const size_t fsize = (size_t)st.st_size;
    void* map = mmap(nullptr, fsize, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (map == MAP_FAILED) return fail(c, "grtLoadPly: mmap of %s failed", path);
    const char* f = static_cast<const char*>(map);
    struct Unmap { void* p; size_t n; ~Unmap() { munmap(p, n); } } unmap{map, fsize};
    // ---- header
    size_t pos = 0;
    auto next_line = [&](std::string& line) -> bool {
        if (pos >= fsize) return false;
        size_t e = pos;
        while (e < fsize && f[e] != '\n') ++e;
        if (e >= fsize) return false;
        line.assign(f + pos, e - pos);
        if (!line.empty() && line.back() == '\r') line.pop_back();
        pos = e + 1;
        return true;
    };
    std::string line;
    if (!next_line(line) || line != "ply") return fail(c, "grtLoadPly: %s is not a PLY file", path);
    bool fmt_ok = false, in_vertex = false, seen_vertex = false, done = false;
    uint64_t n = 0, skip_before = 0, cur_count = 0;
    uint32_t stride = 0, cur_stride = 0;
    uint32_t off[59];
    bool found[59] = {false};
    while (next_line(line)) {
        char a[64] = {0}, b[64] = {0}, d[64] = {0};
        const int k = sscanf(line.c_str(), "%63s %63s %63s", a, b, d);
        if (k <= 0) continue;
        const std::string w = a;
        if (w == "end_header") { done = true; break; }
        if (w == "comment" || w == "obj_info") continue;
        if (w == "format") {
            if (std::string(b) != "binary_little_endian") return fail(c, "grtLoadPly: only binary_little_endian PLY is supported (got %s)", b);
            fmt_ok = true;
        } else if (w == "element") {
            if (!seen_vertex) skip_before += cur_count * (uint64_t)cur_stride;
            in_vertex = std::string(b) == "vertex";
            cur_count = strtoull(d, nullptr, 10);
            cur_stride = 0;
            if (in_vertex) { seen_vertex = true; n = cur_count; }
        } else if (w == "property") {
            if (std::string(b) == "list") {
                if (in_vertex || !seen_vertex) return fail(c, "grtLoadPly: list properties in or before the vertex element are not supported");
                continue;
            }
            const int sz = ply_type_size(b);
            if (!sz) return fail(c, "grtLoadPly: unknown property type %s", b);
            if (in_vertex) {
                for (int q = 0; q < 59; ++q)
                    if (!found[q] && std::string(d) == kPlyProps[q]) {
                        if (sz != 4 || !(std::string(b) == "float" || std::string(b) == "float32"))
                            return fail(c, "grtLoadPly: property %s must be float", d);
                        found[q] = true; off[q] = stride; break;
                    }
                stride += (uint32_t)sz;
            } else {
                cur_stride += (uint32_t)sz;
            }
        }
    }
    if (!done || !fmt_ok || !seen_vertex) return fail(c, "grtLoadPly: malformed header in %s", path);
    for (int q = 0; q < 59; ++q)
        if (!found[q]) return fail(c, "grtLoadPly: vertex property %s not found", kPlyProps[q]);  // happly.h:857 throws here
    const size_t data0 = pos + (size_t)skip_before;
    if (data0 + (size_t)n * stride > fsize) return fail(c, "grtLoadPly: %s is truncated (%llu vertices of %u bytes expected)", path, (unsigned long long)n, stride);
    const unsigned char* rows = reinterpret_cast<const unsigned char*>(f) + data0;
    int rc = grtSetParticlesRaw(c, rows, n, stride, off, alpha_min);
    if (rc) return rc;
    if (n_out) *n_out = n;
    if (center3_out) {  // GaussianData::getCenter (src/GaussianData.cpp:140-152): sequential float sums in vertex order
        float cx = 0.f, cy = 0.f, cz = 0.f;
        for (uint64_t i = 0; i < n; ++i) {
            float v[3];
            memcpy(&v[0], rows + i * stride + off[0], 4); memcpy(&v[1], rows + i * stride + off[1], 4); memcpy(&v[2], rows + i * stride + off[2], 4);
            cx += v[0]; cy += v[1]; cz += v[2];
        }
        const float fn = static_cast<float>(n);
        center3_out[0] = cx / fn; center3_out[1] = cy / fn; center3_out[2] = cz / fn;
    }
    return 0;
}

int grtGetParticles(grtContext* c, void* host_out236, uint64_t n) {
    if (!c) return 1;
    if (n > c->n_particles) return fail(c, "grtGetParticles: %llu requested, %llu resident", (unsigned long long)n, (unsigned long long)c->n_particles);
    if (n && !host_out236) return fail(c, "grtGetParticles: host buffer is NULL");
    DEV(c);
    if (n) CU(c, cudaMemcpy(host_out236, c->d_particles, n * 236ull, cudaMemcpyDeviceToHost));
    return 0;
}

int grtBuildParticleAccel(grtContext* c, void* stream_v, uint64_t* handle_out, void** d_particles_out) {
    if (!c) return 1;
    cudaStream_t stream = (cudaStream_t)stream_v;
    CU(c, cudaSetDevice(c->device));
    // A rebuild (dynamic particle sets: grtSetParticles* + grtBuildParticleAccel per frame) keeps the node / record / SH
    // arrays of the previous accel when they are large enough and takes its temporaries from the context's scratch block.
    c->p_valid = false;
    c->p_live = c->p_nnodes = c->p_levels = 0;
    if (c->ev_valid) CU(c, cudaStreamWaitEvent(stream, c->ev1, 0));  // the node / record arrays are reused: wait for the last launch that read them
    if (handle_out) *handle_out = 0;
    if (d_particles_out) *d_particles_out = c->d_particles;
    c->info.n_particles = c->n_particles;
    c->info.n_live = 0; c->info.n_wide_nodes = 0; c->info.wide_depth_levels = 0; c->info.build_ms = 0;
    if (c->n_particles == 0) return 0;
    const uint32_t n = (uint32_t)c->n_particles;
    if (!c->ev_b0) { CU(c, cudaEventCreate(&c->ev_b0)); CU(c, cudaEventCreate(&c->ev_b1)); }
    CU(c, cudaEventRecord(c->ev_b0, stream));
    c->scratch.begin();
    float4* d_lo = static_cast<float4*>(c->scratch.get(sizeof(float4) * (size_t)n));
    float4* d_hi = static_cast<float4*>(c->scratch.get(sizeof(float4) * (size_t)n));
    const bool own_lo = !d_lo, own_hi = !d_hi;
    if (own_lo) CU(c, cudaMalloc(&d_lo, sizeof(float4) * (size_t)n));
    if (own_hi) CU(c, cudaMalloc(&d_hi, sizeof(float4) * (size_t)n));
    int rc = particles_compute_boxes(c->d_particles, c->d_s, n, d_lo, d_hi, stream, &c->launches);
    LbvhResult res;
    res.reuse_nodes = c->p_nodes;
    res.reuse_cap = c->p_nodes_cap;
    if (!rc) rc = lbvh_build(d_lo, d_hi, n, stream, &res, &c->launches, &c->scratch);
    if (own_lo) cudaFree(d_lo);
    if (own_hi) cudaFree(d_hi);
    if (rc) {
        if (res.perm_owned) cudaFree(res.perm);
        if (res.nodes != c->p_nodes) cudaFree(res.nodes);
        return fail(c, "particle accel build failed: %s", cudaGetErrorString((cudaError_t)rc));
    }
    if (res.nodes && res.nodes != c->p_nodes) {
        cudaFree(c->p_nodes);
        c->p_nodes = res.nodes;
        c->p_nodes_cap = res.nodes_cap ? res.nodes_cap : res.n_nodes;
    }
    c->p_live = res.n_live; c->p_nnodes = res.n_nodes; c->p_levels = res.levels;
    if (res.n_live) {
        if (res.n_live > c->p_recs_cap) {
            cudaFree(c->p_recs); cudaFree(c->p_sh);
            c->p_recs = nullptr; c->p_sh = nullptr; c->p_recs_cap = 0;
            CU(c, cudaMalloc(&c->p_recs, sizeof(PRec) * (size_t)res.n_live));
            CU(c, cudaMalloc(&c->p_sh, sizeof(float4) * SH_F4 * (size_t)res.n_live));
            c->p_recs_cap = res.n_live;
        }
        RC(c, particles_gather(c->d_particles, c->d_s, res.perm, res.n_live, c->p_recs, c->p_sh, stream, &c->launches));
        AccelHeader h{};
        h.magic = ACCEL_MAGIC_PARTICLES; h.nodes = c->p_nodes; h.recs = c->p_recs; h.sh = c->p_sh; h.tris = nullptr;
        h.n_prims = res.n_live; h.n_nodes = res.n_nodes;
        if (!c->d_p_header) CU(c, cudaMalloc(&c->d_p_header, sizeof(AccelHeader)));
        CU(c, cudaMemcpyAsync(c->d_p_header, &h, sizeof(h), cudaMemcpyHostToDevice, stream));
        c->p_valid = true;
    }
    CU(c, cudaEventRecord(c->ev_b1, stream));
    CU(c, cudaStreamSynchronize(stream));
    if (res.perm_owned) cudaFree(res.perm);
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev_b0, c->ev_b1);
    c->info.n_live = res.n_live; c->info.n_wide_nodes = res.n_nodes; c->info.wide_depth_levels = res.levels; c->info.build_ms = ms;
    for (int k = 0; k < 3; ++k) { c->info.scene_lo[k] = res.lo[k]; c->info.scene_hi[k] = res.hi[k]; }
    if (handle_out) *handle_out = c->p_valid ? (uint64_t)(uintptr_t)c->d_p_header : 0;  // 0 when nothing is live
    return 0;
}

// ---------------------------------------------------------------- meshes
int grtSetMeshes(grtContext* c, const grtMeshDesc* meshes, uint32_t n) {
    if (!c) return 1;
    if (n && !meshes) return fail(c, "grtSetMeshes: meshes is NULL");
    // Only validation and a copy of the caller's arrays happen here; the per-vertex transform the reference host does on every
    // update (world vertices of the instance, normals = mat3(transform) * n: src/GaussianTracer.cpp:658-663, 711-794) runs on the
    // device in grtBuildMeshAccel.
    c->h_ov.clear(); c->h_on.clear(); c->h_idx.clear(); c->h_prep.clear();
    uint32_t tri_off = 0;
    for (uint32_t m = 0; m < n; ++m) {
        const grtMeshDesc& d = meshes[m];
        if (d.index_count % 3) return fail(c, "grtSetMeshes: mesh %u index_count %u is not a multiple of 3", m, d.index_count);
        if ((d.vertex_count && (!d.vertices || !d.normals)) || (d.index_count && !d.indices)) return fail(c, "grtSetMeshes: mesh %u has a NULL array", m);
        for (uint32_t k = 0; k < d.index_count; ++k)
            if (d.indices[k] >= d.vertex_count) {
                c->h_ov.clear(); c->h_on.clear(); c->h_idx.clear(); c->h_prep.clear();
                return fail(c, "grtSetMeshes: mesh %u face %u references vertex %u >= %u", m, k / 3, d.indices[k], d.vertex_count);
            }
        MeshPrep p{};
        memcpy(p.T, d.transform, sizeof(p.T));
        p.v_off = (uint32_t)(c->h_ov.size() / 3); p.i_off = (uint32_t)c->h_idx.size(); p.tri_off = tri_off;
        c->h_ov.insert(c->h_ov.end(), d.vertices, d.vertices + 3ull * d.vertex_count);
        c->h_on.insert(c->h_on.end(), d.normals, d.normals + 3ull * d.vertex_count);
        c->h_idx.insert(c->h_idx.end(), d.indices, d.indices + d.index_count);
        tri_off += d.index_count / 3;
        c->h_prep.push_back(p);
    }
    return 0;
}

int grtBuildMeshAccel(grtContext* c, void* stream_v, uint64_t* mesh_handle_out, void** d_meshes_out) {
    if (!c) return 1;
    cudaStream_t stream = (cudaStream_t)stream_v;
    CU(c, cudaSetDevice(c->device));
    // frames in flight may still read the old mesh arrays: the rebuild is ordered behind the context's last launch on the
    // device (no host-side cudaDeviceSynchronize)
    if (c->ev_valid) CU(c, cudaStreamWaitEvent(stream, c->ev1, 0));
    free_mesh_accel(c);
    if (mesh_handle_out) *mesh_handle_out = 0;
    if (d_meshes_out) *d_meshes_out = nullptr;
    c->info.n_mesh_triangles = 0; c->info.n_mesh_wide_nodes = 0; c->info.mesh_build_ms = 0;
    const uint32_t nm = (uint32_t)c->h_prep.size();
    const uint32_t nv = (uint32_t)(c->h_ov.size() / 3), ni = (uint32_t)c->h_idx.size(), nt = ni / 3;
    if (nm == 0) return 0;
    // Moving a primitive rebuilds this accel every frame (updateInstanceTransforms, src/GaussianTracer.cpp:711-794): the
    // persistent arrays come from a pool that is kept between builds, the temporaries from the build scratch block.
    c->mesh_pool.begin();
    auto palloc = [&](size_t bytes, void** out) -> cudaError_t {
        *out = c->mesh_pool.get(bytes ? bytes : 4);
        if (*out) return cudaSuccess;
        cudaError_t e = cudaMalloc(out, bytes ? bytes : 4);
        if (e == cudaSuccess) c->mesh_allocs.push_back(*out);
        return e;
    };
    if (!c->ev_b0) { CU(c, cudaEventCreate(&c->ev_b0)); CU(c, cudaEventCreate(&c->ev_b1)); }
    CU(c, cudaEventRecord(c->ev_b0, stream));
    // persistent: Mesh{faces, vertex_normals}[] exactly as sendGeometryAttributesToDevice leaves it (src/GaussianTracer.cpp:653-709):
    // faces = the index triples, vertex_normals = mat3(transform) * normal
    uint32_t* d_idx = nullptr; float* d_nw = nullptr;
    { void* p = nullptr; CU(c, palloc(sizeof(uint32_t) * (size_t)ni, &p)); d_idx = (uint32_t*)p; }
    { void* p = nullptr; CU(c, palloc(sizeof(float) * 3ull * nv, &p)); d_nw = (float*)p; }
    { void* p = nullptr; CU(c, palloc(sizeof(MeshDev) * nm, &p)); c->d_meshes = (MeshDev*)p; }
    std::vector<MeshDev> hm(nm);
    for (uint32_t m = 0; m < nm; ++m) {
        hm[m].faces = (uint3*)(d_idx + c->h_prep[m].i_off);
        hm[m].vertex_normals = (float3*)(d_nw + 3ull * c->h_prep[m].v_off);
    }
    CU(c, cudaMemcpyAsync(d_idx, c->h_idx.data(), sizeof(uint32_t) * (size_t)ni, cudaMemcpyHostToDevice, stream));
    CU(c, cudaMemcpyAsync(c->d_meshes, hm.data(), sizeof(MeshDev) * nm, cudaMemcpyHostToDevice, stream));
    if (d_meshes_out) *d_meshes_out = c->d_meshes;
    c->scratch.begin();
    struct Tmp {  // scratch block, else cudaMalloc'd for this build only
        void* p = nullptr; bool owned = false;
        ~Tmp() { if (owned) cudaFree(p); }
        cudaError_t alloc(BuildScratch& s, size_t bytes) {
            p = s.get(bytes ? bytes : 4);
            if (p) return cudaSuccess;
            owned = true;
            return cudaMalloc(&p, bytes ? bytes : 4);
        }
    } t_ov, t_on, t_vw, t_prep, t_tri9, t_meta, t_lo, t_hi;
    CU(c, t_ov.alloc(c->scratch, sizeof(float) * 3ull * nv));
    CU(c, t_on.alloc(c->scratch, sizeof(float) * 3ull * nv));
    CU(c, t_vw.alloc(c->scratch, sizeof(float) * 3ull * nv));
    CU(c, t_prep.alloc(c->scratch, sizeof(MeshPrep) * (size_t)nm));
    CU(c, t_tri9.alloc(c->scratch, sizeof(float) * 9ull * nt));
    CU(c, t_meta.alloc(c->scratch, sizeof(uint2) * (size_t)nt));
    CU(c, t_lo.alloc(c->scratch, sizeof(float4) * (size_t)nt));
    CU(c, t_hi.alloc(c->scratch, sizeof(float4) * (size_t)nt));
    float* d_tri9 = (float*)t_tri9.p; uint2* d_meta = (uint2*)t_meta.p; float4 *d_lo = (float4*)t_lo.p, *d_hi = (float4*)t_hi.p;
    CU(c, cudaMemcpyAsync(t_ov.p, c->h_ov.data(), sizeof(float) * 3ull * nv, cudaMemcpyHostToDevice, stream));
    CU(c, cudaMemcpyAsync(t_on.p, c->h_on.data(), sizeof(float) * 3ull * nv, cudaMemcpyHostToDevice, stream));
    CU(c, cudaMemcpyAsync(t_prep.p, c->h_prep.data(), sizeof(MeshPrep) * (size_t)nm, cudaMemcpyHostToDevice, stream));
    int rc = mesh_prepare((const float*)t_ov.p, (const float*)t_on.p, d_idx, (const MeshPrep*)t_prep.p, nm, nv, nt, (float*)t_vw.p, d_nw, d_tri9, d_meta, stream, &c->launches);
    if (!rc && nt == 0) { CU(c, cudaStreamSynchronize(stream)); return 0; }
    if (!rc) rc = mesh_compute_boxes(d_tri9, nt, d_lo, d_hi, stream, &c->launches);
    LbvhResult res;
    res.reuse_nodes = c->m_nodes;
    res.reuse_cap = c->m_nodes_cap;
    if (!rc) rc = lbvh_build(d_lo, d_hi, nt, stream, &res, &c->launches, &c->scratch);
    if (!rc && res.n_live) {
        void* p = nullptr;
        rc = (int)palloc(sizeof(TriRec) * (size_t)res.n_live, &p);
        c->m_tris = (TriRec*)p;
        if (!rc) rc = mesh_gather(d_tri9, d_meta, res.perm, res.n_live, c->m_tris, stream, &c->launches);
    }
    cudaError_t es = cudaStreamSynchronize(stream);
    if (res.perm_owned) cudaFree(res.perm);
    if (res.nodes && res.nodes != c->m_nodes) {
        cudaFree(c->m_nodes);
        c->m_nodes = res.nodes;
        c->m_nodes_cap = res.nodes_cap ? res.nodes_cap : res.n_nodes;
    }
    if (rc || es != cudaSuccess) { free_mesh_accel(c); return fail(c, "mesh accel build failed: %s", cudaGetErrorString(rc ? (cudaError_t)rc : es)); }
    c->m_count = res.n_live; c->m_nnodes = res.n_nodes;
    if (res.n_live) {
        AccelHeader h{};
        h.magic = ACCEL_MAGIC_MESH; h.nodes = c->m_nodes; h.tris = c->m_tris; h.n_prims = res.n_live; h.n_nodes = res.n_nodes;
        void* p = nullptr;
        CU(c, palloc(sizeof(AccelHeader), &p));
        c->d_m_header = (AccelHeader*)p;
        CU(c, cudaMemcpyAsync(c->d_m_header, &h, sizeof(h), cudaMemcpyHostToDevice, stream));
    }
    CU(c, cudaEventRecord(c->ev_b1, stream));
    CU(c, cudaStreamSynchronize(stream));
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev_b0, c->ev_b1);
    c->info.n_mesh_triangles = res.n_live; c->info.n_mesh_wide_nodes = res.n_nodes; c->info.mesh_build_ms = ms;
    if (mesh_handle_out) *mesh_handle_out = (uint64_t)(uintptr_t)c->d_m_header;
    return 0;
}

// ---------------------------------------------------------------- render
static int do_launch(grtContext* c, cudaStream_t stream, const grtParams* p, uint32_t x0, uint32_t y0, uint32_t w, uint32_t h,
                     uint32_t band_rows = 0, uint32_t band_stride = 1, uint32_t band_offset = 0) {
    if (!p) return fail(c, "grtLaunch: host_params is NULL");
    DEV(c);
    if (p->width == 0 || p->height == 0) return 0;
    if (x0 + w > p->width || y0 + h > p->height) return fail(c, "grtLaunch: tile (%u,%u,%u,%u) outside %ux%u", x0, y0, w, h, p->width, p->height);
    if (p->sh_degree_max > 3) return fail(c, "grtLaunch: sh_degree_max %u > 3", p->sh_degree_max);
    // Handles are opaque device pointers owned by this context; anything else is a caller bug
    // (the reference would hand a stale OptixTraversableHandle to optixLaunch and crash).
    if (p->handle != 0 && (p->handle != (uint64_t)(uintptr_t)c->d_p_header || !c->p_valid))
        return fail(c, "grtLaunch: Params.handle is not this context's (current) particle accel");
    if (p->mesh_handle != 0 && p->mesh_handle != (uint64_t)(uintptr_t)c->d_m_header) return fail(c, "grtLaunch: Params.mesh_handle is not this context's mesh accel");
    if (p->mesh_handle != 0 && p->d_meshes != (void*)c->d_meshes) return fail(c, "grtLaunch: Params.d_meshes does not belong to Params.mesh_handle");
    if (!p->output_buffer && !c->dbg_rgb32f) return fail(c, "grtLaunch: Params.output_buffer is NULL");
    LaunchExtras ex{};
    if (p->handle) { ex.p_nodes = c->p_nodes; ex.p_recs = c->p_recs; ex.p_sh = c->p_sh; ex.p_count = c->p_live; }
    if (p->mesh_handle) { ex.m_nodes = c->m_nodes; ex.m_tris = c->m_tris; ex.m_count = c->m_count; }
    ex.x0 = x0; ex.y0 = y0; ex.w = w; ex.h = h;
    ex.rgb32f = c->dbg_rgb32f;
    ex.trace_ids = c->dbg_trace_ids; ex.trace_counts = c->dbg_trace_counts;
    ex.tx0 = c->tx0; ex.ty0 = c->ty0; ex.tw = c->tw; ex.th = c->th; ex.trace_max = c->trace_max;
    ex.counters = c->d_counters;
    if (band_stride > 1u) {
        if (band_rows == 0u || (band_rows & 7u)) return fail(c, "grtLaunchRows: band_rows %u must be a positive multiple of 8 (the packet grid)", band_rows);
        if (band_offset >= band_stride) return fail(c, "grtLaunchRows: band_offset %u >= band_stride %u", band_offset, band_stride);
        ex.band_groups = band_rows / 8u; ex.band_stride = band_stride; ex.band_offset = band_offset;
    } else {
        ex.band_groups = 1u; ex.band_stride = 1u; ex.band_offset = 0u;
    }
    // Every launch gets its own block of scheduler counters out of a ring: launches of one context that are in flight on
    // different streams (tiles or cameras overlapped on one GPU) must not share a tile counter.
    ex.tile_counter = c->d_tile_counter + (size_t)(c->tile_seq++ % TILE_RING) * TILE_SLOT_WORDS;
    CU(c, cudaMemsetAsync(ex.tile_counter, 0, (GRT_SM_REGIONS_HOST ? TILE_SLOT_WORDS : 1u) * sizeof(uint32_t), stream));
    const WavefrontBuffers* wf = nullptr;
    if (ex.m_count != 0u && c->wavefront && p->type != 1 /* NORMAL: rays end at the primitive, one launch does it */) {
        // wavefront buffers: payload of every pixel of the rectangle + two tile lists, kept between launches, grown when a larger
        // rectangle comes (no allocation on the steady-state path)
        const size_t sb = wavefront_state_bytes(ex), lb = wavefront_list_bytes(ex);
        if (sb > c->wf.state_bytes || lb > c->wf.list_bytes || !c->wf.ctrl) {
            CU(c, cudaStreamSynchronize(stream));
            cudaFree(c->wf.state); cudaFree(c->wf.list[0]); cudaFree(c->wf.list[1]);
            c->wf.state = nullptr; c->wf.list[0] = c->wf.list[1] = nullptr; c->wf.state_bytes = c->wf.list_bytes = 0;
            CU(c, cudaMalloc(&c->wf.state, sb));
            CU(c, cudaMalloc(&c->wf.list[0], lb));
            CU(c, cudaMalloc(&c->wf.list[1], lb));
            if (!c->wf.ctrl) CU(c, cudaMalloc(&c->wf.ctrl, WF_CTRL_WORDS * sizeof(uint32_t)));
            c->wf.state_bytes = sb; c->wf.list_bytes = lb;
        }
        if (sb != 0) wf = &c->wf;
    }
    CU(c, cudaEventRecord(c->ev0, stream));
    RC(c, launch_trace(*p, ex, c->counters_on, stream, c->sm_count, &c->smem_attr_set, wf, &c->launches));
    CU(c, cudaEventRecord(c->ev1, stream));
    c->ev_valid = true;
    return 0;
}

int grtLaunch(grtContext* c, void* stream, const grtParams* p) {
    if (!c) return 1;
    if (!p) return fail(c, "grtLaunch: host_params is NULL");
    return do_launch(c, (cudaStream_t)stream, p, 0, 0, p->width, p->height);
}
int grtLaunchTile(grtContext* c, void* stream, const grtParams* p, uint32_t x0, uint32_t y0, uint32_t w, uint32_t h) {
    if (!c) return 1;
    return do_launch(c, (cudaStream_t)stream, p, x0, y0, w, h);
}
int grtLaunchRows(grtContext* c, void* stream, const grtParams* p, uint32_t band_rows, uint32_t band_stride, uint32_t band_offset) {
    if (!c) return 1;
    if (!p) return fail(c, "grtLaunchRows: host_params is NULL");
    if (band_stride == 0u) return fail(c, "grtLaunchRows: band_stride is 0");
    return do_launch(c, (cudaStream_t)stream, p, 0, 0, p->width, p->height, band_rows, band_stride, band_offset);
}
int grtLaunchBatch(grtContext* c, void* stream, const grtParams* arr, uint32_t n) {
    if (!c) return 1;
    if (n && !arr) return fail(c, "grtLaunchBatch: array is NULL");
    for (uint32_t i = 0; i < n; ++i) {
        int rc = do_launch(c, (cudaStream_t)stream, &arr[i], 0, 0, arr[i].width, arr[i].height);
        if (rc) return rc;
    }
    return 0;
}

int grtRenderHost(grtContext* c, const grtParams* p, uint8_t* host_rgb8, uint32_t y0, uint32_t h) {
    if (!c) return 1;
    if (!p || !host_rgb8) return fail(c, "grtRenderHost: bad argument");
    if (h == 0) { y0 = 0; h = p->height; }
    if (y0 + h > p->height) return fail(c, "grtRenderHost: rows [%u,%u) outside height %u", y0, y0 + h, p->height);
    DEV(c);
    const size_t need = (size_t)p->width * p->height * 3;
    if (need > c->frame_bytes) {
        cudaFree(c->d_frame);
        c->d_frame = nullptr; c->frame_bytes = 0;
        CU(c, cudaMalloc(&c->d_frame, need));
        c->frame_bytes = need;
    }
    grtParams q = *p;
    q.output_buffer = c->d_frame;
    if (q.mode_fisheye) CU(c, cudaMemsetAsync(c->d_frame + (size_t)3 * q.width * y0, 0, (size_t)3 * q.width * h, c->own_stream));  // GaussianTracer.cpp:510-513
    int rc = do_launch(c, c->own_stream, &q, 0, y0, q.width, h);
    if (rc) return rc;
    const size_t off = (size_t)3 * q.width * y0;
    CU(c, cudaMemcpyAsync(host_rgb8 + off, c->d_frame + off, (size_t)3 * q.width * h, cudaMemcpyDeviceToHost, c->own_stream));
    CU(c, cudaStreamSynchronize(c->own_stream));
    return 0;
}

// ---------------------------------------------------------------- side channels
int grtSetDebugOutputs(grtContext* c, float* d_rgb32f, uint32_t* d_trace_ids, uint32_t* d_trace_counts, uint32_t tx0, uint32_t ty0, uint32_t tw,
                       uint32_t th, uint32_t trace_max) {
    if (!c) return 1;
    c->dbg_rgb32f = d_rgb32f;
    c->dbg_trace_ids = (tw && th && trace_max) ? d_trace_ids : nullptr;
    c->dbg_trace_counts = d_trace_counts;
    c->tx0 = tx0; c->ty0 = ty0; c->tw = tw; c->th = th; c->trace_max = trace_max;
    return 0;
}
// ---------------------------------------------------------------- multi-GPU: framebuffer gather over NCCL
int grtCommGetUniqueId(uint8_t out128[128]) {
    if (!out128) return 1;
    if (const char* e = nccl_bind()) return fail(nullptr, "grtCommGetUniqueId: %s", e);
    NcclId id;
    int r = g_nccl.GetUniqueId(&id);
    if (r) return fail(nullptr, "ncclGetUniqueId failed: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
    memcpy(out128, id.b, 128);
    return 0;
}
int grtCommInit(grtContext* c, int nranks, int rank, const uint8_t id128[128]) {
    if (!c || !id128) return 1;
    if (nranks < 1 || rank < 0 || rank >= nranks) return fail(c, "grtCommInit: rank %d of %d", rank, nranks);
    if (c->comm) return fail(c, "grtCommInit: this context already has a communicator");
    if (const char* e = nccl_bind()) return fail(c, "grtCommInit: %s", e);
    DEV(c);
    NcclId id;
    memcpy(id.b, id128, 128);
    typedef int (*init_fn)(void**, int, NcclId, int);
    NC(c, ((init_fn)g_nccl.CommInitRank)(&c->comm, nranks, id, rank));
    c->comm_rank = rank; c->comm_nranks = nranks;
    return 0;
}
int grtCommDestroy(grtContext* c) {
    if (!c) return 1;
    if (c->comm) { DEV(c); g_nccl.CommDestroy(c->comm); c->comm = nullptr; }
    c->comm_rank = 0; c->comm_nranks = 1;
    return 0;
}
// Camera-batch partition: every rank contributes `bytes` from d_send; the root receives rank r's block at d_recv_all + r * bytes
// (its own block is copied device-to-device).  One grouped ncclSend/ncclRecv exchange, asynchronous on `stream`.
int grtGatherFrames(grtContext* c, void* stream_v, const void* d_send, size_t bytes, void* d_recv_all, int root) {
    if (!c) return 1;
    if (!c->comm) return fail(c, "grtGatherFrames: no communicator (grtCommInit)");
    if (root < 0 || root >= c->comm_nranks) return fail(c, "grtGatherFrames: root %d of %d", root, c->comm_nranks);
    cudaStream_t stream = (cudaStream_t)stream_v;
    DEV(c);
    if (c->comm_rank == root) {
        if (!d_recv_all) return fail(c, "grtGatherFrames: the root needs a receive buffer");
        NC(c, g_nccl.GroupStart());
        for (int r = 0; r < c->comm_nranks; ++r)
            if (r != root) NC(c, g_nccl.Recv((char*)d_recv_all + (size_t)r * bytes, bytes, /*ncclUint8*/ 1, r, c->comm, stream));
        NC(c, g_nccl.GroupEnd());
        if (d_send && (const char*)d_send != (const char*)d_recv_all + (size_t)root * bytes)
            CU(c, cudaMemcpyAsync((char*)d_recv_all + (size_t)root * bytes, d_send, bytes, cudaMemcpyDeviceToDevice, stream));
    } else {
        NC(c, g_nccl.Send(d_send, bytes, 1, root, c->comm, stream));
    }
    return 0;
}
// Row-band partition of ONE frame (grtLaunchRows with band_stride = nranks, band_offset = rank): every rank sends the bands it
// traced straight out of its frame buffer, the root receives them in place at the same offsets of ITS frame buffer (which
// already holds the root's own bands): no packing, one grouped exchange.
int grtGatherBands(grtContext* c, void* stream_v, void* d_frame, uint32_t width, uint32_t height, uint32_t band_rows, int root) {
    if (!c) return 1;
    if (!c->comm) return fail(c, "grtGatherBands: no communicator (grtCommInit)");
    if (!d_frame || band_rows == 0 || (band_rows & 7u)) return fail(c, "grtGatherBands: bad argument (band_rows %u)", band_rows);
    if (root < 0 || root >= c->comm_nranks) return fail(c, "grtGatherBands: root %d of %d", root, c->comm_nranks);
    cudaStream_t stream = (cudaStream_t)stream_v;
    DEV(c);
    const uint32_t n_bands = (height + band_rows - 1u) / band_rows;
    const size_t row_bytes = (size_t)width * 3u;
    NC(c, g_nccl.GroupStart());
    for (uint32_t b = 0; b < n_bands; ++b) {
        const int owner = (int)(b % (uint32_t)c->comm_nranks);
        const uint32_t y0 = b * band_rows, rows = (y0 + band_rows <= height) ? band_rows : height - y0;
        char* p = (char*)d_frame + (size_t)y0 * row_bytes;
        if (c->comm_rank == root && owner != root) NC(c, g_nccl.Recv(p, rows * row_bytes, 1, owner, c->comm, stream));
        else if (c->comm_rank != root && owner == c->comm_rank) NC(c, g_nccl.Send(p, rows * row_bytes, 1, root, c->comm, stream));
    }
    NC(c, g_nccl.GroupEnd());
    return 0;
}

int grtGetAbiLayout(uint64_t out[32]) {
    if (!out) return 1;
    int k = 0;
    out[k++] = sizeof(grtParams);
    out[k++] = offsetof(grtParams, output_buffer); out[k++] = offsetof(grtParams, width); out[k++] = offsetof(grtParams, height);
    out[k++] = offsetof(grtParams, sh_degree_max); out[k++] = offsetof(grtParams, eye); out[k++] = offsetof(grtParams, U);
    out[k++] = offsetof(grtParams, V); out[k++] = offsetof(grtParams, W); out[k++] = offsetof(grtParams, t_min);
    out[k++] = offsetof(grtParams, t_max); out[k++] = offsetof(grtParams, minTransmittance); out[k++] = offsetof(grtParams, alpha_min);
    out[k++] = offsetof(grtParams, handle); out[k++] = offsetof(grtParams, d_particles); out[k++] = offsetof(grtParams, mesh_handle);
    out[k++] = offsetof(grtParams, mode_fisheye); out[k++] = offsetof(grtParams, d_meshes); out[k++] = offsetof(grtParams, type);
    out[k++] = offsetof(grtParams, traceState);
    out[k++] = 236;                 // bytes per particle record the library reads (GaussianParticle, src/GaussianData.h:12-20)
    out[k++] = sizeof(MeshDev);     // Mesh{faces, vertex_normals} (src/Parameters.h:36-40)
    while (k < 32) out[k++] = 0;
    return 0;
}
int grtSetWavefront(grtContext* c, int enable) { if (!c) return 1; c->wavefront = enable != 0; return 0; }
int grtEnableCounters(grtContext* c, int enable) { if (!c) return 1; c->counters_on = enable != 0; return 0; }
int grtGetCounters(grtContext* c, grtCounters* out, int reset) {
    if (!c || !out) return 1;
    DEV(c);
    CU(c, cudaDeviceSynchronize());
    DeviceCounters d;
    CU(c, cudaMemcpy(&d, c->d_counters, sizeof(d), cudaMemcpyDeviceToHost));
    out->rays = d.rays; out->segments = d.segments; out->rounds = d.rounds; out->hits_gathered = d.hits_gathered;
    out->hits_shaded = d.hits_shaded; out->warp_nodes = d.warp_nodes; out->warp_leaf_prims = d.warp_leaf_prims;
    out->warp_rounds = d.warp_rounds; out->stack_overflow = d.stack_overflow; out->kernel_launches = c->launches;
    out->warp_prims_hit = d.warp_prims_hit; out->warp_consume_iters = d.warp_consume_iters; out->lanes_repass = d.lanes_repass;
    out->order_violations = d.order_violations; out->warp_dfs_fallbacks = d.warp_dfs_fallbacks; out->warp_candidate_iters = d.warp_candidate_iters;
    for (int k = 0; k < 8; ++k) out->dbg[k] = d.dbg[k];
    out->warp_per_ray_passes = d.warp_per_ray_passes;
    out->warp_composite_rounds = d.warp_composite_rounds; out->warp_loop_iters = d.warp_loop_iters;
    out->warp_forced_flushes = d.warp_forced_flushes; out->warp_dfs_iters = d.warp_dfs_iters; out->warp_list_full = d.warp_list_full;
    if (reset) CU(c, cudaMemset(c->d_counters, 0, sizeof(DeviceCounters)));
    return 0;
}
int grtGetBuildInfo(grtContext* c, grtBuildInfo* out) { if (!c || !out) return 1; *out = c->info; return 0; }
int grtLastLaunchMs(grtContext* c, float* ms) {
    if (!c || !ms) return 1;
    if (!c->ev_valid) return fail(c, "grtLastLaunchMs: nothing launched yet");
    DEV(c);
    CU(c, cudaEventSynchronize(c->ev1));
    CU(c, cudaEventElapsedTime(ms, c->ev0, c->ev1));
    return 0;
}

int grtValidateAccel(grtContext* c, int which) {
    if (!c) return 1;
    const WideNode* dn = which ? c->m_nodes : c->p_nodes;
    const uint32_t nn = which ? c->m_nnodes : c->p_nnodes, np = which ? c->m_count : c->p_live;
    if (nn == 0) return np == 0 ? 0 : fail(c, "validate: %u prims but no nodes", np);
    DEV(c);
    CU(c, cudaDeviceSynchronize());
    std::vector<WideNode> nodes(nn);
    CU(c, cudaMemcpy(nodes.data(), dn, sizeof(WideNode) * (size_t)nn, cudaMemcpyDeviceToHost));
    std::vector<uint8_t> seen_prim(np, 0), seen_node(nn, 0);
    struct Item { uint32_t node; float lo[3], hi[3]; };
    std::vector<Item> st;
    st.push_back(Item{0, {-INFINITY, -INFINITY, -INFINITY}, {INFINITY, INFINITY, INFINITY}});
    seen_node[0] = 1;
    uint64_t prim_total = 0;
    while (!st.empty()) {
        Item it = st.back();
        st.pop_back();
        const WideNode& nd = nodes[it.node];
        for (int k = 0; k < WIDE; ++k) {
            const uint32_t ref = nd.child[k];
            if (ref == CHILD_EMPTY) continue;
            const float lo[3] = {nd.lox[k], nd.loy[k], nd.loz[k]}, hi[3] = {nd.hix[k], nd.hiy[k], nd.hiz[k]};
            for (int a = 0; a < 3; ++a) {
                if (!(lo[a] <= hi[a])) return fail(c, "validate: node %u child %d has an inverted/NaN box", it.node, k);
                if (lo[a] < it.lo[a] || hi[a] > it.hi[a]) return fail(c, "validate: node %u child %d box escapes its parent", it.node, k);
            }
            if (ref & CHILD_LEAF) {
                const uint32_t start = ref & 0x0FFFFFFFu, cnt = ((ref >> 28) & 7u) + 1u;
                if (cnt > (uint32_t)LEAF_MAX) return fail(c, "validate: leaf run of %u > LEAF_MAX", cnt);
                for (uint32_t j = start; j < start + cnt; ++j) {
                    if (j >= np) return fail(c, "validate: leaf references record %u >= %u", j, np);
                    if (seen_prim[j]) return fail(c, "validate: record %u referenced twice", j);
                    seen_prim[j] = 1;
                    ++prim_total;
                }
            } else {
                if (ref >= nn) return fail(c, "validate: child node %u >= %u", ref, nn);
                if (seen_node[ref]) return fail(c, "validate: node %u referenced twice", ref);
                seen_node[ref] = 1;
                Item ch{ref, {lo[0], lo[1], lo[2]}, {hi[0], hi[1], hi[2]}};
                st.push_back(ch);
            }
        }
    }
    if (prim_total != np) return fail(c, "validate: %llu of %u records reachable", (unsigned long long)prim_total, np);
    // leaf boxes must contain their records: checked for particles via the record centre (mu)
    if (!which && np) {
        std::vector<PRec> recs(np);
        CU(c, cudaMemcpy(recs.data(), c->p_recs, sizeof(PRec) * (size_t)np, cudaMemcpyDeviceToHost));
        for (uint32_t n = 0; n < nn; ++n)
            for (int k = 0; k < WIDE; ++k) {
                const uint32_t ref = nodes[n].child[k];
                if (ref == CHILD_EMPTY || !(ref & CHILD_LEAF)) continue;
                const uint32_t start = ref & 0x0FFFFFFFu, cnt = ((ref >> 28) & 7u) + 1u;
                for (uint32_t j = start; j < start + cnt; ++j) {
                    const float mu[3] = {recs[j].q2.y, recs[j].q2.z, recs[j].q2.w};
                    const float lo[3] = {nodes[n].lox[k], nodes[n].loy[k], nodes[n].loz[k]}, hi[3] = {nodes[n].hix[k], nodes[n].hiy[k], nodes[n].hiz[k]};
                    for (int a = 0; a < 3; ++a)
                        if (mu[a] < lo[a] || mu[a] > hi[a]) return fail(c, "validate: record %u centre outside its leaf box", j);
                }
            }
    }
    return 0;
}

int grtDebugRadixSort(grtContext* c, uint64_t* host_keys, uint32_t* host_vals, uint64_t n) {
    if (!c) return 1;
    if (n == 0) return 0;
    if (n >= (1ull << 31)) return fail(c, "grtDebugRadixSort: n too large");
    CU(c, cudaSetDevice(c->device));
    uint64_t *k0 = nullptr, *k1 = nullptr; uint32_t *v0 = nullptr, *v1 = nullptr;
    CU(c, cudaMalloc(&k0, 8 * n)); CU(c, cudaMalloc(&k1, 8 * n)); CU(c, cudaMalloc(&v0, 4 * n)); CU(c, cudaMalloc(&v1, 4 * n));
    CU(c, cudaMemcpy(k0, host_keys, 8 * n, cudaMemcpyHostToDevice));
    CU(c, cudaMemcpy(v0, host_vals, 4 * n, cudaMemcpyHostToDevice));
    bool in_tmp = false;
    int rc = radix_sort_pairs(k0, k1, v0, v1, (uint32_t)n, nullptr, &c->launches, &in_tmp);
    if (!rc) {
        rc = (int)cudaMemcpy(host_keys, in_tmp ? k1 : k0, 8 * n, cudaMemcpyDeviceToHost);
        if (!rc) rc = (int)cudaMemcpy(host_vals, in_tmp ? v1 : v0, 4 * n, cudaMemcpyDeviceToHost);
    }
    cudaFree(k0); cudaFree(k1); cudaFree(v0); cudaFree(v1);
    if (rc) return fail(c, "grtDebugRadixSort failed: %s", cudaGetErrorString((cudaError_t)rc));
    return 0;
}

// No counterpart in the reference (see include/grt.h).
int grtSortPairsNoAlloc(grtContext* c, uint64_t* d_keys, uint32_t* d_vals, uint64_t n, void* stream) {
    if (!c) return 1;
    if (n == 0) return 0;
    if (!d_keys || !d_vals) return fail(c, "grtSortPairsNoAlloc: null device pointer");
    const uint32_t max_n = radix_sort_pairs_noalloc_max();
    if (n > max_n)
        return fail(c, "grtSortPairsNoAlloc: n = %llu exceeds the static scratch array's maximum of %u pairs (GRT_NOALLOC_SORT_MAX_N)",
                    (unsigned long long)n, max_n);
    DEV(c);
    const int rc = radix_sort_pairs_noalloc(c->device, d_keys, d_vals, (uint32_t)n, (cudaStream_t)stream, &c->launches);
    if (rc) return fail(c, "grtSortPairsNoAlloc failed: %s", cudaGetErrorString((cudaError_t)rc));
    return 0;
}

}  // extern "C"
