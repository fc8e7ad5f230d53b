/*
 * oracle.cpp -- scalar CPU restatement of the reference's per-ray algorithm.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  PARITY UNPINNED by reference tests (it has none); pinned
 * by analytic KATs and by the reference's own host code built into oracle/_ref (ref_shim.cpp).
 *
 * Build: -O2 -ffp-contract=off (only the explicit fmaf() calls below may fuse; everything else is
 * plain IEEE fp32 in the order written, which is the order of the reference source).
 *
 * What is restated (reference file:line):
 *   raygen bounce loop ............. shaders/tracer.cu:17-110
 *   miss / closesthit .............. shaders/tracer.cu:112-122, 155-187
 *   anyhit k-buffer insertion ...... shaders/tracer.cu:124-153
 *   getRay / getFishEyeRay ......... shaders/tracer.cuh:115-165
 *   getBarycentricNormal ........... shaders/tracer.cuh:167-185
 *   computeResponse ................ shaders/tracer.cuh:187-214 (+ glm mat3_cast gtc/quaternion.inl:47-72)
 *   SHToRadiance / computeRadiance . shaders/tracer.cuh:216-264 (constants src/Parameters.h:10-23)
 *   traceGPs / trace ............... shaders/tracer.cuh:289-373
 *   renderMirror/Normal/Glass ...... shaders/tracer.cuh:396-482
 *   writeOutputBuffer .............. shaders/tracer.cuh:68-73, 484-496
 *   proxy transform ................ src/GaussianTracer.cpp:297-317, 401-420
 *   icosahedron .................... src/geometry/Icosahedron.h:11-38
 *   float3 helpers ................. src/vector_math.h:560-606
 *
 * What is NOT in the reference and is defined here (OptiX is closed source):
 *   optixTrace over the Gaussian IAS is emulated by intersecting the ray with each live particle's
 *   icosahedron proxy; every face crossing (entry and exit) with tmin < t < tmax is offered to the
 *   anyhit rule.  The proxy is intersected in the particle's Gaussian space as 10 slab pairs
 *   ("PROXY-T spec", see proxy_hit_slabs) or, for validation, as 20 world-space triangles.
 */
#include "oracle.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

namespace {

// ----------------------------------------------------------------------------- float3 (vector_math.h)
struct f3 {
    float x, y, z;
};
inline f3 mk(float x, float y, float z) { return f3{x, y, z}; }
inline f3 mk(float s) { return f3{s, s, s}; }
inline f3 operator+(f3 a, f3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
inline f3 operator-(f3 a, f3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline f3 operator-(f3 a) { return mk(-a.x, -a.y, -a.z); }
inline f3 operator*(f3 a, f3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
inline f3 operator*(f3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
inline f3 operator*(float s, f3 a) { return mk(s * a.x, s * a.y, s * a.z); }
inline f3 operator/(f3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
inline f3 operator+(f3 a, float s) { return mk(a.x + s, a.y + s, a.z + s); }
inline f3& operator+=(f3& a, f3 b) { a = a + b; return a; }
inline float dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }           // vector_math.h:572-575
inline f3 cross(f3 a, f3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline float length(f3 v) { return sqrtf(dot(v, v)); }                                // :584-587
inline f3 normalize(f3 v) { float invLen = 1.0f / sqrtf(dot(v, v)); return v * invLen; } // :590-594
inline f3 reflect(f3 i, f3 n) { return i - 2.0f * n * dot(n, i); }                    // :603-606
inline float clampf(float f, float a, float b) { return fmaxf(a, fminf(f, b)); }      // :146-149
inline f3 clamp3(f3 v, float a, float b) { return mk(clampf(v.x, a, b), clampf(v.y, a, b), clampf(v.z, a, b)); }
inline f3 ld3(const float* p) { return mk(p[0], p[1], p[2]); }

inline uint32_t fbits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
// distance in ulps between two finite floats of the same sign region (monotone integer mapping)
inline int64_t ordered(float f) { int32_t i; memcpy(&i, &f, 4); return i < 0 ? (int64_t)INT32_MIN - i : i; }
inline bool ulp_close(float a, float b, int n) {
    if (!std::isfinite(a) || !std::isfinite(b)) return false;
    int64_t d = ordered(a) - ordered(b);
    return (d < 0 ? -d : d) <= n;
}

// ----------------------------------------------------------------------------- constants (tracer.cuh:9-14, Parameters.h:10-23)
constexpr float TRACE_MESH_TMIN = 1e-5f;
constexpr float TRACE_MESH_TMAX = 1e5f;
constexpr uint32_t TIMEOUT_ITERATIONS = 1000;
constexpr uint32_t MAX_BOUNCES = 32;
constexpr float REFRACTION_EPS_SHIFT = 1e-5f;
constexpr uint32_t InvalidParticleId = 0xFFFFFFFFu;
constexpr float InfiniteDistance = 1e20f;

#define SH_C0 0.28209479177387814f
#define SH_C1 0.4886025119029199f
#define SH_C2_0 1.0925484305920792f
#define SH_C2_1 -1.0925484305920792f
#define SH_C2_2 0.31539156525252005f
#define SH_C2_3 -1.0925484305920792f
#define SH_C2_4 0.5462742152960396f
#define SH_C3_0 -0.5900435899266435f
#define SH_C3_1 2.890611442640554f
#define SH_C3_2 -0.4570457994644658f
#define SH_C3_3 0.3731763325901154f
#define SH_C3_4 -0.4570457994644658f
#define SH_C3_5 1.445305721320277f
#define SH_C3_6 -0.5900435899266435f

enum { MIRROR = 0, NORMAL = 1, GLASS = 2 };
enum { TraceLastGaussianPass = 0, TraceGaussianPass = 1, TraceMeshPass = 2, TraceTerminate = 3 };

// ----------------------------------------------------------------------------- glm::mat3_cast (gtc/quaternion.inl:47-72)
// R[c][r]: column c, row r, exactly glm's storage.
struct m3 {
    float c[3][3];
};
inline m3 mat3_cast(const float q[4] /*x,y,z,w*/) {
    const float x = q[0], y = q[1], z = q[2], w = q[3];
    float qxx = x * x, qyy = y * y, qzz = z * z, qxz = x * z, qxy = x * y, qyz = y * z, qwx = w * x, qwy = w * y, qwz = w * z;
    m3 R;
    R.c[0][0] = 1.0f - 2.0f * (qyy + qzz);
    R.c[0][1] = 2.0f * (qxy + qwz);
    R.c[0][2] = 2.0f * (qxz - qwy);
    R.c[1][0] = 2.0f * (qxy - qwz);
    R.c[1][1] = 1.0f - 2.0f * (qxx + qzz);
    R.c[1][2] = 2.0f * (qyz + qwx);
    R.c[2][0] = 2.0f * (qxz + qwy);
    R.c[2][1] = 2.0f * (qyz - qwx);
    R.c[2][2] = 1.0f - 2.0f * (qxx + qyy);
    return R;
}
// glm mat3 * vec3: m[0]*v.x + m[1]*v.y + m[2]*v.z  (type_mat3x3.inl operator*)
inline f3 mul(const m3& m, f3 v) {
    return mk(m.c[0][0] * v.x + m.c[1][0] * v.y + m.c[2][0] * v.z,
              m.c[0][1] * v.x + m.c[1][1] * v.y + m.c[2][1] * v.z,
              m.c[0][2] * v.x + m.c[1][2] * v.y + m.c[2][2] * v.z);
}

// ----------------------------------------------------------------------------- icosahedron (Icosahedron.h:11-38)
void icosahedron(float v[36], uint32_t idx[60]) {
    const float rr = (3 + std::sqrt(5.0f)) / (2 * std::sqrt(3.0f));
    const float ss = 1.0f / rr;
    const float tt = (1.0f + std::sqrt(5.0f)) / (2.0f * rr);
    const float V[12][3] = {{-ss, tt, 0}, {ss, tt, 0}, {-ss, -tt, 0}, {ss, -tt, 0}, {0, -ss, tt}, {0, ss, tt},
                            {0, -ss, -tt}, {0, ss, -tt}, {tt, 0, -ss}, {tt, 0, ss}, {-tt, 0, -ss}, {-tt, 0, ss}};
    const uint32_t I[60] = {0, 11, 5, 0, 5, 1, 0, 1, 7, 0, 7, 10, 0, 10, 11, 1, 5, 9, 5, 11, 4, 11, 10, 2, 10, 7, 6, 7, 1, 8,
                            3, 9, 4, 3, 4, 2, 3, 2, 6, 3, 6, 8, 3, 8, 9, 4, 9, 5, 2, 4, 11, 6, 2, 10, 8, 6, 7, 9, 8, 1};
    memcpy(v, V, sizeof(V));
    memcpy(idx, I, sizeof(I));
}

// ----------------------------------------------------------------------------- per-particle derived data
struct Proxy {
    float A[3][3];   // A[r][c] = (1/scale_r) * R.c[r][c]  == rows of diag(1/scale) * R^T  (tracer.cuh:194-201)
    float mu[3];
    float s;         // sqrtf(2 ln(opacity / alpha_min))   (GaussianTracer.cpp:306)
    uint32_t id;     // PLY vertex index == instanceId     (GaussianTracer.cpp:313)
    float lo[3], hi[3];
    float wv[12][3]; // world-space proxy vertices (triangle mode + AABB)
};

// ---- PROXY-T spec ------------------------------------------------------------------------------
// Shared *definition* (not shared code) with the CUDA kernel: the face-hit distances of a proxy are
//   p   = o - mu                                   (3 fsub)
//   og_r = fma(A[r][2], p.z, fma(A[r][1], p.y, A[r][0]*p.x))
//   dg_r = fma(A[r][2], d.z, fma(A[r][1], d.y, A[r][0]*d.x))
//   10 slabs |n_k . x| <= h_k in Gaussian space, evaluated in this order:
//     k=0..3  n = (1, s1, s2) (unnormalised), (s1,s2) = (+,+),(+,-),(-,+),(-,-), h = sqrt3*s,
//             a = (og.x + s1*og.y) + s2*og.z   (adds only), b likewise from dg
//     k=4,5   n = (0, P, +-Q): a = fma(+-Q, og.z, P*og.y)
//     k=6,7   n = (P, +-Q, 0): a = fma(+-Q, og.y, P*og.x)
//     k=8,9   n = (+-Q, 0, P): a = fma(+-Q, og.x, P*og.z)     h = s
//   per slab: r = 1/b; t1 = (-h - a)*r; t2 = (h - a)*r; te = fmaxf(te, fminf(t1,t2)); tx = fminf(tx, fmaxf(t1,t2))
//   starting from te = -inf, tx = +inf.  The proxy is crossed iff te < tx; its two face hits are te and tx.
constexpr float ICO_P = 0.9341723589627158f;   // (1+sqrt5)/2 / sqrt3
constexpr float ICO_Q = 0.35682208977308993f;  // 2/(1+sqrt5) / sqrt3
constexpr float SQRT3 = 1.7320508075688772f;

inline void slab(float a, float b, float h, float& te, float& tx) {
    float r = 1.0f / b;
    float t1 = (-h - a) * r;
    float t2 = (h - a) * r;
    te = fmaxf(te, fminf(t1, t2));
    tx = fminf(tx, fmaxf(t1, t2));
}

inline bool proxy_hit_slabs(const Proxy& P, f3 o, f3 d, float& te_out, float& tx_out) {
    float px = o.x - P.mu[0], py = o.y - P.mu[1], pz = o.z - P.mu[2];
    float og[3], dg[3];
    for (int r = 0; r < 3; ++r) {
        og[r] = fmaf(P.A[r][2], pz, fmaf(P.A[r][1], py, P.A[r][0] * px));
        dg[r] = fmaf(P.A[r][2], d.z, fmaf(P.A[r][1], d.y, P.A[r][0] * d.x));
    }
    float te = -std::numeric_limits<float>::infinity(), tx = std::numeric_limits<float>::infinity();
    const float h1 = SQRT3 * P.s, h2 = P.s;
    slab((og[0] + og[1]) + og[2], (dg[0] + dg[1]) + dg[2], h1, te, tx);
    slab((og[0] + og[1]) - og[2], (dg[0] + dg[1]) - dg[2], h1, te, tx);
    slab((og[0] - og[1]) + og[2], (dg[0] - dg[1]) + dg[2], h1, te, tx);
    slab((og[0] - og[1]) - og[2], (dg[0] - dg[1]) - dg[2], h1, te, tx);
    slab(fmaf(ICO_Q, og[2], ICO_P * og[1]), fmaf(ICO_Q, dg[2], ICO_P * dg[1]), h2, te, tx);
    slab(fmaf(-ICO_Q, og[2], ICO_P * og[1]), fmaf(-ICO_Q, dg[2], ICO_P * dg[1]), h2, te, tx);
    slab(fmaf(ICO_Q, og[1], ICO_P * og[0]), fmaf(ICO_Q, dg[1], ICO_P * dg[0]), h2, te, tx);
    slab(fmaf(-ICO_Q, og[1], ICO_P * og[0]), fmaf(-ICO_Q, dg[1], ICO_P * dg[0]), h2, te, tx);
    slab(fmaf(ICO_Q, og[0], ICO_P * og[2]), fmaf(ICO_Q, dg[0], ICO_P * dg[2]), h2, te, tx);
    slab(fmaf(-ICO_Q, og[0], ICO_P * og[2]), fmaf(-ICO_Q, dg[0], ICO_P * dg[2]), h2, te, tx);
    te_out = te;
    tx_out = tx;
    return te < tx;
}

// Moller-Trumbore, no culling; returns t,u,v.
inline bool tri_hit(f3 o, f3 d, f3 v0, f3 v1, f3 v2, float& t, float& u, float& v) {
    f3 e1 = v1 - v0, e2 = v2 - v0;
    f3 pv = cross(d, e2);
    float det = dot(e1, pv);
    if (det == 0.0f) return false;
    float inv = 1.0f / det;
    f3 tv = o - v0;
    u = dot(tv, pv) * inv;
    if (u < 0.0f || u > 1.0f) return false;
    f3 qv = cross(tv, e1);
    v = dot(d, qv) * inv;
    if (v < 0.0f || u + v > 1.0f) return false;
    t = dot(e2, qv) * inv;
    return true;
}

uint32_t g_ico_idx[60];
float g_ico_v[36];
struct IcoInit { IcoInit() { icosahedron(g_ico_v, g_ico_idx); } } g_ico_init;

inline int proxy_hit_triangles(const Proxy& P, f3 o, f3 d, float t_out[2]) {
    // validation mode: all 20 world-space triangles; a convex closed proxy yields 0 or 2 crossings
    int n = 0;
    float ts[20];
    for (int f = 0; f < 20; ++f) {
        float t, u, v;
        if (tri_hit(o, d, ld3(P.wv[g_ico_idx[3 * f]]), ld3(P.wv[g_ico_idx[3 * f + 1]]), ld3(P.wv[g_ico_idx[3 * f + 2]]), t, u, v))
            ts[n++] = t;
    }
    if (n < 2) return 0;
    std::sort(ts, ts + n);
    t_out[0] = ts[0];
    t_out[1] = ts[n - 1];  // edge-shared duplicates collapse to first/last
    return 2;
}

bool make_proxy(const OrcParticle& p, uint32_t id, float alpha_min, Proxy& out) {
    // GaussianTracer.cpp:304-311.  opacity <= alpha_min gives s = 0 or NaN: a degenerate instance
    // that is never hit -> culled.
    if (!(p.opacity > alpha_min)) return false;
    float s = sqrtf(2.0f * logf(p.opacity / alpha_min));
    if (!(s > 0.0f) || !std::isfinite(s)) return false;
    m3 R = mat3_cast(p.rotation);
    for (int r = 0; r < 3; ++r) {
        float inv = 1.0f / p.scale[r];
        for (int c = 0; c < 3; ++c) out.A[r][c] = inv * R.c[r][c];
        out.mu[r] = p.position[r];
    }
    out.s = s;
    out.id = id;
    // world vertices: T * (R * S) * v  ==  mu + R * (scale*s * v)
    for (int k = 0; k < 3; ++k) { out.lo[k] = INFINITY; out.hi[k] = -INFINITY; }
    for (int i = 0; i < 12; ++i) {
        f3 sv = mk(p.scale[0] * s * g_ico_v[3 * i], p.scale[1] * s * g_ico_v[3 * i + 1], p.scale[2] * s * g_ico_v[3 * i + 2]);
        f3 w = mul(R, sv) + mk(p.position[0], p.position[1], p.position[2]);
        out.wv[i][0] = w.x; out.wv[i][1] = w.y; out.wv[i][2] = w.z;
        for (int k = 0; k < 3; ++k) { out.lo[k] = fminf(out.lo[k], out.wv[i][k]); out.hi[k] = fmaxf(out.hi[k], out.wv[i][k]); }
    }
    for (int k = 0; k < 3; ++k) {  // conservative padding (the BVH must never cull a slab hit)
        float pad = 1e-4f * (out.hi[k] - out.lo[k]) + 1e-5f * fmaxf(fabsf(out.lo[k]), fabsf(out.hi[k])) + 1e-7f;
        out.lo[k] -= pad; out.hi[k] += pad;
    }
    for (int k = 0; k < 3; ++k) if (!std::isfinite(out.lo[k]) || !std::isfinite(out.hi[k])) return false;
    return true;
}

// ----------------------------------------------------------------------------- CPU BVH (median split; not in the reference)
struct Box { float lo[3], hi[3]; };
struct BvhNode { Box b; uint32_t left, right, first, count; };  // count>0: leaf
struct Bvh {
    std::vector<BvhNode> nodes;
    std::vector<uint32_t> prims;
    void build(const std::vector<Box>& boxes) {
        nodes.clear();
        prims.resize(boxes.size());
        for (size_t i = 0; i < boxes.size(); ++i) prims[i] = (uint32_t)i;
        if (boxes.empty()) return;
        nodes.reserve(boxes.size());
        rec(boxes, 0, (uint32_t)boxes.size());
    }
    uint32_t rec(const std::vector<Box>& boxes, uint32_t first, uint32_t last) {
        uint32_t me = (uint32_t)nodes.size();
        nodes.push_back(BvhNode());
        Box b, cb;
        for (int k = 0; k < 3; ++k) { b.lo[k] = cb.lo[k] = INFINITY; b.hi[k] = cb.hi[k] = -INFINITY; }
        for (uint32_t i = first; i < last; ++i) {
            const Box& p = boxes[prims[i]];
            for (int k = 0; k < 3; ++k) {
                b.lo[k] = fminf(b.lo[k], p.lo[k]); b.hi[k] = fmaxf(b.hi[k], p.hi[k]);
                float c = 0.5f * (p.lo[k] + p.hi[k]);
                cb.lo[k] = fminf(cb.lo[k], c); cb.hi[k] = fmaxf(cb.hi[k], c);
            }
        }
        nodes[me].b = b;
        uint32_t n = last - first;
        if (n <= 4) { nodes[me].first = first; nodes[me].count = n; nodes[me].left = nodes[me].right = 0; return me; }
        int ax = 0;
        float ext = cb.hi[0] - cb.lo[0];
        for (int k = 1; k < 3; ++k) if (cb.hi[k] - cb.lo[k] > ext) { ext = cb.hi[k] - cb.lo[k]; ax = k; }
        uint32_t mid = first + n / 2;
        std::nth_element(prims.begin() + first, prims.begin() + mid, prims.begin() + last, [&](uint32_t a, uint32_t c) {
            return boxes[a].lo[ax] + boxes[a].hi[ax] < boxes[c].lo[ax] + boxes[c].hi[ax];
        });
        nodes[me].count = 0; nodes[me].first = 0;
        uint32_t l = rec(boxes, first, mid);
        uint32_t r = rec(boxes, mid, last);
        nodes[me].left = l; nodes[me].right = r;
        return me;
    }
};
inline bool box_hit(const Box& b, f3 o, f3 inv, float tmin, float tmax, float* t_enter = nullptr) {
    float t0 = tmin, t1 = tmax;
    const float oo[3] = {o.x, o.y, o.z}, ii[3] = {inv.x, inv.y, inv.z};
    for (int k = 0; k < 3; ++k) {
        float a = (b.lo[k] - oo[k]) * ii[k], c = (b.hi[k] - oo[k]) * ii[k];
        // fminf/fmaxf drop NaN (0*inf when the origin lies on a slab plane of an axis-parallel ray)
        t0 = fmaxf(t0, fminf(a, c));
        t1 = fminf(t1, fmaxf(a, c));
    }
    if (t_enter) *t_enter = t0;
    return t0 <= t1 * 1.00001f + 1e-7f;
}

}  // namespace

// ----------------------------------------------------------------------------- scene
struct OrcScene {
    std::vector<OrcParticle> particles;
    std::vector<Proxy> proxies;  // live only
    Bvh bvh;
    float alpha_min_build;
    // meshes (flattened copies)
    struct MeshCopy { std::vector<float> v, n; std::vector<uint32_t> f; };
    std::vector<MeshCopy> meshes;
    struct TriRef { uint32_t mesh, face; };
    std::vector<TriRef> tris;
    Bvh mesh_bvh;
};

namespace {

struct HitPayload { uint32_t particleId; float distance; };

struct Counters {
    uint64_t segments = 0, rounds = 0, hg = 0, hs = 0, nodes = 0, tested = 0;
};

struct PixelCtx {
    const OrcScene* S;
    const OrcRenderArgs* A;
    const OrcParams* prm;
    Counters* cnt;
    uint32_t flags;
    uint32_t hg_pixel;
    uint32_t last_id;  // particle whose face hit defines the current re-launch point
    // trace
    uint32_t* trace_ids;
    uint32_t trace_max, trace_n;
};

// anyhit (tracer.cu:136-153) applied to one candidate face hit.  Returns the (possibly shrunk) tmax.
inline void anyhit(HitPayload* buf, uint32_t K, uint32_t id, float t, float& cur_tmax) {
    HitPayload hit{id, t};
    if (hit.distance < buf[K - 1].distance) {
        for (uint32_t i = 0; i < K; ++i) {  // compareAndSwapHitPayloadValue, strict <
            if (hit.distance < buf[i].distance) std::swap(hit, buf[i]);
        }
        if (buf[K - 1].distance > t) return;  // optixIgnoreIntersection(): tmax unchanged
    }
    cur_tmax = t;  // accepted: traversal tmax shrinks
}

inline void offer(PixelCtx& c, HitPayload* buf, uint32_t K, const Proxy& P, f3 o, f3 d, float tmin, float& cur_tmax) {
    c.cnt->tested++;
    float t2[2];
    if (c.A->use_triangles) {
        if (!proxy_hit_triangles(P, o, d, t2)) return;
    } else {
        if (!proxy_hit_slabs(P, o, d, t2[0], t2[1])) return;
    }
    for (int k = 0; k < 2; ++k) {
        float t = t2[k];
        if (P.id != c.last_id && ulp_close(t, tmin, 4)) c.flags |= ORC_FLAG_TIE;  // another proxy's face hit sitting on the re-launch point
        if (t > tmin && t < cur_tmax) anyhit(buf, K, P.id, t, cur_tmax);
    }
}

// traceGPs (tracer.cuh:289-326) with optixTrace emulated
void traceGPs(PixelCtx& c, HitPayload* buf, uint32_t K, f3 o, f3 d, float tmin, float tmax) {
    for (uint32_t i = 0; i < K; ++i) buf[i] = HitPayload{InvalidParticleId, InfiniteDistance};
    c.cnt->rounds++;
    float cur_tmax = tmax;
    const OrcScene& S = *c.S;
    if (c.A->brute_force) {
        for (const Proxy& P : S.proxies) offer(c, buf, K, P, o, d, tmin, cur_tmax);
        return;
    }
    if (S.bvh.nodes.empty()) return;
    f3 inv = mk(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    uint32_t stack[256];
    int sp = 0;
    if (!box_hit(S.bvh.nodes[0].b, o, inv, tmin, cur_tmax)) return;
    stack[sp++] = 0;
    while (sp) {
        const BvhNode& n = S.bvh.nodes[stack[--sp]];
        c.cnt->nodes++;
        if (n.count) {
            if (!box_hit(n.b, o, inv, tmin, cur_tmax)) continue;  // tmax may have shrunk since the push
            for (uint32_t i = 0; i < n.count; ++i) offer(c, buf, K, S.proxies[S.bvh.prims[n.first + i]], o, d, tmin, cur_tmax);
        } else {
            // near child first so that the k-buffer bound tightens early (any order gives the same K smallest)
            float tl, tr;
            const bool hl = box_hit(S.bvh.nodes[n.left].b, o, inv, tmin, cur_tmax, &tl);
            const bool hr = box_hit(S.bvh.nodes[n.right].b, o, inv, tmin, cur_tmax, &tr);
            if (hl && hr) {
                if (tl <= tr) { stack[sp++] = n.right; stack[sp++] = n.left; }
                else { stack[sp++] = n.left; stack[sp++] = n.right; }
            } else if (hl) stack[sp++] = n.left;
            else if (hr) stack[sp++] = n.right;
        }
    }
}

// computeResponse (tracer.cuh:187-214)
inline float computeResponse(const OrcParticle& gp, f3 o, f3 d) {
    f3 mu = mk(gp.position[0], gp.position[1], gp.position[2]);
    m3 R = mat3_cast(gp.rotation);
    // invCov = inv_s * transpose(R): invCov.c[c][r] = (1/scale_r) * R.c[r][c]
    m3 invCov;
    const float is[3] = {1.0f / gp.scale[0], 1.0f / gp.scale[1], 1.0f / gp.scale[2]};
    for (int cc = 0; cc < 3; ++cc)
        for (int r = 0; r < 3; ++r) invCov.c[cc][r] = is[r] * R.c[r][cc];
    f3 o_g = mul(invCov, o - mu);
    f3 d_g = mul(invCov, d);
    float d_val = -dot(o_g, d_g) / fmaxf(1e-6f, dot(d_g, d_g));
    f3 pos = o + d_val * d;
    f3 p_g = mul(invCov, mu - pos);
    return expf(-0.5f * dot(p_g, p_g));
}

// SHToRadiance + computeRadiance (tracer.cuh:216-264)
inline f3 computeRadiance(const OrcParticle& gp, f3 d, uint32_t deg) {
    f3 sh[16];
    for (int i = 0; i < 16; ++i) sh[i] = mk(gp.sh[3 * i], gp.sh[3 * i + 1], gp.sh[3 * i + 2]);
    f3 L = mk(0.5f) + SH_C0 * sh[0];
    if (deg != 0) {
        float x = d.x, y = d.y, z = d.z;
        L += SH_C1 * (-y * sh[1] + z * sh[2] - x * sh[3]);
        if (deg != 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = x * y, xz = x * z, yz = y * z;
            // ":243" evaluates (2. * zz - xx - yy) in double, then SH_C2_2 * (double) in double, then * float3
            float c6 = (float)((double)SH_C2_2 * (2. * (double)zz - (double)xx - (double)yy));
            L += SH_C2_0 * xy * sh[4] + SH_C2_1 * yz * sh[5] + c6 * sh[6] + SH_C2_3 * xz * sh[7] + SH_C2_4 * (xx - yy) * sh[8];
            if (deg != 2) {
                L += SH_C3_0 * y * (3.0f * xx - yy) * sh[9] + SH_C3_1 * xy * z * sh[10] + SH_C3_2 * y * (4.0f * zz - xx - yy) * sh[11] +
                     SH_C3_3 * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[12] + SH_C3_4 * x * (4.0f * zz - xx - yy) * sh[13] +
                     SH_C3_5 * z * (xx - yy) * sh[14] + SH_C3_6 * x * (xx - 3.0f * yy) * sh[15];
            }
        }
    }
    return mk(fmaxf(L.x, 0.0f), fmaxf(L.y, 0.0f), fmaxf(L.z, 0.0f));
}

struct RayData { f3 radiance; float density; };

// trace (tracer.cuh:328-373)
void trace(PixelCtx& c, RayData& rayData, f3 ray_o, f3 ray_d, float t_min, float t_max) {
    const OrcParams& params = *c.prm;
    const uint32_t K = c.A->K;
    float rayTransmittance = 1.0f - rayData.density;
    constexpr float epsT = 1e-9f;
    float rayLastHitDistance = t_min;
    f3 radiance = mk(0.0f);
    HitPayload buf[64];
    float prev_t = -1.0f;
    c.cnt->segments++;
    c.last_id = InvalidParticleId;
    auto tcheck = [&](float T) {
        if (fabsf(T - params.minTransmittance) <= 5e-8f) c.flags |= ORC_FLAG_TRANSMIT;
        return T > params.minTransmittance;
    };
    while ((rayLastHitDistance <= t_max) && tcheck(rayTransmittance)) {
        traceGPs(c, buf, K, ray_o, ray_d, rayLastHitDistance + epsT, t_max + epsT);
        if (buf[0].particleId == InvalidParticleId) break;
        for (uint32_t i = 0; i < K; ++i) {
            const HitPayload hit = buf[i];
            if ((hit.particleId != InvalidParticleId) && tcheck(rayTransmittance)) {
                const OrcParticle& currParticle = c.S->particles[hit.particleId];
                if (ulp_close(hit.distance, prev_t, 4) || ulp_close(hit.distance, t_max + epsT, 4)) c.flags |= ORC_FLAG_TIE;
                prev_t = hit.distance;
                c.cnt->hg++;
                c.hg_pixel++;
                if (c.trace_ids) { if (c.trace_n < c.trace_max) c.trace_ids[c.trace_n] = hit.particleId; c.trace_n++; }
                if (hit.distance >= rayLastHitDistance) c.last_id = hit.particleId;
                rayLastHitDistance = fmaxf(hit.distance, rayLastHitDistance);
                float hitAlpha = computeResponse(currParticle, ray_o, ray_d);
                hitAlpha = fminf(0.99f, hitAlpha * currParticle.opacity);
                if (fabsf(hitAlpha - params.alpha_min) <= 5e-7f) c.flags |= ORC_FLAG_ALPHA;
                if (params.alpha_min < hitAlpha) {
                    f3 hitRadiance = computeRadiance(currParticle, normalize(ray_d), params.sh_degree_max);
                    radiance += rayTransmittance * hitRadiance * hitAlpha;
                    rayTransmittance *= (1.0f - hitAlpha);
                    c.cnt->hs++;
                }
            }
        }
    }
    rayData.radiance = radiance;
    rayData.density = 1.0f - rayTransmittance;
}

struct RayPayload {
    float t_hit;
    f3 currRayOrigin, currRayDirection, hitNormal;
    unsigned int numBounces;
    f3 accumColor;
    float accumAlpha;
    f3 directLight;
    float blockingRadiance;
};

// refract (tracer.cuh:432-465)
inline void refract(f3& newRayDirction, f3 ray_d, f3 normal, float etai_over_etat, float& t_hit, unsigned int& numBounces) {
    float ri;
    if (dot(ray_d, normal) < 0.0f) {
        ri = 1.0f / etai_over_etat;
    } else {
        ri = etai_over_etat;
        normal = -normal;
    }
    float cos_theta = fminf(dot(-ray_d, normal), 1.0f);
    float sin_theta = sqrtf(1.0f - cos_theta * cos_theta);
    bool cannot_refract = ri * sin_theta > 1.0f;
    if (cannot_refract) {
        f3 reflected_normal = dot(ray_d, normal) < 0.0f ? normal : -normal;
        newRayDirction = reflect(ray_d, reflected_normal);
        numBounces += 1;
    } else {
        f3 r_out_perp = ri * (ray_d + cos_theta * normal);
        f3 r_out_parallel = -sqrtf(fabsf(1.0f - dot(r_out_perp, r_out_perp))) * normal;
        newRayDirction = r_out_perp + r_out_parallel;
        t_hit += REFRACTION_EPS_SHIFT;
    }
}

// closest mesh hit in (TRACE_MESH_TMIN, TRACE_MESH_TMAX), no culling, anyhit disabled (tracer.cuh:266-287)
bool mesh_closest(PixelCtx& c, f3 o, f3 d, uint32_t& mesh, uint32_t& face, float& t_hit, float& bu, float& bv) {
    const OrcScene& S = *c.S;
    if (S.tris.empty()) return false;
    float best = TRACE_MESH_TMAX;
    float second = INFINITY;
    bool found = false;
    auto test = [&](uint32_t ti) {
        const OrcScene::TriRef& tr = S.tris[ti];
        const OrcScene::MeshCopy& M = S.meshes[tr.mesh];
        const uint32_t* f = &M.f[3 * tr.face];
        float t, u, v;
        if (!tri_hit(o, d, ld3(&M.v[3 * f[0]]), ld3(&M.v[3 * f[1]]), ld3(&M.v[3 * f[2]]), t, u, v)) return;
        if (!(t > TRACE_MESH_TMIN)) return;
        if (t < best) { second = best; best = t; mesh = tr.mesh; face = tr.face; bu = u; bv = v; found = true; }
        else if (t < second) second = t;
    };
    if (c.A->brute_force) {
        for (uint32_t i = 0; i < S.tris.size(); ++i) test(i);
    } else {
        f3 inv = mk(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
        uint32_t stack[128];
        int sp = 0;
        stack[sp++] = 0;
        while (sp) {
            const BvhNode& n = S.mesh_bvh.nodes[stack[--sp]];
            if (!box_hit(n.b, o, inv, TRACE_MESH_TMIN, best)) continue;
            if (n.count) for (uint32_t i = 0; i < n.count; ++i) test(S.mesh_bvh.prims[n.first + i]);
            else { stack[sp++] = n.right; stack[sp++] = n.left; }
        }
    }
    // NOTE: neighbouring triangles of a tessellated surface share edges, so "second within 4 ulp" is the
    // normal case on an edge and gives the same t; only flag when the two candidates are different meshes
    // is not tracked -- we simply do not flag here (t identical => same segment end either way).
    t_hit = best;
    return found;
}

void render_pixel(PixelCtx& c, uint32_t px, uint32_t py, f3& rgb_out) {
    const OrcParams& params = *c.prm;
    // raygen (tracer.cu:17-110)
    RayPayload payload;
    payload.t_hit = 0.0f;
    payload.numBounces = 0;
    payload.hitNormal = mk(0.0f);
    payload.accumColor = mk(0.0f);
    payload.accumAlpha = 0.0f;
    payload.blockingRadiance = 0.0f;
    payload.directLight = mk(0.0f);
    RayData rayData;
    rayData.radiance = mk(0.0f);
    rayData.density = 0.0f;
    unsigned int state = TraceGaussianPass;

    float o[3], d[3];
    OrcParams tmp = params;
    if (!orc_get_ray(&tmp, px, py, o, d)) {  // fisheye r > 1: DEFINED as a black pixel
        rgb_out = mk(0.0f);
        return;
    }
    payload.currRayOrigin = ld3(o);
    payload.currRayDirection = ld3(d);

    unsigned int timeout = 0;
    while ((length(payload.currRayDirection) > 0.1f) && (payload.numBounces < MAX_BOUNCES)) {
        const f3 ray_o = payload.currRayOrigin;
        const f3 ray_d = payload.currRayDirection;

        // traceMesh (tracer.cuh:266-287)
        state = TraceMeshPass;
        {
            uint32_t mesh = 0, face = 0;
            float t_hit = 0, bu = 0, bv = 0;
            if (!mesh_closest(c, ray_o, ray_d, mesh, face, t_hit, bu, bv)) {
                // __miss__miss (tracer.cu:112-122)
                payload.currRayOrigin = mk(0.0f);
                payload.currRayDirection = mk(0.0f);
                state = TraceLastGaussianPass;
            } else {
                // __closesthit__closesthit (tracer.cu:155-187)
                unsigned int numBounces = payload.numBounces;
                unsigned int nextState = TraceGaussianPass;
                const OrcScene::MeshCopy& M = c.S->meshes[mesh];
                const uint32_t* f = &M.f[3 * face];
                // getBarycentricNormal (tracer.cuh:167-185)
                f3 n0 = ld3(&M.n[3 * f[0]]), n1 = ld3(&M.n[3 * f[1]]), n2 = ld3(&M.n[3 * f[2]]);
                float w0 = 1.0f - bu - bv, w1 = bu, w2 = bv;
                f3 normal = normalize(w0 * n0 + w1 * n1 + w2 * n2);
                f3 newRayDirection = mk(0.0f);
                if (params.type == MIRROR) {  // renderMirror (tracer.cuh:396-404)
                    newRayDirection = reflect(ray_d, normal);
                    numBounces += 1;
                } else if (params.type == NORMAL) {  // renderNormal (tracer.cuh:406-429)
                    trace(c, rayData, ray_o, ray_d, params.t_min, t_hit);
                    f3 radiance = rayData.radiance;
                    float alpha = rayData.density;
                    payload.accumColor += radiance;
                    payload.accumAlpha += alpha;
                    const f3 normalColor = (normal + 1.0f) / 2.0f;
                    payload.accumColor += normalColor * (1.0f - alpha);
                    payload.accumAlpha += (1.0f - alpha);
                    nextState = TraceTerminate;
                } else if (params.type == GLASS) {  // renderGlass (tracer.cuh:467-482)
                    float n1i = 1.0003f, n2i = 1.5f;
                    refract(newRayDirection, ray_d, normal, n2i / n1i, t_hit, numBounces);
                }
                payload.t_hit = t_hit;
                payload.currRayOrigin = ray_o + t_hit * ray_d;
                payload.currRayDirection = newRayDirection;
                payload.hitNormal = normal;
                payload.numBounces = numBounces;
                state = nextState;
            }
        }
        if (state == TraceTerminate) break;

        if (state == TraceLastGaussianPass) {
            trace(c, rayData, ray_o, ray_d, params.t_min, params.t_max);
            f3 radiance = rayData.radiance;
            float alpha = rayData.density;
            payload.directLight = radiance * alpha;
            payload.accumAlpha = clampf(payload.accumAlpha + alpha, 0.0f, 1.0f);
            state = TraceTerminate;
        } else {
            trace(c, rayData, ray_o, ray_d, params.t_min, payload.t_hit);
            f3 radiance = rayData.radiance;
            float alpha = rayData.density;
            payload.accumColor += mk(1.0f - payload.accumAlpha) * radiance;
            payload.accumAlpha = clampf(payload.accumAlpha + alpha, 0.0f, 1.0f);
            payload.blockingRadiance = clampf(payload.blockingRadiance + alpha, 0.0f, 1.0f);
        }
        payload.accumColor += payload.directLight * (1.0f - payload.blockingRadiance);
        timeout += 1;
        if (timeout > TIMEOUT_ITERATIONS) break;
    }
    rgb_out = payload.accumColor;
}

}  // namespace

// ----------------------------------------------------------------------------- C interface
extern "C" {

OrcScene* orc_scene_create(const OrcParticle* particles, uint64_t n, float alpha_min_build) {
    OrcScene* S = new OrcScene();
    S->particles.assign(particles, particles + n);
    S->alpha_min_build = alpha_min_build;
    S->proxies.reserve(n);
    for (uint64_t i = 0; i < n; ++i) {
        Proxy P;
        if (make_proxy(particles[i], (uint32_t)i, alpha_min_build, P)) S->proxies.push_back(P);
    }
    std::vector<Box> boxes(S->proxies.size());
    for (size_t i = 0; i < boxes.size(); ++i) {
        memcpy(boxes[i].lo, S->proxies[i].lo, 12);
        memcpy(boxes[i].hi, S->proxies[i].hi, 12);
    }
    S->bvh.build(boxes);
    return S;
}
void orc_scene_destroy(OrcScene* s) { delete s; }
uint64_t orc_scene_live_count(const OrcScene* s) { return s->proxies.size(); }

void orc_scene_set_meshes(OrcScene* S, const OrcMesh* meshes, uint32_t n_meshes) {
    S->meshes.clear();
    S->tris.clear();
    std::vector<Box> boxes;
    for (uint32_t m = 0; m < n_meshes; ++m) {
        OrcScene::MeshCopy mc;
        mc.v.assign(meshes[m].vertices, meshes[m].vertices + 3 * (size_t)meshes[m].n_vertices);
        mc.n.assign(meshes[m].normals, meshes[m].normals + 3 * (size_t)meshes[m].n_vertices);
        mc.f.assign(meshes[m].faces, meshes[m].faces + 3 * (size_t)meshes[m].n_faces);
        for (uint32_t f = 0; f < meshes[m].n_faces; ++f) {
            Box b;
            for (int k = 0; k < 3; ++k) { b.lo[k] = INFINITY; b.hi[k] = -INFINITY; }
            for (int j = 0; j < 3; ++j)
                for (int k = 0; k < 3; ++k) {
                    float x = mc.v[3 * mc.f[3 * f + j] + k];
                    b.lo[k] = fminf(b.lo[k], x); b.hi[k] = fmaxf(b.hi[k], x);
                }
            for (int k = 0; k < 3; ++k) {
                float pad = 1e-4f * (b.hi[k] - b.lo[k]) + 1e-5f * fmaxf(fabsf(b.lo[k]), fabsf(b.hi[k])) + 1e-7f;
                b.lo[k] -= pad; b.hi[k] += pad;
            }
            boxes.push_back(b);
            S->tris.push_back(OrcScene::TriRef{m, f});
        }
        S->meshes.push_back(std::move(mc));
    }
    S->mesh_bvh.build(boxes);
}

int orc_render(const OrcScene* S, const OrcRenderArgs* A, OrcStats* stats) {
    const OrcParams& prm = A->params;
    if (A->K < 1 || A->K > 64) return 1;
    uint32_t x0 = A->x0, y0 = A->y0, w = A->w, h = A->h;
    if (w == 0 || h == 0) { x0 = y0 = 0; w = prm.width; h = prm.height; }
    if (x0 + w > prm.width || y0 + h > prm.height) return 2;
    uint32_t nthreads = A->threads ? A->threads : std::max(1u, std::thread::hardware_concurrency());
    std::vector<Counters> cnts(nthreads);
    std::vector<uint64_t> flagged(nthreads, 0);
    // pixel -> trace slot lookup
    std::vector<int32_t> trace_slot;
    if (A->n_trace_pixels && A->trace_ids) {
        trace_slot.assign((size_t)prm.width * prm.height, -1);
        for (uint32_t i = 0; i < A->n_trace_pixels; ++i) trace_slot[A->trace_pixels[i]] = (int32_t)i;
    }
    std::atomic<uint32_t> next_row{0};
    auto t_begin = std::chrono::steady_clock::now();
    auto worker = [&](uint32_t tid) {
        for (;;) {
            uint32_t row = next_row.fetch_add(1);
            if (row >= h) break;
            uint32_t y = y0 + row;
            for (uint32_t x = x0; x < x0 + w; ++x) {
                PixelCtx c;
                c.S = S; c.A = A; c.prm = &prm; c.cnt = &cnts[tid];
                c.flags = 0; c.hg_pixel = 0; c.last_id = InvalidParticleId;
                c.trace_ids = nullptr; c.trace_max = 0; c.trace_n = 0;
                size_t pix = (size_t)y * prm.width + x;
                int32_t slot = trace_slot.empty() ? -1 : trace_slot[pix];
                if (slot >= 0) { c.trace_ids = A->trace_ids + (size_t)slot * A->trace_max; c.trace_max = A->trace_max; }
                f3 rgb;
                render_pixel(c, x, y, rgb);
                if (slot >= 0 && A->trace_counts) A->trace_counts[slot] = c.trace_n;
                if (A->out_rgb32f) { A->out_rgb32f[3 * pix] = rgb.x; A->out_rgb32f[3 * pix + 1] = rgb.y; A->out_rgb32f[3 * pix + 2] = rgb.z; }
                if (A->out_rgb8) {  // writeOutputBuffer (tracer.cuh:484-496)
                    f3 q = clamp3(rgb, 0.0f, 1.0f);
                    A->out_rgb8[3 * pix] = orc_quantize(q.x);
                    A->out_rgb8[3 * pix + 1] = orc_quantize(q.y);
                    A->out_rgb8[3 * pix + 2] = orc_quantize(q.z);
                }
                if (A->out_flags) A->out_flags[pix] = (uint8_t)c.flags;
                if (A->out_hg) A->out_hg[pix] = (uint16_t)std::min<uint32_t>(c.hg_pixel, 65535u);
                if (c.flags) flagged[tid]++;
            }
        }
    };
    std::vector<std::thread> pool;
    for (uint32_t t = 1; t < nthreads; ++t) pool.emplace_back(worker, t);
    worker(0);
    for (auto& t : pool) t.join();
    auto t_end = std::chrono::steady_clock::now();
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        stats->rays = (uint64_t)w * h;
        for (uint32_t t = 0; t < nthreads; ++t) {
            stats->segments += cnts[t].segments; stats->rounds += cnts[t].rounds;
            stats->hits_gathered += cnts[t].hg; stats->hits_shaded += cnts[t].hs;
            stats->nodes_visited += cnts[t].nodes; stats->proxies_tested += cnts[t].tested;
            stats->flagged += flagged[t];
        }
        stats->seconds = std::chrono::duration<double>(t_end - t_begin).count();
        stats->threads = nthreads;
    }
    return 0;
}

float orc_compute_response(const OrcParticle* p, const float o[3], const float d[3]) { return computeResponse(*p, ld3(o), ld3(d)); }

void orc_sh_radiance(const OrcParticle* p, const float d[3], uint32_t deg, float out[3]) {
    f3 L = computeRadiance(*p, ld3(d), deg);
    out[0] = L.x; out[1] = L.y; out[2] = L.z;
}

uint8_t orc_quantize(float x) {  // quantizeUnsigned8Bits (tracer.cuh:68-73)
    x = clampf(x, 0.0f, 1.0f);
    enum { N = (1 << 8) - 1, Np1 = (1 << 8) };
    return (unsigned char)std::min((unsigned int)(x * (float)Np1), (unsigned int)N);
}

int orc_get_ray(const OrcParams* prm, uint32_t ix, uint32_t iy, float o[3], float dd[3]) {
    // raygen passes -U, -V, W (tracer.cu:36-56)
    const f3 U = -ld3(prm->U), V = -ld3(prm->V), W = ld3(prm->W);
    const int width = (int)prm->width, height = (int)prm->height;
    // float2 d = 2.0f * make_float2((x + 0.5)/w, (y + 0.5)/h) - 1.0f   (tracer.cuh:126-130 / 147-151)
    float dx = 2.0f * ((static_cast<float>(ix) + 0.5f) / static_cast<float>(width)) - 1.0f;
    float dy = 2.0f * ((static_cast<float>(iy) + 0.5f) / static_cast<float>(height)) - 1.0f;
    f3 dir;
    if (!prm->mode_fisheye) {  // getRay (tracer.cuh:115-134)
        dir = normalize(dx * U + dy * V + W);
    } else {  // getFishEyeRay (tracer.cuh:136-165)
        float r = sqrtf(dx * dx + dy * dy);
        if (r > 1.0f) return 0;
        float f = 1.0f / sqrtf(2.0f);
        float theta = 2.0f * asinf(r / (2.0f * f));
        float phi = atan2f(dy, dx);
        f3 direction = mk(sinf(theta) * cosf(phi), sinf(theta) * sinf(phi), cosf(theta));
        dir = normalize(direction.x * U + direction.y * V + direction.z * W);
    }
    o[0] = prm->eye[0]; o[1] = prm->eye[1]; o[2] = prm->eye[2];
    dd[0] = dir.x; dd[1] = dir.y; dd[2] = dir.z;
    return 1;
}

int orc_proxy_hit(const OrcParticle* p, float alpha_min_build, const float o[3], const float d[3], int use_triangles, float t_out[2]) {
    Proxy P;
    if (!make_proxy(*p, 0, alpha_min_build, P)) return 0;
    if (use_triangles) return proxy_hit_triangles(P, ld3(o), ld3(d), t_out);
    return proxy_hit_slabs(P, ld3(o), ld3(d), t_out[0], t_out[1]) ? 2 : 0;
}

void orc_icosahedron(float verts[36], uint32_t idx[60]) { icosahedron(verts, idx); }

void orc_refract(const float d[3], const float n[3], float eta, float out_dir[3], float* t_hit, uint32_t* bounces) {
    f3 nd;
    unsigned int b = *bounces;
    refract(nd, ld3(d), ld3(n), eta, *t_hit, b);
    *bounces = b;
    out_dir[0] = nd.x; out_dir[1] = nd.y; out_dir[2] = nd.z;
}

void orc_transform_mesh(const float* verts, const float* normals, uint32_t n, const float T[16], float* vw, float* no) {
    // glm column-major mat4: T[4*c + r].  world = M * vec4(v,1) (what an OptiX instance transform does to the
    // GAS vertices); normals = mat3(transform) * n, NOT inverse-transpose (GaussianTracer.cpp:658-663).
    for (uint32_t i = 0; i < n; ++i) {
        float x = verts[3 * i], y = verts[3 * i + 1], z = verts[3 * i + 2];
        for (int r = 0; r < 3; ++r) vw[3 * i + r] = T[0 + r] * x + T[4 + r] * y + T[8 + r] * z + T[12 + r];
        float nx = normals[3 * i], ny = normals[3 * i + 1], nz = normals[3 * i + 2];
        for (int r = 0; r < 3; ++r) no[3 * i + r] = T[0 + r] * nx + T[4 + r] * ny + T[8 + r] * nz;
    }
}

// ---- batch forms of the known-answer hooks (tests/test_ref_pins.py compares them with the reference-compiled functions of
// oracle/_ref on 1e5 random inputs each)
void orc_compute_response_batch(const OrcParticle* ps, uint64_t n, const uint32_t* pidx, const float* o, const float* d, float* out) {
    for (uint64_t i = 0; i < n; ++i) out[i] = computeResponse(ps[pidx[i]], ld3(o + 3 * i), ld3(d + 3 * i));
}
void orc_sh_radiance_batch(const OrcParticle* ps, uint64_t n, const uint32_t* pidx, const float* d, uint32_t deg, int normalise, float* out) {
    for (uint64_t i = 0; i < n; ++i) {
        f3 dir = ld3(d + 3 * i);
        if (normalise) dir = normalize(dir);  // tracer.cuh:262
        const f3 L = computeRadiance(ps[pidx[i]], dir, deg);
        out[3 * i] = L.x; out[3 * i + 1] = L.y; out[3 * i + 2] = L.z;
    }
}
void orc_quantize_batch(uint64_t n, const float* x, uint8_t* out) {
    for (uint64_t i = 0; i < n; ++i) out[i] = orc_quantize(x[i]);
}
void orc_get_rays(const OrcParams* prm, uint64_t n, const uint32_t* xs, const uint32_t* ys, float* o, float* d, uint8_t* valid) {
    for (uint64_t i = 0; i < n; ++i) {
        o[3 * i] = o[3 * i + 1] = o[3 * i + 2] = d[3 * i] = d[3 * i + 1] = d[3 * i + 2] = NAN;
        valid[i] = (uint8_t)orc_get_ray(prm, xs[i], ys[i], o + 3 * i, d + 3 * i);
    }
}
void orc_refract_batch(uint64_t n, const float* d, const float* nrm, float eta, float* out_dir, float* t_hit_io, uint32_t* bounces_io) {
    for (uint64_t i = 0; i < n; ++i) orc_refract(d + 3 * i, nrm + 3 * i, eta, out_dir + 3 * i, t_hit_io + i, bounces_io + i);
}
void orc_mirror_batch(uint64_t n, const float* d, const float* nrm, float* out_dir, uint32_t* bounces_io) {  // renderMirror, tracer.cuh:396-404
    for (uint64_t i = 0; i < n; ++i) {
        const f3 r = reflect(ld3(d + 3 * i), ld3(nrm + 3 * i));
        out_dir[3 * i] = r.x; out_dir[3 * i + 1] = r.y; out_dir[3 * i + 2] = r.z;
        bounces_io[i] += 1;
    }
}
// Derived proxy data of each particle: A (9 floats, row-major), s, the 12 world-space vertices, the padded box; live = 0 if culled.
void orc_proxy_data(const OrcParticle* ps, uint64_t n, float alpha_min_build, float* A9, float* s_out, float* wv36, float* box6, uint8_t* live) {
    for (uint64_t i = 0; i < n; ++i) {
        Proxy P;
        live[i] = make_proxy(ps[i], (uint32_t)i, alpha_min_build, P) ? 1 : 0;
        if (!live[i]) continue;
        memcpy(A9 + 9 * i, P.A, sizeof(float) * 9);
        s_out[i] = P.s;
        memcpy(wv36 + 36 * i, P.wv, sizeof(float) * 36);
        memcpy(box6 + 6 * i, P.lo, sizeof(float) * 3);
        memcpy(box6 + 6 * i + 3, P.hi, sizeof(float) * 3);
    }
}
// Slab test (PROXY-T) of particle pidx[i] against ray i: hit[i] = 1 and t2[2i], t2[2i+1] = (te, tx) when crossed.
void orc_proxy_hit_batch(const OrcParticle* ps, uint64_t n, const uint32_t* pidx, float alpha_min_build, const float* o, const float* d,
                         float* t2, uint8_t* hit) {
    for (uint64_t i = 0; i < n; ++i) {
        Proxy P;
        hit[i] = 0;
        if (!make_proxy(ps[pidx[i]], pidx[i], alpha_min_build, P)) continue;
        hit[i] = proxy_hit_slabs(P, ld3(o + 3 * i), ld3(d + 3 * i), t2[2 * i], t2[2 * i + 1]) ? 1 : 0;
    }
}
// Moller-Trumbore over an explicit triangle soup (12 vertices, 20 faces per proxy: e.g. the REFERENCE's icosahedron pushed through
// the REFERENCE's instance matrix): nearest and farthest crossing of ray i with soup i.
void orc_soup_hits(uint64_t n, const float* verts36, const uint32_t* idx60, const float* o, const float* d, float* t2, uint8_t* hit) {
    for (uint64_t i = 0; i < n; ++i) {
        const float* v = verts36 + 36 * i;
        int cnt = 0;
        float ts[20];
        for (int f = 0; f < 20; ++f) {
            float t, u, w;
            if (tri_hit(ld3(o + 3 * i), ld3(d + 3 * i), ld3(v + 3 * idx60[3 * f]), ld3(v + 3 * idx60[3 * f + 1]), ld3(v + 3 * idx60[3 * f + 2]), t, u, w)) ts[cnt++] = t;
        }
        hit[i] = cnt >= 2;
        if (cnt >= 2) { std::sort(ts, ts + cnt); t2[2 * i] = ts[0]; t2[2 * i + 1] = ts[cnt - 1]; }
    }
}

}  // extern "C"
