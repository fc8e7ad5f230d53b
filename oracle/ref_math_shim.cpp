/*
 * ref_math_shim.cpp -- C entry points around the REFERENCE's own hot-path math, compiled from where it lies under
 * /root/reference (never copied into this repository):
 *   - the pure functions of shaders/tracer.cuh (quantizeUnsigned8Bits :68-73, getRay / getFishEyeRay :115-165,
 *     computeResponse :187-214, SHToRadiance / computeRadiance :216-264, renderMirror :396-404, refract :432-465),
 *     extracted BY NAME at build time into oracle/_ref/gen/tracer_pure.inc (oracle/extract_ref.py) and compiled against
 *     the stub oracle/ref_stub/optix.h with the reference's own glm, vector_math.h and Parameters.h;
 *   - the proxy transform of GaussianTracer::createGaussianParticlesBVH (src/GaussianTracer.cpp:304-311) and the 3x4
 *     packing of createIAS (:405-409), the normal transform of sendGeometryAttributesToDevice (:658-663), the
 *     primitive placement of createPlane (:580-588), the defaults of initializeParams (:477-486) -- statement ranges
 *     extracted into gen/*.inc;
 *   - src/geometry/Primitives.cpp and src/geometry/Icosahedron.h as they are;
 *   - sizeof / offsetof of the real Params, Mesh and GaussianParticle (src/Parameters.h:36-74, src/GaussianData.h:12-20).
 * Output: part of oracle/_ref/libref_host.so (git-ignored; travels to the GPU box).
 * TEST INFRASTRUCTURE ONLY: pins oracle/oracle.cpp and the host mirror against reference-compiled code.
 */
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <cmath>

#include "Parameters.h"          // reference: includes <optix.h> (our stub), GaussianData.h, geometry/Primitives.h
#include "vector_math.h"         // reference
#include "geometry/Icosahedron.h" // reference

Params params;                    // tracer.cuh declares it __constant__; a plain global on the host

#define __PARAMS__ 1
#include "tracer_pure.inc"       // GENERATED from shaders/tracer.cuh (see header)

extern "C" {

// ---- layouts: numbers come from the reference's own headers
void ref_layout(uint64_t out[32]) {
    int k = 0;
    out[k++] = sizeof(Params);
    out[k++] = offsetof(Params, output_buffer);
    out[k++] = offsetof(Params, width);
    out[k++] = offsetof(Params, height);
    out[k++] = offsetof(Params, sh_degree_max);
    out[k++] = offsetof(Params, eye);
    out[k++] = offsetof(Params, U);
    out[k++] = offsetof(Params, V);
    out[k++] = offsetof(Params, W);
    out[k++] = offsetof(Params, t_min);
    out[k++] = offsetof(Params, t_max);
    out[k++] = offsetof(Params, minTransmittance);
    out[k++] = offsetof(Params, alpha_min);
    out[k++] = offsetof(Params, handle);
    out[k++] = offsetof(Params, d_particles);
    out[k++] = offsetof(Params, mesh_handle);
    out[k++] = offsetof(Params, mode_fisheye);
    out[k++] = offsetof(Params, d_meshes);
    out[k++] = offsetof(Params, type);
    out[k++] = offsetof(Params, traceState);
    out[k++] = sizeof(GaussianParticle);
    out[k++] = offsetof(GaussianParticle, position);
    out[k++] = offsetof(GaussianParticle, scale);
    out[k++] = offsetof(GaussianParticle, rotation);
    out[k++] = offsetof(GaussianParticle, opacity);
    out[k++] = offsetof(GaussianParticle, sh);
    out[k++] = sizeof(Mesh);
    out[k++] = offsetof(Mesh, faces);
    out[k++] = offsetof(Mesh, vertex_normals);
    out[k++] = (uint64_t)MIRROR | ((uint64_t)NORMAL << 8) | ((uint64_t)GLASS << 16);
    out[k++] = MaxNumHitPerTrace;
    out[k++] = ((uint64_t)TIMEOUT_ITERATIONS << 32) | MAX_BOUNCES;
}

// constants of tracer.cuh:9-14 and the SH table of Parameters.h:10-23
void ref_constants(float out[24]) {
    int k = 0;
    out[k++] = TRACE_MESH_TMIN; out[k++] = TRACE_MESH_TMAX; out[k++] = REFRACTION_EPS_SHIFT;
    out[k++] = HitPayload::InfiniteDistance;
    out[k++] = SH_C0; out[k++] = SH_C1;
    out[k++] = SH_C2_0; out[k++] = SH_C2_1; out[k++] = SH_C2_2; out[k++] = SH_C2_3; out[k++] = SH_C2_4;
    out[k++] = SH_C3_0; out[k++] = SH_C3_1; out[k++] = SH_C3_2; out[k++] = SH_C3_3; out[k++] = SH_C3_4; out[k++] = SH_C3_5; out[k++] = SH_C3_6;
}

// initializeParams defaults (src/GaussianTracer.cpp:477-486); alpha_min is the constructor's 0.01f (:45), passed in
void ref_param_defaults(float alpha_min, float out_f[4], int32_t out_i[3]) {
    Params saved = params;
    memset(&params, 0xFF, sizeof(params));
#include "param_defaults.inc"
    out_f[0] = params.t_min; out_f[1] = params.t_max; out_f[2] = params.minTransmittance; out_f[3] = params.alpha_min;
    out_i[0] = (int32_t)params.sh_degree_max; out_i[1] = params.mode_fisheye ? 1 : 0; out_i[2] = params.type;
    params = saved;
}

void ref_quantize(uint64_t n, const float* x, uint8_t* out) {
    for (uint64_t i = 0; i < n; ++i) out[i] = quantizeUnsigned8Bits(x[i]);
}

// Rays for pixels (xs[i], ys[i]); U, V, W are passed exactly as given (raygen passes -U, -V, W: shaders/tracer.cu:36-44,
// the caller does that).  valid[i] = 0 where getFishEyeRay returns without writing a ray (r > 1).
void ref_get_rays(int fisheye, uint64_t n, const uint32_t* xs, const uint32_t* ys, const float U[3], const float V[3], const float W[3],
                  const float eye[3], int width, int height, float* o_out, float* d_out, uint8_t* valid) {
    const float3 u = make_float3(U[0], U[1], U[2]), v = make_float3(V[0], V[1], V[2]), w = make_float3(W[0], W[1], W[2]);
    const float3 e = make_float3(eye[0], eye[1], eye[2]);
    for (uint64_t i = 0; i < n; ++i) {
        const float nan = std::nanf("");
        float3 o = make_float3(nan, nan, nan), d = make_float3(nan, nan, nan);
        const uint3 idx = make_uint3(xs[i], ys[i], 0);
        if (fisheye) getFishEyeRay(idx, u, v, w, e, width, height, o, d);
        else getRay(idx, u, v, w, e, width, height, o, d);
        valid[i] = !(o.x != o.x);
        o_out[3 * i] = o.x; o_out[3 * i + 1] = o.y; o_out[3 * i + 2] = o.z;
        d_out[3 * i] = d.x; d_out[3 * i + 1] = d.y; d_out[3 * i + 2] = d.z;
    }
}

// particles: n_p 236-byte GaussianParticle records; pair i uses particle pidx[i], origin o[i], direction d[i]
void ref_compute_response(const void* particles, uint64_t n, const uint32_t* pidx, const float* o, const float* d, float* out) {
    GaussianParticle* gp = (GaussianParticle*)particles;
    for (uint64_t i = 0; i < n; ++i) {
        const float3 oo = make_float3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), dd = make_float3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
        out[i] = computeResponse(gp[pidx[i]], oo, dd);
    }
}

// computeRadiance(gp, normalize(d)) as trace() calls it (tracer.cuh:262 is inside computeRadiance's caller; the caller passes the
// direction it wants -- normalise=1 reproduces `float3 dir = normalize(ray_d)`)
void ref_sh_radiance(const void* particles, uint64_t n, const uint32_t* pidx, const float* d, uint32_t deg, int normalise, float* out) {
    GaussianParticle* gp = (GaussianParticle*)particles;
    const unsigned saved = params.sh_degree_max;
    params.sh_degree_max = deg;
    for (uint64_t i = 0; i < n; ++i) {
        float3 dd = make_float3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
        if (normalise) dd = normalize(dd);
        const float3 L = computeRadiance(gp[pidx[i]], dd);
        out[3 * i] = L.x; out[3 * i + 1] = L.y; out[3 * i + 2] = L.z;
    }
    params.sh_degree_max = saved;
}

void ref_refract(uint64_t n, const float* d, const float* nrm, float eta, float* out_dir, float* t_hit_io, uint32_t* bounces_io) {
    for (uint64_t i = 0; i < n; ++i) {
        float3 nd = make_float3(0.f, 0.f, 0.f);
        unsigned int nb = bounces_io[i];
        refract(nd, make_float3(d[3 * i], d[3 * i + 1], d[3 * i + 2]), make_float3(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]), eta, t_hit_io[i], nb);
        bounces_io[i] = nb;
        out_dir[3 * i] = nd.x; out_dir[3 * i + 1] = nd.y; out_dir[3 * i + 2] = nd.z;
    }
}

void ref_mirror(uint64_t n, const float* d, const float* nrm, float* out_dir, uint32_t* bounces_io) {
    for (uint64_t i = 0; i < n; ++i) {
        float3 nd = make_float3(0.f, 0.f, 0.f);
        unsigned int nb = bounces_io[i];
        renderMirror(make_float3(d[3 * i], d[3 * i + 1], d[3 * i + 2]), make_float3(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]), nd, nb);
        bounces_io[i] = nb;
        out_dir[3 * i] = nd.x; out_dir[3 * i + 1] = nd.y; out_dir[3 * i + 2] = nd.z;
    }
}

// The 3x4 row-major instance transform OptiX receives for particle i (createGaussianParticlesBVH + createIAS)
void ref_proxy_transforms(const void* particles, uint64_t n, float alpha_min, float* out12) {
    const GaussianParticle* gps = (const GaussianParticle*)particles;
    for (uint64_t i = 0; i < n; ++i) {
        const GaussianParticle& gp = gps[i];
        float opacity = gp.opacity;
#include "proxy_transform.inc"
#include "ias_pack.inc"
        memcpy(out12 + 12 * i, instance_transform, sizeof(float) * 12);
    }
}

void ref_icosahedron(float verts[36], uint32_t idx[60]) {
    Icosahedron ico;
    const std::vector<float3> v = ico.getVertices();
    const std::vector<unsigned int> ix = ico.getIndices();
    for (int i = 0; i < 12; ++i) { verts[3 * i] = v[i].x; verts[3 * i + 1] = v[i].y; verts[3 * i + 2] = v[i].z; }
    for (int i = 0; i < 60; ++i) idx[i] = ix[i];
}

// kind 0 = createPlane, 1 = createSphere.  Call with verts == NULL to get the counts only.
void ref_primitive(int kind, const float position[3], float* verts, float* normals, uint32_t* indices, uint64_t counts[2], float transform_colmajor[16]) {
    Primitives prims;
    const float3 pos = make_float3(position[0], position[1], position[2]);
    Primitive p = kind == 0 ? prims.createPlane(pos) : prims.createSphere(pos);
    counts[0] = p.vertices.size();
    counts[1] = p.indices.size();
    memcpy(transform_colmajor, &p.transform[0][0], sizeof(float) * 16);
    if (!verts) return;
    for (size_t i = 0; i < p.vertices.size(); ++i) {
        verts[3 * i] = p.vertices[i].x; verts[3 * i + 1] = p.vertices[i].y; verts[3 * i + 2] = p.vertices[i].z;
        normals[3 * i] = p.normals[i].x; normals[3 * i + 1] = p.normals[i].y; normals[3 * i + 2] = p.normals[i].z;
    }
    for (size_t i = 0; i < p.indices.size(); ++i) indices[i] = p.indices[i];
}

// vertex_normals[i] = mat3(p.transform) * normal (sendGeometryAttributesToDevice)
void ref_normal_transform(uint64_t n, const float* normals_in, const float transform_colmajor[16], float* out) {
    struct { std::vector<float3> normals; glm::mat4 transform; } p;
    memcpy(&p.transform[0][0], transform_colmajor, sizeof(float) * 16);
    p.normals.resize(n);
    for (uint64_t i = 0; i < n; ++i) p.normals[i] = make_float3(normals_in[3 * i], normals_in[3 * i + 1], normals_in[3 * i + 2]);
    std::vector<float3> vn(n);
    float3* vertex_normals = vn.data();
    for (uint64_t i = 0; i < n; ++i) {
#include "normal_transform.inc"
    }
    for (uint64_t i = 0; i < n; ++i) { out[3 * i] = vn[i].x; out[3 * i + 1] = vn[i].y; out[3 * i + 2] = vn[i].z; }
}

// world-space vertex as OptiX sees it: the instance's 3x4 (createIAS packing of p.transform) applied to the object vertex.
// (OptiX's own arithmetic is closed; this is the plain row-by-row evaluation of that matrix.)
void ref_pack_transform(const float transform_colmajor[16], float out12[12]) {
    glm::mat4 transform;
    memcpy(&transform[0][0], transform_colmajor, sizeof(float) * 16);
#include "ias_pack.inc"
    memcpy(out12, instance_transform, sizeof(float) * 12);
}

// primitive placement 0.75 eye + 0.25 lookat (createPlane / createSphere)
void ref_placement(const float eye[3], const float lookat[3], float out[3]) {
    struct { float3 eye; } params;  // shadows the global on purpose: the fragment reads params.eye
    params.eye = make_float3(eye[0], eye[1], eye[2]);
    const float3 current_lookat = make_float3(lookat[0], lookat[1], lookat[2]);
    float3 cameraPosition = params.eye;
#include "placement.inc"
    out[0] = primitive_position.x; out[1] = primitive_position.y; out[2] = primitive_position.z;
}

// glm transforms the GUI applies to a primitive (config 4): translate * rotateY(deg) * scale, on top of nothing
void ref_trs(const float t[3], float rot_y_deg, float s, float out_colmajor[16]) {
    glm::mat4 m = glm::translate(glm::mat4(1.0f), glm::vec3(t[0], t[1], t[2]));
    m = glm::rotate(m, glm::radians(rot_y_deg), glm::vec3(0.0f, 1.0f, 0.0f));
    m = glm::scale(m, glm::vec3(s, s, s));
    memcpy(out_colmajor, &m[0][0], sizeof(float) * 16);
}
}
