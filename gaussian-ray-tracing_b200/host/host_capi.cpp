// host_capi.cpp -- plain-C entry points over the C++ host mirror so that tests (ctypes) can exercise it without a
// C++ harness: PLY load + activation, camera frame, primitive tessellation, and a whole headless render.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "GaussianTracer.h"

extern "C" {

// -> malloc'd array of 236-byte records (free with grtHostFree); NULL + message on stderr on failure
void* grtHostLoadPly(const char* path, unsigned long long* n_out, float center_out[3]) {
    *n_out = 0;
    try {
        GaussianData gd{std::string(path)};
        const size_t bytes = gd.particles.size() * sizeof(GaussianParticle);
        void* buf = malloc(bytes + 1);
        memcpy(buf, gd.particles.data(), bytes);
        const float3 c = gd.getCenter();
        center_out[0] = c.x; center_out[1] = c.y; center_out[2] = c.z;
        *n_out = gd.particles.size();
        return buf;
    } catch (const std::exception& e) {
        fprintf(stderr, "grtHostLoadPly: %s\n", e.what());
        return nullptr;
    }
}
void grtHostFree(void* p) { free(p); }

void grtHostCameraUVW(const float eye[3], const float lookat[3], const float up[3], float fovY, float aspect, float U[3], float V[3], float W[3]) {
    Camera cam;
    float3 e, l, u;
    e.x = eye[0]; e.y = eye[1]; e.z = eye[2];
    l.x = lookat[0]; l.y = lookat[1]; l.z = lookat[2];
    u.x = up[0]; u.y = up[1]; u.z = up[2];
    cam.setEye(e); cam.setLookat(l); cam.setUp(u); cam.setFovY(fovY); cam.setAspectRatio(aspect);
    float3 a, b, c;
    cam.UVWFrame(a, b, c);
    U[0] = a.x; U[1] = a.y; U[2] = a.z; V[0] = b.x; V[1] = b.y; V[2] = b.z; W[0] = c.x; W[1] = c.y; W[2] = c.z;
}

// kind 0 plane, 1 sphere.  Call with NULL outputs to get the counts, then again with buffers.
int grtHostPrimitive(int kind, const float pos[3], unsigned* n_vertices, unsigned* n_indices, float* vertices, float* normals,
                     unsigned* indices, float transform16[16]) {
    Primitives prims;
    float3 p; p.x = pos[0]; p.y = pos[1]; p.z = pos[2];
    Primitive pr = kind == 0 ? prims.createPlane(p) : prims.createSphere(p);
    *n_vertices = (unsigned)pr.vertices.size();
    *n_indices = (unsigned)pr.indices.size();
    if (vertices) memcpy(vertices, pr.vertices.data(), pr.vertices.size() * sizeof(float3));
    if (normals) memcpy(normals, pr.normals.data(), pr.normals.size() * sizeof(float3));
    if (indices) memcpy(indices, pr.indices.data(), pr.indices.size() * sizeof(unsigned));
    if (transform16) memcpy(transform16, pr.transform.m, sizeof(pr.transform.m));
    return 0;
}

// The reference's start-up + one frame (src/main.cpp:62-100) through the C++ facade; rgb8_out = width*height*3 bytes.
// flags: bit0 fisheye, bit1 createPlane, bit2 createSphere, bit3 device-side PLY ingest.  Returns 0, or 1 with the message in err (if given).
int grtHostRenderPly(const char* ply, unsigned width, unsigned height, unsigned sh_degree, unsigned flags, int mesh_type,
                     unsigned char* rgb8_out, float* kernel_ms_out, char* err, unsigned err_len) {
    try {
        GaussianTracer tracer{std::string(ply), 0, (flags & 8u) != 0};
        tracer.setSize(width, height);
        tracer.initializeOptix();
        CUDAOutputBuffer out(tracer.context(), (int32_t)width, (int32_t)height);
        out.setStream(tracer.stream);
        Camera camera;
        float3 eye; eye.x = 0.f; eye.y = 0.f; eye.z = 3.f;
        float3 up; up.x = 0.f; up.y = 1.f; up.z = 0.f;
        camera.setEye(eye); camera.setLookat(tracer.getGaussianCenter()); camera.setUp(up); camera.setFovY(60.0f);
        bool changed = true;
        tracer.updateCamera(camera, changed);
        tracer.params.sh_degree_max = sh_degree;
        tracer.params.mode_fisheye = (flags & 1u) != 0;
        tracer.setRenderType((unsigned)mesh_type);
        if (flags & 2u) tracer.createPlane();
        if (flags & 4u) tracer.createSphere();
        tracer.render(out);
        const std::vector<uchar3>& px = out.download();
        memcpy(rgb8_out, px.data(), px.size() * sizeof(uchar3));
        if (kernel_ms_out) *kernel_ms_out = tracer.lastKernelMs();
        if (const char* sink = getenv("GRT_HOST_FRAME_SINK")) {  // test hook: also write the frame (".png" or PPM)
            const std::string sp(sink);
            if (sp.size() > 4 && sp.substr(sp.size() - 4) == ".png") out.writePNG(sink); else out.writePPM(sink);
        }
        return 0;
    } catch (const std::exception& e) {
        if (err && err_len) { strncpy(err, e.what(), err_len - 1); err[err_len - 1] = 0; }
        return 1;
    }
}
}
