/*
 * ref_shim.cpp -- C entry points around the REFERENCE's own host code, compiled where it lies under
 * /root/reference (never copied): src/GaussianData.cpp (PLY parse + activation) and src/Camera.cpp
 * (U,V,W frame).  Output: oracle/_ref/libref_host.so (git-ignored; travels to the GPU box).
 * TEST INFRASTRUCTURE ONLY: used to pin our own PLY loader / camera against the reference's code.
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "GaussianData.h"
#include "Camera.h"

static_assert(sizeof(GaussianParticle) == 236, "reference particle record must be 236 bytes");

extern "C" {

// returns a malloc'd array of 236-byte records (caller frees with ref_free)
void* ref_load_ply(const char* path, unsigned long long* n_out, float center_out[3]) {
    try {
        GaussianData gd{std::string(path)};
        *n_out = gd.particles.size();
        void* buf = malloc(gd.particles.size() * sizeof(GaussianParticle) + 1);
        memcpy(buf, gd.particles.data(), gd.particles.size() * sizeof(GaussianParticle));
        float3 c = gd.getCenter();
        center_out[0] = c.x; center_out[1] = c.y; center_out[2] = c.z;
        return buf;
    } catch (const std::exception& e) {
        fprintf(stderr, "ref_load_ply: %s\n", e.what());
        *n_out = 0;
        return nullptr;
    } catch (...) {
        *n_out = 0;
        return nullptr;
    }
}
void ref_free(void* p) { free(p); }

void ref_camera_uvw(const float eye[3], const float lookat[3], const float up[3], float fovY, float aspect,
                    float U[3], float V[3], float W[3]) {
    Camera cam;
    cam.setEye(make_float3(eye[0], eye[1], eye[2]));
    cam.setLookat(make_float3(lookat[0], lookat[1], lookat[2]));
    cam.setUp(make_float3(up[0], up[1], up[2]));
    cam.setFovY(fovY);
    cam.setAspectRatio(aspect);
    float3 u, v, w;
    cam.UVWFrame(u, v, w);
    U[0] = u.x; U[1] = u.y; U[2] = u.z;
    V[0] = v.x; V[1] = v.y; V[2] = v.z;
    W[0] = w.x; W[1] = w.y; W[2] = w.z;
}
}
