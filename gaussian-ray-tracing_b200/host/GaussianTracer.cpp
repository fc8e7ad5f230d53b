#include "GaussianTracer.h"

#include <cstring>
#include <stdexcept>

#include "Exception.h"

GaussianTracer::GaussianTracer(const std::string& filename, int device, bool device_ingest) : m_gsData(filename, device_ingest) {
    stream = nullptr;
    memset(&params, 0, sizeof(params));
    alpha_min = 0.01f;  // GaussianTracer.cpp:45
    current_lookat.x = current_lookat.y = current_lookat.z = 0.0f;
    if (grtCreate(device, &m_ctx) != 0) throw std::runtime_error(std::string("grtCreate failed: ") + grtGetLastError(nullptr));
    if (device_ingest) {  // the constructor must leave count and centre valid (main.cpp:66 asks for the centre before init)
        uint64_t n = 0;
        float cen[3] = {0.f, 0.f, 0.f};
        if (grtLoadPly(m_ctx, filename.c_str(), alpha_min, &n, cen) != 0) {
            const std::string msg = grtGetLastError(m_ctx);
            grtDestroy(m_ctx);
            m_ctx = nullptr;
            throw std::runtime_error(msg);
        }
        float3 c; c.x = cen[0]; c.y = cen[1]; c.z = cen[2];
        m_gsData.setDeviceIngested((size_t)n, c);
    }
    particle_count = m_gsData.getVertexCount();
}

GaussianTracer::~GaussianTracer() {
    if (m_ctx) {
        if (stream) grtStreamDestroy(m_ctx, stream);
        grtDestroy(m_ctx);  // the context owns particles, BVHs, meshes
    }
    delete primitives;
}

void GaussianTracer::setSize(unsigned int width, unsigned int height) {
    params.width = width;
    params.height = height;
}

// Replaces createContext .. createSBT + createGaussianParticlesBVH + initializeParams (GaussianTracer.cpp:72-83).
void GaussianTracer::initializeOptix() {
    if (!m_gsData.deferred()) GRT_CHECK(m_ctx, grtSetParticles(m_ctx, m_gsData.particles.data(), particle_count, alpha_min));
    void* d_particles = nullptr;
    uint64_t handle = 0;
    GRT_CHECK(m_ctx, grtBuildParticleAccel(m_ctx, nullptr, &handle, &d_particles));
    // initializeParams (GaussianTracer.cpp:475-486)
    params.output_buffer = nullptr;
    params.handle = handle;
    params.t_min = 1e-3f;
    params.t_max = 1e5f;
    params.minTransmittance = 0.001f;
    params.alpha_min = alpha_min;
    params.sh_degree_max = 0;
    params.mesh_handle = 0;
    params.mode_fisheye = false;
    params.type = MIRROR;
    params.traceState = nullptr;  // per-pixel state lives in registers
    params.d_particles = d_particles;
    params.d_meshes = nullptr;
    GRT_CHECK(m_ctx, grtStreamCreate(m_ctx, &stream));
}

void GaussianTracer::render(CUDAOutputBuffer& output_buffer) { renderTile(output_buffer, 0, 0, params.width, params.height); }

void GaussianTracer::renderTile(CUDAOutputBuffer& output_buffer, unsigned x0, unsigned y0, unsigned w, unsigned h) {
    if (params.mode_fisheye && w != 0 && h != 0) {  // GaussianTracer.cpp:510-513 clears the frame; a tile clears only its own rows
        unsigned char* base = reinterpret_cast<unsigned char*>(output_buffer.map());
        if (x0 == 0 && w == params.width) {
            GRT_CHECK(m_ctx, grtMemset(m_ctx, base + (size_t)y0 * params.width * sizeof(uchar3), 0, (size_t)h * params.width * sizeof(uchar3), stream));
        } else {
            for (unsigned y = y0; y < y0 + h; ++y)
                GRT_CHECK(m_ctx, grtMemset(m_ctx, base + ((size_t)y * params.width + x0) * sizeof(uchar3), 0, (size_t)w * sizeof(uchar3), stream));
        }
        output_buffer.unmap();
    }
    params.output_buffer = output_buffer.map();
    // every render re-reads the host Params: callers mutate params.mode_fisheye / sh_degree_max directly (gui.cpp:190)
    GRT_CHECK(m_ctx, grtLaunchTile(m_ctx, stream, reinterpret_cast<const grtParams*>(&params), x0, y0, w, h));
    output_buffer.unmap();
    GRT_CHECK(m_ctx, grtStreamSynchronize(m_ctx, stream));  // CUDA_SYNC_CHECK (GaussianTracer.cpp:537)
}

float GaussianTracer::lastKernelMs() {
    float ms = 0.0f;
    GRT_CHECK(m_ctx, grtLastLaunchMs(m_ctx, &ms));
    return ms;
}

void GaussianTracer::updateCamera(Camera& camera, bool& camera_changed) {
    if (!camera_changed) return;
    camera_changed = false;
    camera.setAspectRatio(static_cast<float>(params.width) / static_cast<float>(params.height));
    params.eye = camera.eye();
    camera.UVWFrame(params.U, params.V, params.W);
    current_lookat = camera.lookat();
}

void GaussianTracer::setRenderType(unsigned int renderType) { params.type = (int)renderType; }

float3 GaussianTracer::primitivePosition() const {
    const float cw = 0.75f, gw = 1.0f - cw;
    float3 p;
    p.x = current_lookat.x * gw + params.eye.x * cw;
    p.y = current_lookat.y * gw + params.eye.y * cw;
    p.z = current_lookat.z * gw + params.eye.z * cw;
    return p;
}

void GaussianTracer::rebuildMeshes() {
    std::vector<Primitive>& ps = primitives->getPrimitives();
    std::vector<grtMeshDesc> descs(ps.size());
    for (size_t i = 0; i < ps.size(); ++i) {
        grtMeshDesc& d = descs[i];
        d.vertices = reinterpret_cast<const float*>(ps[i].vertices.data());
        d.normals = reinterpret_cast<const float*>(ps[i].normals.data());
        d.indices = ps[i].indices.data();
        d.vertex_count = (uint32_t)ps[i].vertices.size();
        d.index_count = (uint32_t)ps[i].indices.size();
        memcpy(d.transform, ps[i].transform.m, sizeof(d.transform));
    }
    GRT_CHECK(m_ctx, grtSetMeshes(m_ctx, descs.data(), (uint32_t)descs.size()));
    uint64_t mh = 0;
    void* dm = nullptr;
    GRT_CHECK(m_ctx, grtBuildMeshAccel(m_ctx, stream, &mh, &dm));
    params.mesh_handle = mh;
    params.d_meshes = reinterpret_cast<Mesh*>(dm);
}

void GaussianTracer::createPlane() { primitives->createPlane(primitivePosition()); rebuildMeshes(); }
void GaussianTracer::createSphere() { primitives->createSphere(primitivePosition()); rebuildMeshes(); }
void GaussianTracer::createLoadMesh(std::string filename) { primitives->createLoadMesh(filename, primitivePosition()); rebuildMeshes(); }

// gizmo move (GaussianTracer.cpp:711-794): the caller's Primitive carries the new transform
void GaussianTracer::updateInstanceTransforms(Primitive& p) {
    for (Primitive& q : primitives->getPrimitives())
        if (q.instanceIndex == p.instanceIndex) q.transform = p.transform;
    rebuildMeshes();
}

void GaussianTracer::removePrimitive() {  // GaussianTracer.cpp:553-566
    primitives->clearPrimitives();
    rebuildMeshes();  // zero meshes: handle 0, d_meshes NULL
}
