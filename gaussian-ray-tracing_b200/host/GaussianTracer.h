// GaussianTracer.h -- host facade with the reference's surface (src/GaussianTracer.h:20-111).  Same method names,
// same public fields (params, stream, primitives), same call order
//     ctor -> setSize -> initializeOptix -> { updateCamera -> render }*
// The body calls libgrt (include/grt.h) instead of OptiX: initializeOptix() builds the device LBVH.
#pragma once
#include <string>
#include <vector>

#include "CUDAOutputBuffer.h"
#include "Camera.h"
#include "GaussianData.h"
#include "Parameters.h"
#include "Primitives.h"
#include "grt.h"

enum ReflectionPrimitiveType { PLANE = 0, SPHERE, CUSTOM };

class GaussianTracer {
public:
    // device_ingest (addition): the PLY is activated on the GPU (grtLoadPly) instead of by GaussianData on the host
    GaussianTracer(const std::string& filename, int device = 0, bool device_ingest = false);
    ~GaussianTracer();

    void initializeOptix();
    void render(CUDAOutputBuffer& output_buffer);
    void updateCamera(Camera& camera, bool& camera_changed);
    void updateInstanceTransforms(Primitive& p);
    std::vector<Primitive>& getPrimitives() { return primitives->getPrimitives(); }
    void removePrimitive();
    void setRenderType(unsigned int renderType);
    void setSize(unsigned int width, unsigned int height);
    float3 getGaussianCenter() { return m_gsData.getCenter(); }

    Params params;
    void* stream;  // CUstream in the reference

    Primitives* primitives = new Primitives();
    void createPlane();
    void createSphere();
    void createLoadMesh(std::string filename);

    // additions (not in the reference): access for headless callers and multi-GPU tiling
    grtContext* context() { return m_ctx; }
    void renderTile(CUDAOutputBuffer& output_buffer, unsigned x0, unsigned y0, unsigned w, unsigned h);
    float lastKernelMs();
    const GaussianData& data() const { return m_gsData; }

private:
    float3 primitivePosition() const;  // 0.75 eye + 0.25 lookat (GaussianTracer.cpp:580-588)
    void rebuildMeshes();              // createGAS/createIAS/buildAccelationStructure + sendGeometryAttributesToDevice

    grtContext* m_ctx = nullptr;
    GaussianData m_gsData;
    size_t particle_count;
    float alpha_min;
    float3 current_lookat;
};
