// Parameters.h -- launch-parameter ABI of the hot path, mirror of the reference's src/Parameters.h:36-93.
// `Params` keeps the reference's field names and its exact 144-byte layout (it IS grtParams, include/grt.h);
// OptixTraversableHandle becomes an opaque 64-bit handle returned by libgrt.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector_types.h>

#include "grt.h"

struct Mesh {  // src/Parameters.h:36-40
    uint3* faces;
    float3* vertex_normals;
};

struct Params {  // src/Parameters.h:42-74
    uchar3* output_buffer;
    unsigned int width;
    unsigned int height;
    unsigned int sh_degree_max;
    float3 eye;
    float3 U;
    float3 V;
    float3 W;
    float t_min;
    float t_max;
    float minTransmittance;
    float alpha_min;
    uint64_t handle;
    void* d_particles;
    uint64_t mesh_handle;
    bool mode_fisheye;
    Mesh* d_meshes;
    int type;
    unsigned int* traceState;
};
static_assert(sizeof(Params) == 144 && sizeof(grtParams) == 144, "Params must stay 144 bytes");
static_assert(offsetof(Params, eye) == offsetof(grtParams, eye) && offsetof(Params, W) == offsetof(grtParams, W) &&
                  offsetof(Params, alpha_min) == offsetof(grtParams, alpha_min) && offsetof(Params, handle) == offsetof(grtParams, handle) &&
                  offsetof(Params, d_particles) == offsetof(grtParams, d_particles) &&
                  offsetof(Params, mesh_handle) == offsetof(grtParams, mesh_handle) &&
                  offsetof(Params, mode_fisheye) == offsetof(grtParams, mode_fisheye) && offsetof(Params, d_meshes) == offsetof(grtParams, d_meshes) &&
                  offsetof(Params, type) == offsetof(grtParams, type) && offsetof(Params, traceState) == offsetof(grtParams, traceState),
              "Params and grtParams must agree field by field");

enum MeshType { MIRROR = 0, NORMAL = 1, GLASS = 2 };                                                     // src/Parameters.h:80-85
enum TraceState { TraceGaussianPass = 0, TraceMeshPass, TraceLastGaussianPass, TraceTerminate };          // src/Parameters.h:87-93
