// GaussianData.h -- PLY -> activated particle records; same surface as the reference's src/GaussianData.h:12-42.
// The reader is our own (single pass over the vertex element straight into the 236-byte records; the reference
// copies 59 whole columns through happly vectors first), the arithmetic is the reference's (src/GaussianData.cpp:97-131).
#pragma once
#include <string>
#include <vector>
#include <vector_types.h>

struct GaussianParticle {  // 236 bytes, byte-compatible with src/GaussianData.h:12-20 (glm::vec3, glm::vec3, glm::quat x,y,z,w)
    float position[3];
    float scale[3];     // exp(scale_k)
    float rotation[4];  // normalised, memory order x,y,z,w  (rot_0 of the file is w)
    float opacity;      // sigmoid(opacity)
    float3 sh[16];      // sh[0] = f_dc, sh[k] = (f_rest[k-1], f_rest[k+14], f_rest[k+29])
};
static_assert(sizeof(GaussianParticle) == 236, "GaussianParticle must stay 236 bytes");

class GaussianData {
public:
    GaussianData(const std::string& filename);  // throws std::runtime_error on a missing file / property, like happly
    // addition: deferred = true reads nothing on the host; GaussianTracer then ingests the file on the device
    // (grtLoadPly: mmap + CUDA activation kernel) and reports count and centre back through setDeviceIngested().
    GaussianData(const std::string& filename, bool deferred);
    ~GaussianData();
    size_t getVertexCount() const { return m_deferred ? m_count : particles.size(); }
    float3 getCenter();
    std::vector<GaussianParticle> particles;  // empty in deferred mode (the records live on the device: grtGetParticles)
    bool deferred() const { return m_deferred; }
    const std::string& filename() const { return m_filename; }
    void setDeviceIngested(size_t count, float3 center) { m_count = count; m_center = center; }

private:
    void loadPly();
    std::string m_filename;
    bool m_deferred = false;
    size_t m_count = 0;
    float3 m_center{};
};
