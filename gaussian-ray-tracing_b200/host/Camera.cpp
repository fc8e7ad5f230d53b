#include "Camera.h"

#include <cmath>

namespace {
inline float3 f3(float x, float y, float z) { float3 r; r.x = x; r.y = y; r.z = z; return r; }
inline float3 cross3(const float3& a, const float3& b) { return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline float len3(const float3& v) { return sqrtf(v.x * v.x + v.y * v.y + v.z * v.z); }
inline float3 scaled(const float3& v, float s) { return f3(v.x * s, v.y * s, v.z * s); }
inline float3 unit(const float3& v) { return scaled(v, 1.0f / len3(v)); }  // vector_math.h:590-594 semantics
}  // namespace

Camera::Camera() : m_eye(f3(0, 0, 0)), m_lookat(f3(0, 0, 0)), m_up(f3(0, 0, 0)), m_fovY(0.0f), m_aspectRatio(1.0f) {}

void Camera::UVWFrame(float3& U, float3& V, float3& W) const {
    W = f3(m_lookat.x - m_eye.x, m_lookat.y - m_eye.y, m_lookat.z - m_eye.z);
    const float wlen = len3(W);
    U = unit(cross3(W, m_up));
    V = unit(cross3(U, W));
    const float vlen = wlen * tanf(0.5f * m_fovY * 3.14159265358979323846f / 180.0f);
    V = scaled(V, vlen);
    U = scaled(U, vlen * m_aspectRatio);
}
