// Camera.h -- look-at camera, same surface as the reference's src/Camera.h:9-64 (the glm view/projection helpers
// only feed the GUI gizmo and are left out).
#pragma once
#include <vector_types.h>

class Camera {
public:
    Camera();
    const float3& eye() const { return m_eye; }
    const float3& lookat() const { return m_lookat; }
    void setEye(const float3& val) { m_eye = val; }
    void setLookat(const float3& val) { m_lookat = val; }
    void setUp(const float3& val) { m_up = val; }
    void setFovY(float val) { m_fovY = val; }
    void setAspectRatio(float val) { m_aspectRatio = val; }
    void setMoveSpeed(const float& val) { m_moveSpeed = val; }
    // length(W) = focal length, |V| = |W| tan(fovY/2), |U| = |V| aspect  (src/Camera.cpp:3-13)
    void UVWFrame(float3& U, float3& V, float3& W) const;

private:
    float3 m_eye, m_lookat, m_up;
    float m_fovY;
    float m_aspectRatio;
    float m_moveSpeed = 1.0f;
};
