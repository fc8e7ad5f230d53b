// Primitives.h -- reflection primitives (plane, UV sphere), same surface as src/geometry/Primitives.h:14-65.
// glm::mat4 is replaced by Mat4: 16 floats in glm's memory order (column-major, m[col*4+row]).
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include <vector_types.h>

struct Mat4 {
    float m[16];
    static Mat4 identity();
    static Mat4 translate(float3 p);
    static Mat4 rotateY(float degrees);
    static Mat4 scale(float s);
    Mat4 operator*(const Mat4& b) const;  // glm semantics: (A*B) v = A (B v)
};

struct Primitive {
    size_t index;
    std::string type;
    size_t instanceIndex;
    std::vector<float3> vertices;
    std::vector<unsigned int> indices;
    std::vector<float3> normals;
    size_t vertex_count;
    Mat4 transform;
    uint64_t gas;  // unused: the mesh accel is built for all primitives at once by libgrt
};

class Primitives {
public:
    Primitive createPlane(float3 position);   // 0.3 x 0.5 quad, normal +z           (Primitives.cpp:6-61)
    Primitive createSphere(float3 position);  // r = 0.3, 180 x 90 tessellation      (Primitives.cpp:63-140)
    Primitive createLoadMesh(std::string filename, float3 position);  // OBJ loading is out of scope: throws
    std::vector<Primitive>& getPrimitives() { return m_primitives; }
    void clearPrimitives() { m_primitives.clear(); numberOfMesh = numberOfPlane = numberOfSphere = numberOfLoaded = 0; }
    size_t& getMeshCount() { return numberOfMesh; }
    size_t& getPlaneCount() { return numberOfPlane; }
    size_t& getSphereCount() { return numberOfSphere; }
    size_t& getLoadedCount() { return numberOfLoaded; }

private:
    Mat4 getInitialTransform(float3 position);
    std::vector<Primitive> m_primitives;
    size_t numberOfMesh = 0, numberOfPlane = 0, numberOfSphere = 0, numberOfLoaded = 0;
};
