#include "Primitives.h"

#include <cmath>
#include <stdexcept>

namespace {
const float kPi = 3.14159265358979323846f;
inline float3 f3(float x, float y, float z) { float3 r; r.x = x; r.y = y; r.z = z; return r; }

// Two triangles per cell of a (rows x cols)-vertex grid: (ll, lr, ur) and (ur, ul, ll), the winding the reference
// uses for both the plane and the sphere (Primitives.cpp:38-51, 117-129).
void gridIndices(unsigned rows, unsigned cols, std::vector<unsigned int>& out) {
    out.reserve(out.size() + 6u * (rows - 1) * (cols - 1));
    for (unsigned r = 0; r + 1 < rows; ++r)
        for (unsigned c = 0; c + 1 < cols; ++c) {
            const unsigned ll = r * cols + c, lr = ll + 1, ul = ll + cols, ur = ul + 1;
            const unsigned tri[6] = {ll, lr, ur, ur, ul, ll};
            out.insert(out.end(), tri, tri + 6);
        }
}
}  // namespace

Mat4 Mat4::identity() {
    Mat4 r{};
    r.m[0] = r.m[5] = r.m[10] = r.m[15] = 1.0f;
    return r;
}
Mat4 Mat4::translate(float3 p) {
    Mat4 r = identity();
    r.m[12] = p.x; r.m[13] = p.y; r.m[14] = p.z;
    return r;
}
Mat4 Mat4::rotateY(float degrees) {
    const float a = degrees * kPi / 180.0f, c = cosf(a), s = sinf(a);
    Mat4 r = identity();
    r.m[0] = c; r.m[2] = -s;   // column 0 = (c, 0, -s)
    r.m[8] = s; r.m[10] = c;   // column 2 = (s, 0, c)
    return r;
}
Mat4 Mat4::scale(float s) {
    Mat4 r = identity();
    r.m[0] = r.m[5] = r.m[10] = s;
    return r;
}
Mat4 Mat4::operator*(const Mat4& b) const {
    Mat4 r{};
    for (int c = 0; c < 4; ++c)
        for (int row = 0; row < 4; ++row) {
            float acc = 0.0f;
            for (int k = 0; k < 4; ++k) acc += m[k * 4 + row] * b.m[c * 4 + k];
            r.m[c * 4 + row] = acc;
        }
    return r;
}

Mat4 Primitives::getInitialTransform(float3 position) { return Mat4::translate(position); }  // Primitives.cpp:204-216

Primitive Primitives::createPlane(float3 position) {
    Primitive p;
    const float width = 0.3f, height = 0.5f;
    const float3 corner = f3(-width * 0.5f, -height * 0.5f, 0.0f);
    for (unsigned j = 0; j < 2; ++j)
        for (unsigned i = 0; i < 2; ++i) {
            p.vertices.push_back(f3(corner.x + float(i) * width, corner.y + float(j) * height, corner.z + 0.0f));
            p.normals.push_back(f3(0.0f, 0.0f, 1.0f));
        }
    gridIndices(2, 2, p.indices);
    p.index = numberOfPlane++;
    p.type = "Plane";
    p.instanceIndex = numberOfMesh++;
    p.vertex_count = p.vertices.size();
    p.transform = getInitialTransform(position);
    p.gas = 0;
    m_primitives.push_back(p);
    return p;
}

Primitive Primitives::createSphere(float3 position) {
    Primitive p;
    const unsigned tessU = 180, tessV = 90;
    const float radius = 0.3f;
    const float phiStep = 2.0f * kPi / (float)tessU, thetaStep = kPi / (float)(tessV - 1);
    p.vertices.reserve((tessU + 1) * tessV);
    p.normals.reserve((tessU + 1) * tessV);
    for (unsigned lat = 0; lat < tessV; ++lat) {  // rings from +y downwards
        const float theta = (float)lat * thetaStep, st = sinf(theta), ct = cosf(theta);
        for (unsigned lon = 0; lon <= tessU; ++lon) {  // tessU + 1 vertices per ring (seam duplicated)
            const float phi = (float)lon * phiStep;
            const float3 n = f3(cosf(phi) * st, ct, sinf(phi) * st);
            p.normals.push_back(n);
            p.vertices.push_back(f3(n.x * radius, n.y * radius, n.z * radius));
        }
    }
    gridIndices(tessV, tessU + 1, p.indices);
    p.index = numberOfSphere++;
    p.type = "Sphere";
    p.instanceIndex = numberOfMesh++;
    p.vertex_count = p.vertices.size();
    p.transform = getInitialTransform(position);
    p.gas = 0;
    m_primitives.push_back(p);
    return p;
}

Primitive Primitives::createLoadMesh(std::string filename, float3) {
    // The reference loads OBJ files through tinyobjloader, whose submodule is empty in the reference tree; OBJ
    // meshes are in none of the benchmark configurations.
    throw std::runtime_error("createLoadMesh(" + filename + "): OBJ loading is out of scope of this build");
}
