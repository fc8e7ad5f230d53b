/* Stub <tiny_obj_loader.h> (the reference's third_party/tinyobjloader submodule is empty): only the names
 * Primitives::createLoadMesh mentions, so that src/geometry/Primitives.cpp compiles.  ParseFromFile always fails;
 * OBJ loading is out of scope.  TEST INFRASTRUCTURE ONLY. */
#pragma once
#include <string>
#include <vector>
namespace tinyobj {
typedef float real_t;
struct index_t { int vertex_index, normal_index, texcoord_index; };
struct mesh_t { std::vector<index_t> indices; std::vector<unsigned char> num_face_vertices; };
struct shape_t { std::string name; mesh_t mesh; };
struct attrib_t { std::vector<real_t> vertices, normals, texcoords; };
struct ObjReaderConfig {};
class ObjReader {
public:
    bool ParseFromFile(const std::string&, const ObjReaderConfig&) { err_ = "tinyobjloader is not vendored"; return false; }
    const std::string& Error() const { return err_; }
    const std::string& Warning() const { return warn_; }
    const attrib_t& GetAttrib() const { return attrib_; }
    const std::vector<shape_t>& GetShapes() const { return shapes_; }
private:
    std::string err_, warn_; attrib_t attrib_; std::vector<shape_t> shapes_;
};
}
