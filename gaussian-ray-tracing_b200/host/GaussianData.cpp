#include "GaussianData.h"

#include <cmath>
#include <cstdint>
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>

namespace {
struct Prop {
    std::string name;
    int type;  // 0 char 1 uchar 2 short 3 ushort 4 int 5 uint 6 float 7 double
    size_t offset;
};
const char* kTypeNames[][2] = {{"char", "int8"}, {"uchar", "uint8"}, {"short", "int16"}, {"ushort", "uint16"},
                               {"int", "int32"}, {"uint", "uint32"}, {"float", "float32"}, {"double", "float64"}};
const size_t kTypeSize[] = {1, 1, 2, 2, 4, 4, 4, 8};

int typeOf(const std::string& t) {
    for (int i = 0; i < 8; ++i)
        if (t == kTypeNames[i][0] || t == kTypeNames[i][1]) return i;
    throw std::runtime_error("PLY: unknown property type '" + t + "'");
}
template <class T> T rd(const unsigned char* p, bool swap) {
    unsigned char b[sizeof(T)];
    for (size_t i = 0; i < sizeof(T); ++i) b[i] = p[swap ? sizeof(T) - 1 - i : i];
    T v;
    memcpy(&v, b, sizeof(T));
    return v;
}
float asFloat(const unsigned char* p, int type, bool swap) {
    switch (type) {
        case 0: return (float)rd<int8_t>(p, swap);
        case 1: return (float)rd<uint8_t>(p, swap);
        case 2: return (float)rd<int16_t>(p, swap);
        case 3: return (float)rd<uint16_t>(p, swap);
        case 4: return (float)rd<int32_t>(p, swap);
        case 5: return (float)rd<uint32_t>(p, swap);
        case 6: return rd<float>(p, swap);
        default: return (float)rd<double>(p, swap);
    }
}
}  // namespace

GaussianData::GaussianData(const std::string& filename) : m_filename(filename) {
    printf("Loading PLY file: %s\n", filename.c_str());
    loadPly();
}
GaussianData::GaussianData(const std::string& filename, bool deferred) : m_filename(filename), m_deferred(deferred) {
    printf("Loading PLY file: %s\n", filename.c_str());
    if (!deferred) loadPly();
}
GaussianData::~GaussianData() {}

namespace {
struct File {  // C stdio on purpose: this library is also dlopen()ed into Python processes
    FILE* f;
    explicit File(const char* p) : f(fopen(p, "rb")) {}
    ~File() { if (f) fclose(f); }
    bool line(std::string& out) {
        out.clear();
        int c;
        while ((c = fgetc(f)) != EOF && c != '\n') out.push_back((char)c);
        if (!out.empty() && out.back() == '\r') out.pop_back();
        return c != EOF || !out.empty();
    }
};
std::vector<std::string> words(const std::string& s) {
    std::vector<std::string> w;
    size_t i = 0;
    while (i < s.size()) {
        while (i < s.size() && isspace((unsigned char)s[i])) ++i;
        size_t j = i;
        while (j < s.size() && !isspace((unsigned char)s[j])) ++j;
        if (j > i) w.push_back(s.substr(i, j - i));
        i = j;
    }
    return w;
}
}  // namespace

void GaussianData::loadPly() {
    File file(m_filename.c_str());
    FILE* f = file.f;
    if (!f) throw std::runtime_error("PLY: could not open file " + m_filename);
    std::string line;
    file.line(line);
    if (line.substr(0, 3) != "ply") throw std::runtime_error("PLY: bad magic in " + m_filename);
    enum { ASCII, LE, BE } fmt = LE;
    size_t count = 0;
    bool inVertex = false, seenVertex = false, ended = false;
    std::vector<Prop> props;
    size_t stride = 0;
    // elements that precede "vertex" would have to be skipped; 3DGS / 3DGRUT files have vertex first
    while (file.line(line)) {
        const std::vector<std::string> w = words(line);
        if (w.empty()) continue;
        if (w[0] == "format" && w.size() >= 2) {
            if (w[1] == "ascii") fmt = ASCII;
            else if (w[1] == "binary_little_endian") fmt = LE;
            else if (w[1] == "binary_big_endian") fmt = BE;
            else throw std::runtime_error("PLY: unknown format " + w[1]);
        } else if (w[0] == "element" && w.size() >= 3) {
            const size_t n = (size_t)strtoull(w[2].c_str(), nullptr, 10);
            if (w[1] == "vertex") {
                if (seenVertex) throw std::runtime_error("PLY: duplicate vertex element");
                inVertex = seenVertex = true;
                count = n;
            } else {
                if (!seenVertex && n != 0) throw std::runtime_error("PLY: element '" + w[1] + "' before vertex is not supported");
                inVertex = false;
            }
        } else if (w[0] == "property" && inVertex && w.size() >= 3) {
            if (w[1] == "list") throw std::runtime_error("PLY: list property in the vertex element");
            Prop p{w[2], typeOf(w[1]), stride};
            stride += kTypeSize[p.type];
            props.push_back(p);
        } else if (w[0] == "end_header") {
            ended = true;
            break;
        }
    }
    if (!ended) throw std::runtime_error("PLY: header without end_header");
    if (!seenVertex) throw std::runtime_error("PLY: no vertex element");
    // the 59 properties GaussianData::parse reads (src/GaussianData.cpp:29-92); anything else (nx,ny,nz) is ignored
    std::vector<std::string> need = {"x", "y", "z", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3", "opacity",
                                     "f_dc_0", "f_dc_1", "f_dc_2"};
    for (int k = 0; k < 45; ++k) need.push_back("f_rest_" + std::to_string(k));
    std::vector<int> col(need.size(), -1);
    for (size_t i = 0; i < need.size(); ++i) {
        for (size_t j = 0; j < props.size(); ++j)
            if (props[j].name == need[i]) col[i] = (int)j;
        if (col[i] < 0) throw std::runtime_error("PLY parser: element vertex does not have property " + need[i]);  // happly.h:857
    }
    printf("Number of vertices: %zu\n", count);
    particles.resize(count);
    std::vector<unsigned char> row(stride);
    std::vector<float> v(need.size());
    const bool swap = (fmt == BE);
    for (size_t i = 0; i < count; ++i) {
        if (fmt == ASCII) {
            std::vector<double> all(props.size());
            for (size_t j = 0; j < props.size(); ++j)
                if (fscanf(f, "%lf", &all[j]) != 1) throw std::runtime_error("PLY: truncated ascii vertex data");
            for (size_t k = 0; k < need.size(); ++k) v[k] = (float)all[col[k]];
        } else {
            if (fread(row.data(), 1, stride, f) != stride) throw std::runtime_error("PLY: truncated binary vertex data");
            for (size_t k = 0; k < need.size(); ++k) v[k] = asFloat(row.data() + props[col[k]].offset, props[col[k]].type, swap);
        }
        GaussianParticle& p = particles[i];
        p.position[0] = v[0]; p.position[1] = v[1]; p.position[2] = v[2];
        p.scale[0] = expf(v[3]); p.scale[1] = expf(v[4]); p.scale[2] = expf(v[5]);
        const float norm = sqrtf(v[6] * v[6] + v[7] * v[7] + v[8] * v[8] + v[9] * v[9]);
        p.rotation[3] = v[6] / norm;  // w
        p.rotation[0] = v[7] / norm;
        p.rotation[1] = v[8] / norm;
        p.rotation[2] = v[9] / norm;
        p.opacity = 1.0f / (1.0f + expf(-v[10]));
        p.sh[0].x = v[11]; p.sh[0].y = v[12]; p.sh[0].z = v[13];
        const float* fr = &v[14];
        for (int k = 1; k < 16; ++k) { p.sh[k].x = fr[k - 1]; p.sh[k].y = fr[k + 14]; p.sh[k].z = fr[k + 29]; }
    }
}

float3 GaussianData::getCenter() {  // src/GaussianData.cpp:140-152: sequential float sums
    if (m_deferred) return m_center;  // computed the same way by grtLoadPly
    float3 c; c.x = c.y = c.z = 0.0f;
    for (auto& p : particles) { c.x += p.position[0]; c.y += p.position[1]; c.z += p.position[2]; }
    const float n = static_cast<float>(particles.size());
    c.x /= n; c.y /= n; c.z /= n;
    return c;
}
