#include "CUDAOutputBuffer.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "Exception.h"

CUDAOutputBuffer::CUDAOutputBuffer(grtContext* ctx, int32_t width, int32_t height) : m_ctx(ctx) { resize(width, height); }

CUDAOutputBuffer::~CUDAOutputBuffer() {
    if (m_device_pixels) grtFree(m_ctx, m_device_pixels);
}

void CUDAOutputBuffer::resize(int32_t width, int32_t height) {
    if (width < 1 || height < 1) throw std::runtime_error("CUDAOutputBuffer: size must be at least 1x1");
    if (m_width == width && m_height == height) return;
    if (m_device_pixels) GRT_CHECK(m_ctx, grtFree(m_ctx, m_device_pixels));
    m_device_pixels = nullptr;
    m_width = width;
    m_height = height;
    void* p = nullptr;
    GRT_CHECK(m_ctx, grtMalloc(m_ctx, (size_t)width * height * sizeof(uchar3), &p));
    m_device_pixels = reinterpret_cast<uchar3*>(p);
    m_host_pixels.assign((size_t)width * height, uchar3{0, 0, 0});
}

uchar3* CUDAOutputBuffer::map() { return m_device_pixels; }
void CUDAOutputBuffer::unmap() {}

const std::vector<uchar3>& CUDAOutputBuffer::download() {
    GRT_CHECK(m_ctx, grtMemcpyDtoH(m_ctx, m_host_pixels.data(), m_device_pixels, m_host_pixels.size() * sizeof(uchar3), m_stream));
    GRT_CHECK(m_ctx, grtStreamSynchronize(m_ctx, m_stream));
    return m_host_pixels;
}

void CUDAOutputBuffer::writePPM(const char* path) {
    download();
    FILE* f = fopen(path, "wb");
    if (!f) throw std::runtime_error(std::string("cannot write ") + path);
    fprintf(f, "P6\n%d %d\n255\n", m_width, m_height);
    fwrite(m_host_pixels.data(), sizeof(uchar3), m_host_pixels.size(), f);
    fclose(f);
}

// Headless frame sink (replaces Display.cpp's GL blit): an RGB8 PNG without any compression library --
// zlib "stored" blocks, CRC-32 and Adler-32 computed here.  Row 0 first (launch y = 0), like writePPM.
namespace {
uint32_t crc32_update(uint32_t c, const unsigned char* p, size_t n) {
    static uint32_t table[256];
    static bool init = false;
    if (!init) {
        for (uint32_t i = 0; i < 256; ++i) {
            uint32_t v = i;
            for (int k = 0; k < 8; ++k) v = (v & 1u) ? 0xEDB88320u ^ (v >> 1) : v >> 1;
            table[i] = v;
        }
        init = true;
    }
    for (size_t i = 0; i < n; ++i) c = table[(c ^ p[i]) & 0xFFu] ^ (c >> 8);
    return c;
}
void put_be32(std::vector<unsigned char>& v, uint32_t x) {
    v.push_back((unsigned char)(x >> 24)); v.push_back((unsigned char)(x >> 16)); v.push_back((unsigned char)(x >> 8)); v.push_back((unsigned char)x);
}
void write_chunk(FILE* f, const char* type, const std::vector<unsigned char>& data) {
    std::vector<unsigned char> head;
    put_be32(head, (uint32_t)data.size());
    fwrite(head.data(), 1, 4, f);
    fwrite(type, 1, 4, f);
    if (!data.empty()) fwrite(data.data(), 1, data.size(), f);
    uint32_t c = crc32_update(0xFFFFFFFFu, reinterpret_cast<const unsigned char*>(type), 4);
    if (!data.empty()) c = crc32_update(c, data.data(), data.size());
    std::vector<unsigned char> tail;
    put_be32(tail, c ^ 0xFFFFFFFFu);
    fwrite(tail.data(), 1, 4, f);
}
}  // namespace

void CUDAOutputBuffer::writePNG(const char* path) {
    const std::vector<uchar3>& px = download();
    FILE* f = fopen(path, "wb");
    if (!f) throw std::runtime_error(std::string("cannot write ") + path);
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    fwrite(sig, 1, 8, f);
    std::vector<unsigned char> ihdr;
    put_be32(ihdr, (uint32_t)m_width); put_be32(ihdr, (uint32_t)m_height);
    ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);  // 8-bit RGB, no interlace
    write_chunk(f, "IHDR", ihdr);
    // raw scanlines: filter byte 0 + RGB row
    const size_t row = (size_t)m_width * 3 + 1;
    std::vector<unsigned char> raw(row * (size_t)m_height);
    for (int y = 0; y < m_height; ++y) {
        raw[y * row] = 0;
        memcpy(&raw[y * row + 1], &px[(size_t)y * m_width], (size_t)m_width * 3);
    }
    std::vector<unsigned char> z;
    z.push_back(0x78); z.push_back(0x01);  // zlib header, no compression
    uint32_t a = 1, b = 0;                  // Adler-32
    for (size_t off = 0; off < raw.size() || raw.empty();) {
        const size_t n = std::min<size_t>(65535, raw.size() - off);
        z.push_back(off + n >= raw.size() ? 1 : 0);  // BFINAL, BTYPE = 00 (stored)
        z.push_back((unsigned char)(n & 0xFF)); z.push_back((unsigned char)(n >> 8));
        z.push_back((unsigned char)(~n & 0xFF)); z.push_back((unsigned char)((~n >> 8) & 0xFF));
        z.insert(z.end(), raw.begin() + off, raw.begin() + off + n);
        for (size_t i = 0; i < n; ++i) { a = (a + raw[off + i]) % 65521u; b = (b + a) % 65521u; }
        off += n;
        if (raw.empty()) break;
    }
    put_be32(z, (b << 16) | a);
    write_chunk(f, "IDAT", z);
    write_chunk(f, "IEND", std::vector<unsigned char>());
    fclose(f);
}
