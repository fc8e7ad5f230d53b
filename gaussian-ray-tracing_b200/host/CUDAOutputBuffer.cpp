#include "CUDAOutputBuffer.h"

#include <cstdio>
#include <stdexcept>

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
