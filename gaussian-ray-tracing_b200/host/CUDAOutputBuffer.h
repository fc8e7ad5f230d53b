// CUDAOutputBuffer.h -- headless stand-in for the reference's GL-interop buffer (src/CUDAOutputBuffer.h:12-40).
// Same map/unmap/width/height/resize/setStream surface; the pixels live in plain device memory owned through
// libgrt (no OpenGL on the box), getPBO() returns 0, and download() copies the frame to the host.
#pragma once
#include <cstdint>
#include <vector>
#include <vector_types.h>

#include "grt.h"

class CUDAOutputBuffer {
public:
    CUDAOutputBuffer(grtContext* ctx, int32_t width, int32_t height);
    ~CUDAOutputBuffer();
    void setStream(void* stream) { m_stream = stream; }
    void resize(int32_t width, int32_t height);
    uchar3* map();
    void unmap();
    int32_t width() const { return m_width; }
    int32_t height() const { return m_height; }
    unsigned int getPBO() { return 0u; }
    const std::vector<uchar3>& download();            // synchronous device -> host copy
    void writePPM(const char* path);                  // binary P6, row 0 first (launch y = 0)
    void writePNG(const char* path);                  // 8-bit RGB PNG (stored deflate blocks, no external library)

private:
    grtContext* m_ctx;
    void* m_stream = nullptr;
    uchar3* m_device_pixels = nullptr;
    std::vector<uchar3> m_host_pixels;
    int32_t m_width = 0, m_height = 0;
};
