// ingest.cu -- PLY vertex rows -> activated 236-byte GaussianParticle records, on the device.
//
// Replaces the per-vertex loop of GaussianData::parse (src/GaussianData.cpp:97-131) and the proxy-size loop of
// createGaussianParticlesBVH (src/GaussianTracer.cpp:304-306).  The reference copies 59 property columns through
// std::vectors on the host (happly) and activates them one vertex at a time; here the raw vertex payload of the file
// (any float-property order, extra properties allowed) is uploaded as it lies in the file and one kernel produces
//   * the reference-layout records (pos, exp(scale), normalised quaternion x,y,z,w, sigmoid(opacity), 16 x float3 SH) and
//   * the proxy in-sphere radius s = sqrtf(2 ln(opacity / alpha_min)), 0 for particles that can never be hit.
// HBM-streaming: stride bytes in, 240 bytes out per vertex; rows are staged through shared memory so that both the
// global loads and the global stores are coalesced whatever the property order is.
//
// Numerics: exp / log are evaluated in double and rounded once, i.e. the correctly rounded float.  A host libm's expf is
// only faithfully rounded (glibc: <= 0.502 ulp), so ~0.05 % of scale/opacity values differ by one ulp from the host route
// (measured in tests/test_gpu_ingest.py) -- the reference's own MSVC and Linux builds differ in the same way.
// Everything else is the reference's fp32 operation order with un-fused IEEE operators (bit-exact).
#include "grt_internal.h"

namespace grt {

constexpr int INGEST_THREADS = 128;

__device__ __forceinline__ float load_f32(const unsigned char* p, bool aligned) {
    if (aligned) return *reinterpret_cast<const float*>(p);
    const uint32_t v = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
    return __uint_as_float(v);
}

__global__ void __launch_bounds__(INGEST_THREADS) particles_activate_kernel(const unsigned char* __restrict__ rows, const PlyLayout L,
                                                                            uint32_t n, float alpha_min, float* __restrict__ out59,
                                                                            float* __restrict__ d_s) {
    extern __shared__ __align__(16) unsigned char s_in[];                              // INGEST_THREADS x stride bytes
    float* s_out = reinterpret_cast<float*>(s_in + (((size_t)INGEST_THREADS * L.stride + 15) & ~(size_t)15));  // INGEST_THREADS x 59 floats
    const uint32_t v0 = blockIdx.x * INGEST_THREADS;
    const uint32_t nv = min((uint32_t)INGEST_THREADS, n - v0);
    const size_t in_bytes = (size_t)nv * L.stride;
    const unsigned char* src = rows + (size_t)v0 * L.stride;
    // coalesced copy in: 16-byte chunks when the tile start is 16-byte aligned, bytes otherwise
    if (((reinterpret_cast<uintptr_t>(src) | in_bytes) & 15u) == 0) {
        const uint4* s4 = reinterpret_cast<const uint4*>(src);
        uint4* d4 = reinterpret_cast<uint4*>(s_in);
        for (size_t k = threadIdx.x; k < in_bytes / 16; k += INGEST_THREADS) d4[k] = __ldg(s4 + k);
    } else {
        for (size_t k = threadIdx.x; k < in_bytes; k += INGEST_THREADS) s_in[k] = __ldg(src + k);
    }
    __syncthreads();
    const uint32_t t = threadIdx.x;
    if (t < nv) {
        const unsigned char* row = s_in + (size_t)t * L.stride;
        const bool al = L.aligned != 0;
        float p[59];
#pragma unroll
        for (int k = 0; k < 59; ++k) p[k] = load_f32(row + L.off[k], al);
        float* o = s_out + t * 59;
        o[0] = p[0]; o[1] = p[1]; o[2] = p[2];
        // GaussianData.cpp:101-103
        o[3] = (float)exp((double)p[3]); o[4] = (float)exp((double)p[4]); o[5] = (float)exp((double)p[5]);
        // GaussianData.cpp:104-111: quat(w = rot_0, x, y, z), glm memory order x,y,z,w
        const float nrm = __fsqrt_rn(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(p[6], p[6]), __fmul_rn(p[7], p[7])), __fmul_rn(p[8], p[8])),
                                               __fmul_rn(p[9], p[9])));
        o[6] = __fdiv_rn(p[7], nrm); o[7] = __fdiv_rn(p[8], nrm); o[8] = __fdiv_rn(p[9], nrm); o[9] = __fdiv_rn(p[6], nrm);
        // GaussianData.cpp:112
        const float opacity = __fdiv_rn(1.0f, __fadd_rn(1.0f, (float)exp(-(double)p[10])));
        o[10] = opacity;
        o[11] = p[11]; o[12] = p[12]; o[13] = p[13];
        // GaussianData.cpp:113-128: sh[k] = (f_rest[k-1], f_rest[k+14], f_rest[k+29])
#pragma unroll
        for (int k = 1; k < 16; ++k) {
            o[11 + 3 * k + 0] = p[14 + k - 1];
            o[11 + 3 * k + 1] = p[14 + k + 14];
            o[11 + 3 * k + 2] = p[14 + k + 29];
        }
        // GaussianTracer.cpp:304-306; opacity <= alpha_min -> the instance can never be hit
        float s = 0.0f;
        if (opacity > alpha_min) {
            s = __fsqrt_rn(__fmul_rn(2.0f, (float)log((double)__fdiv_rn(opacity, alpha_min))));
            if (!(s > 0.0f) || !isfinite(s)) s = 0.0f;
        }
        d_s[v0 + t] = s;
    }
    __syncthreads();
    // coalesced copy out (records are 236 bytes: float granularity)
    float* dst = out59 + (size_t)v0 * 59;
    for (uint32_t k = threadIdx.x; k < nv * 59u; k += INGEST_THREADS) dst[k] = s_out[k];
}

int particles_activate(const unsigned char* d_rows, const PlyLayout& L, uint32_t n, float alpha_min, float* d_out59, float* d_s,
                       cudaStream_t stream, unsigned long long* launches) {
    if (!n) return 0;
    const size_t smem = (((size_t)INGEST_THREADS * L.stride + 15) & ~(size_t)15) + (size_t)INGEST_THREADS * 59 * sizeof(float);
    if (smem > 200 * 1024) return (int)cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(particles_activate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    particles_activate_kernel<<<(n + INGEST_THREADS - 1) / INGEST_THREADS, INGEST_THREADS, smem, stream>>>(d_rows, L, n, alpha_min, d_out59, d_s);
    ++*launches;
    return (int)cudaGetLastError();
}

}  // namespace grt
