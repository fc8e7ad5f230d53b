// build.cu -- on-device acceleration-structure build for sm_100a.
//
// Replaces the reference's createGaussianParticlesBVH / createGAS / createIAS / optixAccelBuild
// (src/GaussianTracer.cpp:297-473): there the host loops over N particles building 80-byte OptixInstance
// records and OptiX builds a two-level BVH on RT-core hardware.  B200 has no RT cores, so this file builds a
// software BVH entirely on the device:
//   proxy boxes -> 63-bit Morton codes -> hand-written LSD radix sort (no CUB/Thrust) -> Karras 2012
//   binary hierarchy -> bottom-up atomic refit -> level-synchronous collapse into 8-wide nodes whose
//   leaf children are runs of <= LEAF_MAX Morton-consecutive records.
// All kernels are one-time (per scene) and HBM-bound streaming passes; none is on the per-frame path.
#include <algorithm>
#include <cfloat>
#include <cstdio>
#include <cstdlib>
#include <utility>
#include <vector>

#include <mutex>

#include "grt_internal.h"

namespace grt {

#define CK(x)                              \
    do {                                   \
        cudaError_t e_ = (x);              \
        if (e_ != cudaSuccess) return (int)e_; \
    } while (0)

static inline unsigned int cdiv(unsigned long long a, unsigned int b) { return (unsigned int)((a + b - 1) / b); }

// ------------------------------------------------------------------------------------------- radix sort
// LSD, 8 bits per pass, stable.  Per pass: (1) per-tile digit histograms + global digit totals, (2) one warp per digit turns
// its row of tile counts into global output offsets, (3) every tile re-reads its keys and scatters them.  A tile is
// RS_THREADS x RS_ITEMS keys, warp w of the block owning the w-th contiguous slice of it.
#ifndef GRT_RS_ITEMS
#define GRT_RS_ITEMS 8
#endif
constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = GRT_RS_ITEMS;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;
constexpr int RS_WARPS = RS_THREADS / 32;

// ghist[digit][tile] = count, gtot[digit] += count (gtot zeroed by the host before the pass)
__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(const uint64_t* __restrict__ keys, uint32_t n, int shift,
                                                             uint32_t* __restrict__ ghist, uint32_t* __restrict__ gtot, uint32_t nblocks) {
    __shared__ uint32_t h[256];
    const uint32_t tid = threadIdx.x;
    h[tid] = 0;
    __syncthreads();
    const uint64_t base = (uint64_t)blockIdx.x * RS_TILE;
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        uint64_t i = base + (uint64_t)r * RS_THREADS + tid;
        if (i < n) atomicAdd(&h[(uint32_t)(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    const uint32_t v = h[tid];
    ghist[(uint64_t)tid * nblocks + blockIdx.x] = v;
    if (v) atomicAdd(&gtot[tid], v);
}

// one warp per digit: exclusive scan of the digit's row of tile counts, offset by the total of all smaller digits
__global__ void __launch_bounds__(RS_THREADS) rs_scan_kernel(uint32_t* __restrict__ ghist, const uint32_t* __restrict__ gtot, uint32_t nblocks) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t d = blockIdx.x * RS_WARPS + (threadIdx.x >> 5);
    uint32_t* row = ghist + (uint64_t)d * nblocks;
    constexpr int PRE = 16;  // rows of up to 32 * PRE tiles are fetched with all loads in flight at once
    uint32_t v[PRE];
    if (nblocks <= 32u * PRE) {
#pragma unroll
        for (int c = 0; c < PRE; ++c) v[c] = (c * 32u + lane < nblocks) ? row[c * 32u + lane] : 0u;
    }
    uint32_t base = 0;
    for (uint32_t j = lane; j < d; j += 32u) base += gtot[j];
#pragma unroll
    for (int o = 16; o; o >>= 1) base += __shfl_xor_sync(0xFFFFFFFFu, base, o);
    if (nblocks <= 32u * PRE) {
#pragma unroll
        for (int c = 0; c < PRE; ++c) {
            if (c * 32u < nblocks) {
                uint32_t inc = v[c];
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
                    if (lane >= (uint32_t)o) inc += t;
                }
                if (c * 32u + lane < nblocks) row[c * 32u + lane] = base + inc - v[c];
                base += __shfl_sync(0xFFFFFFFFu, inc, 31);
            }
        }
        return;
    }
    for (uint32_t b0 = 0; b0 < nblocks; b0 += 32u) {
        const uint32_t x = (b0 + lane < nblocks) ? row[b0 + lane] : 0u;
        uint32_t inc = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
            if (lane >= (uint32_t)o) inc += t;
        }
        if (b0 + lane < nblocks) row[b0 + lane] = base + inc - x;
        base += __shfl_sync(0xFFFFFFFFu, inc, 31);
    }
}

// Scatter of one tile.  The keys are first put into digit order inside shared memory (stable: warp slice, round, lane), then
// written out by consecutive threads, so that the keys of one digit -- which are consecutive in the output -- leave as
// coalesced runs instead of 32 scattered 8-byte stores per warp.
__global__ void __launch_bounds__(RS_THREADS) rs_scatter_kernel(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                                uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                                                uint32_t n, int shift, const uint32_t* __restrict__ ghist, uint32_t nblocks) {
    __shared__ uint32_t cnt[RS_WARPS][256];  // per-warp digit counts, then the warp's start inside the tile's digit run
    __shared__ uint32_t delta[256];          // global start of the tile's digit run minus its start inside the sorted tile
    __shared__ uint32_t wsum[RS_WARPS];
    __shared__ uint64_t skey[RS_TILE];
    __shared__ uint32_t sval[RS_TILE];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) cnt[w][tid] = 0;
    __syncthreads();
    const uint64_t tile0 = (uint64_t)blockIdx.x * RS_TILE;
    const uint32_t nv = (uint32_t)min((uint64_t)RS_TILE, (uint64_t)n - tile0);
    const uint32_t base = warp * (32 * RS_ITEMS);
    uint64_t key[RS_ITEMS];
    uint32_t val[RS_ITEMS], off[RS_ITEMS];
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        const uint32_t i = base + r * 32u + lane;
        key[r] = 0; val[r] = 0;
        if (i < nv) { key[r] = keys_in[tile0 + i]; val[r] = vals_in[tile0 + i]; }
    }
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        const bool valid = base + r * 32u + lane < nv;
        // invalid lanes get a private pseudo-digit so they only match themselves
        const uint32_t dgt = valid ? ((uint32_t)(key[r] >> shift) & 255u) : (256u + lane);
        const uint32_t peers = __match_any_sync(0xFFFFFFFFu, dgt);
        const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
        uint32_t old = 0;
        if (valid && rank == 0) { old = cnt[warp][dgt]; cnt[warp][dgt] = old + __popc(peers); }
        old = __shfl_sync(0xFFFFFFFFu, old, __ffs(peers) - 1);
        off[r] = old + rank;
        __syncwarp();
    }
    __syncthreads();
    {   // digit tid: the warps' shares of the digit become offsets inside the digit's run; the runs are laid out by an
        // exclusive scan of the digit totals over the block
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) { const uint32_t c = cnt[w][tid]; cnt[w][tid] = run; run += c; }
        uint32_t inc = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
            if (lane >= (uint32_t)o) inc += t;
        }
        if (lane == 31u) wsum[warp] = inc;
        __syncthreads();
        uint32_t lstart = inc - run;
        for (uint32_t w = 0; w < warp; ++w) lstart += wsum[w];
        delta[tid] = ghist[(uint64_t)tid * nblocks + blockIdx.x] - lstart;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) cnt[w][tid] += lstart;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        if (base + r * 32u + lane < nv) {
            const uint32_t lp = cnt[warp][(uint32_t)(key[r] >> shift) & 255u] + off[r];
            skey[lp] = key[r];
            sval[lp] = val[r];
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        const uint32_t i = r * RS_THREADS + tid;
        if (i < nv) {
            const uint64_t k = skey[i];
            const uint32_t pos = i + delta[(uint32_t)(k >> shift) & 255u];
            keys_out[pos] = k;
            vals_out[pos] = sval[i];
        }
    }
}

void BuildScratch::begin() {
    if (want > cap) {
        if (base) cudaFree(base);
        base = nullptr;
        cap = 0;
        const size_t ask = want + want / 8 + (1u << 20);
        if (cudaMalloc(&base, ask) == cudaSuccess) cap = ask; else { base = nullptr; (void)cudaGetLastError(); }
    }
    off = 0;
    want = 0;
}
void* BuildScratch::get(size_t bytes) {
    bytes = (bytes + 255) & ~(size_t)255;
    want += bytes;
    if (off + bytes > cap) return nullptr;
    void* p = base + off;
    off += bytes;
    return p;
}
void BuildScratch::release() {
    if (base) cudaFree(base);
    base = nullptr;
    cap = off = want = 0;
}

int radix_sort_pairs(uint64_t* d_keys, uint64_t* d_keys_tmp, uint32_t* d_vals, uint32_t* d_vals_tmp, uint32_t n,
                     cudaStream_t stream, unsigned long long* launches, bool* result_in_tmp, BuildScratch* scratch, int passes) {
    *result_in_tmp = (passes & 1) != 0;
    if (n == 0) return 0;
    const uint32_t nblocks = cdiv(n, RS_TILE);
    const size_t hist_bytes = sizeof(uint32_t) * (256ull * nblocks + 8ull * 256ull);  // tile counts + 8 passes of digit totals
    uint32_t* d_hist = scratch ? static_cast<uint32_t*>(scratch->get(hist_bytes)) : nullptr;
    const bool own_hist = d_hist == nullptr;
    if (own_hist) CK(cudaMalloc(&d_hist, hist_bytes));
    uint32_t* d_tot = d_hist + 256ull * nblocks;
    CK(cudaMemsetAsync(d_tot, 0, sizeof(uint32_t) * 8 * 256, stream));
    uint64_t* kin = d_keys; uint64_t* kout = d_keys_tmp;
    uint32_t* vin = d_vals; uint32_t* vout = d_vals_tmp;
    for (int pass = 0; pass < passes; ++pass) {   // (keys that only use their low 8 * passes bits need no more passes)
        const int shift = pass * 8;
        rs_hist_kernel<<<nblocks, RS_THREADS, 0, stream>>>(kin, n, shift, d_hist, d_tot + 256 * pass, nblocks);
        rs_scan_kernel<<<256 / RS_WARPS, RS_THREADS, 0, stream>>>(d_hist, d_tot + 256 * pass, nblocks);
        rs_scatter_kernel<<<nblocks, RS_THREADS, 0, stream>>>(kin, vin, kout, vout, n, shift, d_hist, nblocks);
        *launches += 3;
        std::swap(kin, kout);
        std::swap(vin, vout);
    }
    // an even number of passes: the result is back in the original buffers
    if (own_hist) {
        cudaError_t e = cudaStreamSynchronize(stream);
        cudaFree(d_hist);
        if (e != cudaSuccess) return (int)e;
    }
    CK(cudaGetLastError());
    return 0;
}

// ------------------------------------------------------------------------------------------- allocation-free sort
// The same eight LSD passes as radix_sort_pairs, for callers that must not allocate (CUDA graph capture, serving loops): the
// ping-pong key / value buffers and the histograms live in ONE file-scope device array instead of cudaMalloc'd or BuildScratch
// memory.  Every GPU that loads the library holds the array until the process exits; every call shares it (calls that overlap
// on the GPU race) and it keeps the last call's contents -- the digit totals, the only part that has to start at zero, are
// cleared by each call.
constexpr uint32_t NOALLOC_SORT_MAX_N = GRT_NOALLOC_SORT_MAX_N;
constexpr size_t NA_KEYS_OFF = 0;
constexpr size_t NA_VALS_OFF = NA_KEYS_OFF + sizeof(uint64_t) * (size_t)NOALLOC_SORT_MAX_N;
constexpr size_t NA_HIST_OFF = NA_VALS_OFF + sizeof(uint32_t) * (size_t)NOALLOC_SORT_MAX_N;
constexpr size_t NA_HIST_BYTES = sizeof(uint32_t) * (256ull * ((NOALLOC_SORT_MAX_N + RS_TILE - 1) / RS_TILE) + 8ull * 256ull);
constexpr size_t NA_TOTAL = NA_HIST_OFF + NA_HIST_BYTES;   // 12.5 MiB + 520 KiB at 2^20 pairs
static_assert(NA_VALS_OFF % 256 == 0 && NA_HIST_OFF % 256 == 0, "sub-buffers stay 256-byte aligned");
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_TOTAL];

// In host code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address is looked up once per
// device, with that device current, and cached.
constexpr int NA_MAX_DEVICES = 64;
static std::mutex g_noalloc_mu;
static unsigned char* g_noalloc_ptr[NA_MAX_DEVICES] = {};

int radix_sort_pairs_noalloc(int device, uint64_t* d_keys, uint32_t* d_vals, uint32_t n, cudaStream_t stream, unsigned long long* launches) {
    if (n == 0) return 0;
    if (n > NOALLOC_SORT_MAX_N || device < 0 || device >= NA_MAX_DEVICES) return (int)cudaErrorInvalidValue;  // (the caller words the error)
    unsigned char* base = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (!g_noalloc_ptr[device]) {
            CK(cudaSetDevice(device));
            void* p = nullptr;
            CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
            g_noalloc_ptr[device] = static_cast<unsigned char*>(p);
        }
        base = g_noalloc_ptr[device];
    }
    uint64_t* k_tmp = reinterpret_cast<uint64_t*>(base + NA_KEYS_OFF);
    uint32_t* v_tmp = reinterpret_cast<uint32_t*>(base + NA_VALS_OFF);
    uint32_t* d_hist = reinterpret_cast<uint32_t*>(base + NA_HIST_OFF);
    const uint32_t nblocks = cdiv(n, RS_TILE);
    uint32_t* d_tot = d_hist + 256ull * nblocks;
    CK(cudaMemsetAsync(d_tot, 0, sizeof(uint32_t) * 8 * 256, stream));   // the digit totals are accumulated with atomics
    uint64_t* kin = d_keys; uint64_t* kout = k_tmp;
    uint32_t* vin = d_vals; uint32_t* vout = v_tmp;
    for (int pass = 0; pass < 8; ++pass) {
        const int shift = pass * 8;
        rs_hist_kernel<<<nblocks, RS_THREADS, 0, stream>>>(kin, n, shift, d_hist, d_tot + 256 * pass, nblocks);
        rs_scan_kernel<<<256 / RS_WARPS, RS_THREADS, 0, stream>>>(d_hist, d_tot + 256 * pass, nblocks);
        rs_scatter_kernel<<<nblocks, RS_THREADS, 0, stream>>>(kin, vin, kout, vout, n, shift, d_hist, nblocks);
        *launches += 3;
        std::swap(kin, kout);
        std::swap(vin, vout);
    }
    // 8 passes: the result is back in the caller's buffers
    CK(cudaGetLastError());
    return 0;
}
uint32_t radix_sort_pairs_noalloc_max() { return NOALLOC_SORT_MAX_N; }

// ------------------------------------------------------------------------------------------- bounds + Morton
__device__ __forceinline__ uint32_t enc_f(float f) {
    uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ __forceinline__ float dec_f(uint32_t e) {
    uint32_t b = (e & 0x80000000u) ? (e & 0x7FFFFFFFu) : ~e;
#ifdef __CUDA_ARCH__
    return __uint_as_float(b);
#else
    float f; memcpy(&f, &b, 4); return f;
#endif
}

// bounds[0..2] = encoded min of box centres, bounds[3..5] = encoded max, bounds[6] = live count
// (grid-stride, one set of atomics per block: min / max / count do not depend on the order)
__global__ void __launch_bounds__(256) bounds_kernel(const float4* __restrict__ lo, const float4* __restrict__ hi, uint32_t n,
                                                     uint32_t* __restrict__ bounds) {
    __shared__ float sm[8][6];
    __shared__ uint32_t sl[8];
    float c[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, C[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    uint32_t live = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = lo[i], b = hi[i];
        if (a.x <= b.x) {
            live++;
            const float m[3] = {0.5f * (a.x + b.x), 0.5f * (a.y + b.y), 0.5f * (a.z + b.z)};
#pragma unroll
            for (int k = 0; k < 3; ++k) { c[k] = fminf(c[k], m[k]); C[k] = fmaxf(C[k], m[k]); }
        }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            c[k] = fminf(c[k], __shfl_xor_sync(0xFFFFFFFFu, c[k], o));
            C[k] = fmaxf(C[k], __shfl_xor_sync(0xFFFFFFFFu, C[k], o));
        }
        live += __shfl_xor_sync(0xFFFFFFFFu, live, o);
    }
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { sm[warp][k] = c[k]; sm[warp][3 + k] = C[k]; }
        sl[warp] = live;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t nl = 0;
        for (int w = 0; w < 8; ++w) {
            nl += sl[w];
#pragma unroll
            for (int k = 0; k < 3; ++k) { c[k] = fminf(c[k], sm[w][k]); C[k] = fmaxf(C[k], sm[w][3 + k]); }
        }
        if (nl) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                atomicMin(&bounds[k], enc_f(c[k]));
                atomicMax(&bounds[3 + k], enc_f(C[k]));
            }
            atomicAdd(&bounds[6], nl);
        }
    }
}

__device__ __forceinline__ uint64_t expand21(uint64_t x) {
    x &= 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

// GRT_MORTON_SIZE_EVERY = k > 0: extended Morton code (Vinkler, Bittner, Havran 2017) -- after every k levels of x,y,z bits one
// bit of the box's diagonal (relative to the scene's) is inserted, so that primitives much larger than their neighbours
// split off high in the tree instead of inflating the boxes of every node down to their leaf.  0 = plain 63-bit code.
#ifndef GRT_MORTON_PASSES_DEFAULT
#define GRT_MORTON_PASSES_DEFAULT 0   // 0: by size (4 or 6)
#endif
#ifndef GRT_MORTON_SIZE_EVERY
#define GRT_MORTON_SIZE_EVERY 0
#endif

// drop = 21 - bits per axis kept: the key then has 3 * (21 - drop) bits and needs fewer radix passes (dead boxes keep the all-ones
// key, which is larger than any live key in the bits the passes look at as long as those are more than the key has).
__global__ void morton_kernel(const float4* __restrict__ lo, const float4* __restrict__ hi, uint32_t n, const uint32_t* __restrict__ bounds,
                              uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t drop) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 a = lo[i], b = hi[i];
    uint64_t key = 0xFFFFFFFFFFFFFFFFull;  // dead boxes sort last
    if (a.x <= b.x) {
        const float c[3] = {0.5f * (a.x + b.x), 0.5f * (a.y + b.y), 0.5f * (a.z + b.z)};
        uint64_t q[3];
        float ext2 = 0.0f;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float mn = dec_f(bounds[k]), mx = dec_f(bounds[3 + k]);
            const float ext = fmaxf(mx - mn, 1e-30f);
            ext2 += ext * ext;
            float u = (c[k] - mn) / ext;
            u = fminf(fmaxf(u, 0.0f), 1.0f);
            q[k] = (uint64_t)min(2097151u, (uint32_t)(u * 2097152.0f));
        }
        if (GRT_MORTON_SIZE_EVERY == 0) {
            key = (expand21(q[0] >> drop) << 2) | (expand21(q[1] >> drop) << 1) | expand21(q[2] >> drop);
        } else {
            const float dx = b.x - a.x, dy = b.y - a.y, dz = b.z - a.z;
            const float rel = fminf(sqrtf((dx * dx + dy * dy + dz * dz) / ext2), 1.0f);
            const uint32_t qs = min(2097151u, (uint32_t)(rel * 2097152.0f));
            key = 0;
            int bits = 0, lvl = 20, sb = 20;
            while (bits < 63) {
                for (int r = 0; r < GRT_MORTON_SIZE_EVERY && bits < 63; ++r, --lvl) {
#pragma unroll
                    for (int k = 0; k < 3; ++k)
                        if (bits < 63) { key = (key << 1) | ((q[k] >> lvl) & 1ull); ++bits; }
                }
                if (bits < 63) { key = (key << 1) | ((uint64_t)(qs >> sb) & 1ull); ++bits; --sb; }
            }
        }
    }
    keys[i] = key;
    vals[i] = i;
}

// ------------------------------------------------------------------------------------------- Karras hierarchy
// Unified node ids: internal i in [0, n-1), leaf j as (n-1)+j.
__device__ __forceinline__ int delta_fn(const uint64_t* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clzll((long long)(a ^ b));
}

// A binary node is one 48-byte record (children, leaf range, box): whoever visits a node touches two sectors, not six arrays.
struct BinNode {
    uint32_t left, right, first, last;
    float4 lo, hi;
};

__global__ void karras_kernel(const uint64_t* __restrict__ keys, int n, BinNode* __restrict__ bn, uint32_t* __restrict__ parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (delta_fn(keys, n, i, i + 1) - delta_fn(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta_fn(keys, n, i, i - d);
    int lmax = 2;
    while (delta_fn(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta_fn(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta_fn(keys, n, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) >> 1;
        if (delta_fn(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const uint32_t l_id = (lo == gamma) ? (uint32_t)(n - 1 + gamma) : (uint32_t)gamma;
    const uint32_t r_id = (hi == gamma + 1) ? (uint32_t)(n - 1 + gamma + 1) : (uint32_t)(gamma + 1);
    *reinterpret_cast<uint4*>(&bn[i]) = make_uint4(l_id, r_id, (uint32_t)lo, (uint32_t)hi);
    parent[l_id] = (uint32_t)i;
    parent[r_id] = (uint32_t)i;
    if (i == 0) parent[0] = 0xFFFFFFFFu;
}

// leafbox[2j], leafbox[2j+1] = box of the j-th record in Morton order
__global__ void leaf_boxes_kernel(const float4* __restrict__ lo, const float4* __restrict__ hi, const uint32_t* __restrict__ perm, uint32_t n,
                                  float4* __restrict__ leafbox) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint32_t i = perm[j];
    leafbox[2u * j] = lo[i];
    leafbox[2u * j + 1u] = hi[i];
}

__global__ void refit_kernel(uint32_t n, const uint32_t* __restrict__ parent, BinNode* bn, const float4* __restrict__ leafbox, uint32_t* flags) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    uint32_t p = parent[n - 1 + j];
    while (p != 0xFFFFFFFFu) {
        __threadfence();
        if (atomicAdd(&flags[p], 1u) == 0u) return;  // first child to arrive: the sibling will finish the job
        __threadfence();
        const uint32_t l = bn[p].left, r = bn[p].right;
        const float4* pl = l >= n - 1 ? leafbox + 2u * (l - (n - 1)) : &bn[l].lo;
        const float4* pr = r >= n - 1 ? leafbox + 2u * (r - (n - 1)) : &bn[r].lo;
        const float4 la = __ldcg(pl), lb = __ldcg(pl + 1), ra = __ldcg(pr), rb = __ldcg(pr + 1);
        __stcg(&bn[p].lo, make_float4(fminf(la.x, ra.x), fminf(la.y, ra.y), fminf(la.z, ra.z), 0.f));
        __stcg(&bn[p].hi, make_float4(fmaxf(lb.x, rb.x), fmaxf(lb.y, rb.y), fmaxf(lb.z, rb.z), 0.f));
        p = parent[p];
    }
}

// ------------------------------------------------------------------------------------------- collapse to 8-wide
struct BinTree {
    const BinNode* bn;
    const float4* leafbox;
    uint32_t n;
    // box of a unified node id (internal i in [0, n-1), leaf j as (n-1)+j): lo at [0], hi at [1]
    __device__ __forceinline__ const float4* box(uint32_t id) const { return id >= n - 1 ? leafbox + 2u * (id - (n - 1)) : &bn[id].lo; }
};

__device__ __forceinline__ void write_child(WideNode* nd, int k, float4 lo, float4 hi, uint32_t ref) {
    nd->lox[k] = lo.x; nd->loy[k] = lo.y; nd->loz[k] = lo.z;
    nd->hix[k] = hi.x; nd->hiy[k] = hi.y; nd->hiz[k] = hi.z;
    nd->child[k] = ref;
    nd->pad[k] = 0;
}

// What the collapse needs to know about a candidate child: fetched once when the child appears (all loads independent),
// so that opening a child costs one round of memory latency instead of one per candidate per step.
struct Slot {
    uint32_t id, cnt, first, l, r;
    float area;
};
__device__ __forceinline__ Slot load_slot(const BinTree& T, uint32_t id) {
    Slot s;
    s.id = id;
    if (id >= T.n - 1) {
        s.cnt = 1u; s.first = id - (T.n - 1); s.l = s.r = 0u; s.area = -1.0f;
        return s;
    }
    const uint4 h = *reinterpret_cast<const uint4*>(&T.bn[id]);
    const float4 a = T.bn[id].lo, d = T.bn[id].hi;
    const uint32_t f = h.z, la = h.w;
    s.l = h.x;
    s.r = h.y;
    const float ex = d.x - a.x, ey = d.y - a.y, ez = d.z - a.z;
    s.area = ex * ey + ey * ez + ez * ex;
    s.cnt = la - f + 1u;
    s.first = f;
    return s;
}
__device__ __forceinline__ Slot leaf_slot(const BinTree& T, uint32_t j) {
    Slot s;
    s.id = T.n - 1 + j; s.cnt = 1u; s.first = j; s.l = s.r = 0u; s.area = -1.0f;
    return s;
}

// One level: work items (binary node, wide-node index) come from q_in, the items of the next level go to q_out.  The item
// counts live on the device (lvl[0] in, lvl[1] out) so that several levels can be queued without a host round trip; a wide
// node whose index does not fit into `cap` is counted but not written (the host then repeats the collapse into a larger array).
__global__ void __launch_bounds__(128) collapse_kernel(BinTree T, const uint2* __restrict__ q_in, uint2* __restrict__ q_out,
                                                       uint32_t* __restrict__ lvl, uint32_t* __restrict__ node_counter,
                                                       WideNode* __restrict__ nodes, uint32_t cap) {
    const uint32_t n_in = lvl[0];
    uint32_t* const n_out = lvl + 1;
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < n_in; t += gridDim.x * blockDim.x) {
        const uint32_t b = q_in[t].x, w = q_in[t].y;
        Slot sl[WIDE];
        int ns = 2;
        {
            const uint32_t l = T.bn[b].left, r = T.bn[b].right;
            sl[0] = load_slot(T, l);
            sl[1] = load_slot(T, r);
        }
        // Open the largest-area child until the node is full.  A subtree of more than WIDE records is split into its
        // two binary children; a smaller one is only opened if ALL its records fit into the free slots (it is then
        // flattened into single-record leaf children), so that the bottom of the tree is made of well-filled nodes
        // instead of chains of 2- and 3-child nodes.
        while (ns < WIDE) {
            int best = -1;
            float bestA = -1.0f;
#pragma unroll
            for (int k = 0; k < WIDE; ++k) {
                if (k >= ns) continue;
                const uint32_t c = sl[k].cnt;
                if (c <= (uint32_t)LEAF_MAX) continue;  // (covers single records: cnt 1)
                if (LEAF_MAX == 1 && c <= (uint32_t)WIDE && ns - 1 + (int)c > WIDE) continue;
                if (sl[k].area > bestA) { bestA = sl[k].area; best = k; }
            }
            if (best < 0) break;
            Slot o;
#pragma unroll
            for (int k = 0; k < WIDE; ++k) if (k == best) o = sl[k];
            Slot na, nb;
            bool flat = LEAF_MAX == 1 && o.cnt <= (uint32_t)WIDE;
            if (!flat) {
                na = load_slot(T, o.l);
                nb = load_slot(T, o.r);
            } else {
                na = leaf_slot(T, o.first);
            }
#pragma unroll
            for (int k = 0; k < WIDE; ++k) if (k == best) sl[k] = na;
            if (!flat) {
#pragma unroll
                for (int k = 0; k < WIDE; ++k) if (k == ns) sl[k] = nb;
                ++ns;
            } else {
                for (uint32_t j = 1; j < o.cnt; ++j) {
                    const Slot lf = leaf_slot(T, o.first + j);
#pragma unroll
                    for (int k = 0; k < WIDE; ++k) if (k == ns) sl[k] = lf;
                    ++ns;
                }
            }
        }
        // children that stay subtrees become work items of the next level: one reservation per node, not per child
        uint32_t m = 0;
#pragma unroll
        for (int k = 0; k < WIDE; ++k) m += (k < ns && sl[k].cnt > (uint32_t)LEAF_MAX) ? 1u : 0u;
        uint32_t wi = 0, qo = 0;
        if (m) {
            wi = atomicAdd(node_counter, m);
            qo = atomicAdd(n_out, m);
        }
        const bool fits = w < cap;
        WideNode* nd = &nodes[fits ? w : 0u];
#pragma unroll
        for (int k = 0; k < WIDE; ++k) {
            if (k < ns) {
                const uint32_t id = sl[k].id;
                uint32_t ref;
                if (sl[k].cnt <= (uint32_t)LEAF_MAX) {
                    ref = CHILD_LEAF | ((sl[k].cnt - 1u) << 28) | sl[k].first;
                } else {
                    q_out[qo++] = make_uint2(id, wi);
                    ref = wi++;
                }
                if (fits) { const float4* bx = T.box(id); write_child(nd, k, bx[0], bx[1], ref); }
            } else if (fits) {
                write_child(nd, k, make_float4(INFINITY, INFINITY, INFINITY, 0.f), make_float4(-INFINITY, -INFINITY, -INFINITY, 0.f), CHILD_EMPTY);
            }
        }
    }
}

// n_live <= LEAF_MAX: one node, one leaf child holding everything
__global__ void tiny_root_kernel(const float4* __restrict__ lo, const float4* __restrict__ hi, const uint32_t* __restrict__ perm, uint32_t n,
                                 WideNode* __restrict__ nodes) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    float4 a = make_float4(INFINITY, INFINITY, INFINITY, 0.f), b = make_float4(-INFINITY, -INFINITY, -INFINITY, 0.f);
    for (uint32_t j = 0; j < n; ++j) {
        const float4 l = lo[perm[j]], h = hi[perm[j]];
        a.x = fminf(a.x, l.x); a.y = fminf(a.y, l.y); a.z = fminf(a.z, l.z);
        b.x = fmaxf(b.x, h.x); b.y = fmaxf(b.y, h.y); b.z = fmaxf(b.z, h.z);
    }
    write_child(&nodes[0], 0, a, b, CHILD_LEAF | ((n - 1u) << 28) | 0u);
    for (int k = 1; k < WIDE; ++k)
        write_child(&nodes[0], k, make_float4(INFINITY, INFINITY, INFINITY, 0.f), make_float4(-INFINITY, -INFINITY, -INFINITY, 0.f), CHILD_EMPTY);
}

// GRT_BUILD_TIMING=1: event-timed phases of lbvh_build on stderr (diagnostic; the events themselves cost a few microseconds)
struct PhaseTimer {
    bool on = false;
    cudaStream_t stream = nullptr;
    std::vector<std::pair<const char*, cudaEvent_t>> marks;
    explicit PhaseTimer(cudaStream_t s) : stream(s) {
        const char* e = getenv("GRT_BUILD_TIMING");
        on = e && e[0] == '1';
        mark("start");
    }
    void mark(const char* name) {
        if (!on) return;
        cudaEvent_t ev;
        cudaEventCreate(&ev);
        cudaEventRecord(ev, stream);
        marks.emplace_back(name, ev);
    }
    ~PhaseTimer() {
        if (!on) return;
        cudaStreamSynchronize(stream);
        fprintf(stderr, "[lbvh_build]");
        for (size_t i = 1; i < marks.size(); ++i) {
            float ms = 0;
            cudaEventElapsedTime(&ms, marks[i - 1].second, marks[i].second);
            fprintf(stderr, " %s %.1f us |", marks[i].first, ms * 1e3f);
        }
        fprintf(stderr, "\n");
        for (auto& m : marks) cudaEventDestroy(m.second);
    }
};

struct DevBuf {  // a build temporary: from the context's scratch block, else cudaMalloc'd and freed on scope exit
    void* p = nullptr;
    bool owned = false;
    BuildScratch* scratch = nullptr;
    ~DevBuf() { if (p && owned) cudaFree(p); }
    cudaError_t alloc(size_t bytes) {
        p = scratch ? scratch->get(bytes ? bytes : 1) : nullptr;
        if (p) return cudaSuccess;
        owned = true;
        return cudaMalloc(&p, bytes ? bytes : 1);
    }
    template <class T> T* as() { return reinterpret_cast<T*>(p); }
};

int lbvh_build(const float4* d_lo, const float4* d_hi, uint32_t n, cudaStream_t stream, LbvhResult* out, unsigned long long* launches,
               BuildScratch* scratch) {
    WideNode* const reuse_nodes = out->reuse_nodes;
    const uint32_t reuse_cap = out->reuse_cap;
    *out = LbvhResult();
    if (n == 0) return 0;
    if (n >= (1u << 28)) return (int)cudaErrorInvalidValue;
    const int TB = 256;
    PhaseTimer pt(stream);
    DevBuf bounds, keys, keys2, vals2;
    bounds.scratch = keys.scratch = keys2.scratch = vals2.scratch = scratch;
    CK(bounds.alloc(8 * sizeof(uint32_t)));
    const uint32_t h_init[8] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u, 0u, 0u};
    CK(cudaMemcpyAsync(bounds.p, h_init, sizeof(h_init), cudaMemcpyHostToDevice, stream));
    bounds_kernel<<<std::min(cdiv(n, 256), 148u * 8u), 256, 0, stream>>>(d_lo, d_hi, n, bounds.as<uint32_t>());
    ++*launches;
    CK(keys.alloc(sizeof(uint64_t) * n));
    CK(keys2.alloc(sizeof(uint64_t) * n));
    CK(vals2.alloc(sizeof(uint32_t) * n));
    uint32_t* d_perm = scratch ? static_cast<uint32_t*>(scratch->get(sizeof(uint32_t) * n)) : nullptr;
    out->perm_owned = d_perm == nullptr;
    if (!d_perm) CK(cudaMalloc(&d_perm, sizeof(uint32_t) * n));
    out->perm = d_perm;
    // Radix passes (of 8 bits) over the Morton keys: 8 = the full 63-bit code; 6 = 15 bits per axis, 4 = 10 bits per axis (records that
    // share a cell keep their input order; Karras splits equal keys by index).  GRT_MORTON_PASSES overrides.
    static const int env_passes = [] { const char* e = getenv("GRT_MORTON_PASSES"); return e ? atoi(e) : 0; }();
    // Default: 6 passes (measured at 1M: the very same tree as with 8, 0.99 -> 0.94 ms), 4 for sets of up to 2^17 boxes (1024 cells per
    // axis are plenty for them; 100k: 0.49 -> 0.41 ms).
    const int passes = (env_passes == 4 || env_passes == 6 || env_passes == 8) ? env_passes : GRT_MORTON_PASSES_DEFAULT != 0 ? GRT_MORTON_PASSES_DEFAULT : n <= (1u << 17) ? 4 : 6;
    const uint32_t drop = GRT_MORTON_SIZE_EVERY != 0 ? 0u : passes == 8 ? 0u : passes == 6 ? 6u : 11u;
    morton_kernel<<<cdiv(n, TB), TB, 0, stream>>>(d_lo, d_hi, n, bounds.as<uint32_t>(), keys.as<uint64_t>(), d_perm, drop);
    ++*launches;
    pt.mark("bounds+morton");
    bool in_tmp = false;
    int rc = radix_sort_pairs(keys.as<uint64_t>(), keys2.as<uint64_t>(), d_perm, vals2.as<uint32_t>(), n, stream, launches, &in_tmp, scratch,
                              GRT_MORTON_SIZE_EVERY != 0 ? 8 : passes);
    if (rc) return rc;
    uint32_t h_bounds[8];
    CK(cudaMemcpyAsync(h_bounds, bounds.p, sizeof(h_bounds), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    pt.mark("sort+sync");
    const uint32_t nl = h_bounds[6];
    out->n_live = nl;
    for (int k = 0; k < 3; ++k) { out->lo[k] = dec_f(h_bounds[k]); out->hi[k] = dec_f(h_bounds[3 + k]); }
    if (nl == 0) return 0;

    if (nl <= (uint32_t)LEAF_MAX) {
        WideNode* nodes = reuse_cap >= 1 ? reuse_nodes : nullptr;
        if (!nodes) CK(cudaMalloc(&nodes, sizeof(WideNode)));
        tiny_root_kernel<<<1, 32, 0, stream>>>(d_lo, d_hi, d_perm, nl, nodes);
        ++*launches;
        CK(cudaStreamSynchronize(stream));
        out->nodes = nodes;
        out->n_nodes = 1;
        out->levels = 1;
        return 0;
    }

    DevBuf bn, leafbox, parent, flags, qa, qb, ctr, wide_tmp;
    for (DevBuf* b : {&bn, &leafbox, &parent, &flags, &qa, &qb, &ctr, &wide_tmp}) b->scratch = scratch;
    CK(bn.alloc(sizeof(BinNode) * (size_t)(nl - 1)));
    CK(leafbox.alloc(32ull * nl));
    CK(parent.alloc(4ull * (2ull * nl - 1)));
    CK(flags.alloc(4ull * (nl - 1)));
    CK(cudaMemsetAsync(flags.p, 0, 4ull * (nl - 1), stream));
    karras_kernel<<<cdiv(nl - 1, TB), TB, 0, stream>>>(keys.as<uint64_t>(), (int)nl, bn.as<BinNode>(), parent.as<uint32_t>());
    leaf_boxes_kernel<<<cdiv(nl, TB), TB, 0, stream>>>(d_lo, d_hi, d_perm, nl, leafbox.as<float4>());
    refit_kernel<<<cdiv(nl, TB), TB, 0, stream>>>(nl, parent.as<uint32_t>(), bn.as<BinNode>(), leafbox.as<float4>(), flags.as<uint32_t>());
    *launches += 3;
    pt.mark("karras+leaf+refit");

    // collapse, level by level.  Upper bound on wide nodes: every wide node but the root is a binary
    // internal node with > LEAF_MAX prims, and two such nodes never share a leaf run, so <= nl.
    // The levels are queued COLLAPSE_CHUNK at a time with their item counts on the device; the nodes are written
    // straight into the caller's array when it is likely to be large enough (checked afterwards), else into scratch.
    const uint32_t max_nodes = nl;
    constexpr int COLLAPSE_MAX_CHUNK = 16;
    // levels queued per host round trip: an 8-wide tree over nl records is about log8(nl) + 6 levels deep
    int chunk = 7;
    for (uint32_t m = nl; m > 1u; m >>= 3) ++chunk;
    const int COLLAPSE_CHUNK = std::min(chunk, COLLAPSE_MAX_CHUNK);
    CK(qa.alloc(sizeof(uint2) * (size_t)max_nodes));
    CK(qb.alloc(sizeof(uint2) * (size_t)max_nodes));
    CK(ctr.alloc((COLLAPSE_MAX_CHUNK + 2) * sizeof(uint32_t)));  // [0] = node counter, [1..] = items per level of the chunk
    BinTree T{bn.as<BinNode>(), leafbox.as<float4>(), nl};
    const uint32_t grid = std::min(cdiv(max_nodes, 128), 148u * 8u);
    uint32_t n_nodes = 0, levels = 0;
    WideNode* target = (reuse_nodes && reuse_cap >= nl / 6u + 16u) ? reuse_nodes : nullptr;
    uint32_t cap = target ? reuse_cap : max_nodes;
    for (int attempt = 0; attempt < 2; ++attempt) {
        if (!target) {
            CK(wide_tmp.alloc(sizeof(WideNode) * (size_t)max_nodes));
            target = wide_tmp.as<WideNode>();
            cap = max_nodes;
        }
        const uint2 root_item = make_uint2(0u, 0u);
        CK(cudaMemcpyAsync(qa.p, &root_item, sizeof(uint2), cudaMemcpyHostToDevice, stream));
        uint32_t h_ctr[COLLAPSE_MAX_CHUNK + 2] = {1u, 1u};  // one node (the root) allocated, one item at level 0
        CK(cudaMemcpyAsync(ctr.p, h_ctr, sizeof(h_ctr), cudaMemcpyHostToDevice, stream));
        uint2* qi = qa.as<uint2>();
        uint2* qo = qb.as<uint2>();
        levels = 0;
        for (;;) {
            for (int l = 0; l < COLLAPSE_CHUNK; ++l) {
                collapse_kernel<<<grid, 128, 0, stream>>>(T, qi, qo, ctr.as<uint32_t>() + 1 + l, ctr.as<uint32_t>(), target, cap);
                std::swap(qi, qo);
            }
            *launches += COLLAPSE_CHUNK;
            CK(cudaMemcpyAsync(h_ctr, ctr.p, sizeof(h_ctr), cudaMemcpyDeviceToHost, stream));
            CK(cudaStreamSynchronize(stream));
            for (int l = 0; l < COLLAPSE_CHUNK; ++l) levels += h_ctr[1 + l] != 0u;
            if (h_ctr[1 + COLLAPSE_CHUNK] == 0u) break;
            if (levels > 4096) return (int)cudaErrorUnknown;
            uint32_t next[COLLAPSE_MAX_CHUNK + 2] = {h_ctr[0], h_ctr[1 + COLLAPSE_CHUNK]};
            CK(cudaMemcpyAsync(ctr.p, next, sizeof(next), cudaMemcpyHostToDevice, stream));
        }
        n_nodes = h_ctr[0];
        if (n_nodes <= cap) break;
        target = nullptr;  // the caller's array was too small after all: once more, into scratch
    }
    CK(cudaGetLastError());
    pt.mark("collapse+sync");
    WideNode* nodes = target;
    if (target != reuse_nodes) {
        nodes = n_nodes <= reuse_cap ? reuse_nodes : nullptr;
        if (!nodes) {
            out->nodes_cap = n_nodes + n_nodes / 8u + 16u;  // headroom: a slightly different particle set rebuilds in place
            CK(cudaMalloc(&nodes, sizeof(WideNode) * (size_t)out->nodes_cap));
        }
        CK(cudaMemcpyAsync(nodes, target, sizeof(WideNode) * (size_t)n_nodes, cudaMemcpyDeviceToDevice, stream));
        CK(cudaStreamSynchronize(stream));
    }
    out->nodes = nodes;
    out->n_nodes = n_nodes;
    out->levels = levels;
    return 0;
}

// ------------------------------------------------------------------------------------------- particles
// Raw record = reference GaussianParticle, 59 floats: pos[0..2] scale[3..5] rot x,y,z,w [6..9] opacity[10] sh[11..58]
constexpr int RAW_FLOATS = 59;

struct M3 { float c[3][3]; };  // c[col][row], glm storage

// glm::mat3_cast (third_party/glm/gtc/quaternion.inl:47-72), un-fused IEEE ops in the source order so that
// host (oracle) and device produce bit-identical matrices.
__device__ __forceinline__ M3 mat3_cast_rn(float x, float y, float z, float w) {
    const float qxx = __fmul_rn(x, x), qyy = __fmul_rn(y, y), qzz = __fmul_rn(z, z);
    const float qxz = __fmul_rn(x, z), qxy = __fmul_rn(x, y), qyz = __fmul_rn(y, z);
    const float qwx = __fmul_rn(w, x), qwy = __fmul_rn(w, y), qwz = __fmul_rn(w, z);
    M3 R;
    R.c[0][0] = __fsub_rn(1.0f, __fmul_rn(2.0f, __fadd_rn(qyy, qzz)));
    R.c[0][1] = __fmul_rn(2.0f, __fadd_rn(qxy, qwz));
    R.c[0][2] = __fmul_rn(2.0f, __fsub_rn(qxz, qwy));
    R.c[1][0] = __fmul_rn(2.0f, __fsub_rn(qxy, qwz));
    R.c[1][1] = __fsub_rn(1.0f, __fmul_rn(2.0f, __fadd_rn(qxx, qzz)));
    R.c[1][2] = __fmul_rn(2.0f, __fadd_rn(qyz, qwx));
    R.c[2][0] = __fmul_rn(2.0f, __fadd_rn(qxz, qwy));
    R.c[2][1] = __fmul_rn(2.0f, __fsub_rn(qyz, qwx));
    R.c[2][2] = __fsub_rn(1.0f, __fmul_rn(2.0f, __fadd_rn(qxx, qyy)));
    return R;
}

// icosahedron of src/geometry/Icosahedron.h:13-29 (unit in-sphere)
__constant__ float c_ico[12][3] = {
    {-0.66158453f, 1.0704663f, 0.f}, {0.66158453f, 1.0704663f, 0.f}, {-0.66158453f, -1.0704663f, 0.f}, {0.66158453f, -1.0704663f, 0.f},
    {0.f, -0.66158453f, 1.0704663f}, {0.f, 0.66158453f, 1.0704663f}, {0.f, -0.66158453f, -1.0704663f}, {0.f, 0.66158453f, -1.0704663f},
    {1.0704663f, 0.f, -0.66158453f}, {1.0704663f, 0.f, 0.66158453f}, {-1.0704663f, 0.f, -0.66158453f}, {-1.0704663f, 0.f, 0.66158453f}};

__global__ void particle_boxes_kernel(const float* __restrict__ raw, const float* __restrict__ s_arr, uint32_t n, float4* __restrict__ lo,
                                      float4* __restrict__ hi) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* p = raw + (size_t)i * RAW_FLOATS;
    const float s = s_arr[i];
    float4 a = make_float4(1.f, 1.f, 1.f, 0.f), b = make_float4(-1.f, -1.f, -1.f, 0.f);  // dead: lo > hi
    if (s > 0.0f) {
        const M3 R = mat3_cast_rn(p[6], p[7], p[8], p[9]);
        const float sx = p[3] * s, sy = p[4] * s, sz = p[5] * s;
        float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int v = 0; v < 12; ++v) {
            const float vx = sx * c_ico[v][0], vy = sy * c_ico[v][1], vz = sz * c_ico[v][2];
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const float wv = p[r] + (R.c[0][r] * vx + R.c[1][r] * vy + R.c[2][r] * vz);
                mn[r] = fminf(mn[r], wv);
                mx[r] = fmaxf(mx[r], wv);
            }
        }
        bool ok = true;
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            // conservative padding: the slab test of the trace kernel runs in Gaussian space, so the world-space
            // box must absorb both roundings.  1e-4 of the extent is ~3 orders above fp32 noise.
            const float pad = 1e-4f * (mx[r] - mn[r]) + 1e-5f * fmaxf(fabsf(mn[r]), fabsf(mx[r])) + 1e-7f;
            mn[r] -= pad;
            mx[r] += pad;
            ok = ok && isfinite(mn[r]) && isfinite(mx[r]);
        }
        if (ok) {
            a = make_float4(mn[0], mn[1], mn[2], 0.f);
            b = make_float4(mx[0], mx[1], mx[2], 0.f);
        }
    }
    lo[i] = a;
    hi[i] = b;
}

__global__ void particle_gather_kernel(const float* __restrict__ raw, const float* __restrict__ s_arr, const uint32_t* __restrict__ perm,
                                       uint32_t n_live, PRec* __restrict__ recs, float4* __restrict__ sh) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t i = j < n_live ? perm[j] : 0u;
    // SH coefficients: the warp copies the 32 x 48 floats of its particles with consecutive lanes on consecutive floats, so
    // that both the reads (within a 192-byte row segment) and the writes (one contiguous 6 KB block) are coalesced
    {
        const uint32_t j0 = j - lane;
        const uint32_t cnt = min(32u, n_live > j0 ? n_live - j0 : 0u) * (uint32_t)(SH_F4 * 4);
        float* o = reinterpret_cast<float*>(sh + (size_t)j0 * SH_F4);
#pragma unroll 4
        for (uint32_t e = lane; e < 32u * (uint32_t)(SH_F4 * 4); e += 32u) {
            const uint32_t src = __shfl_sync(0xFFFFFFFFu, i, e / (uint32_t)(SH_F4 * 4));
            if (e < cnt) o[e] = __ldg(raw + (size_t)src * RAW_FLOATS + 11u + e % (uint32_t)(SH_F4 * 4));
        }
    }
    if (j >= n_live) return;
    const float* p = raw + (size_t)i * RAW_FLOATS;
    const M3 R = mat3_cast_rn(p[6], p[7], p[8], p[9]);
    float A[3][3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const float inv = __fdiv_rn(1.0f, p[3 + r]);  // inv_s[r][r] = 1.0f / gp.scale (tracer.cuh:196-199)
#pragma unroll
        for (int c = 0; c < 3; ++c) A[r][c] = __fmul_rn(inv, R.c[r][c]);  // (inv_s * RT): row r of A = column r of R, scaled
    }
    PRec rec;
    rec.q0 = make_float4(A[0][0], A[0][1], A[0][2], A[1][0]);
    rec.q1 = make_float4(A[1][1], A[1][2], A[2][0], A[2][1]);
    rec.q2 = make_float4(A[2][2], p[0], p[1], p[2]);
    rec.q3 = make_float4(s_arr[i], p[10], __uint_as_float(i), 0.f);
    recs[j] = rec;
}

int particles_compute_boxes(const void* d_particles236, const float* d_s, uint32_t n, float4* d_lo, float4* d_hi, cudaStream_t stream,
                            unsigned long long* launches) {
    if (!n) return 0;
    particle_boxes_kernel<<<cdiv(n, 128), 128, 0, stream>>>((const float*)d_particles236, d_s, n, d_lo, d_hi);
    ++*launches;
    return (int)cudaGetLastError();
}

int particles_gather(const void* d_particles236, const float* d_s, const uint32_t* d_perm, uint32_t n_live, PRec* d_recs, float4* d_sh,
                     cudaStream_t stream, unsigned long long* launches) {
    if (!n_live) return 0;
    particle_gather_kernel<<<cdiv(n_live, 128), 128, 0, stream>>>((const float*)d_particles236, d_s, d_perm, n_live, d_recs, d_sh);
    ++*launches;
    return (int)cudaGetLastError();
}

// ------------------------------------------------------------------------------------------- meshes
__global__ void tri_boxes_kernel(const float* __restrict__ tri9, uint32_t n, float4* __restrict__ lo, float4* __restrict__ hi) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* t = tri9 + (size_t)i * 9;
    float mn[3], mx[3];
    bool ok = true;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        mn[k] = fminf(t[k], fminf(t[3 + k], t[6 + k]));
        mx[k] = fmaxf(t[k], fmaxf(t[3 + k], t[6 + k]));
        const float pad = 1e-4f * (mx[k] - mn[k]) + 1e-5f * fmaxf(fabsf(mn[k]), fabsf(mx[k])) + 1e-7f;
        mn[k] -= pad;
        mx[k] += pad;
        ok = ok && isfinite(mn[k]) && isfinite(mx[k]);
    }
    lo[i] = ok ? make_float4(mn[0], mn[1], mn[2], 0.f) : make_float4(1.f, 1.f, 1.f, 0.f);
    hi[i] = ok ? make_float4(mx[0], mx[1], mx[2], 0.f) : make_float4(-1.f, -1.f, -1.f, 0.f);
}

__global__ void tri_gather_kernel(const float* __restrict__ tri9, const uint2* __restrict__ meta, const uint32_t* __restrict__ perm, uint32_t n,
                                  TriRec* __restrict__ out) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint32_t i = perm[j];
    const float* t = tri9 + (size_t)i * 9;
    const uint2 m = meta[i];
    TriRec r;
    r.a = make_float4(t[0], t[1], t[2], __uint_as_float(m.x));
    r.b = make_float4(__fsub_rn(t[3], t[0]), __fsub_rn(t[4], t[1]), __fsub_rn(t[5], t[2]), __uint_as_float(m.y));
    r.c = make_float4(__fsub_rn(t[6], t[0]), __fsub_rn(t[7], t[1]), __fsub_rn(t[8], t[2]), 0.f);
    out[j] = r;
}

// Reflection primitives arrive in OBJECT space with a glm 4x4 each (src/geometry/Primitives.h:14-29).  What the reference does on
// the host per update -- the instance transform OptiX applies to the GAS vertices, and vertex_normals[i] = mat3(transform) * n
// (src/GaussianTracer.cpp:658-663, NOT the inverse transpose) -- runs here on the device, in glm's un-fused evaluation order
// (column by column: (T[0]*x + T[1]*y) + T[2]*z (+ T[3])), so that the bits equal the oracle's host evaluation.
__global__ void mesh_vertices_kernel(const float* __restrict__ ov, const float* __restrict__ on, const MeshPrep* __restrict__ prep, uint32_t n_meshes,
                                     uint32_t n_verts, float* __restrict__ vw, float* __restrict__ nw) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_verts) return;
    uint32_t m = 0;
    while (m + 1 < n_meshes && i >= prep[m + 1].v_off) ++m;
    const float* T = prep[m].T;  // glm column-major: T[4*col + row]
    const float x = ov[3 * i], y = ov[3 * i + 1], z = ov[3 * i + 2];
    const float nx = on[3 * i], ny = on[3 * i + 1], nz = on[3 * i + 2];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        vw[3 * i + r] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(T[0 + r], x), __fmul_rn(T[4 + r], y)), __fmul_rn(T[8 + r], z)), T[12 + r]);
        nw[3 * i + r] = __fadd_rn(__fadd_rn(__fmul_rn(T[0 + r], nx), __fmul_rn(T[4 + r], ny)), __fmul_rn(T[8 + r], nz));
    }
}
__global__ void mesh_triangles_kernel(const float* __restrict__ vw, const uint32_t* __restrict__ idx, const MeshPrep* __restrict__ prep, uint32_t n_meshes,
                                      uint32_t n_tris, float* __restrict__ tri9, uint2* __restrict__ meta) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tris) return;
    uint32_t m = 0;
    while (m + 1 < n_meshes && t >= prep[m + 1].tri_off) ++m;
    const uint32_t f = t - prep[m].tri_off;
    const uint32_t* ix = idx + prep[m].i_off + 3u * f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float* v = vw + 3u * (prep[m].v_off + ix[c]);
        tri9[9 * (size_t)t + 3 * c] = v[0]; tri9[9 * (size_t)t + 3 * c + 1] = v[1]; tri9[9 * (size_t)t + 3 * c + 2] = v[2];
    }
    meta[t] = make_uint2(m, f);
}
int mesh_prepare(const float* d_ov, const float* d_on, const uint32_t* d_idx, const MeshPrep* d_prep, uint32_t n_meshes, uint32_t n_verts,
                 uint32_t n_tris, float* d_vw, float* d_nw, float* d_tri9, uint2* d_meta, cudaStream_t stream, unsigned long long* launches) {
    if (n_verts) { mesh_vertices_kernel<<<cdiv(n_verts, 128), 128, 0, stream>>>(d_ov, d_on, d_prep, n_meshes, n_verts, d_vw, d_nw); ++*launches; }
    if (n_tris) { mesh_triangles_kernel<<<cdiv(n_tris, 128), 128, 0, stream>>>(d_vw, d_idx, d_prep, n_meshes, n_tris, d_tri9, d_meta); ++*launches; }
    return (int)cudaGetLastError();
}

int mesh_compute_boxes(const float* d_tri9, uint32_t n, float4* d_lo, float4* d_hi, cudaStream_t stream, unsigned long long* launches) {
    if (!n) return 0;
    tri_boxes_kernel<<<cdiv(n, 128), 128, 0, stream>>>(d_tri9, n, d_lo, d_hi);
    ++*launches;
    return (int)cudaGetLastError();
}
int mesh_gather(const float* d_tri9, const uint2* d_meta, const uint32_t* d_perm, uint32_t n, TriRec* out, cudaStream_t stream,
                unsigned long long* launches) {
    if (!n) return 0;
    tri_gather_kernel<<<cdiv(n, 128), 128, 0, stream>>>(d_tri9, d_meta, d_perm, n, out);
    ++*launches;
    return (int)cudaGetLastError();
}

}  // namespace grt
