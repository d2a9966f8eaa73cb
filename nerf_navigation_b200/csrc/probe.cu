// probe.cu — L2-gather roofline probe ("the builder must measure L2_gather_peak on the box").
// Independent pseudo-random loads of 4 / 8 / 16 bytes from a resident buffer (the hash table's size), all SMs, many
// loads in flight per thread.  bench.py times it with CUDA events and reports useful bytes/s — the denominator the
// field-query kernels' gather rate is compared with.  Not on the product path.
#include "common.cuh"

template <typename T>
__device__ __forceinline__ uint32_t probe_fold(T v);
template <>
__device__ __forceinline__ uint32_t probe_fold<uint32_t>(uint32_t v) { return v; }
template <>
__device__ __forceinline__ uint32_t probe_fold<uint2>(uint2 v) { return v.x ^ v.y; }
template <>
__device__ __forceinline__ uint32_t probe_fold<uint4>(uint4 v) { return v.x ^ v.y ^ v.z ^ v.w; }

// pattern 0: every lane independent (uniform-random points, hashed levels)
// pattern 1: the 8 loads of one group are the 8 corners of a hashed cell (x^y*p1^z*p2), i.e. 4 aligned-or-not pairs
template <typename T, int UNROLL>
__global__ void __launch_bounds__(256)
k_probe_gather(const T *__restrict__ buf, uint32_t rows, int iters, int pattern, uint32_t *__restrict__ sink) {
    uint32_t state = (blockIdx.x * blockDim.x + threadIdx.x) * 747796405u + 2891336453u;
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it) {
        uint32_t idx[UNROLL];
        if (pattern == 0) {
#pragma unroll
            for (int k = 0; k < UNROLL; ++k) {
                state = state * 747796405u + 2891336453u;
                uint32_t h = ((state >> ((state >> 28u) + 4u)) ^ state) * 277803737u;
                idx[k] = __umulhi((h >> 22u) ^ h, rows);
            }
        } else {
            state = state * 747796405u + 2891336453u;
            const uint32_t h = ((state >> ((state >> 28u) + 4u)) ^ state) * 277803737u;
            const uint32_t cx = h & 1023u, cy = (h >> 10) & 1023u, cz = (h >> 20) & 1023u;
#pragma unroll
            for (int k = 0; k < UNROLL; ++k) {
                const uint32_t x = cx + (k & 1), y = cy + ((k >> 1) & 1), z = cz + ((k >> 2) & 1);
                idx[k] = (x ^ (y * 2654435761u) ^ (z * 805459861u)) % rows;
            }
        }
#pragma unroll
        for (int k = 0; k < UNROLL; ++k) acc ^= probe_fold<T>(__ldg(buf + idx[k]));
    }
    if (acc == 0x9e3779b9u) sink[0] = acc;  // practically never; keeps the loads alive
}

extern "C" int ngp_probe_gather(const void *buf, int64_t buf_bytes, int elem_bytes, int iters, int pattern,
                                int blocks, uint32_t *sink, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(buf, "ngp_probe_gather");
    NGP_REQUIRE(buf && sink && buf_bytes > 0 && iters > 0 && blocks > 0, NGP_ERR_ARG, "ngp_probe_gather: bad argument");
    const uint32_t rows = (uint32_t)(buf_bytes / elem_bytes);
    cudaStream_t st = (cudaStream_t)stream;
    switch (elem_bytes) {
        case 4: k_probe_gather<uint32_t, 8><<<blocks, 256, 0, st>>>((const uint32_t *)buf, rows, iters, pattern, sink); break;
        case 8: k_probe_gather<uint2, 8><<<blocks, 256, 0, st>>>((const uint2 *)buf, rows, iters, pattern, sink); break;
        case 16: k_probe_gather<uint4, 8><<<blocks, 256, 0, st>>>((const uint4 *)buf, rows, iters, pattern, sink); break;
        default: NGP_REQUIRE(false, NGP_ERR_UNSUPPORTED, "ngp_probe_gather: elem_bytes must be 4, 8 or 16");
    }
    NGP_CHECK_LAUNCH("ngp_probe_gather");
    return NGP_OK;
}
