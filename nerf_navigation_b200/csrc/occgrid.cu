// occgrid.cu — occupancy-grid maintenance for the cuda_ray paths.
// Reference: NeRFRenderer.update_extra_state, full-sweep branch (nerf/renderer.py:457-486) + the EMA-max / mean / packbits
// tail (:519-530).  The reference builds meshgrids, Morton indices and jittered positions with a dozen torch ops per
// cascade and scatters through an index tensor; here
//   k_occ_positions: one thread per cell IN MORTON ORDER writes the jittered query position of (cascade, cell)
//                    (cell coordinates = morton3D_invert(index), so sigma comes back already in grid order: no scatter),
//   [the fused field query, field.cu, evaluates sigma at those positions]
//   k_occ_update:    grid = max(grid * decay, sigma * density_scale) where grid >= 0 (cells marked -1 stay untrained,
//                    renderer.py:519-520), with per-block partial sums of max(grid, 0) for the mean density.
// Jitter: counter-based (PCG hash of seed, cell) uniform in [-1, 1) half cells; the reference draws torch.rand_like, which
// no other implementation can reproduce, so only the distribution is kept.  jitter = 0 gives the cell centres exactly.
#include "common.cuh"
#include "march.cuh"

#define OCC_THREADS 256

__device__ __forceinline__ uint32_t occ_hash(uint32_t v) {   // PCG-RXS-M-XS 32
    uint32_t s = v * 747796405u + 2891336453u;
    uint32_t w = ((s >> ((s >> 28u) + 4u)) ^ s) * 277803737u;
    return (w >> 22u) ^ w;
}

__global__ void __launch_bounds__(OCC_THREADS)
k_occ_positions(int cascade, uint32_t G, float bound, float jitter, uint32_t seed, float *__restrict__ xyz) {
    const uint32_t idx = blockIdx.x * OCC_THREADS + threadIdx.x;
    const uint32_t n = G * G * G;
    if (idx >= n) return;
    const uint32_t c[3] = {ngp_morton_compact(idx), ngp_morton_compact(idx >> 1), ngp_morton_compact(idx >> 2)};
    const float b = fminf((float)(1 << cascade), bound);       // renderer.py:474
    const float half = b / (float)G;                            // :475
    uint32_t h = occ_hash(seed ^ (idx * 3u + (uint32_t)cascade * 0x9e3779b9u));
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        // :470 `2 * coords.float() / (G - 1) - 1`: torch divides a tensor by a scalar as a product with its reciprocal
        const float u = __fmul_rn(2.0f * (float)c[a], __frcp_rn((float)(G - 1))) - 1.0f;
        h = occ_hash(h + (uint32_t)a);
        const float r = (float)(h >> 8) * (1.0f / 8388608.0f) - 1.0f;                  // [-1, 1)
        xyz[(size_t)idx * 3 + a] = fmaf(jitter * r, half, u * (b - half));              // :477-479
    }
}

__global__ void __launch_bounds__(OCC_THREADS)
k_occ_update(float *__restrict__ grid, const float *__restrict__ sigma, int64_t n, float decay, float density_scale,
             float *__restrict__ block_sums) {
    __shared__ float warp_sums[OCC_THREADS / 32];
    const int64_t i = (int64_t)blockIdx.x * OCC_THREADS + threadIdx.x;
    float v = 0.0f;
    if (i < n) {
        float g = grid[i];
        const float s = sigma[i] * density_scale;                                       // :483
        if (g >= 0.0f && s >= 0.0f) g = fmaxf(g * decay, s);                            // :519-520
        grid[i] = g;
        v = fmaxf(g, 0.0f);                                                             // :521 clamp(min=0)
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f;
        for (int w = 0; w < OCC_THREADS / 32; ++w) t += warp_sums[w];
        block_sums[blockIdx.x] = t;
    }
}

extern "C" int ngp_occ_grid_positions(int cascade, uint32_t G, float bound, float jitter, uint32_t seed, float *xyz,
                                      ngp_stream_t stream) {
    NGP_REQUIRE(xyz, NGP_ERR_ARG, "ngp_occ_grid_positions: null pointer");
    NGP_REQUIRE_DEVICE(xyz, "ngp_occ_grid_positions");
    NGP_REQUIRE(G >= 2 && G <= 1024 && (G & (G - 1)) == 0 && cascade >= 0 && cascade < 8 && bound > 0.0f, NGP_ERR_ARG,
                "ngp_occ_grid_positions: G must be a power of two in [2, 1024], cascade in [0, 8)");
    const uint32_t n = G * G * G;
    k_occ_positions<<<(n + OCC_THREADS - 1) / OCC_THREADS, OCC_THREADS, 0, (cudaStream_t)stream>>>(cascade, G, bound, jitter,
                                                                                                   seed, xyz);
    NGP_CHECK_LAUNCH("ngp_occ_grid_positions");
    return NGP_OK;
}

extern "C" int64_t ngp_occ_grid_update_blocks(int64_t n) { return ngp_div_up(n, OCC_THREADS); }

extern "C" int ngp_occ_grid_update(float *grid, const float *sigma, int64_t n, float decay, float density_scale,
                                   float *block_sums, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(grid, "ngp_occ_grid_update");
    if (n == 0) return NGP_OK;
    NGP_REQUIRE(grid && sigma && block_sums, NGP_ERR_ARG, "ngp_occ_grid_update: null pointer");
    k_occ_update<<<(unsigned)ngp_div_up(n, OCC_THREADS), OCC_THREADS, 0, (cudaStream_t)stream>>>(grid, sigma, n, decay,
                                                                                                 density_scale, block_sums);
    NGP_CHECK_LAUNCH("ngp_occ_grid_update");
    return NGP_OK;
}
