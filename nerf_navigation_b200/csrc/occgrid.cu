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

// ------------------------------------------------------------------------------------------------ partial refresh
// update_extra_state after its 16th call (renderer.py:488-517): per cascade, N = G^3 / 4 uniformly random cells plus N cells
// drawn (with replacement) from the currently occupied ones (grid > 0) are re-queried; the others keep their value.
//   k_occ_count / k_occ_scan / k_occ_compact: ordered list of the occupied cells (block counts, one-block exclusive scan,
//       ordered write: the list — hence the draw — is reproducible run to run, unlike an atomics-based compaction);
//   k_occ_partial_select: cell index + jittered position of the 2N queries from a counter-based RNG;
//   [the fused field query evaluates sigma at the positions]
//   k_occ_scatter_max: tmp[cell] = max over the queries that hit the cell (the reference's index_put keeps an arbitrary
//       one of the duplicates; the maximum is one of them, and deterministic), then k_occ_update as for the full sweep.
__global__ void __launch_bounds__(OCC_THREADS)
k_occ_count(const float *__restrict__ grid, uint32_t n, uint32_t *__restrict__ block_counts) {
    const uint32_t i = blockIdx.x * OCC_THREADS + threadIdx.x;
    const int c = __syncthreads_count(i < n && grid[i] > 0.0f);
    if (threadIdx.x == 0) block_counts[blockIdx.x] = (uint32_t)c;
}

// exclusive scan of nb block counts in place (one block), total -> *total
__global__ void __launch_bounds__(1024)
k_occ_scan(uint32_t *__restrict__ counts, uint32_t nb, uint32_t *__restrict__ total) {
    __shared__ uint32_t part[1024];
    const uint32_t per = (nb + 1023u) / 1024u, lo = threadIdx.x * per, hi = min(nb, lo + per);
    uint32_t sum = 0;
    for (uint32_t i = lo; i < hi; ++i) sum += counts[i];
    part[threadIdx.x] = sum;
    __syncthreads();
    for (uint32_t o = 1; o < 1024; o <<= 1) {
        const uint32_t v = threadIdx.x >= o ? part[threadIdx.x - o] : 0u;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    uint32_t run = part[threadIdx.x] - sum;
    for (uint32_t i = lo; i < hi; ++i) {
        const uint32_t c = counts[i];
        counts[i] = run;
        run += c;
    }
    if (threadIdx.x == 1023) *total = part[1023];
}

__global__ void __launch_bounds__(OCC_THREADS)
k_occ_compact(const float *__restrict__ grid, uint32_t n, const uint32_t *__restrict__ block_offsets,
              uint32_t *__restrict__ list) {
    __shared__ uint32_t warp_base[OCC_THREADS / 32];
    const uint32_t i = blockIdx.x * OCC_THREADS + threadIdx.x;
    const bool occ = i < n && grid[i] > 0.0f;
    const unsigned m = __ballot_sync(0xffffffffu, occ);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) warp_base[warp] = (uint32_t)__popc(m);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t run = block_offsets[blockIdx.x];
        for (int w = 0; w < OCC_THREADS / 32; ++w) {
            const uint32_t c = warp_base[w];
            warp_base[w] = run;
            run += c;
        }
    }
    __syncthreads();
    if (occ) list[warp_base[warp] + __popc(m & ((1u << lane) - 1u))] = i;
}

__global__ void __launch_bounds__(OCC_THREADS)
k_occ_partial_select(int cascade, uint32_t G, float bound, float jitter, uint32_t seed, uint32_t N,
                     const uint32_t *__restrict__ occ_list, const uint32_t *__restrict__ n_occ_ptr,
                     int32_t *__restrict__ cell_idx, float *__restrict__ xyz) {
    const uint32_t q = blockIdx.x * OCC_THREADS + threadIdx.x;
    if (q >= 2u * N) return;
    const uint32_t n_occ = *n_occ_ptr;
    uint32_t h = occ_hash(seed ^ (q * 0x9e3779b9u + (uint32_t)cascade));
    uint32_t idx;
    if (q < N) {                                  // torch.randint(0, G, (N, 3)) -> morton3D (renderer.py:494-495)
        uint32_t c[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            h = occ_hash(h + 0x632be5abu + (uint32_t)a);
            c[a] = (uint32_t)(((uint64_t)h * G) >> 32);
        }
        idx = ngp_morton3D(c[0], c[1], c[2]);
    } else if (n_occ > 0u) {                      // occ_indices[randint(0, n_occ, [N])] (:497-499)
        h = occ_hash(h + 0x85ebca6bu);
        idx = occ_list[(uint32_t)(((uint64_t)h * n_occ) >> 32)];
    } else {                                      // nothing occupied (the reference's randint(0, 0) raises): no query
        cell_idx[q] = -1;
        xyz[(size_t)q * 3] = xyz[(size_t)q * 3 + 1] = xyz[(size_t)q * 3 + 2] = 0.0f;
        return;
    }
    cell_idx[q] = (int32_t)idx;
    const uint32_t c[3] = {ngp_morton_compact(idx), ngp_morton_compact(idx >> 1), ngp_morton_compact(idx >> 2)};
    const float b = fminf((float)(1 << cascade), bound);
    const float half = b / (float)G;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float u = __fmul_rn(2.0f * (float)c[a], __frcp_rn((float)(G - 1))) - 1.0f;     // :505
        h = occ_hash(h + 0xc2b2ae35u + (uint32_t)a);
        const float r = (float)(h >> 8) * (1.0f / 8388608.0f) - 1.0f;
        xyz[(size_t)q * 3 + a] = fmaf(jitter * r, half, u * (b - half));                     // :509-511
    }
}

__global__ void __launch_bounds__(OCC_THREADS)
k_occ_fill(float *__restrict__ p, int64_t n, float v) {
    const int64_t i = (int64_t)blockIdx.x * OCC_THREADS + threadIdx.x;
    if (i < n) p[i] = v;
}

__global__ void __launch_bounds__(OCC_THREADS)
k_occ_scatter_max(float *__restrict__ tmp, const int32_t *__restrict__ cell_idx, const float *__restrict__ sigma, uint32_t nq,
                  float density_scale) {
    const uint32_t q = blockIdx.x * OCC_THREADS + threadIdx.x;
    if (q >= nq) return;
    const int32_t c = cell_idx[q];
    if (c < 0) return;
    const float s = sigma[q] * density_scale;            // :513-514
    if (!(s >= 0.0f)) return;                            // a negative / NaN density never enters the grid (:524)
    atomicMax(reinterpret_cast<int *>(tmp) + c, __float_as_int(s));   // non-negative floats order like their bit patterns
}

// ------------------------------------------------------------------------------------------------ mark_untrained_grid
// renderer.py:381-444: a cell no training camera sees is marked -1 (never refreshed, never occupied).  One thread per
// (cascade, cell in Morton order) walks the poses (cam2world [B,4,4]): cam = (cell centre - t) R, visible iff z > 0 and
// |x| < cx / fx * z + 2 half, |y| < cy / fy * z + 2 half.
__global__ void __launch_bounds__(OCC_THREADS)
k_occ_mark_untrained(const float *__restrict__ poses, int B, float sx, float sy, uint32_t C, uint32_t G, float bound,
                     float *__restrict__ grid) {
    const uint32_t n = G * G * G;
    const uint32_t i = blockIdx.x * OCC_THREADS + threadIdx.x;
    if (i >= n * C) return;
    const uint32_t cas = i / n, idx = i - cas * n;
    const uint32_t c[3] = {ngp_morton_compact(idx), ngp_morton_compact(idx >> 1), ngp_morton_compact(idx >> 2)};
    const float b = fminf((float)(1 << cas), bound);
    const float half = b / (float)G;
    float w[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) w[a] = __fmul_rn(__fmul_rn(2.0f * (float)c[a], __frcp_rn((float)(G - 1))) - 1.0f, b - half);   // :413, 420
    const float h2 = half * 2.0f;
    bool seen = false;
    for (int p = 0; p < B && !seen; ++p) {
        const float *P = poses + (size_t)p * 16;
        const float d0 = w[0] - __ldg(P + 3), d1 = w[1] - __ldg(P + 7), d2 = w[2] - __ldg(P + 11);          // :428
        const float x = fmaf(d2, __ldg(P + 8), fmaf(d1, __ldg(P + 4), d0 * __ldg(P + 0)));                  // :429 (row vector @ R)
        const float y = fmaf(d2, __ldg(P + 9), fmaf(d1, __ldg(P + 5), d0 * __ldg(P + 1)));
        const float z = fmaf(d2, __ldg(P + 10), fmaf(d1, __ldg(P + 6), d0 * __ldg(P + 2)));
        seen = z > 0.0f && fabsf(x) < __fadd_rn(__fmul_rn(sx, z), h2) && fabsf(y) < __fadd_rn(__fmul_rn(sy, z), h2);   // :432-435
    }
    if (!seen) grid[i] = -1.0f;                                                                             // :442
}

// ------------------------------------------------------------------------------------------------ A* occupancy
// nav/quad_plot.py:64-83: density on a side^3 lattice over [lo, hi]^3, MaxPool3d(k), > thresh.
//   k_lattice_positions: torch.linspace's own arithmetic (start + i step for the first half, end - (n-1-i) step for the
//       second) in meshgrid 'ij' order;  [fused field query]  k_pool_thresh: max over each k^3 block, thresholded.
__global__ void __launch_bounds__(OCC_THREADS)
k_lattice_positions(float lo0, float lo1, float lo2, float hi0, float hi1, float hi2, uint32_t side, float *__restrict__ xyz) {
    const uint32_t i = blockIdx.x * OCC_THREADS + threadIdx.x;
    if (i >= side * side * side) return;
    const uint32_t c[3] = {i / (side * side), (i / side) % side, i % side};
    const float lo[3] = {lo0, lo1, lo2}, hi[3] = {hi0, hi1, hi2};
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float step = __fdiv_rn(hi[a] - lo[a], (float)(side - 1));
        xyz[(size_t)i * 3 + a] = c[a] < side / 2 ? fmaf(step, (float)c[a], lo[a]) : hi[a] - step * (float)(side - 1 - c[a]);
    }
}

__global__ void __launch_bounds__(OCC_THREADS)
k_pool_thresh(const float *__restrict__ sigma, uint32_t side, uint32_t k, float thresh, uint8_t *__restrict__ out) {
    const uint32_t g = side / k;
    const uint32_t o = blockIdx.x * OCC_THREADS + threadIdx.x;
    if (o >= g * g * g) return;
    const uint32_t c[3] = {o / (g * g), (o / g) % g, o % g};
    float m = -3.402823466e+38f;
    for (uint32_t a = 0; a < k; ++a)
        for (uint32_t b = 0; b < k; ++b)
            for (uint32_t d = 0; d < k; ++d)
                m = fmaxf(m, __ldg(sigma + ((size_t)(c[0] * k + a) * side + (c[1] * k + b)) * side + (c[2] * k + d)));
    out[o] = m > thresh ? 1 : 0;
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

extern "C" int64_t ngp_occ_grid_partial_workspace_bytes(uint32_t G) {
    const int64_t n = (int64_t)G * G * G;
    return (ngp_div_up(n, OCC_THREADS) + 4 + n) * 4;   // block counts, total, occupied-cell list
}

extern "C" int ngp_occ_grid_partial_select(const float *grid, int cascade, uint32_t G, float bound, float jitter, uint32_t seed,
                                           uint32_t N, void *workspace, int32_t *cell_idx, float *xyz, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(grid, "ngp_occ_grid_partial_select");
    NGP_REQUIRE(grid && workspace && cell_idx && xyz, NGP_ERR_ARG, "ngp_occ_grid_partial_select: null pointer");
    NGP_REQUIRE(G >= 2 && G <= 1024 && (G & (G - 1)) == 0 && cascade >= 0 && cascade < 8 && bound > 0.0f && N >= 1, NGP_ERR_ARG,
                "ngp_occ_grid_partial_select: G must be a power of two in [2, 1024], cascade in [0, 8), N >= 1");
    const uint32_t n = G * G * G, nb = (n + OCC_THREADS - 1) / OCC_THREADS;
    uint32_t *counts = (uint32_t *)workspace, *total = counts + nb, *list = counts + nb + 4;
    cudaStream_t st = (cudaStream_t)stream;
    k_occ_count<<<nb, OCC_THREADS, 0, st>>>(grid, n, counts);
    k_occ_scan<<<1, 1024, 0, st>>>(counts, nb, total);
    k_occ_compact<<<nb, OCC_THREADS, 0, st>>>(grid, n, counts, list);
    k_occ_partial_select<<<(2 * N + OCC_THREADS - 1) / OCC_THREADS, OCC_THREADS, 0, st>>>(cascade, G, bound, jitter, seed, N, list,
                                                                                         total, cell_idx, xyz);
    NGP_CHECK_LAUNCH("ngp_occ_grid_partial_select");
    return NGP_OK;
}

extern "C" int ngp_occ_grid_scatter_update(float *grid, const int32_t *cell_idx, const float *sigma, uint32_t nq, int64_t n,
                                           float decay, float density_scale, float *tmp, float *block_sums,
                                           ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(grid, "ngp_occ_grid_scatter_update");
    NGP_REQUIRE(grid && cell_idx && sigma && tmp && block_sums, NGP_ERR_ARG, "ngp_occ_grid_scatter_update: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    k_occ_fill<<<(unsigned)ngp_div_up(n, OCC_THREADS), OCC_THREADS, 0, st>>>(tmp, n, -1.0f);
    if (nq) k_occ_scatter_max<<<(nq + OCC_THREADS - 1) / OCC_THREADS, OCC_THREADS, 0, st>>>(tmp, cell_idx, sigma, nq, density_scale);
    k_occ_update<<<(unsigned)ngp_div_up(n, OCC_THREADS), OCC_THREADS, 0, st>>>(grid, tmp, n, decay, 1.0f, block_sums);
    NGP_CHECK_LAUNCH("ngp_occ_grid_scatter_update");
    return NGP_OK;
}

extern "C" int ngp_occ_grid_mark_untrained(const float *poses, int B, float fx, float fy, float cx, float cy, uint32_t C,
                                           uint32_t G, float bound, float *grid, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(grid, "ngp_occ_grid_mark_untrained");
    NGP_REQUIRE(poses && grid && B >= 0, NGP_ERR_ARG, "ngp_occ_grid_mark_untrained: null pointer / bad B");
    NGP_REQUIRE(G >= 2 && G <= 1024 && (G & (G - 1)) == 0 && C >= 1 && C <= 8 && bound > 0.0f, NGP_ERR_ARG,
                "ngp_occ_grid_mark_untrained: G must be a power of two in [2, 1024], C in [1, 8]");
    const uint32_t n = G * G * G * C;
    // cx / fx and cy / fy are python floats in the reference (double division), rounded when they meet the fp32 tensor
    const float sx = (float)((double)cx / (double)fx), sy = (float)((double)cy / (double)fy);
    k_occ_mark_untrained<<<(n + OCC_THREADS - 1) / OCC_THREADS, OCC_THREADS, 0, (cudaStream_t)stream>>>(poses, B, sx, sy, C, G,
                                                                                                        bound, grid);
    NGP_CHECK_LAUNCH("ngp_occ_grid_mark_untrained");
    return NGP_OK;
}

extern "C" int ngp_lattice_positions(const float *lo_host, const float *hi_host, uint32_t side, float *xyz, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(xyz, "ngp_lattice_positions");
    NGP_REQUIRE(lo_host && hi_host && xyz && side >= 2 && side <= 1024, NGP_ERR_ARG, "ngp_lattice_positions: bad arguments");
    const uint32_t n = side * side * side;
    k_lattice_positions<<<(n + OCC_THREADS - 1) / OCC_THREADS, OCC_THREADS, 0, (cudaStream_t)stream>>>(
        lo_host[0], lo_host[1], lo_host[2], hi_host[0], hi_host[1], hi_host[2], side, xyz);
    NGP_CHECK_LAUNCH("ngp_lattice_positions");
    return NGP_OK;
}

extern "C" int ngp_maxpool_threshold(const float *sigma, uint32_t side, uint32_t k, float thresh, uint8_t *out,
                                     ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(sigma, "ngp_maxpool_threshold");
    NGP_REQUIRE(sigma && out && k >= 1 && side >= k, NGP_ERR_ARG, "ngp_maxpool_threshold: bad arguments");
    const uint32_t g = side / k, n = g * g * g;
    k_pool_thresh<<<(n + OCC_THREADS - 1) / OCC_THREADS, OCC_THREADS, 0, (cudaStream_t)stream>>>(sigma, side, k, thresh, out);
    NGP_CHECK_LAUNCH("ngp_maxpool_threshold");
    return NGP_OK;
}
