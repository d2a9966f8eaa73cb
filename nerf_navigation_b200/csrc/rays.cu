// rays.cu — ray ops behind raymarching.*: AABB near/far, Morton codes, occupancy bit packing, occupancy-grid ray
// marching (training: count -> scan -> write, deterministic; inference: alive-ray stepping) and front-to-back
// compositing forward/backward.  Reference semantics: raymarching/src/raymarching.cu (line ranges cited per kernel).
// Unlike the reference (atomicAdd slot reservation, raymarching.cu:409-410) sample offsets are an exclusive scan of the
// per-ray counts in ray order, so `rays` rows and offsets are deterministic.
#include <float.h>

#include "common.cuh"
#include "march.cuh"

#define RAY_THREADS 128

// ---------------------------------------------------------------- near/far  (raymarching.cu:93-147)
__global__ void __launch_bounds__(RAY_THREADS)
k_near_far(const float *__restrict__ rays_o, const float *__restrict__ rays_d, const float *__restrict__ aabb, int64_t N,
           float min_near, float *__restrict__ nears, float *__restrict__ fars) {
    const int64_t n = (int64_t)blockIdx.x * RAY_THREADS + threadIdx.x;
    if (n >= N) return;
    const float o[3] = {rays_o[n * 3], rays_o[n * 3 + 1], rays_o[n * 3 + 2]};
    const float d[3] = {rays_d[n * 3], rays_d[n * 3 + 1], rays_d[n * 3 + 2]};
    float near, far;
    ngp_near_far_ray(o, d, aabb, min_near, near, far);
    nears[n] = near;
    fars[n] = far;
}

extern "C" int ngp_near_far_from_aabb(const float *rays_o, const float *rays_d, const float *aabb, int64_t N,
                                      float min_near, float *nears, float *fars, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(rays_o, "ngp_near_far_from_aabb");
    if (N == 0) return NGP_OK;
    NGP_REQUIRE(rays_o && rays_d && aabb && nears && fars, NGP_ERR_ARG, "ngp_near_far_from_aabb: null pointer");
    k_near_far<<<(unsigned)ngp_div_up(N, RAY_THREADS), RAY_THREADS, 0, (cudaStream_t)stream>>>(rays_o, rays_d, aabb, N,
                                                                                              min_near, nears, fars);
    NGP_CHECK_LAUNCH("ngp_near_far_from_aabb");
    return NGP_OK;
}

// ---------------------------------------------------------------- morton  (raymarching.cu:216-262)
__global__ void k_morton3D(const int32_t *__restrict__ coords, int64_t N, int32_t *__restrict__ indices) {
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    indices[n] = (int32_t)ngp_morton3D((uint32_t)coords[n * 3], (uint32_t)coords[n * 3 + 1], (uint32_t)coords[n * 3 + 2]);
}
__global__ void k_morton3D_invert(const int32_t *__restrict__ indices, int64_t N, int32_t *__restrict__ coords) {
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const uint32_t v = (uint32_t)indices[n];
    coords[n * 3] = (int32_t)ngp_morton_compact(v);
    coords[n * 3 + 1] = (int32_t)ngp_morton_compact(v >> 1);
    coords[n * 3 + 2] = (int32_t)ngp_morton_compact(v >> 2);
}
extern "C" int ngp_morton3D(const int32_t *coords, int64_t N, int32_t *indices, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(coords, "ngp_morton3D");
    if (N == 0) return NGP_OK;
    NGP_REQUIRE(coords && indices, NGP_ERR_ARG, "ngp_morton3D: null pointer");
    k_morton3D<<<(unsigned)ngp_div_up(N, 256), 256, 0, (cudaStream_t)stream>>>(coords, N, indices);
    NGP_CHECK_LAUNCH("ngp_morton3D");
    return NGP_OK;
}
extern "C" int ngp_morton3D_invert(const int32_t *indices, int64_t N, int32_t *coords, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(indices, "ngp_morton3D_invert");
    if (N == 0) return NGP_OK;
    NGP_REQUIRE(coords && indices, NGP_ERR_ARG, "ngp_morton3D_invert: null pointer");
    k_morton3D_invert<<<(unsigned)ngp_div_up(N, 256), 256, 0, (cudaStream_t)stream>>>(indices, N, coords);
    NGP_CHECK_LAUNCH("ngp_morton3D_invert");
    return NGP_OK;
}

// ---------------------------------------------------------------- packbits  (raymarching.cu:269-291)
// one thread per output byte; its 8 floats are two aligned float4 loads (32 B = one sector), stores coalesce.
__global__ void k_packbits(const float *__restrict__ grid, int64_t N, float thresh, uint8_t *__restrict__ bitfield) {
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const float4 a = __ldg(reinterpret_cast<const float4 *>(grid) + n * 2);
    const float4 b = __ldg(reinterpret_cast<const float4 *>(grid) + n * 2 + 1);
    uint32_t bits = 0;
    bits |= (a.x > thresh) ? 1u : 0u;
    bits |= (a.y > thresh) ? 2u : 0u;
    bits |= (a.z > thresh) ? 4u : 0u;
    bits |= (a.w > thresh) ? 8u : 0u;
    bits |= (b.x > thresh) ? 16u : 0u;
    bits |= (b.y > thresh) ? 32u : 0u;
    bits |= (b.z > thresh) ? 64u : 0u;
    bits |= (b.w > thresh) ? 128u : 0u;
    bitfield[n] = (uint8_t)bits;
}
extern "C" int ngp_packbits(const float *grid, int64_t N, float thresh, uint8_t *bitfield, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(grid, "ngp_packbits");
    if (N == 0) return NGP_OK;
    NGP_REQUIRE(grid && bitfield, NGP_ERR_ARG, "ngp_packbits: null pointer");
    NGP_REQUIRE(((uintptr_t)grid & 15) == 0, NGP_ERR_ARG, "ngp_packbits: grid must be 16-byte aligned");
    k_packbits<<<(unsigned)ngp_div_up(N, 256), 256, 0, (cudaStream_t)stream>>>(grid, N, thresh, bitfield);
    NGP_CHECK_LAUNCH("ngp_packbits");
    return NGP_OK;
}

// ---------------------------------------------------------------- training marcher  (raymarching.cu:313-484)
__device__ __forceinline__ MarchRay load_ray(const float *__restrict__ rays_o, const float *__restrict__ rays_d, int64_t n) {
    MarchRay r;
    r.ox = rays_o[n * 3]; r.oy = rays_o[n * 3 + 1]; r.oz = rays_o[n * 3 + 2];
    r.dx = rays_d[n * 3]; r.dy = rays_d[n * 3 + 1]; r.dz = rays_d[n * 3 + 2];
    r.rdx = 1.0f / r.dx; r.rdy = 1.0f / r.dy; r.rdz = 1.0f / r.dz;
    return r;
}

// pass 1: per-ray sample count (and jittered start t0 when perturb).  SEGS: the pass also leaves the ray's RUNS — maximal
// stretches of consecutive samples, (t of the first sample, length) — in segs[run * N + ray] (coalesced across the rays of a
// warp) and their number in nseg[ray].  Inside a run t_{k+1} = t_k + clamp(t_k dt_gamma, dt_min, dt_max) and nothing else,
// so the write pass (k_march_replay) rebuilds the samples from the runs without looking at the occupancy grid again.  A ray
// with more than MARCH_SEG_CAP runs keeps its count in nseg and is marched again by the write pass.
#ifndef MARCH_SEG_CAP
#define MARCH_SEG_CAP 16
#endif
template <bool SEGS>
__global__ void __launch_bounds__(RAY_THREADS)
k_march_count(const float *__restrict__ rays_o, const float *__restrict__ rays_d, MarchCfg cfg, uint32_t max_steps,
              int64_t N, const float *__restrict__ nears, const float *__restrict__ fars, uint32_t perturb,
              int32_t *__restrict__ counts, float *__restrict__ t_first, int32_t *__restrict__ nseg,
              float2 *__restrict__ segs) {
    // the marcher iteration of march.cuh in its cheaper, value-identical form (ngp_march_cell_fast / ngp_march_exit_fast:
    // levels from the exponent bits, min(2^level, bound) and its reciprocal from tables, no Morton decode for the exit)
    __shared__ float mipb[8], miprb[8];
    if (threadIdx.x < 8) {
        mipb[threadIdx.x] = fminf((float)(1 << threadIdx.x), cfg.bound);
        miprb[threadIdx.x] = 1.0f / mipb[threadIdx.x];
    }
    __syncthreads();
    const int64_t n = (int64_t)blockIdx.x * RAY_THREADS + threadIdx.x;
    if (n >= N) return;
    const MarchRay r = load_ray(rays_o, rays_d, n);
    float t = nears[n];
    if (perturb) t = fmaf(cfg.dt_min, ngp_pcg_jitter(42ull, (uint64_t)n), t);  // :352-355, seed :489
    const float far = fars[n];
    uint32_t steps = 0, run_start = 0;
    int32_t runs = 0;
    bool in_run = false;
    float tf = t, run_t = t;
#pragma unroll 1
    while (t < far && steps < max_steps) {
        float dt;
        int lvl;
        uint32_t nxyz;
        const uint32_t cell = ngp_march_cell_fast(r, cfg, miprb, t, dt, lvl, nxyz);
        if ((__ldg(cfg.grid + (cell >> 3)) >> (cell & 7u)) & 1u) {
            if (steps == 0) tf = t;   // the march is memoryless in t: pass 2 resumes here, skipping the leading empty space
            if (SEGS && !in_run) {
                in_run = true;
                run_t = t;
                run_start = steps;
            }
            ++steps;
            t += dt;
        } else {
            if (SEGS && in_run) {
                if (runs < MARCH_SEG_CAP) segs[(int64_t)runs * N + n] = make_float2(run_t, __int_as_float((int)(steps - run_start)));
                ++runs;
                in_run = false;
            }
            const float tt = ngp_march_exit_fast(r, cfg, mipb, t, nxyz, lvl);
            if (cfg.lvl_dt >= 0) {    // raymarching.cu:398-402; dt_gamma == 0: the step is dt_min
                do {
                    t += cfg.dt_min;
                } while (t < tt);
            } else {
                do {
                    t = ngp_march_advance(t, cfg);
                } while (t < tt);
            }
        }
    }
    if (SEGS) {
        if (in_run) {
            if (runs < MARCH_SEG_CAP) segs[(int64_t)runs * N + n] = make_float2(run_t, __int_as_float((int)(steps - run_start)));
            ++runs;
        }
        nseg[n] = runs;
    }
    counts[n] = (int32_t)steps;
    t_first[n] = tf;
}

// exclusive scan of counts -> offsets (3 kernels: per-block totals, scan of totals, per-block scan + base)
#define SCAN_THREADS 256
#define SCAN_ITEMS 8
#define SCAN_TILE (SCAN_THREADS * SCAN_ITEMS)

__device__ __forceinline__ int32_t block_exclusive_scan(int32_t v, int32_t *total) {
    __shared__ int32_t warp_sums[SCAN_THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int32_t u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int32_t w = lane < SCAN_THREADS / 32 ? warp_sums[lane] : 0;
#pragma unroll
        for (int o = 1; o < SCAN_THREADS / 32; o <<= 1) {
            const int32_t u = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += u;
        }
        if (lane < SCAN_THREADS / 32) warp_sums[lane] = w;
    }
    __syncthreads();
    const int32_t base = warp > 0 ? warp_sums[warp - 1] : 0;
    *total = warp_sums[SCAN_THREADS / 32 - 1];
    __syncthreads();
    return base + inc - v;
}

__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_tile_totals(const int32_t *__restrict__ counts, int64_t N, int32_t *__restrict__ tile_totals) {
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    int32_t s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i)
        if (base + i < N) s += counts[base + i];
    int32_t total;
    block_exclusive_scan(s, &total);
    if (threadIdx.x == 0) tile_totals[blockIdx.x] = total;
}

// single block: in-place exclusive scan of tile totals, adds the incoming counter value, bumps the counters
__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_totals(int32_t *__restrict__ tile_totals, int64_t n_tiles, int32_t *__restrict__ counter, int64_t N) {
    __shared__ int32_t carry;
    if (threadIdx.x == 0) carry = counter[0];
    __syncthreads();
    for (int64_t start = 0; start < n_tiles; start += SCAN_THREADS) {
        const int64_t i = start + threadIdx.x;
        const int32_t v = i < n_tiles ? tile_totals[i] : 0;
        int32_t total;
        const int32_t ex = block_exclusive_scan(v, &total);
        const int32_t c = carry;
        if (i < n_tiles) tile_totals[i] = c + ex;
        __syncthreads();
        if (threadIdx.x == 0) carry = c + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        counter[0] = carry;          // total samples (reference: atomicAdd(counter, num_steps))
        counter[1] += (int32_t)N;    // rays processed   (atomicAdd(counter + 1, 1))
    }
}

__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_write_rays(const int32_t *__restrict__ counts, const int32_t *__restrict__ tile_totals, int64_t N,
                  int32_t *__restrict__ rays) {
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    int32_t c[SCAN_ITEMS];
    int32_t s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        c[i] = base + i < N ? counts[base + i] : 0;
        s += c[i];
    }
    int32_t total;
    int32_t off = block_exclusive_scan(s, &total) + tile_totals[blockIdx.x];
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        if (base + i < N) {
            rays[(base + i) * 3] = (int32_t)(base + i);
            rays[(base + i) * 3 + 1] = off;
            rays[(base + i) * 3 + 2] = c[i];
        }
        off += c[i];
    }
}

// pass 2: write samples at the scanned offsets
__global__ void __launch_bounds__(RAY_THREADS)
k_march_write(const float *__restrict__ rays_o, const float *__restrict__ rays_d, MarchCfg cfg, int64_t N, int64_t M,
              const float *__restrict__ nears, const float *__restrict__ fars, uint32_t perturb,
              const int32_t *__restrict__ rays, const float *__restrict__ t_first, float *__restrict__ xyzs,
              float *__restrict__ dirs, float *__restrict__ deltas) {
    const int64_t n = (int64_t)blockIdx.x * RAY_THREADS + threadIdx.x;
    if (n >= N) return;
    const int64_t off = rays[n * 3 + 1];
    const uint32_t cnt = (uint32_t)rays[n * 3 + 2];
    if (cnt == 0 || off + cnt >= M) return;  // :419-420
    const MarchRay r = load_ray(rays_o, rays_d, n);
    float t = nears[n];
    if (perturb) t = fmaf(cfg.dt_min, ngp_pcg_jitter(42ull, (uint64_t)n), t);
    float last_t = t;      // the first delta is measured from the ray's start (:470-473)
    t = t_first[n];        // resume at the first occupied sample found by pass 1
    const float far = fars[n];
    uint32_t step = 0;
    float px, py, pz, dt;
    while (t < far && step < cnt) {
        if (!ngp_march_step(r, cfg, t, px, py, pz, dt)) continue;
        const int64_t k = off + step;
        xyzs[k * 3] = t;   // the sample's parameter: k_march_expand turns it into the position (and writes dirs) coalesced
        t += dt;
        *reinterpret_cast<float2 *>(deltas + k * 2) = make_float2(dt, t - last_t);
        last_t = t;
        ++step;
    }
}

// pass 3 of the thread-per-ray form: one WARP per ray turns the parameters left in xyzs[k][0] into positions and writes the
// directions, lanes = consecutive samples (coalesced).  The thread-per-ray kernel above used to write eight scattered
// floats per sample (5.5 ms for the estimator's 7.9x10^7 samples); now it writes three and this kernel streams the rest.
// Same expression as the marcher's own position (ngp_march_probe): clamp(fma(t, d, o)).
__global__ void __launch_bounds__(256)
k_march_expand(const float *__restrict__ rays_o, const float *__restrict__ rays_d, float bound, int64_t N, int64_t M,
               const int32_t *__restrict__ rays, float *__restrict__ xyzs, float *__restrict__ dirs) {
    const int64_t n = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (n >= N) return;
    const int64_t off = rays[n * 3 + 1];
    const uint32_t cnt = (uint32_t)rays[n * 3 + 2];
    if (cnt == 0 || off + cnt >= M) return;
    const MarchRay r = load_ray(rays_o, rays_d, n);
    for (uint32_t s = lane; s < cnt; s += 32) {
        const int64_t k = off + s;
        const float t = xyzs[k * 3];
        xyzs[k * 3] = ngp_clampf(fmaf(t, r.dx, r.ox), -bound, bound);
        xyzs[k * 3 + 1] = ngp_clampf(fmaf(t, r.dy, r.oy), -bound, bound);
        xyzs[k * 3 + 2] = ngp_clampf(fmaf(t, r.dz, r.oz), -bound, bound);
        dirs[k * 3] = r.dx; dirs[k * 3 + 1] = r.dy; dirs[k * 3 + 2] = r.dz;
    }
}

// Write pass from the runs of k_march_count<true>: one WARP per ray.  Phase 1: lane j replays run j — t_{k+1} = t_k +
// dt(t_k), the marcher's own recurrence, no occupancy lookups — and leaves the parameters t_k in the warp's shared-memory
// strip (in xyzs[k][0] for a ray with more than MARCH_REPLAY_STRIP samples); a ray with more runs than MARCH_SEG_CAP is
// marched again by the whole warp (march_ray_warp, the small-batch write pass).  Phase 2: lanes = consecutive samples, everything else follows
// from t_k alone and is stored coalesced: dt_k = dt(t_k), the position clamp(o + t_k d), the direction, and
// deltas_k = (dt_k, (t_k + dt_k) - (t_{k-1} + dt_{k-1})) with the ray's start for k = 0 — the marcher's t - last_t
// (:470-473), since inside a run t_{k-1} + dt_{k-1} IS t_k.  Same float operations per sample as k_march_write: same bits.
#define MARCH_REPLAY_STRIP 1024
template <bool WRITE>
__device__ __forceinline__ uint32_t march_ray_warp(const MarchRay &r, const MarchCfg &cfg, float t_cur, float far,
                                                   uint32_t limit, float last_t, int64_t off, float *__restrict__ xyzs,
                                                   float *__restrict__ dirs, float *__restrict__ deltas, float &t_first);
__global__ void __launch_bounds__(256)
k_march_replay(const float *__restrict__ rays_o, const float *__restrict__ rays_d, MarchCfg cfg, int64_t N, int64_t M,
               const float *__restrict__ nears, const float *__restrict__ fars, uint32_t perturb,
               const int32_t *__restrict__ rays, const int32_t *__restrict__ nseg, const float2 *__restrict__ segs,
               float *xyzs, float *__restrict__ dirs, float *__restrict__ deltas) {
    __shared__ float strips[8][MARCH_REPLAY_STRIP];
    const int64_t n = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (n >= N) return;
    const int64_t off = rays[n * 3 + 1];
    const uint32_t cnt = (uint32_t)rays[n * 3 + 2];
    if (cnt == 0 || off + cnt >= M) return;  // :419-420
    const MarchRay r = load_ray(rays_o, rays_d, n);
    float t0 = nears[n];
    if (perturb) t0 = fmaf(cfg.dt_min, ngp_pcg_jitter(42ull, (uint64_t)n), t0);
    const bool in_strip = cnt <= MARCH_REPLAY_STRIP;
    float *strip = strips[(threadIdx.x >> 5) & 7];
    const int ns = nseg[n];
    if (ns <= MARCH_SEG_CAP) {
        float t = 0.0f;
        int len = 0;
        if (lane < ns) {
            const float2 sg = __ldg(segs + (int64_t)lane * N + n);
            t = sg.x;
            len = __float_as_int(sg.y);
        }
        int pre = len;   // inclusive scan of the run lengths
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, pre, o);
            if (lane >= o) pre += v;
        }
        const int s0 = pre - len;
        if (in_strip) {
            for (int i = 0; i < len; ++i) {
                strip[s0 + i] = t;
                t += ngp_clampf(t * cfg.dt_gamma, cfg.dt_min, cfg.dt_max);
            }
        } else {
            for (int i = 0; i < len; ++i) {
                xyzs[(off + s0 + i) * 3] = t;
                t += ngp_clampf(t * cfg.dt_gamma, cfg.dt_min, cfg.dt_max);
            }
        }
    } else {   // (warp-uniform) more runs than the workspace holds: the warp marches the ray again, 32 probes at a time, and
        float tf;   // writes the samples itself — the small-batch write pass; run 0 starts at the first occupied sample
        march_ray_warp<true>(r, cfg, __ldg(segs + n).x, fars[n], cnt, t0, off, xyzs, dirs, deltas, tf);
        return;
    }
    __syncwarp();   // the parameters written above are read by other lanes below
    float carry = t0;   // t + dt of the sample before this block of 32 (the ray's start before the first)
    for (uint32_t base = 0; base < cnt; base += 32) {
        const uint32_t s = base + lane;
        const bool in = s < cnt;
        const int64_t k = off + s;
        const float t = !in ? 0.0f : (in_strip ? strip[s] : xyzs[k * 3]);
        const float dt = ngp_clampf(t * cfg.dt_gamma, cfg.dt_min, cfg.dt_max);
        const float tn = t + dt;
        float prev = __shfl_up_sync(0xffffffffu, tn, 1);
        if (lane == 0) prev = carry;
        carry = __shfl_sync(0xffffffffu, tn, 31);
        if (in) {
            xyzs[k * 3] = ngp_clampf(fmaf(t, r.dx, r.ox), -cfg.bound, cfg.bound);
            xyzs[k * 3 + 1] = ngp_clampf(fmaf(t, r.dy, r.oy), -cfg.bound, cfg.bound);
            xyzs[k * 3 + 2] = ngp_clampf(fmaf(t, r.dz, r.oz), -cfg.bound, cfg.bound);
            dirs[k * 3] = r.dx; dirs[k * 3 + 1] = r.dy; dirs[k * 3 + 2] = r.dz;
            *reinterpret_cast<float2 *>(deltas + k * 2) = make_float2(dt, tn - prev);
        }
    }
}

// ---- warp-per-ray variants for small ray batches (training step: 4096 rays; the estimator's real use: 1024 rays).
// With one thread per ray a batch of a few thousand rays is pure latency: < 1 warp per SM walking ~10^3 dependent
// steps.  The parameter sequence t_k of a ray does not depend on the occupancy (march.cuh: ngp_march_advance), so the 32
// lanes of a warp probe 32 consecutive t_k of ONE ray at once and the warp then replays the reference's sequential
// decisions (sample / skip to the voxel exit) on the probe results with ballots and shuffles — same visits, same
// samples, bit for bit.
#ifndef MARCH_WARP_MAX_RAYS
#define MARCH_WARP_MAX_RAYS 8192
#endif

template <bool WRITE>
__device__ __forceinline__ uint32_t march_ray_warp(const MarchRay &r, const MarchCfg &cfg, float t_cur, float far,
                                                   uint32_t limit, float last_t, int64_t off, float *__restrict__ xyzs,
                                                   float *__restrict__ dirs, float *__restrict__ deltas, float &t_first) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    uint32_t steps = 0;
    bool done = false;
    while (!done && t_cur < far && steps < limit) {
        float w = t_cur;                       // lane i: the i-th successor of t_cur
        for (int j = 0; j < lane; ++j) w = ngp_march_advance(w, cfg);
        const float wn = ngp_march_advance(w, cfg);
        float px = 0.f, py = 0.f, pz = 0.f, dt = 0.f, tt = w;
        bool occ = false;
        if (w < far) occ = ngp_march_probe(r, cfg, w, px, py, pz, dt, tt);
        const unsigned occ_mask = __ballot_sync(FULL, occ);
        const float w_window_end = __shfl_sync(FULL, wn, 31);
        float t_next = w_window_end;
        int my_s = -1;
        float my_last = 0.0f;
        int c = 0;                              // lane whose t the reference marcher is at (warp-uniform)
        while (c < 32) {
            const float tw = __shfl_sync(FULL, w, c);
            if (!(tw < far) || steps >= limit) {
                done = true;
                break;
            }
            if ((occ_mask >> c) & 1u) {         // sample here, go to the next t
                if (lane == c) {
                    my_s = (int)steps;
                    my_last = last_t;
                }
                if (steps == 0) t_first = tw;
                last_t = __shfl_sync(FULL, wn, c);
                ++steps;
                ++c;
            } else {                            // empty: first t of the sequence at or past the voxel exit
                const float ttc = __shfl_sync(FULL, tt, c);
                const unsigned ahead = c >= 31 ? 0u : (FULL << (c + 1));
                const unsigned m = __ballot_sync(FULL, w >= ttc) & ahead;
                if (m) {
                    c = __ffs(m) - 1;
                } else {                        // the voxel ends beyond this window
                    float tc = w_window_end;
                    while (tc < ttc) tc = ngp_march_advance(tc, cfg);
                    t_next = tc;
                    break;
                }
            }
        }
        if (WRITE && my_s >= 0) {
            const int64_t k = off + my_s;
            xyzs[k * 3] = px; xyzs[k * 3 + 1] = py; xyzs[k * 3 + 2] = pz;
            dirs[k * 3] = r.dx; dirs[k * 3 + 1] = r.dy; dirs[k * 3 + 2] = r.dz;
            deltas[k * 2] = dt;
            deltas[k * 2 + 1] = wn - my_last;
        }
        t_cur = t_next;
    }
    return steps;
}

__global__ void __launch_bounds__(RAY_THREADS)
k_march_count_warp(const float *__restrict__ rays_o, const float *__restrict__ rays_d, MarchCfg cfg, uint32_t max_steps,
                   int64_t N, const float *__restrict__ nears, const float *__restrict__ fars, uint32_t perturb,
                   int32_t *__restrict__ counts, float *__restrict__ t_first) {
    const int64_t n = ((int64_t)blockIdx.x * RAY_THREADS + threadIdx.x) >> 5;
    if (n >= N) return;
    const MarchRay r = load_ray(rays_o, rays_d, n);
    float t = nears[n];
    if (perturb) t = fmaf(cfg.dt_min, ngp_pcg_jitter(42ull, (uint64_t)n), t);
    float tf = t;
    const uint32_t steps = march_ray_warp<false>(r, cfg, t, fars[n], max_steps, t, 0, nullptr, nullptr, nullptr, tf);
    if ((threadIdx.x & 31) == 0) {
        counts[n] = (int32_t)steps;
        t_first[n] = tf;
    }
}

__global__ void __launch_bounds__(RAY_THREADS)
k_march_write_warp(const float *__restrict__ rays_o, const float *__restrict__ rays_d, MarchCfg cfg, int64_t N, int64_t M,
                   const float *__restrict__ nears, const float *__restrict__ fars, uint32_t perturb,
                   const int32_t *__restrict__ rays, const float *__restrict__ t_first, float *__restrict__ xyzs,
                   float *__restrict__ dirs, float *__restrict__ deltas) {
    const int64_t n = ((int64_t)blockIdx.x * RAY_THREADS + threadIdx.x) >> 5;
    if (n >= N) return;
    const int64_t off = rays[n * 3 + 1];
    const uint32_t cnt = (uint32_t)rays[n * 3 + 2];
    if (cnt == 0 || off + cnt >= M) return;  // :419-420
    const MarchRay r = load_ray(rays_o, rays_d, n);
    float t = nears[n];
    if (perturb) t = fmaf(cfg.dt_min, ngp_pcg_jitter(42ull, (uint64_t)n), t);
    float tf = 0.0f;
    march_ray_warp<true>(r, cfg, t_first[n], fars[n], cnt, t, off, xyzs, dirs, deltas, tf);
}

// workspace: counts [N], tile totals [tiles + 8], t_first [N]; for batches marched one thread per ray also nseg [N] and,
// 8-byte aligned, segs [MARCH_SEG_CAP][N] (only the runs that exist are touched)
static int64_t march_ws_head_words(int64_t N) { return 2 * N + ngp_div_up(N, SCAN_TILE) + 8; }
static int64_t march_ws_segs_offset(int64_t N) { return ((march_ws_head_words(N) + N) * 4 + 7) & ~(int64_t)7; }
// NGP_B200_MARCH_WRITE=remarch: the write pass marches the grid again (k_march_write + k_march_expand, the form before
// the runs existed) — kept as the cross-check of k_march_replay (tests/test_rays_gpu.py compares the two bit for bit)
static bool march_write_remarch() {
    const char *e = getenv("NGP_B200_MARCH_WRITE");
    return e && strcmp(e, "remarch") == 0;
}
extern "C" int64_t ngp_march_rays_train_workspace_bytes(int64_t N) {
    if (N <= MARCH_WARP_MAX_RAYS) return march_ws_head_words(N) * (int64_t)sizeof(int32_t);
    return march_ws_segs_offset(N) + (int64_t)MARCH_SEG_CAP * N * (int64_t)sizeof(float2);
}

extern "C" int ngp_march_rays_train(const float *rays_o, const float *rays_d, const uint8_t *grid, float bound,
                                    float dt_gamma, uint32_t max_steps, int64_t N, uint32_t C, uint32_t H, int64_t M,
                                    const float *nears, const float *fars, float *xyzs, float *dirs, float *deltas,
                                    int32_t *rays, int32_t *counter, uint32_t perturb, void *workspace,
                                    int64_t workspace_bytes, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(rays_o, "ngp_march_rays_train");
    if (N == 0) return NGP_OK;
    NGP_REQUIRE(rays_o && rays_d && grid && nears && fars && rays && counter, NGP_ERR_ARG,
                "ngp_march_rays_train: null pointer");
    NGP_REQUIRE(M == 0 || (xyzs && dirs && deltas), NGP_ERR_ARG, "ngp_march_rays_train: null sample buffers");
    NGP_REQUIRE(((uintptr_t)deltas & 7u) == 0, NGP_ERR_ARG, "ngp_march_rays_train: deltas must be 8-byte aligned");
    NGP_REQUIRE(C >= 1 && C <= 8 && H >= 1 && H <= 1024, NGP_ERR_ARG, "ngp_march_rays_train: bad cascade/grid size");
    NGP_REQUIRE(workspace && workspace_bytes >= ngp_march_rays_train_workspace_bytes(N), NGP_ERR_ARG,
                "ngp_march_rays_train: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    int32_t *counts = (int32_t *)workspace;
    int32_t *tile_totals = counts + N;
    const int64_t n_tiles = ngp_div_up(N, SCAN_TILE);
    float *t_first = (float *)(tile_totals + n_tiles + 8);
    const MarchCfg cfg = ngp_march_cfg(grid, bound, dt_gamma, max_steps, C, H);
    const unsigned gr = (unsigned)ngp_div_up(N, RAY_THREADS);
    const bool by_warp = N <= MARCH_WARP_MAX_RAYS;
    const unsigned gw = (unsigned)ngp_div_up(N * 32, RAY_THREADS);
    int32_t *nseg = (int32_t *)workspace + march_ws_head_words(N);
    float2 *segs = (float2 *)((char *)workspace + march_ws_segs_offset(N));
    if (by_warp)
        k_march_count_warp<<<gw, RAY_THREADS, 0, st>>>(rays_o, rays_d, cfg, max_steps, N, nears, fars, perturb, counts, t_first);
    else
        k_march_count<true><<<gr, RAY_THREADS, 0, st>>>(rays_o, rays_d, cfg, max_steps, N, nears, fars, perturb, counts, t_first,
                                                        nseg, segs);
    k_scan_tile_totals<<<(unsigned)n_tiles, SCAN_THREADS, 0, st>>>(counts, N, tile_totals);
    k_scan_totals<<<1, SCAN_THREADS, 0, st>>>(tile_totals, n_tiles, counter, N);
    k_scan_write_rays<<<(unsigned)n_tiles, SCAN_THREADS, 0, st>>>(counts, tile_totals, N, rays);
    if (M > 0) {
        if (by_warp)
            k_march_write_warp<<<gw, RAY_THREADS, 0, st>>>(rays_o, rays_d, cfg, N, M, nears, fars, perturb, rays, t_first, xyzs, dirs, deltas);
        else if (march_write_remarch()) {
            k_march_write<<<gr, RAY_THREADS, 0, st>>>(rays_o, rays_d, cfg, N, M, nears, fars, perturb, rays, t_first, xyzs, dirs, deltas);
            k_march_expand<<<(unsigned)ngp_div_up(N * 32, 256), 256, 0, st>>>(rays_o, rays_d, bound, N, M, rays, xyzs, dirs);
        } else
            k_march_replay<<<(unsigned)ngp_div_up(N * 32, 256), 256, 0, st>>>(rays_o, rays_d, cfg, N, M, nears, fars, perturb, rays,
                                                                              nseg, segs, xyzs, dirs, deltas);
    }
    NGP_CHECK_LAUNCH("ngp_march_rays_train");
    return NGP_OK;
}

// Second half of a two-call use: after ngp_march_rays_train(..., M = 0, ...) has counted, scanned and filled `rays` /
// `counter` (the host reads counter[0] to size the sample buffers exactly), write the samples without marching the count
// pass again.  `rays` and `workspace` must be the ones of that count call.
extern "C" int ngp_march_rays_train_write(const float *rays_o, const float *rays_d, const uint8_t *grid, float bound,
                                          float dt_gamma, uint32_t max_steps, int64_t N, uint32_t C, uint32_t H, int64_t M,
                                          const float *nears, const float *fars, float *xyzs, float *dirs, float *deltas,
                                          const int32_t *rays, uint32_t perturb, const void *workspace, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(rays_o, "ngp_march_rays_train_write");
    if (N == 0 || M == 0) return NGP_OK;
    NGP_REQUIRE(rays_o && rays_d && grid && nears && fars && rays && xyzs && dirs && deltas && workspace, NGP_ERR_ARG,
                "ngp_march_rays_train_write: null pointer");
    NGP_REQUIRE(C >= 1 && C <= 8 && H >= 1 && H <= 1024, NGP_ERR_ARG, "ngp_march_rays_train_write: bad cascade/grid size");
    NGP_REQUIRE(((uintptr_t)deltas & 7u) == 0, NGP_ERR_ARG, "ngp_march_rays_train_write: deltas must be 8-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    const int32_t *counts = (const int32_t *)workspace;
    const int64_t n_tiles = ngp_div_up(N, SCAN_TILE);
    const float *t_first = (const float *)(counts + N + n_tiles + 8);
    const MarchCfg cfg = ngp_march_cfg(grid, bound, dt_gamma, max_steps, C, H);
    if (N <= MARCH_WARP_MAX_RAYS)
        k_march_write_warp<<<(unsigned)ngp_div_up(N * 32, RAY_THREADS), RAY_THREADS, 0, st>>>(rays_o, rays_d, cfg, N, M, nears, fars,
                                                                                              perturb, rays, t_first, xyzs, dirs, deltas);
    else if (march_write_remarch()) {
        k_march_write<<<(unsigned)ngp_div_up(N, RAY_THREADS), RAY_THREADS, 0, st>>>(rays_o, rays_d, cfg, N, M, nears, fars, perturb,
                                                                                    rays, t_first, xyzs, dirs, deltas);
        k_march_expand<<<(unsigned)ngp_div_up(N * 32, 256), 256, 0, st>>>(rays_o, rays_d, bound, N, M, rays, xyzs, dirs);
    } else
        k_march_replay<<<(unsigned)ngp_div_up(N * 32, 256), 256, 0, st>>>(
            rays_o, rays_d, cfg, N, M, nears, fars, perturb, rays, (const int32_t *)workspace + march_ws_head_words(N),
            (const float2 *)((const char *)workspace + march_ws_segs_offset(N)), xyzs, dirs, deltas);
    NGP_CHECK_LAUNCH("ngp_march_rays_train_write");
    return NGP_OK;
}

// ---------------------------------------------------------------- training marcher, gradient to the rays
// EXTENSION (absent in the reference, whose cuda_ray path returns no pose gradient): the sample
// abscissae t_k are constants w.r.t. the pose, x_k = clamp(o + t_k d, -bound, bound), dirs_k = d, exactly what the torch
// sampler differentiates (renderer.py:159-160):
//   dL/do = sum_k m_k * dL/dx_k,   dL/dd = sum_k t_k (m_k * dL/dx_k) + sum_k dL/ddir_k,   m_k = per-axis "not clamped".
// t_k is rebuilt from the emitted deltas: t_k + dt_k = t_start + cumsum(deltas[:,1])_k.
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// One WARP per ray: lanes = 32 consecutive samples (coalesced loads), the running t is a warp scan of deltas[:, 1], the
// sums over the ray's samples are warp reductions.
__global__ void __launch_bounds__(256)
k_march_train_bwd_warp(const float *__restrict__ grad_xyzs, const float *__restrict__ grad_dirs,
                       const float *__restrict__ rays_o, const float *__restrict__ rays_d, const float *__restrict__ nears,
                       const float *__restrict__ deltas, const int32_t *__restrict__ rays, float bound, float dt_min,
                       uint32_t perturb, int64_t N, int64_t M, float *__restrict__ grad_o, float *__restrict__ grad_d) {
    const int64_t i = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (i >= N) return;
    const int64_t n = rays[i * 3], off = rays[i * 3 + 1];
    const uint32_t cnt = (uint32_t)rays[i * 3 + 2];
    float go[3] = {0.f, 0.f, 0.f}, gd[3] = {0.f, 0.f, 0.f};
    if (cnt != 0 && off + cnt < M) {
        const float o[3] = {rays_o[n * 3], rays_o[n * 3 + 1], rays_o[n * 3 + 2]};
        const float d[3] = {rays_d[n * 3], rays_d[n * 3 + 1], rays_d[n * 3 + 2]};
        float tc = nears[n];
        if (perturb) tc = fmaf(dt_min, ngp_pcg_jitter(42ull, (uint64_t)n), tc);
        for (uint32_t base = 0; base < cnt; base += 32) {
            const uint32_t s = base + lane;
            const bool in = s < cnt;
            const int64_t k = off + s;
            const float2 dl = in ? __ldg(reinterpret_cast<const float2 *>(deltas) + k) : make_float2(0.0f, 0.0f);
            float te = dl.y;   // inclusive scan -> t_k + dt_k - t_start of the block
#pragma unroll
            for (int sh = 1; sh < 32; sh <<= 1) {
                const float v = __shfl_up_sync(0xffffffffu, te, sh);
                if (lane >= sh) te += v;
            }
            te += tc;
            const float t = te - dl.x;
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const float p = fmaf(t, d[a], o[a]);
                const float g = (in && p >= -bound && p <= bound) ? __ldg(grad_xyzs + k * 3 + a) : 0.0f;
                go[a] += g;
                gd[a] = fmaf(t, g, gd[a]);
                if (grad_dirs && in) gd[a] += __ldg(grad_dirs + k * 3 + a);
            }
            tc = __shfl_sync(0xffffffffu, te, 31);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float so = warp_sum(go[a]), sd = warp_sum(gd[a]);
        if (lane == 0) {
            grad_o[n * 3 + a] = so;
            grad_d[n * 3 + a] = sd;
        }
    }
}

extern "C" int ngp_march_rays_train_backward(const float *grad_xyzs, const float *grad_dirs, const float *rays_o,
                                             const float *rays_d, const float *nears, const float *deltas,
                                             const int32_t *rays, float bound, uint32_t max_steps, uint32_t perturb,
                                             int64_t N, int64_t M, float *grad_rays_o, float *grad_rays_d,
                                             ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(grad_xyzs, "ngp_march_rays_train_backward");
    if (N == 0) return NGP_OK;
    NGP_REQUIRE(rays_o && rays_d && nears && rays && grad_rays_o && grad_rays_d, NGP_ERR_ARG,
                "ngp_march_rays_train_backward: null pointer");
    NGP_REQUIRE(M == 0 || (grad_xyzs && deltas), NGP_ERR_ARG, "ngp_march_rays_train_backward: null sample buffers");
    const float dt_min = 2.0f * 1.7320508075688772f / (float)max_steps;
    k_march_train_bwd_warp<<<(unsigned)ngp_div_up(N * 32, 256), 256, 0, (cudaStream_t)stream>>>(
        grad_xyzs, grad_dirs, rays_o, rays_d, nears, deltas, rays, bound, dt_min, perturb, N, M, grad_rays_o, grad_rays_d);
    NGP_CHECK_LAUNCH("ngp_march_rays_train_backward");
    return NGP_OK;
}

// ---------------------------------------------------------------- compositing, training  (raymarching.cu:505-582, 606-688)
// __expf as in the reference (fast intrinsic) so alpha matches its CUDA build.
__global__ void __launch_bounds__(RAY_THREADS)
k_composite_train_fwd(const float *__restrict__ sigmas, const float *__restrict__ rgbs, const float *__restrict__ deltas,
                      const int32_t *__restrict__ rays, int64_t M, int64_t N, float *__restrict__ weights_sum,
                      float *__restrict__ depth, float *__restrict__ image) {
    const int64_t n = (int64_t)blockIdx.x * RAY_THREADS + threadIdx.x;
    if (n >= N) return;
    const uint32_t id = (uint32_t)rays[n * 3], off = (uint32_t)rays[n * 3 + 1], cnt = (uint32_t)rays[n * 3 + 2];
    float T = 1.0f, r = 0.0f, g = 0.0f, b = 0.0f, ws = 0.0f, t = 0.0f, d = 0.0f;
    if (cnt != 0 && (int64_t)off + cnt < M) {
        for (uint32_t s = 0; s < cnt; ++s) {
            const int64_t k = (int64_t)off + s;
            const float2 dl = __ldg(reinterpret_cast<const float2 *>(deltas) + k);
            const float alpha = 1.0f - __expf(-__ldg(sigmas + k) * dl.x);
            const float w = alpha * T;
            r = fmaf(w, __ldg(rgbs + k * 3), r);
            g = fmaf(w, __ldg(rgbs + k * 3 + 1), g);
            b = fmaf(w, __ldg(rgbs + k * 3 + 2), b);
            t += dl.y;
            d = fmaf(w, t, d);
            ws += w;
            T *= 1.0f - alpha;
            if (T < 1e-4f) break;
        }
    }
    weights_sum[id] = ws;
    depth[id] = d;
    image[id * 3] = r; image[id * 3 + 1] = g; image[id * 3 + 2] = b;
}

// One WARP per ray (lanes = 32 consecutive samples, coalesced loads; warp scans for the transmittance and the running t,
// warp sums for the accumulators): the thread-per-ray loop above reads four strided streams per lane.
__global__ void __launch_bounds__(256)
k_composite_train_fwd_warp(const float *__restrict__ sigmas, const float *__restrict__ rgbs, const float *__restrict__ deltas,
                           const int32_t *__restrict__ rays, int64_t M, int64_t N, float *__restrict__ weights_sum,
                           float *__restrict__ depth, float *__restrict__ image) {
    const int64_t n = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (n >= N) return;
    const uint32_t id = (uint32_t)__ldg(rays + n * 3), off = (uint32_t)__ldg(rays + n * 3 + 1), cnt = (uint32_t)__ldg(rays + n * 3 + 2);
    // rgbs are read as the block's 96 consecutive floats (flat element lane + 32 j, j = 0..2: three coalesced loads instead
    // of three stride-3 ones); element f belongs to sample f / 3, whose weight comes by shuffle, and to component f % 3 =
    // (lane + 2 j) % 3, so the three per-lane partial sums are sorted into r, g, b once, after the loop.
    const int sj0 = lane / 3, sj1 = (lane + 32) / 3, sj2 = (lane + 64) / 3;
    float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f, wsum = 0.0f, dsum = 0.0f;
    if (cnt != 0 && (int64_t)off + cnt < M) {
        float Tc = 1.0f, tc = 0.0f;
        for (uint32_t base = 0; base < cnt; base += 32) {
            const uint32_t s = base + lane, left = cnt - base;
            const bool in = s < cnt;
            const int64_t k = (int64_t)off + s;
            const float *cb = rgbs + ((int64_t)off + base) * 3;
            const float v0 = (uint32_t)sj0 < left ? __ldg(cb + lane) : 0.0f, v1 = (uint32_t)sj1 < left ? __ldg(cb + lane + 32) : 0.0f,
                        v2 = (uint32_t)sj2 < left ? __ldg(cb + lane + 64) : 0.0f;
            const float2 dl = in ? __ldg(reinterpret_cast<const float2 *>(deltas) + k) : make_float2(0.0f, 0.0f);
            const float alpha = in ? 1.0f - __expf(-__ldg(sigmas + k) * dl.x) : 0.0f;
            float P = 1.0f - alpha, t = dl.y;   // inclusive scans: transmittance after the sample, t of the sample
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const float vp = __shfl_up_sync(0xffffffffu, P, o), vt = __shfl_up_sync(0xffffffffu, t, o);
                if (lane >= o) {
                    P *= vp;
                    t += vt;
                }
            }
            float Pb = __shfl_up_sync(0xffffffffu, P, 1);
            if (lane == 0) Pb = 1.0f;
            const float w = alpha * (Tc * Pb), Tafter = Tc * P;
            t += tc;
            const unsigned brk = __ballot_sync(0xffffffffu, in && Tafter < 1e-4f);
            const int last = brk ? __ffs(brk) - 1 : 31;   // the sample that saw T < 1e-4 is still accumulated (:560-570)
            const bool use = in && lane <= last;
            const unsigned um = __ballot_sync(0xffffffffu, use);   // samples past the end never touch the sums (a NaN there stays out)
            const float wu = use ? w : 0.0f;
            const float w0 = __shfl_sync(0xffffffffu, wu, sj0), w1 = __shfl_sync(0xffffffffu, wu, sj1), w2 = __shfl_sync(0xffffffffu, wu, sj2);
            if ((um >> sj0) & 1u) a0 = fmaf(w0, v0, a0);
            if ((um >> sj1) & 1u) a1 = fmaf(w1, v1, a1);
            if ((um >> sj2) & 1u) a2 = fmaf(w2, v2, a2);
            if (use) {
                dsum = fmaf(w, t, dsum);
                wsum += w;
            }
            if (brk) break;
            Tc = __shfl_sync(0xffffffffu, Tafter, 31);
            tc = __shfl_sync(0xffffffffu, t, 31);
        }
    }
    const int c0 = lane % 3;   // component of a0; a1 holds (c0 + 2) % 3, a2 holds (c0 + 1) % 3
    const float r = warp_sum(c0 == 0 ? a0 : (c0 == 1 ? a1 : a2));
    const float g = warp_sum(c0 == 1 ? a0 : (c0 == 2 ? a1 : a2));
    const float b = warp_sum(c0 == 2 ? a0 : (c0 == 0 ? a1 : a2));
    const float ws = warp_sum(wsum), d = warp_sum(dsum);
    if (lane == 0) {
        weights_sum[id] = ws;
        depth[id] = d;
        image[id * 3] = r; image[id * 3 + 1] = g; image[id * 3 + 2] = b;
    }
}

__global__ void __launch_bounds__(RAY_THREADS)
k_composite_train_bwd(const float *__restrict__ grad_ws, const float *__restrict__ grad_image,
                      const float *__restrict__ sigmas, const float *__restrict__ rgbs, const float *__restrict__ deltas,
                      const int32_t *__restrict__ rays, const float *__restrict__ weights_sum,
                      const float *__restrict__ image, int64_t M, int64_t N, float *__restrict__ grad_sigmas,
                      float *__restrict__ grad_rgbs) {
    const int64_t n = (int64_t)blockIdx.x * RAY_THREADS + threadIdx.x;
    if (n >= N) return;
    const uint32_t id = (uint32_t)rays[n * 3], off = (uint32_t)rays[n * 3 + 1], cnt = (uint32_t)rays[n * 3 + 2];
    if (cnt == 0 || (int64_t)off + cnt >= M) return;
    const float gr = grad_image[id * 3], gg = grad_image[id * 3 + 1], gb = grad_image[id * 3 + 2], gws = grad_ws[id];
    const float rf = image[id * 3], gf = image[id * 3 + 1], bf = image[id * 3 + 2], wsf = weights_sum[id];
    float T = 1.0f, r = 0.0f, g = 0.0f, b = 0.0f;
    for (uint32_t s = 0; s < cnt; ++s) {
        const int64_t k = (int64_t)off + s;
        const float dl0 = __ldg(deltas + k * 2);
        const float c0 = __ldg(rgbs + k * 3), c1 = __ldg(rgbs + k * 3 + 1), c2 = __ldg(rgbs + k * 3 + 2);
        const float alpha = 1.0f - __expf(-__ldg(sigmas + k) * dl0);
        const float w = alpha * T;
        r = fmaf(w, c0, r);
        g = fmaf(w, c1, g);
        b = fmaf(w, c2, b);
        T *= 1.0f - alpha;
        if (T < 1e-4f) break;  // before writing this sample's gradients (:658-661)
        grad_rgbs[k * 3] = gr * w; grad_rgbs[k * 3 + 1] = gg * w; grad_rgbs[k * 3 + 2] = gb * w;
        grad_sigmas[k] = dl0 * (gr * (T * c0 - (rf - r)) + gg * (T * c1 - (gf - g)) + gb * (T * c2 - (bf - b)) +
                                gws * (1.0f - wsf));
    }
}

// The same backward with one WARP per ray: lanes = 32 consecutive samples, so every load and store is coalesced (the
// thread-per-ray loop above reads three strided streams per lane: 5.4 ms for the estimator's 7.9x10^7 samples against
// 0.6 ms of HBM traffic).  The running transmittance and colour become warp scans (prefix product / prefix sums carried
// from one 32-sample block to the next): same formulas, summation order within a block differs by a few ulp.
__global__ void __launch_bounds__(256)
k_composite_train_bwd_warp(const float *__restrict__ grad_ws, const float *__restrict__ grad_image,
                           const float *__restrict__ sigmas, const float *__restrict__ rgbs, const float *__restrict__ deltas,
                           const int32_t *__restrict__ rays, const float *__restrict__ weights_sum,
                           const float *__restrict__ image, int64_t M, int64_t N, float *__restrict__ grad_sigmas,
                           float *__restrict__ grad_rgbs) {
    const int64_t n = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (n >= N) return;
    const uint32_t id = (uint32_t)__ldg(rays + n * 3), off = (uint32_t)__ldg(rays + n * 3 + 1), cnt = (uint32_t)__ldg(rays + n * 3 + 2);
    if (cnt == 0 || (int64_t)off + cnt >= M) return;
    const float gr = __ldg(grad_image + id * 3), gg = __ldg(grad_image + id * 3 + 1), gb = __ldg(grad_image + id * 3 + 2);
    const float gws = __ldg(grad_ws + id);
    const float rf = __ldg(image + id * 3), gf = __ldg(image + id * 3 + 1), bf = __ldg(image + id * 3 + 2);
    const float wsf = __ldg(weights_sum + id);
    float Tc = 1.0f, rc = 0.0f, gc = 0.0f, bc = 0.0f;   // carried in from the previous blocks
    for (uint32_t base = 0; base < cnt; base += 32) {
        const uint32_t s = base + lane;
        const bool in = s < cnt;
        const int64_t k = (int64_t)off + s;
        const float dl0 = in ? __ldg(deltas + k * 2) : 0.0f;
        const float c0 = in ? __ldg(rgbs + k * 3) : 0.0f, c1 = in ? __ldg(rgbs + k * 3 + 1) : 0.0f, c2 = in ? __ldg(rgbs + k * 3 + 2) : 0.0f;
        const float alpha = in ? 1.0f - __expf(-__ldg(sigmas + k) * dl0) : 0.0f;
        float P = 1.0f - alpha;   // -> inclusive prefix product: transmittance after the sample / carried-in T
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const float v = __shfl_up_sync(0xffffffffu, P, o);
            if (lane >= o) P *= v;
        }
        float Pb = __shfl_up_sync(0xffffffffu, P, 1);
        if (lane == 0) Pb = 1.0f;
        const float Tafter = Tc * P, w = alpha * (Tc * Pb);
        float r = w * c0, g = w * c1, b = w * c2;   // -> inclusive prefix sums
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const float vr = __shfl_up_sync(0xffffffffu, r, o), vg = __shfl_up_sync(0xffffffffu, g, o),
                        vb = __shfl_up_sync(0xffffffffu, b, o);
            if (lane >= o) {
                r += vr; g += vg; b += vb;
            }
        }
        r += rc; g += gc; b += bc;
        const unsigned brk = __ballot_sync(0xffffffffu, in && Tafter < 1e-4f);   // :658-661: that sample writes nothing, the ray ends
        const int first = brk ? __ffs(brk) - 1 : 32;
        if (in && lane < first) {
            grad_rgbs[k * 3] = gr * w; grad_rgbs[k * 3 + 1] = gg * w; grad_rgbs[k * 3 + 2] = gb * w;
            grad_sigmas[k] = dl0 * (gr * (Tafter * c0 - (rf - r)) + gg * (Tafter * c1 - (gf - g)) + gb * (Tafter * c2 - (bf - b)) +
                                    gws * (1.0f - wsf));
        }
        if (brk) break;
        Tc = __shfl_sync(0xffffffffu, Tafter, 31);
        rc = __shfl_sync(0xffffffffu, r, 31); gc = __shfl_sync(0xffffffffu, g, 31); bc = __shfl_sync(0xffffffffu, b, 31);
    }
}

extern "C" int ngp_composite_rays_train_forward(const float *sigmas, const float *rgbs, const float *deltas,
                                                const int32_t *rays, int64_t M, int64_t N, float *weights_sum,
                                                float *depth, float *image, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(sigmas, "ngp_composite_rays_train_forward");
    if (N == 0) return NGP_OK;
    NGP_REQUIRE(rays && weights_sum && depth && image, NGP_ERR_ARG, "ngp_composite_rays_train_forward: null pointer");
    NGP_REQUIRE(M == 0 || (sigmas && rgbs && deltas), NGP_ERR_ARG, "ngp_composite_rays_train_forward: null samples");
    static int serial = -1;   // NGP_B200_COMPOSITE_BWD=serial also selects the thread-per-ray forward
    if (serial < 0) {
        const char *e = getenv("NGP_B200_COMPOSITE_BWD");
        serial = (e && strcmp(e, "serial") == 0) ? 1 : 0;
    }
    if (serial)
        k_composite_train_fwd<<<(unsigned)ngp_div_up(N, RAY_THREADS), RAY_THREADS, 0, (cudaStream_t)stream>>>(
            sigmas, rgbs, deltas, rays, M, N, weights_sum, depth, image);
    else
        k_composite_train_fwd_warp<<<(unsigned)ngp_div_up(N * 32, 256), 256, 0, (cudaStream_t)stream>>>(
            sigmas, rgbs, deltas, rays, M, N, weights_sum, depth, image);
    NGP_CHECK_LAUNCH("ngp_composite_rays_train_forward");
    return NGP_OK;
}

extern "C" int ngp_composite_rays_train_backward(const float *grad_weights_sum, const float *grad_image,
                                                 const float *sigmas, const float *rgbs, const float *deltas,
                                                 const int32_t *rays, const float *weights_sum, const float *image,
                                                 int64_t M, int64_t N, float *grad_sigmas, float *grad_rgbs,
                                                 ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(grad_weights_sum, "ngp_composite_rays_train_backward");
    if (N == 0) return NGP_OK;
    NGP_REQUIRE(grad_weights_sum && grad_image && rays && weights_sum && image, NGP_ERR_ARG,
                "ngp_composite_rays_train_backward: null pointer");
    NGP_REQUIRE(M == 0 || (sigmas && rgbs && deltas && grad_sigmas && grad_rgbs), NGP_ERR_ARG,
                "ngp_composite_rays_train_backward: null samples");
    static int serial = -1;   // NGP_B200_COMPOSITE_BWD=serial: the thread-per-ray kernel (the reference's summation order)
    if (serial < 0) {
        const char *e = getenv("NGP_B200_COMPOSITE_BWD");
        serial = (e && strcmp(e, "serial") == 0) ? 1 : 0;
    }
    if (serial)
        k_composite_train_bwd<<<(unsigned)ngp_div_up(N, RAY_THREADS), RAY_THREADS, 0, (cudaStream_t)stream>>>(
            grad_weights_sum, grad_image, sigmas, rgbs, deltas, rays, weights_sum, image, M, N, grad_sigmas, grad_rgbs);
    else
        k_composite_train_bwd_warp<<<(unsigned)ngp_div_up(N * 32, 256), 256, 0, (cudaStream_t)stream>>>(
            grad_weights_sum, grad_image, sigmas, rgbs, deltas, rays, weights_sum, image, M, N, grad_sigmas, grad_rgbs);
    NGP_CHECK_LAUNCH("ngp_composite_rays_train_backward");
    return NGP_OK;
}

// ---------------------------------------------------------------- inference marcher / compositor  (:706-814, :828-913)
__global__ void __launch_bounds__(RAY_THREADS)
k_march_rays(int64_t n_alive, uint32_t n_step, const int32_t *__restrict__ rays_alive, const float *__restrict__ rays_t,
             const float *__restrict__ rays_o, const float *__restrict__ rays_d, MarchCfg cfg,
             const float *__restrict__ fars, float *__restrict__ xyzs, float *__restrict__ dirs,
             float *__restrict__ deltas, uint32_t perturb) {
    const int64_t n = (int64_t)blockIdx.x * RAY_THREADS + threadIdx.x;
    if (n >= n_alive) return;
    const int32_t id = rays_alive[n];
    const MarchRay r = load_ray(rays_o, rays_d, id);
    float t = rays_t[id];
    if (perturb) t = fmaf(cfg.dt_min, ngp_pcg_jitter((uint64_t)perturb, (uint64_t)n), t);  // :745-748, seed :819
    float last_t = t;
    const float far = fars[id];
    uint32_t step = 0;
    float px, py, pz, dt;
    while (t < far && step < n_step) {
        if (!ngp_march_step(r, cfg, t, px, py, pz, dt)) continue;
        const int64_t k = n * n_step + step;
        xyzs[k * 3] = px; xyzs[k * 3 + 1] = py; xyzs[k * 3 + 2] = pz;
        dirs[k * 3] = r.dx; dirs[k * 3 + 1] = r.dy; dirs[k * 3 + 2] = r.dz;
        t += dt;
        deltas[k * 2] = dt;
        deltas[k * 2 + 1] = t - last_t;
        last_t = t;
        ++step;
    }
}

__global__ void __launch_bounds__(RAY_THREADS)
k_composite_rays(int64_t n_alive, uint32_t n_step, int32_t *__restrict__ rays_alive, float *__restrict__ rays_t,
                 const float *__restrict__ sigmas, const float *__restrict__ rgbs, const float *__restrict__ deltas,
                 float *__restrict__ weights_sum, float *__restrict__ depth, float *__restrict__ image) {
    const int64_t n = (int64_t)blockIdx.x * RAY_THREADS + threadIdx.x;
    if (n >= n_alive) return;
    const int32_t id = rays_alive[n];
    float t = rays_t[id], ws = weights_sum[id], d = depth[id];
    float r = image[id * 3], g = image[id * 3 + 1], b = image[id * 3 + 2];
    uint32_t step = 0;
    while (step < n_step) {
        const int64_t k = n * n_step + step;
        const float dl0 = deltas[k * 2];
        if (dl0 == 0.0f) break;
        const float alpha = 1.0f - __expf(-sigmas[k] * dl0);
        const float T = 1.0f - ws;
        const float w = alpha * T;
        ws += w;
        t += deltas[k * 2 + 1];
        d = fmaf(w, t, d);
        r = fmaf(w, rgbs[k * 3], r);
        g = fmaf(w, rgbs[k * 3 + 1], g);
        b = fmaf(w, rgbs[k * 3 + 2], b);
        if ((double)T < 1e-4) break;  // double literal in the reference (:880)
        ++step;
    }
    if (step < n_step) rays_alive[n] = -1;
    else rays_t[id] = t;
    weights_sum[id] = ws;
    depth[id] = d;
    image[id * 3] = r; image[id * 3 + 1] = g; image[id * 3 + 2] = b;
}

extern "C" int ngp_march_rays(int64_t n_alive, uint32_t n_step, const int32_t *rays_alive, const float *rays_t,
                              const float *rays_o, const float *rays_d, float bound, float dt_gamma, uint32_t max_steps,
                              uint32_t C, uint32_t H, const uint8_t *grid, const float *nears, const float *fars,
                              float *xyzs, float *dirs, float *deltas, uint32_t perturb, ngp_stream_t stream) {
    if (n_alive == 0) return NGP_OK;
    NGP_REQUIRE_DEVICE(rays_alive, "ngp_march_rays");
    (void)nears;
    NGP_REQUIRE(rays_alive && rays_t && rays_o && rays_d && grid && fars && xyzs && dirs && deltas, NGP_ERR_ARG,
                "ngp_march_rays: null pointer");
    NGP_REQUIRE(C >= 1 && C <= 8 && H >= 1 && H <= 1024, NGP_ERR_ARG, "ngp_march_rays: bad cascade/grid size");
    const MarchCfg cfg = ngp_march_cfg(grid, bound, dt_gamma, max_steps, C, H);
    k_march_rays<<<(unsigned)ngp_div_up(n_alive, RAY_THREADS), RAY_THREADS, 0, (cudaStream_t)stream>>>(
        n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, cfg, fars, xyzs, dirs, deltas, perturb);
    NGP_CHECK_LAUNCH("ngp_march_rays");
    return NGP_OK;
}

extern "C" int ngp_composite_rays(int64_t n_alive, uint32_t n_step, int32_t *rays_alive, float *rays_t,
                                  const float *sigmas, const float *rgbs, const float *deltas, float *weights_sum,
                                  float *depth, float *image, ngp_stream_t stream) {
    if (n_alive == 0) return NGP_OK;
    NGP_REQUIRE_DEVICE(rays_alive, "ngp_composite_rays");
    NGP_REQUIRE(rays_alive && rays_t && sigmas && rgbs && deltas && weights_sum && depth && image, NGP_ERR_ARG,
                "ngp_composite_rays: null pointer");
    k_composite_rays<<<(unsigned)ngp_div_up(n_alive, RAY_THREADS), RAY_THREADS, 0, (cudaStream_t)stream>>>(
        n_alive, n_step, rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth, image);
    NGP_CHECK_LAUNCH("ngp_composite_rays");
    return NGP_OK;
}

// ------------------------------------------------------------------------------------------------ pose -> rays closure
// The estimator's measurement (nav/estimator_helpers.py:293-327) builds ALL H*W rays of one pose with get_rays
// (nerf/utils.py:52-116) and then indexes the ~1k sampled pixels; here only the sampled pixels' rays are generated, for a
// batch of pose hypotheses at once, with the same conventions: pixel centres at +0.5, camera-space direction
// ((i - cx) / fx, (j - cy) / fy, 1) normalised, rays_d = dir @ R^T, rays_o = t (pose = cam2world [R | t]).
//   k_pose_rays_fwd: one thread per (hypothesis, pixel);
//   k_pose_rays_bwd: dL/d(pose) = [sum_r grad_d (x) dir_r | sum_r grad_o]: one CTA per hypothesis, fixed-order tree over its
//       rays (deterministic), so the estimator step has no eager elementwise / matmul launches around the marcher.
__device__ __forceinline__ void pose_pixel_dir(int64_t pix, int W, float fx, float fy, float cx, float cy, float (&d)[3]) {
    const float i = (float)(pix % W) + 0.5f, j = (float)(pix / W) + 0.5f;
    const float x = __fdiv_rn(i - cx, fx), y = __fdiv_rn(j - cy, fy);
    const float n = sqrtf(fmaf(x, x, fmaf(y, y, 1.0f)));
    d[0] = __fdiv_rn(x, n); d[1] = __fdiv_rn(y, n); d[2] = __fdiv_rn(1.0f, n);
}

__global__ void __launch_bounds__(RAY_THREADS)
k_pose_rays_fwd(const float *__restrict__ poses, const int64_t *__restrict__ pix, int P, int R, int W, float fx, float fy,
                float cx, float cy, float *__restrict__ rays_o, float *__restrict__ rays_d) {
    const int64_t q = (int64_t)blockIdx.x * RAY_THREADS + threadIdx.x;
    if (q >= (int64_t)P * R) return;
    const int p = (int)(q / R), r = (int)(q - (int64_t)p * R);
    float d[3];
    pose_pixel_dir(pix[r], W, fx, fy, cx, cy, d);
    const float *M = poses + (size_t)p * 16;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        rays_d[q * 3 + a] = fmaf(d[2], __ldg(M + 4 * a + 2), fmaf(d[1], __ldg(M + 4 * a + 1), d[0] * __ldg(M + 4 * a)));
        rays_o[q * 3 + a] = __ldg(M + 4 * a + 3);
    }
}

__global__ void __launch_bounds__(256)
k_pose_rays_bwd(const float *__restrict__ grad_o, const float *__restrict__ grad_d, const int64_t *__restrict__ pix, int R, int W,
                float fx, float fy, float cx, float cy, float *__restrict__ grad_poses) {
    __shared__ float part[8][12];
    const int p = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float acc[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) acc[k] = 0.0f;
    for (int r = threadIdx.x; r < R; r += 256) {
        float d[3];
        pose_pixel_dir(pix[r], W, fx, fy, cx, cy, d);
        const int64_t q = (int64_t)p * R + r;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float gd = grad_d ? grad_d[q * 3 + a] : 0.0f;
            acc[4 * a + 0] = fmaf(gd, d[0], acc[4 * a + 0]);
            acc[4 * a + 1] = fmaf(gd, d[1], acc[4 * a + 1]);
            acc[4 * a + 2] = fmaf(gd, d[2], acc[4 * a + 2]);
            acc[4 * a + 3] += grad_o ? grad_o[q * 3 + a] : 0.0f;
        }
    }
#pragma unroll
    for (int k = 0; k < 12; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
        if (lane == 0) part[warp][k] = acc[k];
    }
    __syncthreads();
    if (threadIdx.x < 16) {
        float v = 0.0f;
        if (threadIdx.x < 12)
            for (int w = 0; w < 8; ++w) v += part[w][threadIdx.x];
        grad_poses[(size_t)p * 16 + threadIdx.x] = v;      // row 3 of the pose takes no gradient
    }
}

extern "C" int ngp_pose_rays_forward(const float *poses, const int64_t *pix, int P, int R, int W, float fx, float fy, float cx,
                                     float cy, float *rays_o, float *rays_d, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(poses, "ngp_pose_rays_forward");
    if ((int64_t)P * R == 0) return NGP_OK;
    NGP_REQUIRE(poses && pix && rays_o && rays_d && P > 0 && R > 0 && W > 0, NGP_ERR_ARG, "ngp_pose_rays_forward: bad arguments");
    k_pose_rays_fwd<<<(unsigned)ngp_div_up((int64_t)P * R, RAY_THREADS), RAY_THREADS, 0, (cudaStream_t)stream>>>(
        poses, pix, P, R, W, fx, fy, cx, cy, rays_o, rays_d);
    NGP_CHECK_LAUNCH("ngp_pose_rays_forward");
    return NGP_OK;
}

extern "C" int ngp_pose_rays_backward(const float *grad_o, const float *grad_d, const int64_t *pix, int P, int R, int W, float fx,
                                      float fy, float cx, float cy, float *grad_poses, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(grad_poses, "ngp_pose_rays_backward");
    if (P == 0) return NGP_OK;
    NGP_REQUIRE((grad_o || grad_d) && pix && grad_poses && P > 0 && R >= 0 && W > 0, NGP_ERR_ARG,
                "ngp_pose_rays_backward: bad arguments");
    k_pose_rays_bwd<<<P, 256, 0, (cudaStream_t)stream>>>(grad_o, grad_d, pix, R, W, fx, fy, cx, cy, grad_poses);
    NGP_CHECK_LAUNCH("ngp_pose_rays_backward");
    return NGP_OK;
}

// ---------------------------------------------------------------- 12-state -> cam2world pose  (estimator closure)
// nav/estimator_helpers.py:293-327 + nav/math_utils.py:26-38, 159-174: R = I + sin(a) K + (1 - cos(a)) K K with
// a = |v|, K = skew(v / (1e-10 + a)), v = state[6:9]; pose = [flip_yz (rot_x(pi/2) R) neg_yz | flip_yz p], p = state[0:3].
// One thread per hypothesis, evaluated in double and rounded once (a few dozen flops per state); the backward is the
// analytic derivative of the same expression (v = 0: zero gradient, as autograd gives through the 1e-10).
// c, s = cos / sin of the float32 pi / 2 the reference evaluates.
struct PoseMat {
    double m[3][3];
};
__device__ __forceinline__ void state_rotation(const double v[3], double c, double s, PoseMat &K, PoseMat &R, double &ang, double &sn,
                                               double &cs) {
    ang = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    const double inv = 1.0 / (1e-10 + ang);
    const double k[3] = {v[0] * inv, v[1] * inv, v[2] * inv};
    const double Km[3][3] = {{0.0, -k[2], k[1]}, {k[2], 0.0, -k[0]}, {-k[1], k[0], 0.0}};
    sn = sin(ang);
    cs = cos(ang);
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double kk = 0.0;
            for (int l = 0; l < 3; ++l) kk += Km[i][l] * Km[l][j];
            K.m[i][j] = Km[i][j];
            R.m[i][j] = (i == j ? 1.0 : 0.0) + sn * Km[i][j] + (1.0 - cs) * kk;
        }
    (void)c; (void)s;
}

__global__ void k_state_to_pose_fwd(const float *__restrict__ states, int P, float c, float s, float *__restrict__ poses) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const float *st = states + (int64_t)p * 12;
    const double v[3] = {st[6], st[7], st[8]};
    PoseMat K, R;
    double ang, sn, cs;
    state_rotation(v, c, s, K, R, ang, sn, cs);
    // XR = rot_x R: row 0 = R0, row 1 = c R1 - s R2, row 2 = s R1 + c R2
    double XR[3][3];
    for (int j = 0; j < 3; ++j) {
        XR[0][j] = R.m[0][j];
        XR[1][j] = (double)c * R.m[1][j] - (double)s * R.m[2][j];
        XR[2][j] = (double)s * R.m[1][j] + (double)c * R.m[2][j];
    }
    // flip_yz A: rows (A1, A2, A0); A neg_yz: columns 1, 2 negated
    float *out = poses + (int64_t)p * 16;
    const int src[3] = {1, 2, 0};
    for (int i = 0; i < 3; ++i) {
        out[i * 4 + 0] = (float)XR[src[i]][0];
        out[i * 4 + 1] = (float)-XR[src[i]][1];
        out[i * 4 + 2] = (float)-XR[src[i]][2];
        out[i * 4 + 3] = st[src[i]];   // flip_yz p = (p1, p2, p0)
    }
    out[12] = 0.0f; out[13] = 0.0f; out[14] = 0.0f; out[15] = 1.0f;
}

__global__ void k_state_to_pose_bwd(const float *__restrict__ states, const float *__restrict__ grad_poses, int P, float c, float s,
                                    float *__restrict__ grad_states) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const float *st = states + (int64_t)p * 12;
    const float *G = grad_poses + (int64_t)p * 16;
    float *gs = grad_states + (int64_t)p * 12;
    for (int i = 0; i < 12; ++i) gs[i] = 0.0f;   // velocity and angular velocity take no gradient from the measurement
    const int src[3] = {1, 2, 0};
    // translation: pose[i][3] = p[src[i]]
    for (int i = 0; i < 3; ++i) gs[src[i]] = G[i * 4 + 3];
    // dL/dXR[src[i]][j] = G[i][j] * (j == 0 ? 1 : -1)
    double gXR[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) gXR[src[i]][j] = (double)G[i * 4 + j] * (j == 0 ? 1.0 : -1.0);
    // dL/dR = rot_x^T dL/dXR
    double gR[3][3];
    for (int j = 0; j < 3; ++j) {
        gR[0][j] = gXR[0][j];
        gR[1][j] = (double)c * gXR[1][j] + (double)s * gXR[2][j];
        gR[2][j] = -(double)s * gXR[1][j] + (double)c * gXR[2][j];
    }
    const double v[3] = {st[6], st[7], st[8]};
    PoseMat K, R;
    double ang, sn, cs;
    state_rotation(v, c, s, K, R, ang, sn, cs);
    if (!(ang > 0.0)) return;   // v = 0: K = 0, sin = 0, 1 - cos = 0 -> every path to v carries a zero factor
    // R = I + sn K + (1 - cs) K K
    double KK[3][3], gK[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            KK[i][j] = 0.0;
            for (int l = 0; l < 3; ++l) KK[i][j] += K.m[i][l] * K.m[l][j];
        }
    double g_ang = 0.0;   // explicit dependence on the angle: dR/da = cs K + sn K K
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) g_ang += gR[i][j] * (cs * K.m[i][j] + sn * KK[i][j]);
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {   // d(K K)/dK: gR K^T + K^T gR
            double a = 0.0, b = 0.0;
            for (int l = 0; l < 3; ++l) {
                a += gR[i][l] * K.m[j][l];
                b += K.m[l][i] * gR[l][j];
            }
            gK[i][j] = sn * gR[i][j] + (1.0 - cs) * (a + b);
        }
    const double gk[3] = {gK[2][1] - gK[1][2], gK[0][2] - gK[2][0], gK[1][0] - gK[0][1]};
    // k = v / (1e-10 + a), a = |v|
    const double den = 1e-10 + ang;
    const double gkv = gk[0] * v[0] + gk[1] * v[1] + gk[2] * v[2];
    for (int j = 0; j < 3; ++j) gs[6 + j] = (float)(gk[j] / den - gkv * v[j] / (ang * den * den) + g_ang * v[j] / ang);
}

extern "C" int ngp_state_to_pose_forward(const float *states, int P, float cos_half_pi, float sin_half_pi, float *poses,
                                         ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(states, "ngp_state_to_pose_forward");
    if (P == 0) return NGP_OK;
    NGP_REQUIRE(states && poses && P > 0, NGP_ERR_ARG, "ngp_state_to_pose_forward: null pointer / bad P");
    k_state_to_pose_fwd<<<(unsigned)ngp_div_up((int64_t)P, 128), 128, 0, (cudaStream_t)stream>>>(states, P, cos_half_pi, sin_half_pi,
                                                                                                poses);
    NGP_CHECK_LAUNCH("ngp_state_to_pose_forward");
    return NGP_OK;
}

extern "C" int ngp_state_to_pose_backward(const float *states, const float *grad_poses, int P, float cos_half_pi, float sin_half_pi,
                                          float *grad_states, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(states, "ngp_state_to_pose_backward");
    if (P == 0) return NGP_OK;
    NGP_REQUIRE(states && grad_poses && grad_states && P > 0, NGP_ERR_ARG, "ngp_state_to_pose_backward: null pointer / bad P");
    k_state_to_pose_bwd<<<(unsigned)ngp_div_up((int64_t)P, 128), 128, 0, (cudaStream_t)stream>>>(states, grad_poses, P, cos_half_pi,
                                                                                                sin_half_pi, grad_states);
    NGP_CHECK_LAUNCH("ngp_state_to_pose_backward");
    return NGP_OK;
}
