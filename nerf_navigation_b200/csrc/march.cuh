// march.cuh — occupancy-grid marcher shared by the standalone ray ops and the fused render kernels.
// Reference semantics: raymarching/src/raymarching.cu:44-83 (mip level, Morton), :361-403 (one marcher iteration).
// Arithmetic notes (bit-exact sample counts are the target): explicit fmaf where nvcc contracts the reference
// expression under its default -fmad=true; the cell index product `0.5 * (x/mip_bound + 1) * H` is evaluated in
// fp64 in the reference (double literal) — for power-of-two H the fp32 product is exact and identical, otherwise the
// fp64 path is taken.
#pragma once
#include <stdint.h>

__host__ __device__ __forceinline__ uint32_t ngp_expand_bits(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
__host__ __device__ __forceinline__ uint32_t ngp_morton3D(uint32_t x, uint32_t y, uint32_t z) {
    return ngp_expand_bits(x) | (ngp_expand_bits(y) << 1) | (ngp_expand_bits(z) << 2);
}
__host__ __device__ __forceinline__ uint32_t ngp_morton_compact(uint32_t x) {
    x &= 0x49249249u;
    x = (x | (x >> 2)) & 0xc30c30c3u;
    x = (x | (x >> 4)) & 0x0f00f00fu;
    x = (x | (x >> 8)) & 0xff0000ffu;
    x = (x | (x >> 16)) & 0x0000ffffu;
    return x;
}

struct MarchRay {
    float ox, oy, oz, dx, dy, dz, rdx, rdy, rdz;
};

struct MarchCfg {
    float bound, dt_gamma, dt_min, dt_max, rH, Hf;
    uint32_t C, H, H3;
    int h_pow2;
    int lvl_dt;   // dt_gamma == 0: the (then constant) mip level of dt = dt_min, raymarching.cu:51-56; -1 otherwise
    const uint8_t *grid;
};

__host__ __device__ __forceinline__ MarchCfg ngp_march_cfg(const uint8_t *grid, float bound, float dt_gamma,
                                                           uint32_t max_steps, uint32_t C, uint32_t H) {
    MarchCfg c;
    c.bound = bound;
    c.dt_gamma = dt_gamma;
    c.dt_min = 2.0f * 1.7320508075688772f / (float)max_steps;                   // raymarching.cu:347
    c.dt_max = 2.0f * 1.7320508075688772f * (float)(1u << (C - 1)) / (float)H;  // :348
    c.rH = 1.0f / (float)H;
    c.Hf = (float)H;
    c.C = C;
    c.H = H;
    c.H3 = H * H * H;
    c.h_pow2 = (H & (H - 1)) == 0;
    c.lvl_dt = -1;
    if (dt_gamma == 0.0f) {   // clamp(t * 0, dt_min, dt_max) = dt_min for every finite t
        int e;
        frexpf(c.dt_min * c.Hf * 0.5f, &e);
        c.lvl_dt = e < 0 ? 0 : (e > (int)C - 1 ? (int)C - 1 : e);
    }
    c.grid = grid;
    return c;
}

__device__ __forceinline__ float ngp_clampf(float v, float lo, float hi) { return fminf(hi, fmaxf(lo, v)); }

// frexpf exponent clamped to [0, C-1]  (raymarching.cu:44-56)
__device__ __forceinline__ int ngp_mip_level(float v, uint32_t C) {
    int e;
    frexpf(v, &e);
    return min((int)C - 1, max(0, e));
}

__device__ __forceinline__ int ngp_cell_coord(float p, float mip_rbound, const MarchCfg &c) {
    const float u = fmaf(p, mip_rbound, 1.0f);
    float v;
    if (c.h_pow2) v = 0.5f * u * c.Hf;
    else v = (float)(0.5 * (double)u * (double)c.H);
    return (int)ngp_clampf(v, 0.0f, c.Hf - 1.0f);
}

// Probe of the marcher at parameter t0 (no state change): occupancy of the cell holding the sample, the sample position
// and step, and for an empty cell the parameter `tt` at which the ray leaves the voxel (raymarching.cu:361-396).
__device__ __forceinline__ bool ngp_march_probe(const MarchRay &r, const MarchCfg &c, float t0, float &px, float &py,
                                                float &pz, float &dt, float &tt) {
    const float x = ngp_clampf(fmaf(t0, r.dx, r.ox), -c.bound, c.bound);
    const float y = ngp_clampf(fmaf(t0, r.dy, r.oy), -c.bound, c.bound);
    const float z = ngp_clampf(fmaf(t0, r.dz, r.oz), -c.bound, c.bound);
    dt = ngp_clampf(t0 * c.dt_gamma, c.dt_min, c.dt_max);
    const float mx = fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z)));
    const int level = max(ngp_mip_level(mx, c.C), ngp_mip_level(dt * c.Hf * 0.5f, c.C));
    const float mip_bound = fminf((float)(1 << level), c.bound);
    const float mip_rbound = 1.0f / mip_bound;
    const int nx = ngp_cell_coord(x, mip_rbound, c);
    const int ny = ngp_cell_coord(y, mip_rbound, c);
    const int nz = ngp_cell_coord(z, mip_rbound, c);
    const uint32_t index = (uint32_t)level * c.H3 + ngp_morton3D((uint32_t)nx, (uint32_t)ny, (uint32_t)nz);
    const bool occ = (__ldg(c.grid + (index >> 3)) >> (index & 7u)) & 1u;
    px = x; py = y; pz = z;
    tt = t0;
    if (occ) return true;
    const float tx = fmaf(fmaf(((float)nx + 0.5f + 0.5f * copysignf(1.0f, r.dx)) * c.rH, 2.0f, -1.0f), mip_bound, -x) * r.rdx;
    const float ty = fmaf(fmaf(((float)ny + 0.5f + 0.5f * copysignf(1.0f, r.dy)) * c.rH, 2.0f, -1.0f), mip_bound, -y) * r.rdy;
    const float tz = fmaf(fmaf(((float)nz + 0.5f + 0.5f * copysignf(1.0f, r.dz)) * c.rH, 2.0f, -1.0f), mip_bound, -z) * r.rdz;
    tt = t0 + fmaxf(0.0f, fminf(tx, fminf(ty, tz)));
    return false;
}

// The probe in two halves, for marchers that keep several rays' occupancy loads in flight per thread (render5.cu): the
// bit address of the cell that holds the sample at t0, then — only for an empty cell, after the load came back — the voxel
// exit.  Every expression is the one ngp_march_probe evaluates (the position is recomputed from (o, d, t0), the cell
// coordinates are decoded from the Morton index), so both forms give the same bits.
__device__ __forceinline__ uint32_t ngp_march_cell(const MarchRay &r, const MarchCfg &c, float t0, float &dt) {
    const float x = ngp_clampf(fmaf(t0, r.dx, r.ox), -c.bound, c.bound);
    const float y = ngp_clampf(fmaf(t0, r.dy, r.oy), -c.bound, c.bound);
    const float z = ngp_clampf(fmaf(t0, r.dz, r.oz), -c.bound, c.bound);
    dt = ngp_clampf(t0 * c.dt_gamma, c.dt_min, c.dt_max);
    const float mx = fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z)));
    const int level = max(ngp_mip_level(mx, c.C), ngp_mip_level(dt * c.Hf * 0.5f, c.C));
    const float mip_rbound = 1.0f / fminf((float)(1 << level), c.bound);
    const int nx = ngp_cell_coord(x, mip_rbound, c);
    const int ny = ngp_cell_coord(y, mip_rbound, c);
    const int nz = ngp_cell_coord(z, mip_rbound, c);
    return (uint32_t)level * c.H3 + ngp_morton3D((uint32_t)nx, (uint32_t)ny, (uint32_t)nz);
}
// frexpf exponent of a finite v >= 0 from its bit pattern (zero / subnormal give a negative value, clamped to 0 like the
// negative exponents frexpf returns for them)
__device__ __forceinline__ int ngp_mip_level_bits(float v, uint32_t C) {
    const int e = (int)((__float_as_uint(v) >> 23) & 0xffu) - 126;
    return min((int)C - 1, max(0, e));
}
// fp64 form of the cell coordinate for grids whose H is not a power of two, out of line so that the common fp32 form
// does not carry its (predicated-off) double-precision instructions
static __device__ __noinline__ float ngp_cell_coord_f64(float u, uint32_t H) { return (float)(0.5 * (double)u * (double)H); }
__device__ __forceinline__ int ngp_cell_coord_fast(float p, float mip_rbound, const MarchCfg &c) {
    const float u = fmaf(p, mip_rbound, 1.0f);
    const float v = c.h_pow2 ? 0.5f * u * c.Hf : ngp_cell_coord_f64(u, c.H);
    return (int)ngp_clampf(v, 0.0f, c.Hf - 1.0f);
}
// The same cell address with fewer instructions (render5.cu): the level of the position from the exponent bits; dt and its
// level from the configuration when dt_gamma == 0; min(2^level, bound) and its reciprocal from tables filled with the
// expressions above (mipb[l] = fminf((float)(1 << l), bound), miprb[l] = 1.0f / mipb[l]); the cell coordinates are also
// returned packed (nx | ny << 10 | nz << 20) so that the exit needs no Morton decode.  Identical values.
__device__ __forceinline__ uint32_t ngp_march_cell_fast(const MarchRay &r, const MarchCfg &c, const float *miprb, float t0,
                                                        float &dt, int &level, uint32_t &nxyz) {
    const float x = ngp_clampf(fmaf(t0, r.dx, r.ox), -c.bound, c.bound);
    const float y = ngp_clampf(fmaf(t0, r.dy, r.oy), -c.bound, c.bound);
    const float z = ngp_clampf(fmaf(t0, r.dz, r.oz), -c.bound, c.bound);
    int lvl_dt = c.lvl_dt;
    if (lvl_dt < 0) {
        dt = ngp_clampf(t0 * c.dt_gamma, c.dt_min, c.dt_max);
        lvl_dt = ngp_mip_level(dt * c.Hf * 0.5f, c.C);
    } else {
        dt = c.dt_min;
    }
    const float mx = fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z)));
    level = max(ngp_mip_level_bits(mx, c.C), lvl_dt);
    const float mip_rbound = miprb[level];
    const int nx = ngp_cell_coord_fast(x, mip_rbound, c);
    const int ny = ngp_cell_coord_fast(y, mip_rbound, c);
    const int nz = ngp_cell_coord_fast(z, mip_rbound, c);
    nxyz = (uint32_t)nx | ((uint32_t)ny << 10) | ((uint32_t)nz << 20);
    return (uint32_t)level * c.H3 + ngp_morton3D((uint32_t)nx, (uint32_t)ny, (uint32_t)nz);
}
// r.rdx/rdy/rdz must be set
__device__ __forceinline__ float ngp_march_exit_fast(const MarchRay &r, const MarchCfg &c, const float *mipb, float t0,
                                                     uint32_t nxyz, int level) {
    const int nx = (int)(nxyz & 1023u), ny = (int)((nxyz >> 10) & 1023u), nz = (int)(nxyz >> 20);
    const float x = ngp_clampf(fmaf(t0, r.dx, r.ox), -c.bound, c.bound);
    const float y = ngp_clampf(fmaf(t0, r.dy, r.oy), -c.bound, c.bound);
    const float z = ngp_clampf(fmaf(t0, r.dz, r.oz), -c.bound, c.bound);
    const float mip_bound = mipb[level];
    const float tx = fmaf(fmaf(((float)nx + 0.5f + 0.5f * copysignf(1.0f, r.dx)) * c.rH, 2.0f, -1.0f), mip_bound, -x) * r.rdx;
    const float ty = fmaf(fmaf(((float)ny + 0.5f + 0.5f * copysignf(1.0f, r.dy)) * c.rH, 2.0f, -1.0f), mip_bound, -y) * r.rdy;
    const float tz = fmaf(fmaf(((float)nz + 0.5f + 0.5f * copysignf(1.0f, r.dz)) * c.rH, 2.0f, -1.0f), mip_bound, -z) * r.rdz;
    return t0 + fmaxf(0.0f, fminf(tx, fminf(ty, tz)));
}
// r.rdx/rdy/rdz must be set
__device__ __forceinline__ float ngp_march_exit(const MarchRay &r, const MarchCfg &c, float t0, uint32_t index) {
    const uint32_t level = index / c.H3, m = index - level * c.H3;
    const int nx = (int)ngp_morton_compact(m), ny = (int)ngp_morton_compact(m >> 1), nz = (int)ngp_morton_compact(m >> 2);
    const float x = ngp_clampf(fmaf(t0, r.dx, r.ox), -c.bound, c.bound);
    const float y = ngp_clampf(fmaf(t0, r.dy, r.oy), -c.bound, c.bound);
    const float z = ngp_clampf(fmaf(t0, r.dz, r.oz), -c.bound, c.bound);
    const float mip_bound = fminf((float)(1 << level), c.bound);
    const float tx = fmaf(fmaf(((float)nx + 0.5f + 0.5f * copysignf(1.0f, r.dx)) * c.rH, 2.0f, -1.0f), mip_bound, -x) * r.rdx;
    const float ty = fmaf(fmaf(((float)ny + 0.5f + 0.5f * copysignf(1.0f, r.dy)) * c.rH, 2.0f, -1.0f), mip_bound, -y) * r.rdy;
    const float tz = fmaf(fmaf(((float)nz + 0.5f + 0.5f * copysignf(1.0f, r.dz)) * c.rH, 2.0f, -1.0f), mip_bound, -z) * r.rdz;
    return t0 + fmaxf(0.0f, fminf(tx, fminf(ty, tz)));
}

// One step of the parameter sequence of a ray: t_{k+1} = t_k + clamp(t_k dt_gamma, dt_min, dt_max).  Both branches of the
// marcher advance by this recurrence (occupied: t += dt, :478; empty: the do-while of :398-402), so the sequence a ray
// visits does not depend on the occupancy — only which of its elements are visited does.
__device__ __forceinline__ float ngp_march_advance(float t, const MarchCfg &c) {
    return t + ngp_clampf(t * c.dt_gamma, c.dt_min, c.dt_max);
}

// One marcher iteration at parameter t.  Occupied: returns true, sample position in (px,py,pz), step in dt, t unchanged.
// Empty: returns false with t advanced past the voxel (raymarching.cu:393-402).
__device__ __forceinline__ bool ngp_march_step(const MarchRay &r, const MarchCfg &c, float &t, float &px, float &py,
                                               float &pz, float &dt) {
    float tt;
    if (ngp_march_probe(r, c, t, px, py, pz, dt, tt)) return true;
    float tc = t;
    do {
        tc = ngp_march_advance(tc, c);
    } while (tc < tt);
    t = tc;
    return false;
}

// AABB slab test of one ray (raymarching.cu:93-147): near clamped to min_near; a miss returns FLT_MAX for both.
__device__ __forceinline__ void ngp_near_far_ray(const float (&o)[3], const float (&d)[3], const float *__restrict__ aabb,
                                                 float min_near, float &near, float &far) {
    float tn = 0.0f, tf = 0.0f;
    bool miss = false;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float r = 1.0f / d[a];
        float lo = (__ldg(aabb + a) - o[a]) * r, hi = (__ldg(aabb + a + 3) - o[a]) * r;
        if (lo > hi) { const float t = lo; lo = hi; hi = t; }
        if (a == 0) { tn = lo; tf = hi; continue; }
        if (!miss) {
            if (tn > hi || lo > tf) miss = true;
            else {
                if (lo > tn) tn = lo;
                if (hi < tf) tf = hi;
            }
        }
    }
    if (miss) { near = 3.402823466e+38f; far = 3.402823466e+38f; return; }
    if (tn < min_near) tn = min_near;
    near = tn;
    far = tf;
}

// ---- PCG32 (O'Neill) as used by raymarching/src/pcg32.h: seed(:57-72), advance(:149-170), next_float(:107-116)
struct NgpPcg32 {
    uint64_t state, inc;
};
__device__ __forceinline__ uint32_t ngp_pcg_next(NgpPcg32 &r) {
    const uint64_t old = r.state;
    r.state = old * 0x5851f42d4c957f2dULL + r.inc;
    const uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    const uint32_t rot = (uint32_t)(old >> 59u);
    return (xs >> rot) | (xs << ((~rot + 1u) & 31));
}
__device__ __forceinline__ float ngp_pcg_jitter(uint64_t seed, uint64_t n) {
    NgpPcg32 r;
    r.state = 0;
    r.inc = (1ull << 1) | 1ull;
    ngp_pcg_next(r);
    r.state += seed;
    ngp_pcg_next(r);
    uint64_t cur_mult = 0x5851f42d4c957f2dULL, cur_plus = r.inc, acc_mult = 1, acc_plus = 0, delta = n;
    while (delta > 0) {
        if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
        cur_plus = (cur_mult + 1) * cur_plus;
        cur_mult *= cur_mult;
        delta >>= 1;
    }
    r.state = acc_mult * r.state + acc_plus;
    return __uint_as_float((ngp_pcg_next(r) >> 9) | 0x3f800000u) - 1.0f;
}
