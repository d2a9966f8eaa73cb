// common.cuh — shared device/host helpers for libngp_b200 (sm_100a only).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/ngp_b200.h"

#define NGP_SM_COUNT 148

void ngp_set_error(const char *fmt, ...);

#define NGP_REQUIRE(cond, code, ...)      \
    do {                                  \
        if (!(cond)) {                    \
            ngp_set_error(__VA_ARGS__);   \
            return (code);                \
        }                                 \
    } while (0)

#define NGP_CHECK_LAUNCH(name)                                                          \
    do {                                                                                \
        cudaError_t e__ = cudaGetLastError();                                           \
        if (e__ != cudaSuccess) {                                                       \
            ngp_set_error("%s: launch failed: %s", name, cudaGetErrorString(e__));      \
            return NGP_ERR_CUDA;                                                        \
        }                                                                               \
    } while (0)

// The caller's CURRENT device is part of the contract (include/ngp_b200.h): every entry point checks that its first
// device pointer lives on it (one cudaPointerGetAttributes, ~1 us; ngp_set_device_check(0) switches it off).
int ngp_device_mismatch(const void *p, const char *fn);
#define NGP_REQUIRE_DEVICE(p, fn)                      \
    do {                                               \
        const int rc__ = ngp_device_mismatch((p), (fn)); \
        if (rc__) return rc__;                         \
    } while (0)

static inline int64_t ngp_div_up(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------------------------------------
// Hash-grid level geometry (reference semantics: gridencoder/src/gridencoder.cu:35-72, 126-127)
// ---------------------------------------------------------------------------------------------
struct LevelMeta {
    float scale;         // exp2f(level*S)*H - 1
    uint32_t rows;       // hashmap_size of the level
    uint32_t row_offset; // first row of the level in the table
    uint32_t m1, m2;     // dense strides for dims 1, 2 (0 when the stride walk stopped before that dim)
    uint32_t hashed;     // 1: use the xor-prime hash
    uint32_t pow2mask;   // rows-1 when rows is a power of two, else 0
};

__device__ __forceinline__ float ngp_level_scale(uint32_t level, float S, uint32_t H) {
    // same expression as the reference kernel; the product with H and the -1 contract to one fma there
    return fmaf(exp2f((float)level * S), (float)H, -1.0f);
}

template <int D>
__device__ __forceinline__ LevelMeta ngp_level_meta(uint32_t level, const int32_t *__restrict__ offsets, float S,
                                                    uint32_t H, uint32_t gridtype, int align_corners) {
    LevelMeta m;
    m.scale = ngp_level_scale(level, S, H);
    const uint32_t o0 = (uint32_t)offsets[level], o1 = (uint32_t)offsets[level + 1];
    m.rows = o1 - o0;
    m.row_offset = o0;
    const uint32_t resolution = (uint32_t)ceilf(m.scale) + 1u;
    const uint32_t step = align_corners ? resolution : resolution + 1u;
    uint32_t stride = 1;
    uint32_t mm[3] = {0, 0, 0};
#pragma unroll
    for (int d = 0; d < D; ++d) {
        if (stride <= m.rows) {
            mm[d] = stride;
            stride *= step;
        }
    }
    m.m1 = mm[1];
    m.m2 = mm[2];
    m.hashed = (gridtype == 0 && stride > m.rows) ? 1u : 0u;
    m.pow2mask = ((m.rows & (m.rows - 1u)) == 0u) ? m.rows - 1u : 0u;
    return m;
}

__device__ __forceinline__ uint32_t ngp_wrap(uint32_t idx, const LevelMeta &m) {
    if (m.pow2mask) return idx & m.pow2mask;
    return idx < m.rows ? idx : idx % m.rows;
}

// row index of the corner (x, y, z) — level-local
__device__ __forceinline__ uint32_t ngp_row3(uint32_t x, uint32_t y, uint32_t z, const LevelMeta &m) {
    uint32_t idx = m.hashed ? (x ^ (y * 2654435761u) ^ (z * 805459861u)) : (x + y * m.m1 + z * m.m2);
    return ngp_wrap(idx, m);
}
__device__ __forceinline__ uint32_t ngp_row2(uint32_t x, uint32_t y, const LevelMeta &m) {
    uint32_t idx = m.hashed ? (x ^ (y * 2654435761u)) : (x + y * m.m1);
    return ngp_wrap(idx, m);
}

// table element loads (read-only path), C channels starting at row
template <typename T>
__device__ __forceinline__ float ngp_ld1(const T *p);
template <>
__device__ __forceinline__ float ngp_ld1<float>(const float *p) { return __ldg(p); }
template <>
__device__ __forceinline__ float ngp_ld1<__half>(const __half *p) { return __half2float(__ldg(p)); }

__device__ __forceinline__ float2 ngp_ld2(const float *p) { return __ldg(reinterpret_cast<const float2 *>(p)); }
__device__ __forceinline__ float2 ngp_ld2(const __half *p) {
    return __half22float2(__ldg(reinterpret_cast<const __half2 *>(p)));
}
