// grid.cu — standalone multiresolution hash-grid encoder (forward, input-gradient by re-gather).
// Reference semantics: gridencoder/src/gridencoder.cu:75-224 (forward + dy_dx), :317-343 (input backward).
// B200 design: one thread per point walks all levels (x read once, level metadata in shared memory),
// corners are independent read-only loads so 2^D gathers per level are in flight per thread; the backward
// re-gathers the corners instead of round-tripping a [B, L*D*C] dy_dx tensor through HBM.
#include <stdarg.h>
#include <string.h>

#include <stdlib.h>

#include "common.cuh"

// ---------------------------------------------------------------- error state (one TU owns it)
static thread_local char g_err[512] = "";
void ngp_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
extern "C" const char *ngp_last_error(void) { return g_err; }
extern "C" int ngp_abi_version(void) { return 2; }

// device contract check (common.cuh): -1 = read NGP_B200_CHECK_DEVICE on first use (default on)
static int g_device_check = -1;
extern "C" int ngp_set_device_check(int on) {
    const int prev = g_device_check < 0 ? 1 : g_device_check;
    if (on >= 0) g_device_check = on ? 1 : 0;
    return prev;
}
int ngp_device_mismatch(const void *p, const char *fn) {
    if (g_device_check < 0) {
        const char *e = getenv("NGP_B200_CHECK_DEVICE");
        g_device_check = (e && e[0] == '0') ? 0 : 1;
    }
    if (!g_device_check || !p) return NGP_OK;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return NGP_OK;
    }
    if (a.type != cudaMemoryTypeDevice && a.type != cudaMemoryTypeManaged) {
        ngp_set_error("%s: expected a device pointer (got host / unregistered memory)", fn);
        return NGP_ERR_ARG;
    }
    int dev = 0;
    cudaGetDevice(&dev);
    if (a.type == cudaMemoryTypeDevice && a.device != dev) {
        ngp_set_error("%s: buffers live on device %d but the current device is %d (select the device before the call)", fn,
                      a.device, dev);
        return NGP_ERR_ARG;
    }
    return NGP_OK;
}

#define GRID_MAX_L 32
#define GRID_THREADS 256

__global__ void k_level_scales(float *scales, int L, float S, uint32_t H) {
    const int l = threadIdx.x;
    if (l < L) scales[l] = ngp_level_scale((uint32_t)l, S, H);
}

extern "C" int ngp_grid_level_scales(float *scales, int L, float S, uint32_t H, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(scales, "ngp_grid_level_scales");
    NGP_REQUIRE(scales && L > 0 && L <= GRID_MAX_L, NGP_ERR_ARG, "ngp_grid_level_scales: bad arguments");
    k_level_scales<<<1, GRID_MAX_L, 0, (cudaStream_t)stream>>>(scales, L, S, H);
    NGP_CHECK_LAUNCH("ngp_grid_level_scales");
    return NGP_OK;
}

template <typename T>
__device__ __forceinline__ void st1(T *p, float v);
template <>
__device__ __forceinline__ void st1<float>(float *p, float v) { *p = v; }
template <>
__device__ __forceinline__ void st1<__half>(__half *p, float v) { *p = __float2half(v); }

// cell position of a point at one level; returns false if any coordinate is outside [0,1]
template <int D>
__device__ __forceinline__ void locate(const float (&x)[D], const LevelMeta &m, int align_corners, uint32_t (&cell)[D],
                                       float (&frac)[D]) {
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const float p = fmaf(x[d], m.scale, align_corners ? 0.0f : 0.5f);
        cell[d] = (uint32_t)floorf(p);
        frac[d] = p - (float)cell[d];
    }
}

template <int D>
__device__ __forceinline__ uint32_t corner_row(const uint32_t (&cell)[D], int corner, const LevelMeta &m) {
    if (D == 3)
        return ngp_row3(cell[0] + (corner & 1), cell[1] + ((corner >> 1) & 1), cell[D - 1] + ((corner >> 2) & 1), m);
    return ngp_row2(cell[0] + (corner & 1), cell[1] + ((corner >> 1) & 1), m);
}

template <typename T, int D, int C>
__global__ void __launch_bounds__(GRID_THREADS)
k_grid_forward(const float *__restrict__ inputs, const T *__restrict__ table, const int32_t *__restrict__ offsets,
               T *__restrict__ outputs, int out_layout, int64_t B, int L, float S, uint32_t H, int calc_grad,
               T *__restrict__ dy_dx, uint32_t gridtype, int align_corners, uint32_t *__restrict__ indices) {
    __shared__ LevelMeta meta[GRID_MAX_L];
    if (threadIdx.x < L) meta[threadIdx.x] = ngp_level_meta<D>(threadIdx.x, offsets, S, H, gridtype, align_corners);
    __syncthreads();
    const int64_t b = (int64_t)blockIdx.x * GRID_THREADS + threadIdx.x;
    if (b >= B) return;
    float x[D];
    bool oob = false;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        x[d] = inputs[b * D + d];
        oob |= (x[d] < 0.0f || x[d] > 1.0f);
    }
    constexpr int NC = 1 << D;
    for (int l = 0; l < L; ++l) {
        const LevelMeta m = meta[l];
        T *o = out_layout == 0 ? outputs + ((int64_t)l * B + b) * C : outputs + (b * L + l) * C;
        T *g = (calc_grad && dy_dx) ? dy_dx + ((b * L + l) * D) * C : nullptr;
        uint32_t *ix = indices ? indices + (b * L + l) * NC : nullptr;
        if (oob) {  // gridencoder.cu:99-123
#pragma unroll
            for (int c = 0; c < C; ++c) st1(o + c, 0.0f);
            if (g)
                for (int k = 0; k < D * C; ++k) st1(g + k, 0.0f);
            if (ix)
                for (int k = 0; k < NC; ++k) ix[k] = 0xFFFFFFFFu;
            continue;
        }
        uint32_t cell[D];
        float frac[D];
        locate<D>(x, m, align_corners, cell, frac);
        const T *lvl = table + (int64_t)m.row_offset * C;
        float val[NC][C];
        uint32_t rows[NC];
#pragma unroll
        for (int k = 0; k < NC; ++k) rows[k] = corner_row<D>(cell, k, m);
#pragma unroll
        for (int k = 0; k < NC; ++k) {
            if (C == 2) {
                const float2 v = ngp_ld2(lvl + (int64_t)rows[k] * 2);
                val[k][0] = v.x;
                val[k][C - 1] = v.y;
            } else {
#pragma unroll
                for (int c = 0; c < C; ++c) val[k][c] = ngp_ld1(lvl + (int64_t)rows[k] * C + c);
            }
        }
        float acc[C];
#pragma unroll
        for (int c = 0; c < C; ++c) acc[c] = 0.0f;
#pragma unroll
        for (int k = 0; k < NC; ++k) {
            float w = 1.0f;
#pragma unroll
            for (int d = 0; d < D; ++d) w *= (k >> d) & 1 ? frac[d] : 1.0f - frac[d];
#pragma unroll
            for (int c = 0; c < C; ++c) acc[c] = fmaf(w, val[k][c], acc[c]);
        }
#pragma unroll
        for (int c = 0; c < C; ++c) st1(o + c, acc[c]);
        if (ix) {
#pragma unroll
            for (int k = 0; k < NC; ++k) ix[k] = rows[k];
        }
        if (g) {  // gridencoder.cu:180-223
#pragma unroll
            for (int gd = 0; gd < D; ++gd) {
                float ga[C];
#pragma unroll
                for (int c = 0; c < C; ++c) ga[c] = 0.0f;
#pragma unroll
                for (int k = 0; k < (1 << (D - 1)); ++k) {
                    float w = m.scale;
                    int corner = 0;
#pragma unroll
                    for (int nd = 0; nd < D - 1; ++nd) {
                        const int d = nd >= gd ? nd + 1 : nd;
                        const int bit = (k >> nd) & 1;
                        w *= bit ? frac[d] : 1.0f - frac[d];
                        corner |= bit << d;
                    }
#pragma unroll
                    for (int c = 0; c < C; ++c) {
                        // for a half table the reference subtracts in half precision; values here are exact floats
                        ga[c] = fmaf(w, val[corner | (1 << gd)][c] - val[corner][c], ga[c]);
                    }
                }
#pragma unroll
                for (int c = 0; c < C; ++c) st1(g + gd * C + c, ga[c]);
            }
        }
    }
}

// dL/dx01[b,d] = sum_{l,c} grad[b,l,c] * dy_dx[b,l,d,c], corners re-gathered (no stored dy_dx)
template <typename T, int D, int C>
__global__ void __launch_bounds__(GRID_THREADS)
k_grid_backward_input(const T *__restrict__ grad, int grad_layout, const float *__restrict__ inputs,
                      const T *__restrict__ table, const int32_t *__restrict__ offsets, float *__restrict__ grad_inputs,
                      int64_t B, int L, float S, uint32_t H, uint32_t gridtype, int align_corners) {
    __shared__ LevelMeta meta[GRID_MAX_L];
    if (threadIdx.x < L) meta[threadIdx.x] = ngp_level_meta<D>(threadIdx.x, offsets, S, H, gridtype, align_corners);
    __syncthreads();
    const int64_t b = (int64_t)blockIdx.x * GRID_THREADS + threadIdx.x;
    if (b >= B) return;
    float x[D], gx[D];
    bool oob = false;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        x[d] = inputs[b * D + d];
        gx[d] = 0.0f;
        oob |= (x[d] < 0.0f || x[d] > 1.0f);
    }
    constexpr int NC = 1 << D;
    if (!oob) {
        for (int l = 0; l < L; ++l) {
            const LevelMeta m = meta[l];
            const T *gp = grad_layout == 0 ? grad + ((int64_t)l * B + b) * C : grad + (b * L + l) * C;
            float gl[C];
#pragma unroll
            for (int c = 0; c < C; ++c) gl[c] = ngp_ld1(gp + c);
            uint32_t cell[D];
            float frac[D];
            locate<D>(x, m, align_corners, cell, frac);
            const T *lvl = table + (int64_t)m.row_offset * C;
            float val[NC][C];
#pragma unroll
            for (int k = 0; k < NC; ++k) {
                const uint32_t row = corner_row<D>(cell, k, m);
                if (C == 2) {
                    const float2 v = ngp_ld2(lvl + (int64_t)row * 2);
                    val[k][0] = v.x;
                    val[k][C - 1] = v.y;
                } else {
#pragma unroll
                    for (int c = 0; c < C; ++c) val[k][c] = ngp_ld1(lvl + (int64_t)row * C + c);
                }
            }
#pragma unroll
            for (int gd = 0; gd < D; ++gd) {
                float ga[C];
#pragma unroll
                for (int c = 0; c < C; ++c) ga[c] = 0.0f;
#pragma unroll
                for (int k = 0; k < (1 << (D - 1)); ++k) {
                    float w = m.scale;
                    int corner = 0;
#pragma unroll
                    for (int nd = 0; nd < D - 1; ++nd) {
                        const int d = nd >= gd ? nd + 1 : nd;
                        const int bit = (k >> nd) & 1;
                        w *= bit ? frac[d] : 1.0f - frac[d];
                        corner |= bit << d;
                    }
#pragma unroll
                    for (int c = 0; c < C; ++c) ga[c] = fmaf(w, val[corner | (1 << gd)][c] - val[corner][c], ga[c]);
                }
#pragma unroll
                for (int c = 0; c < C; ++c) gx[gd] = fmaf(gl[c], ga[c], gx[gd]);
            }
        }
    }
#pragma unroll
    for (int d = 0; d < D; ++d) grad_inputs[b * D + d] = gx[d];
}

template <typename T, int D>
static int launch_forward_c(int C, const float *inputs, const void *table, const int32_t *offsets, void *outputs,
                            int out_layout, int64_t B, int L, float S, uint32_t H, int calc, void *dy_dx,
                            uint32_t gridtype, int align, uint32_t *indices, cudaStream_t st) {
    const unsigned grid = (unsigned)ngp_div_up(B, GRID_THREADS);
#define LAUNCH_FWD(CC)                                                                                                  \
    k_grid_forward<T, D, CC><<<grid, GRID_THREADS, 0, st>>>(inputs, (const T *)table, offsets, (T *)outputs, out_layout, \
                                                            B, L, S, H, calc, (T *)dy_dx, gridtype, align, indices)
    switch (C) {
        case 1: LAUNCH_FWD(1); break;
        case 2: LAUNCH_FWD(2); break;
        case 4: LAUNCH_FWD(4); break;
        case 8: LAUNCH_FWD(8); break;
        default: ngp_set_error("GridEncoder: C must be 1, 2, 4 or 8 (got %d)", C); return NGP_ERR_UNSUPPORTED;
    }
#undef LAUNCH_FWD
    return NGP_OK;
}

extern "C" int ngp_grid_encode_forward(const float *inputs, const void *embeddings, int table_dtype,
                                       const int32_t *offsets, void *outputs, int out_layout, int64_t B, int D, int C,
                                       int L, float S, uint32_t H, int calc_grad_inputs, void *dy_dx, uint32_t gridtype,
                                       int align_corners, uint32_t *indices, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(inputs, "ngp_grid_encode_forward");
    if (B == 0) return NGP_OK;
    NGP_REQUIRE(inputs && embeddings && offsets && outputs, NGP_ERR_ARG, "ngp_grid_encode_forward: null pointer");
    NGP_REQUIRE(L > 0 && L <= GRID_MAX_L, NGP_ERR_ARG, "ngp_grid_encode_forward: L must be in [1,%d]", GRID_MAX_L);
    NGP_REQUIRE(D == 2 || D == 3, NGP_ERR_UNSUPPORTED, "GridEncoder: D must be 2 or 3 (got %d)", D);
    NGP_REQUIRE(out_layout == 0 || out_layout == 1, NGP_ERR_ARG, "ngp_grid_encode_forward: bad out_layout");
    cudaStream_t st = (cudaStream_t)stream;
    int rc;
    if (table_dtype == NGP_DTYPE_F32)
        rc = D == 3 ? launch_forward_c<float, 3>(C, inputs, embeddings, offsets, outputs, out_layout, B, L, S, H,
                                                 calc_grad_inputs, dy_dx, gridtype, align_corners, indices, st)
                    : launch_forward_c<float, 2>(C, inputs, embeddings, offsets, outputs, out_layout, B, L, S, H,
                                                 calc_grad_inputs, dy_dx, gridtype, align_corners, indices, st);
    else if (table_dtype == NGP_DTYPE_F16)
        rc = D == 3 ? launch_forward_c<__half, 3>(C, inputs, embeddings, offsets, outputs, out_layout, B, L, S, H,
                                                  calc_grad_inputs, dy_dx, gridtype, align_corners, indices, st)
                    : launch_forward_c<__half, 2>(C, inputs, embeddings, offsets, outputs, out_layout, B, L, S, H,
                                                  calc_grad_inputs, dy_dx, gridtype, align_corners, indices, st);
    else {
        ngp_set_error("ngp_grid_encode_forward: table_dtype must be 0 (f32) or 1 (f16)");
        return NGP_ERR_UNSUPPORTED;
    }
    if (rc != NGP_OK) return rc;
    NGP_CHECK_LAUNCH("ngp_grid_encode_forward");
    return NGP_OK;
}

template <typename T, int D>
static int launch_bwd_input_c(int C, const void *grad, int grad_layout, const float *inputs, const void *table,
                              const int32_t *offsets, float *grad_inputs, int64_t B, int L, float S, uint32_t H,
                              uint32_t gridtype, int align, cudaStream_t st) {
    const unsigned grid = (unsigned)ngp_div_up(B, GRID_THREADS);
#define LAUNCH_BWD(CC)                                                                                             \
    k_grid_backward_input<T, D, CC><<<grid, GRID_THREADS, 0, st>>>((const T *)grad, grad_layout, inputs,           \
                                                                   (const T *)table, offsets, grad_inputs, B, L, S, \
                                                                   H, gridtype, align)
    switch (C) {
        case 1: LAUNCH_BWD(1); break;
        case 2: LAUNCH_BWD(2); break;
        case 4: LAUNCH_BWD(4); break;
        case 8: LAUNCH_BWD(8); break;
        default: ngp_set_error("GridEncoder: C must be 1, 2, 4 or 8 (got %d)", C); return NGP_ERR_UNSUPPORTED;
    }
#undef LAUNCH_BWD
    return NGP_OK;
}

extern "C" int ngp_grid_encode_backward_input(const void *grad, int grad_layout, const float *inputs,
                                              const void *embeddings, int table_dtype, const int32_t *offsets,
                                              float *grad_inputs, int64_t B, int D, int C, int L, float S, uint32_t H,
                                              uint32_t gridtype, int align_corners, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(grad, "ngp_grid_encode_backward_input");
    if (B == 0) return NGP_OK;
    NGP_REQUIRE(grad && inputs && embeddings && offsets && grad_inputs, NGP_ERR_ARG,
                "ngp_grid_encode_backward_input: null pointer");
    NGP_REQUIRE(L > 0 && L <= GRID_MAX_L, NGP_ERR_ARG, "ngp_grid_encode_backward_input: L out of range");
    NGP_REQUIRE(D == 2 || D == 3, NGP_ERR_UNSUPPORTED, "GridEncoder: D must be 2 or 3 (got %d)", D);
    cudaStream_t st = (cudaStream_t)stream;
    int rc;
    if (table_dtype == NGP_DTYPE_F32)
        rc = D == 3 ? launch_bwd_input_c<float, 3>(C, grad, grad_layout, inputs, embeddings, offsets, grad_inputs, B, L,
                                                   S, H, gridtype, align_corners, st)
                    : launch_bwd_input_c<float, 2>(C, grad, grad_layout, inputs, embeddings, offsets, grad_inputs, B, L,
                                                   S, H, gridtype, align_corners, st);
    else if (table_dtype == NGP_DTYPE_F16)
        rc = D == 3 ? launch_bwd_input_c<__half, 3>(C, grad, grad_layout, inputs, embeddings, offsets, grad_inputs, B,
                                                    L, S, H, gridtype, align_corners, st)
                    : launch_bwd_input_c<__half, 2>(C, grad, grad_layout, inputs, embeddings, offsets, grad_inputs, B,
                                                    L, S, H, gridtype, align_corners, st);
    else {
        ngp_set_error("ngp_grid_encode_backward_input: bad table_dtype");
        return NGP_ERR_UNSUPPORTED;
    }
    if (rc != NGP_OK) return rc;
    NGP_CHECK_LAUNCH("ngp_grid_encode_backward_input");
    return NGP_OK;
}
