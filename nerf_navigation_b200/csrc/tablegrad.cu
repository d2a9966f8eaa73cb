// tablegrad.cu — hash-table gradient as a deterministic sorted segmented reduction (no float atomics).
// Reference: kernel_grid_backward (gridencoder/src/gridencoder.cu:227-314) scatters w*grad with atomicAdd, so its
// sums depend on the (arbitrary) arrival order.  Here every contribution (point b, level l, corner k) gets the key
// `global table row`; a stable LSD radix sort (CUB, plumbing) orders contributions by row while keeping them in
// ascending (b,l,k) order inside a row; one thread per row segment then adds its contributions in that fixed order.
// Result: bit-identical gradients run to run, and only rows that were touched are written (others are memset 0).
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

#define TG_THREADS 256
#define TG_MAX_L 32

template <int D>
__device__ __forceinline__ bool tg_locate(const float *__restrict__ inputs, int64_t b, const LevelMeta &m, int align,
                                          uint32_t (&cell)[D], float (&frac)[D]) {
    bool oob = false;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const float x = inputs[b * D + d];
        oob |= (x < 0.0f || x > 1.0f);
        const float p = fmaf(x, m.scale, align ? 0.0f : 0.5f);
        cell[d] = (uint32_t)floorf(p);
        frac[d] = p - (float)cell[d];
    }
    return !oob;
}

template <int D>
__device__ __forceinline__ uint32_t tg_row(const uint32_t (&cell)[D], int k, const LevelMeta &m) {
    if (D == 3) return ngp_row3(cell[0] + (k & 1), cell[1] + ((k >> 1) & 1), cell[D - 1] + ((k >> 2) & 1), m);
    return ngp_row2(cell[0] + (k & 1), cell[1] + ((k >> 1) & 1), m);
}

// one thread per (point, level): emits 2^D keys; payload = contribution id
template <int D>
__global__ void __launch_bounds__(TG_THREADS)
k_tg_keys(const float *__restrict__ inputs, const int32_t *__restrict__ offsets, int64_t B, int L, float S, uint32_t H,
          uint32_t gridtype, int align, uint32_t sentinel, uint32_t *__restrict__ keys, uint32_t *__restrict__ ids) {
    __shared__ LevelMeta meta[TG_MAX_L];
    if (threadIdx.x < L) meta[threadIdx.x] = ngp_level_meta<D>(threadIdx.x, offsets, S, H, gridtype, align);
    __syncthreads();
    const int64_t t = (int64_t)blockIdx.x * TG_THREADS + threadIdx.x;
    if (t >= B * L) return;
    const int64_t b = t / L;
    const int l = (int)(t - b * L);
    const LevelMeta m = meta[l];
    uint32_t cell[D];
    float frac[D];
    const bool ok = tg_locate<D>(inputs, b, m, align, cell, frac);
    constexpr int NC = 1 << D;
#pragma unroll
    for (int k = 0; k < NC; ++k) {
        keys[t * NC + k] = ok ? m.row_offset + tg_row<D>(cell, k, m) : sentinel;
        ids[t * NC + k] = (uint32_t)(t * NC + k);
    }
}

// one thread per sorted element; segment heads add their segment in sorted (= ascending contribution id) order
template <typename G, int D, int C>
__global__ void __launch_bounds__(TG_THREADS)
k_tg_reduce(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ ids, int64_t n, uint32_t sentinel,
            const G *__restrict__ grad, int grad_layout, const float *__restrict__ inputs,
            const int32_t *__restrict__ offsets, int64_t B, int L, float S, uint32_t H, uint32_t gridtype, int align,
            float *__restrict__ grad_table) {
    __shared__ LevelMeta meta[TG_MAX_L];
    if (threadIdx.x < L) meta[threadIdx.x] = ngp_level_meta<D>(threadIdx.x, offsets, S, H, gridtype, align);
    __syncthreads();
    const int64_t i = (int64_t)blockIdx.x * TG_THREADS + threadIdx.x;
    if (i >= n) return;
    const uint32_t key = keys[i];
    if (key == sentinel) return;
    if (i > 0 && keys[i - 1] == key) return;  // not a segment head
    constexpr int NC = 1 << D;
    float acc[C];
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] = 0.0f;
    for (int64_t j = i; j < n && keys[j] == key; ++j) {
        const uint32_t id = ids[j];
        const int k = (int)(id % NC);
        const int64_t t = id / NC;
        const int64_t b = t / L;
        const int l = (int)(t - b * L);
        uint32_t cell[D];
        float frac[D];
        tg_locate<D>(inputs, b, meta[l], align, cell, frac);
        float w = 1.0f;
#pragma unroll
        for (int d = 0; d < D; ++d) w *= (k >> d) & 1 ? frac[d] : 1.0f - frac[d];
        const G *g = grad_layout == 0 ? grad + ((int64_t)l * B + b) * C : grad + (b * L + l) * C;
#pragma unroll
        for (int c = 0; c < C; ++c) acc[c] = __fadd_rn(acc[c], __fmul_rn(w, (float)g[c]));  // product then add, as the atomics do
    }
#pragma unroll
    for (int c = 0; c < C; ++c) grad_table[(int64_t)key * C + c] = acc[c];
}

static int tg_bits(uint32_t v) {
    int b = 0;
    while ((1ull << b) <= v) ++b;
    return b;
}

extern "C" int64_t ngp_grid_backward_table_workspace_bytes(int64_t B, int D, int L) {
    const int64_t n = B * L * (1ll << D);
    // 2 key buffers + 2 payload buffers + radix-sort scratch (histograms / tile descriptors: well under n/4 + 16 MiB)
    return 4 * n * 4 + n / 4 + (16ll << 20);
}

template <typename G, int D>
static int tg_launch_reduce(int C, const uint32_t *keys, const uint32_t *ids, int64_t n, uint32_t sentinel,
                            const void *grad, int grad_layout, const float *inputs, const int32_t *offsets, int64_t B,
                            int L, float S, uint32_t H, uint32_t gridtype, int align, float *gt, cudaStream_t st) {
    const unsigned grid = (unsigned)ngp_div_up(n, TG_THREADS);
#define TG_RED(CC)                                                                                                       \
    k_tg_reduce<G, D, CC><<<grid, TG_THREADS, 0, st>>>(keys, ids, n, sentinel, (const G *)grad, grad_layout, inputs, \
                                                       offsets, B, L, S, H, gridtype, align, gt)
    switch (C) {
        case 1: TG_RED(1); break;
        case 2: TG_RED(2); break;
        case 4: TG_RED(4); break;
        case 8: TG_RED(8); break;
        default: ngp_set_error("GridEncoder: C must be 1, 2, 4 or 8 (got %d)", C); return NGP_ERR_UNSUPPORTED;
    }
#undef TG_RED
    return NGP_OK;
}

extern "C" int ngp_grid_encode_backward_table(const void *grad, int grad_layout, int grad_dtype, const float *inputs,
                                              const int32_t *offsets, const int32_t *offsets_host,
                                              float *grad_embeddings, int64_t B, int D, int C, int L, float S,
                                              uint32_t H, uint32_t gridtype, int align_corners, void *workspace,
                                              int64_t workspace_bytes, ngp_stream_t stream) {
    NGP_REQUIRE(grad && inputs && offsets && offsets_host && grad_embeddings, NGP_ERR_ARG,
                "ngp_grid_encode_backward_table: null pointer");
    NGP_REQUIRE(L > 0 && L <= TG_MAX_L, NGP_ERR_ARG, "ngp_grid_encode_backward_table: L out of range");
    NGP_REQUIRE(D == 2 || D == 3, NGP_ERR_UNSUPPORTED, "GridEncoder: D must be 2 or 3 (got %d)", D);
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t total_rows = (uint32_t)offsets_host[L];
    cudaError_t e = cudaMemsetAsync(grad_embeddings, 0, (size_t)total_rows * C * sizeof(float), st);
    NGP_REQUIRE(e == cudaSuccess, NGP_ERR_CUDA, "ngp_grid_encode_backward_table: memset: %s", cudaGetErrorString(e));
    if (B == 0) return NGP_OK;
    const int64_t n = B * L * (1ll << D);
    NGP_REQUIRE(n < (1ll << 31), NGP_ERR_UNSUPPORTED, "ngp_grid_encode_backward_table: B*L*2^D must be < 2^31");
    NGP_REQUIRE(workspace && workspace_bytes >= ngp_grid_backward_table_workspace_bytes(B, D, L), NGP_ERR_ARG,
                "ngp_grid_encode_backward_table: workspace too small");
    uint32_t *keys_a = (uint32_t *)workspace, *keys_b = keys_a + n, *ids_a = keys_b + n, *ids_b = ids_a + n;
    void *cub_tmp = (void *)(ids_b + n);
    size_t cub_avail = (size_t)workspace_bytes - (size_t)(4 * n * 4);
    const uint32_t sentinel = total_rows;
    const unsigned kgrid = (unsigned)ngp_div_up(B * L, TG_THREADS);
    if (D == 3)
        k_tg_keys<3><<<kgrid, TG_THREADS, 0, st>>>(inputs, offsets, B, L, S, H, gridtype, align_corners, sentinel, keys_a, ids_a);
    else
        k_tg_keys<2><<<kgrid, TG_THREADS, 0, st>>>(inputs, offsets, B, L, S, H, gridtype, align_corners, sentinel, keys_a, ids_a);
    NGP_CHECK_LAUNCH("ngp_grid_encode_backward_table(keys)");
    cub::DoubleBuffer<uint32_t> dk(keys_a, keys_b), dv(ids_a, ids_b);
    size_t need = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, need, dk, dv, (int)n, 0, tg_bits(sentinel), st);
    NGP_REQUIRE(need <= cub_avail, NGP_ERR_ARG, "ngp_grid_encode_backward_table: sort scratch %zu > %zu", need, cub_avail);
    e = cub::DeviceRadixSort::SortPairs(cub_tmp, need, dk, dv, (int)n, 0, tg_bits(sentinel), st);
    NGP_REQUIRE(e == cudaSuccess, NGP_ERR_CUDA, "ngp_grid_encode_backward_table: sort: %s", cudaGetErrorString(e));
    int rc;
    if (grad_dtype == NGP_DTYPE_F32)
        rc = D == 3 ? tg_launch_reduce<float, 3>(C, dk.Current(), dv.Current(), n, sentinel, grad, grad_layout, inputs,
                                                 offsets, B, L, S, H, gridtype, align_corners, grad_embeddings, st)
                    : tg_launch_reduce<float, 2>(C, dk.Current(), dv.Current(), n, sentinel, grad, grad_layout, inputs,
                                                 offsets, B, L, S, H, gridtype, align_corners, grad_embeddings, st);
    else if (grad_dtype == NGP_DTYPE_F16)
        rc = D == 3 ? tg_launch_reduce<__half, 3>(C, dk.Current(), dv.Current(), n, sentinel, grad, grad_layout, inputs,
                                                  offsets, B, L, S, H, gridtype, align_corners, grad_embeddings, st)
                    : tg_launch_reduce<__half, 2>(C, dk.Current(), dv.Current(), n, sentinel, grad, grad_layout, inputs,
                                                  offsets, B, L, S, H, gridtype, align_corners, grad_embeddings, st);
    else {
        ngp_set_error("ngp_grid_encode_backward_table: bad grad_dtype");
        return NGP_ERR_UNSUPPORTED;
    }
    if (rc != NGP_OK) return rc;
    NGP_CHECK_LAUNCH("ngp_grid_encode_backward_table(reduce)");
    return NGP_OK;
}
