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

// Segmented reduction of the sorted contributions, in two fixed-shape levels (bit-reproducible: the association depends
// only on positions in the sorted order, and the sort is stable):
//   level 1 (k_tg_reduce): one thread per sorted element computes w * grad; a block of 256 consecutive elements does a
//     segmented inclusive scan (warp shuffles, then an in-order carry across its 8 warps).  A run that starts and ends
//     inside the block is written to the table; a run that crosses the block's edges leaves a per-block partial
//     (`lead`: continues from the previous block; `trail`: starts here and continues into the next).
//   level 2 (k_tg_stitch): one thread per block whose trailing run continues adds the following blocks' lead partials
//     in block order.  A row hit by every sample (e.g. the zero padding of march_rays_train) is thus summed as
//     256-element trees + a short serial pass instead of one thread walking 10^5 elements.
struct TgPartial {
    uint32_t trail_key;   // key of the trailing run (valid iff trail_open)
    uint32_t lead_cont;   // the lead run covers the whole block and continues into the next block
    uint32_t trail_open;  // the trailing run starts in this block and continues into the next
    uint32_t pad;
};

template <typename G, int D, int C>
__global__ void __launch_bounds__(TG_THREADS)
k_tg_reduce(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ ids, int64_t n, uint32_t sentinel,
            const G *__restrict__ grad, int grad_layout, const float *__restrict__ inputs,
            const int32_t *__restrict__ offsets, int64_t B, int L, float S, uint32_t H, uint32_t gridtype, int align,
            float *__restrict__ grad_table, TgPartial *__restrict__ part, float *__restrict__ lead_sum,
            float *__restrict__ trail_sum) {
    __shared__ LevelMeta meta[TG_MAX_L];
    __shared__ uint32_t w_first[TG_THREADS / 32], w_last[TG_THREADS / 32];
    __shared__ float w_sum[TG_THREADS / 32][C];
    if (threadIdx.x < L) meta[threadIdx.x] = ngp_level_meta<D>(threadIdx.x, offsets, S, H, gridtype, align);
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t block0 = (int64_t)blockIdx.x * TG_THREADS;
    const int64_t i = block0 + threadIdx.x;
    const bool in = i < n;
    const uint32_t key = in ? keys[i] : 0xffffffffu;   // past-the-end elements form their own run (never written)
    constexpr int NC = 1 << D;
    float val[C];
#pragma unroll
    for (int c = 0; c < C; ++c) val[c] = 0.0f;
    if (in && key != sentinel) {
        const uint32_t id = ids[i];
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
        for (int c = 0; c < C; ++c) val[c] = __fmul_rn(w, (float)g[c]);  // product then add, as the atomics do
    }
    // warp-level segmented inclusive scan (keys are sorted: equal keys are contiguous)
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t ku = __shfl_up_sync(0xffffffffu, key, o);
        const bool take = lane >= o && ku == key;
#pragma unroll
        for (int c = 0; c < C; ++c) {
            const float vu = __shfl_up_sync(0xffffffffu, val[c], o);
            if (take) val[c] = __fadd_rn(val[c], vu);
        }
    }
    if (lane == 0) w_first[warp] = key;
    if (lane == 31) {
        w_last[warp] = key;
#pragma unroll
        for (int c = 0; c < C; ++c) w_sum[warp][c] = val[c];
    }
    __syncthreads();
    // carry from the previous warps of the block for threads in their warp's leading run (fixed order: nearest first)
    if (key == w_first[warp]) {
        for (int w = warp - 1; w >= 0 && w_last[w] == key; --w) {
#pragma unroll
            for (int c = 0; c < C; ++c) val[c] = __fadd_rn(val[c], w_sum[w][c]);
            if (w_first[w] != key) break;
        }
    }
    if (!in || key == sentinel) {
        if (threadIdx.x == TG_THREADS - 1) part[blockIdx.x] = TgPartial{0u, 0u, 0u, 0u};
        if (threadIdx.x == 0) {  // a block starting in the sentinel / past-the-end zone has no lead run worth keeping
#pragma unroll
            for (int c = 0; c < C; ++c) lead_sum[(int64_t)blockIdx.x * C + c] = 0.0f;
        }
        return;
    }
    const bool last_in_block = threadIdx.x == TG_THREADS - 1 || i + 1 == n;
    const uint32_t next_key = i + 1 < n ? keys[i + 1] : 0xffffffffu;
    const bool run_end_here = last_in_block || next_key != key;
    if (!run_end_here) return;
    const uint32_t block_first = w_first[0];
    const bool from_prev = key == block_first && block0 > 0 && keys[block0 - 1] == key;   // run continues from block-1
    const bool to_next = last_in_block && next_key == key;                                // run continues into block+1
    if (from_prev) {
#pragma unroll
        for (int c = 0; c < C; ++c) lead_sum[(int64_t)blockIdx.x * C + c] = val[c];
    } else if (key == block_first) {
#pragma unroll
        for (int c = 0; c < C; ++c) lead_sum[(int64_t)blockIdx.x * C + c] = 0.0f;
    }
    if (last_in_block) {
        TgPartial p;
        p.trail_key = key;
        p.lead_cont = (from_prev && to_next) ? 1u : 0u;
        p.trail_open = (!from_prev && to_next) ? 1u : 0u;
        p.pad = 0u;
        part[blockIdx.x] = p;
        if (p.trail_open) {
#pragma unroll
            for (int c = 0; c < C; ++c) trail_sum[(int64_t)blockIdx.x * C + c] = val[c];
        }
    }
    if (!from_prev && !to_next) {
#pragma unroll
        for (int c = 0; c < C; ++c) grad_table[(int64_t)key * C + c] = val[c];
    }
}

// level 2: stitch the runs that cross block edges, in block order
template <int C>
__global__ void __launch_bounds__(TG_THREADS)
k_tg_stitch(const TgPartial *__restrict__ part, const float *__restrict__ lead_sum, const float *__restrict__ trail_sum,
            int64_t nblocks, float *__restrict__ grad_table) {
    const int64_t b = (int64_t)blockIdx.x * TG_THREADS + threadIdx.x;
    if (b >= nblocks) return;
    const TgPartial p = part[b];
    if (!p.trail_open) return;
    float acc[C];
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] = trail_sum[b * C + c];
    for (int64_t q = b + 1; q < nblocks; ++q) {
#pragma unroll
        for (int c = 0; c < C; ++c) acc[c] = __fadd_rn(acc[c], lead_sum[q * C + c]);
        if (!part[q].lead_cont) break;
    }
#pragma unroll
    for (int c = 0; c < C; ++c) grad_table[(int64_t)p.trail_key * C + c] = acc[c];
}

// ------------------------------------------------------------------------------------------------ atomic variant
// The reference's own method (kernel_grid_backward, gridencoder.cu:227-314): scatter w * grad with float atomics, sums in
// arrival order (NOT bit-reproducible).  Kept as an opt-in for throughput: no keys, no sort, and (point, level) pairs
// whose upstream gradient is exactly zero (march_rays_train's zero padding, terminated rays) issue nothing.  One thread
// per (point, level), level-major, with runs of neighbouring lanes that share a cell pre-summed in registers; C = 2 / 4
// rows go out as one vector RED (red.global.add.v2/v4.f32).
template <typename G, int D, int C>
__global__ void __launch_bounds__(TG_THREADS)
k_tg_atomic(const G *__restrict__ grad, int grad_layout, const float *__restrict__ inputs,
            const int32_t *__restrict__ offsets, int64_t B, int L, float S, uint32_t H, uint32_t gridtype, int align,
            float *__restrict__ grad_table) {
    __shared__ LevelMeta meta[TG_MAX_L];
    if (threadIdx.x < L) meta[threadIdx.x] = ngp_level_meta<D>(threadIdx.x, offsets, S, H, gridtype, align);
    __syncthreads();
    // level-major: the lanes of a warp are consecutive points at one level.  Consecutive samples of a ray fall into the
    // same cell on the coarse and middle levels, so runs of lanes with equal cells are summed in registers first
    // (segmented warp reduction) and only the run heads issue atomics.
    const int64_t t = (int64_t)blockIdx.x * TG_THREADS + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const bool in = t < B * L;
    const int l = in ? (int)(t / B) : 0;
    const int64_t b = in ? t - (int64_t)l * B : 0;
    constexpr int NC = 1 << D;
    float gv[C];
    bool any = false;
    if (in) {
        const G *g = grad_layout == 0 ? grad + ((int64_t)l * B + b) * C : grad + (b * L + l) * C;
#pragma unroll
        for (int c = 0; c < C; ++c) {
            gv[c] = (float)g[c];
            any |= gv[c] != 0.0f;
        }
    }
    const LevelMeta m = meta[l];
    uint32_t cell[D];
    float frac[D];
    bool live = in && any;
    if (live) live = tg_locate<D>(inputs, b, m, align, cell, frac);
    if (!live) {
#pragma unroll
        for (int d = 0; d < D; ++d) {
            cell[d] = 0xffffffffu;
            frac[d] = 0.0f;
        }
    }
    // run heads: first lane, or (level, cell, liveness) differs from the previous lane
    bool same_prev = lane > 0;
    {
        const int lp = __shfl_up_sync(0xffffffffu, l, 1);
        const int vp = __shfl_up_sync(0xffffffffu, (int)live, 1);
        same_prev = same_prev && lp == l && vp == (int)live;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const uint32_t cp = __shfl_up_sync(0xffffffffu, cell[d], 1);   // every lane shuffles: no short-circuit
            same_prev = same_prev && cp == cell[d];
        }
    }
    const unsigned heads = __ballot_sync(0xffffffffu, !same_prev);
    float val[NC][C];
#pragma unroll
    for (int k = 0; k < NC; ++k) {
        float w = 1.0f;
#pragma unroll
        for (int d = 0; d < D; ++d) w *= (k >> d) & 1 ? frac[d] : 1.0f - frac[d];
#pragma unroll
        for (int c = 0; c < C; ++c) val[k][c] = live ? __fmul_rn(w, gv[c]) : 0.0f;
    }
    if (__popc(heads) <= 24) {   // enough merging to pay for the shuffles (warp-uniform decision)
        const int seg = __popc(heads & (0xffffffffu >> (31 - lane)));   // run id of this lane
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int oseg = __shfl_down_sync(0xffffffffu, seg, o);
            const bool take = lane + o < 32 && oseg == seg;
#pragma unroll
            for (int k = 0; k < NC; ++k)
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    const float v = __shfl_down_sync(0xffffffffu, val[k][c], o);
                    if (take) val[k][c] += v;
                }
        }
        if (same_prev) live = false;   // only run heads write
    }
    if (!live) return;
#pragma unroll
    for (int k = 0; k < NC; ++k) {
        float *dst = grad_table + ((int64_t)m.row_offset + tg_row<D>(cell, k, m)) * C;
        if (C == 2) {
            atomicAdd(reinterpret_cast<float2 *>(dst), make_float2(val[k][0], val[k][1]));
        } else if (C == 4) {
            atomicAdd(reinterpret_cast<float4 *>(dst), make_float4(val[k][0], val[k][1], val[k][2], val[k][3]));
        } else {
#pragma unroll
            for (int c = 0; c < C; ++c) atomicAdd(dst + c, val[k][c]);
        }
    }
}

template <typename G, int D>
static int tg_launch_atomic(int C, const void *grad, int grad_layout, const float *inputs, const int32_t *offsets, int64_t B,
                            int L, float S, uint32_t H, uint32_t gridtype, int align, float *gt, cudaStream_t st) {
    const unsigned grid = (unsigned)ngp_div_up(B * L, TG_THREADS);
#define TG_ATOM(CC)                                                                                                        \
    k_tg_atomic<G, D, CC><<<grid, TG_THREADS, 0, st>>>((const G *)grad, grad_layout, inputs, offsets, B, L, S, H, gridtype, \
                                                       align, gt)
    switch (C) {
        case 1: TG_ATOM(1); break;
        case 2: TG_ATOM(2); break;
        case 4: TG_ATOM(4); break;
        case 8: TG_ATOM(8); break;
        default: ngp_set_error("GridEncoder: C must be 1, 2, 4 or 8 (got %d)", C); return NGP_ERR_UNSUPPORTED;
    }
#undef TG_ATOM
    return NGP_OK;
}

extern "C" int ngp_grid_encode_backward_table_atomic(const void *grad, int grad_layout, int grad_dtype, const float *inputs,
                                                     const int32_t *offsets, int64_t total_rows_,
                                                     float *grad_embeddings, int64_t B, int D, int C, int L, float S,
                                                     uint32_t H, uint32_t gridtype, int align_corners,
                                                     ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(grad, "ngp_grid_encode_backward_table_atomic");
    NGP_REQUIRE(grad && inputs && offsets && total_rows_ > 0 && total_rows_ < (1ll << 32) && grad_embeddings, NGP_ERR_ARG,
                "ngp_grid_encode_backward_table_atomic: null pointer");
    NGP_REQUIRE(L > 0 && L <= TG_MAX_L, NGP_ERR_ARG, "ngp_grid_encode_backward_table_atomic: L out of range");
    NGP_REQUIRE(D == 2 || D == 3, NGP_ERR_UNSUPPORTED, "GridEncoder: D must be 2 or 3 (got %d)", D);
    NGP_REQUIRE(((uintptr_t)grad_embeddings & 15u) == 0, NGP_ERR_ARG,
                "ngp_grid_encode_backward_table_atomic: grad_embeddings must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(grad_embeddings, 0, (size_t)total_rows_ * C * sizeof(float), st);
    NGP_REQUIRE(e == cudaSuccess, NGP_ERR_CUDA, "ngp_grid_encode_backward_table_atomic: memset: %s", cudaGetErrorString(e));
    if (B == 0) return NGP_OK;
    int rc;
    if (grad_dtype == NGP_DTYPE_F32)
        rc = D == 3 ? tg_launch_atomic<float, 3>(C, grad, grad_layout, inputs, offsets, B, L, S, H, gridtype, align_corners, grad_embeddings, st)
                    : tg_launch_atomic<float, 2>(C, grad, grad_layout, inputs, offsets, B, L, S, H, gridtype, align_corners, grad_embeddings, st);
    else if (grad_dtype == NGP_DTYPE_F16)
        rc = D == 3 ? tg_launch_atomic<__half, 3>(C, grad, grad_layout, inputs, offsets, B, L, S, H, gridtype, align_corners, grad_embeddings, st)
                    : tg_launch_atomic<__half, 2>(C, grad, grad_layout, inputs, offsets, B, L, S, H, gridtype, align_corners, grad_embeddings, st);
    else {
        ngp_set_error("ngp_grid_encode_backward_table_atomic: bad grad_dtype");
        return NGP_ERR_UNSUPPORTED;
    }
    if (rc != NGP_OK) return rc;
    NGP_CHECK_LAUNCH("ngp_grid_encode_backward_table_atomic");
    return NGP_OK;
}

// ------------------------------------------------------------------------------------------------ fixed-point atomics
// Bit-reproducible WITHOUT a sort: float atomics depend on arrival order because float addition is not associative; integer
// addition is.  Every contribution is scaled by one power of two (chosen from max|grad| and the row-count bound so that no
// sum can overflow), rounded to int64 and added with 64-bit integer atomics; any arrival order gives the same 64-bit sums.
// A last pass converts back.  Resolution: 2^-(61 - ceil(log2(2^D B))) of max|grad| per contribution (2^-40 for 3x10^5
// samples), far below fp32's own rounding of the sum.  Same thread mapping and run pre-summing as k_tg_atomic.
struct TgFixedHdr {
    unsigned int max_bits;   // float bits of max |grad| (atomicMax on the bit pattern of non-negative floats)
    int shift;               // contributions are scaled by 2^shift
    unsigned int nonfinite;  // an upstream gradient was inf / NaN: the result is all NaN (what GradScaler's inf check looks for)
    unsigned int pad;
};

template <typename G>
__global__ void __launch_bounds__(TG_THREADS) k_tg_absmax(const G *__restrict__ grad, int64_t n, TgFixedHdr *hdr) {
    float m = 0.0f;
    bool bad = false;
    for (int64_t i = (int64_t)blockIdx.x * TG_THREADS + threadIdx.x; i < n; i += (int64_t)gridDim.x * TG_THREADS) {
        const float v = fabsf((float)grad[i]);
        if (!(v <= 3.402823466e38f)) bad = true;   // inf or NaN: integers cannot carry it, the finish pass writes NaN
        else m = v > m ? v : m;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0.0f) atomicMax(&hdr->max_bits, __float_as_uint(m));
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(&hdr->nonfinite, 1u);
}

__global__ void k_tg_fixed_shift(TgFixedHdr *hdr, int count_bits) {
    const unsigned int b = hdr->max_bits;
    const int e = (int)((b >> 23) & 0xff) - 127;          // max < 2^(e+1)   (denormal max: e = -127, still a bound)
    int sh = 61 - (e + 1) - count_bits;
    hdr->shift = sh > 120 ? 120 : (sh < -120 ? -120 : sh);
}

template <typename G, int D, int C>
__global__ void __launch_bounds__(TG_THREADS)
k_tg_fixed(const G *__restrict__ grad, int grad_layout, const float *__restrict__ inputs, const int32_t *__restrict__ offsets,
           int64_t B, int L, float S, uint32_t H, uint32_t gridtype, int align, const TgFixedHdr *__restrict__ hdr,
           unsigned long long *__restrict__ acc) {
    __shared__ LevelMeta meta[TG_MAX_L];
    if (threadIdx.x < L) meta[threadIdx.x] = ngp_level_meta<D>(threadIdx.x, offsets, S, H, gridtype, align);
    __syncthreads();
    const float scale = __uint_as_float((uint32_t)(hdr->shift + 127) << 23);   // 2^shift, |shift| <= 120
    const int64_t t = (int64_t)blockIdx.x * TG_THREADS + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const bool in = t < B * L;
    const int l = in ? (int)(t / B) : 0;
    const int64_t b = in ? t - (int64_t)l * B : 0;
    constexpr int NC = 1 << D;
    float gv[C];
    bool any = false;
    if (in) {
        const G *g = grad_layout == 0 ? grad + ((int64_t)l * B + b) * C : grad + (b * L + l) * C;
#pragma unroll
        for (int c = 0; c < C; ++c) {
            gv[c] = (float)g[c];
            any |= gv[c] != 0.0f;
        }
    }
    const LevelMeta m = meta[l];
    uint32_t cell[D];
    float frac[D];
    bool live = in && any;
    if (live) live = tg_locate<D>(inputs, b, m, align, cell, frac);
    if (!live) {
#pragma unroll
        for (int d = 0; d < D; ++d) {
            cell[d] = 0xffffffffu;
            frac[d] = 0.0f;
        }
    }
    bool same_prev = lane > 0;
    {
        const int lp = __shfl_up_sync(0xffffffffu, l, 1);
        const int vp = __shfl_up_sync(0xffffffffu, (int)live, 1);
        same_prev = same_prev && lp == l && vp == (int)live;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const uint32_t cp = __shfl_up_sync(0xffffffffu, cell[d], 1);
            same_prev = same_prev && cp == cell[d];
        }
    }
    const unsigned heads = __ballot_sync(0xffffffffu, !same_prev);
    long long val[NC][C];   // the integers are what is summed, in registers too: the result cannot depend on the grouping
#pragma unroll
    for (int k = 0; k < NC; ++k) {
        float w = 1.0f;
#pragma unroll
        for (int d = 0; d < D; ++d) w *= (k >> d) & 1 ? frac[d] : 1.0f - frac[d];
#pragma unroll
        for (int c = 0; c < C; ++c) val[k][c] = live ? __float2ll_rn(__fmul_rn(__fmul_rn(w, gv[c]), scale)) : 0ll;   // (inf / NaN: see nonfinite)
    }
    if (__popc(heads) <= 24) {
        const int seg = __popc(heads & (0xffffffffu >> (31 - lane)));
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int oseg = __shfl_down_sync(0xffffffffu, seg, o);
            const bool take = lane + o < 32 && oseg == seg;
#pragma unroll
            for (int k = 0; k < NC; ++k)
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    const long long v = __shfl_down_sync(0xffffffffu, val[k][c], o);
                    if (take) val[k][c] += v;
                }
        }
        if (same_prev) live = false;
    }
    if (!live) return;
#pragma unroll
    for (int k = 0; k < NC; ++k) {
        unsigned long long *dst = acc + ((int64_t)m.row_offset + tg_row<D>(cell, k, m)) * C;
#pragma unroll
        for (int c = 0; c < C; ++c)
            if (val[k][c] != 0) atomicAdd(dst + c, (unsigned long long)val[k][c]);
    }
}

__global__ void __launch_bounds__(TG_THREADS)
k_tg_fixed_finish(const long long *__restrict__ acc, const TgFixedHdr *__restrict__ hdr, float *__restrict__ out, int64_t n) {
    const double inv = __longlong_as_double((long long)(1023 - hdr->shift) << 52);   // 2^-shift
    const bool bad = hdr->nonfinite != 0u;
    for (int64_t i = (int64_t)blockIdx.x * TG_THREADS + threadIdx.x; i < n; i += (int64_t)gridDim.x * TG_THREADS)
        out[i] = bad ? __int_as_float(0x7fc00000) : (float)((double)acc[i] * inv);
}

extern "C" int64_t ngp_grid_backward_table_fixed_workspace_bytes(int64_t total_rows, int C) {
    return 256 + total_rows * C * 8;
}

template <typename G, int D>
static int tg_launch_fixed(int C, const void *grad, int grad_layout, const float *inputs, const int32_t *offsets, int64_t B,
                           int L, float S, uint32_t H, uint32_t gridtype, int align, const TgFixedHdr *hdr,
                           unsigned long long *acc, cudaStream_t st) {
    const unsigned grid = (unsigned)ngp_div_up(B * L, TG_THREADS);
#define TG_FIX(CC)                                                                                                        \
    k_tg_fixed<G, D, CC><<<grid, TG_THREADS, 0, st>>>((const G *)grad, grad_layout, inputs, offsets, B, L, S, H, gridtype, \
                                                      align, hdr, acc)
    switch (C) {
        case 1: TG_FIX(1); break;
        case 2: TG_FIX(2); break;
        case 4: TG_FIX(4); break;
        case 8: TG_FIX(8); break;
        default: ngp_set_error("GridEncoder: C must be 1, 2, 4 or 8 (got %d)", C); return NGP_ERR_UNSUPPORTED;
    }
#undef TG_FIX
    return NGP_OK;
}

extern "C" int ngp_grid_encode_backward_table_fixed(const void *grad, int grad_layout, int grad_dtype, const float *inputs,
                                                    const int32_t *offsets, int64_t total_rows_, float *grad_embeddings,
                                                    int64_t B, int D, int C, int L, float S, uint32_t H, uint32_t gridtype,
                                                    int align_corners, void *workspace, int64_t workspace_bytes,
                                                    ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(grad, "ngp_grid_encode_backward_table_fixed");
    NGP_REQUIRE(grad && inputs && offsets && total_rows_ > 0 && total_rows_ < (1ll << 32) && grad_embeddings, NGP_ERR_ARG,
                "ngp_grid_encode_backward_table_fixed: null pointer");
    NGP_REQUIRE(L > 0 && L <= TG_MAX_L, NGP_ERR_ARG, "ngp_grid_encode_backward_table_fixed: L out of range");
    NGP_REQUIRE(D == 2 || D == 3, NGP_ERR_UNSUPPORTED, "GridEncoder: D must be 2 or 3 (got %d)", D);
    NGP_REQUIRE(grad_dtype == NGP_DTYPE_F32 || grad_dtype == NGP_DTYPE_F16, NGP_ERR_UNSUPPORTED,
                "ngp_grid_encode_backward_table_fixed: bad grad_dtype");
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t n_out = total_rows_ * C;
    if (B == 0) {
        cudaError_t e0 = cudaMemsetAsync(grad_embeddings, 0, (size_t)n_out * sizeof(float), st);
        NGP_REQUIRE(e0 == cudaSuccess, NGP_ERR_CUDA, "ngp_grid_encode_backward_table_fixed: memset: %s", cudaGetErrorString(e0));
        return NGP_OK;
    }
    NGP_REQUIRE(workspace && ((uintptr_t)workspace & 15u) == 0 &&
                    workspace_bytes >= ngp_grid_backward_table_fixed_workspace_bytes(total_rows_, C),
                NGP_ERR_ARG, "ngp_grid_encode_backward_table_fixed: workspace too small or misaligned");
    TgFixedHdr *hdr = (TgFixedHdr *)workspace;
    unsigned long long *acc = (unsigned long long *)((char *)workspace + 256);
    cudaError_t e = cudaMemsetAsync(workspace, 0, (size_t)(256 + n_out * 8), st);
    NGP_REQUIRE(e == cudaSuccess, NGP_ERR_CUDA, "ngp_grid_encode_backward_table_fixed: memset: %s", cudaGetErrorString(e));
    const int64_t n_in = B * L * C;
    const unsigned rgrid = (unsigned)(ngp_div_up(n_in, TG_THREADS) < 148 * 8 ? ngp_div_up(n_in, TG_THREADS) : 148 * 8);
    if (grad_dtype == NGP_DTYPE_F32) k_tg_absmax<float><<<rgrid, TG_THREADS, 0, st>>>((const float *)grad, n_in, hdr);
    else k_tg_absmax<__half><<<rgrid, TG_THREADS, 0, st>>>((const __half *)grad, n_in, hdr);
    int count_bits = 0;   // a row receives at most 2^D contributions per point
    while ((1ll << count_bits) < (B << D)) ++count_bits;
    k_tg_fixed_shift<<<1, 1, 0, st>>>(hdr, count_bits);
    int rc;
    if (grad_dtype == NGP_DTYPE_F32)
        rc = D == 3 ? tg_launch_fixed<float, 3>(C, grad, grad_layout, inputs, offsets, B, L, S, H, gridtype, align_corners, hdr, acc, st)
                    : tg_launch_fixed<float, 2>(C, grad, grad_layout, inputs, offsets, B, L, S, H, gridtype, align_corners, hdr, acc, st);
    else
        rc = D == 3 ? tg_launch_fixed<__half, 3>(C, grad, grad_layout, inputs, offsets, B, L, S, H, gridtype, align_corners, hdr, acc, st)
                    : tg_launch_fixed<__half, 2>(C, grad, grad_layout, inputs, offsets, B, L, S, H, gridtype, align_corners, hdr, acc, st);
    if (rc != NGP_OK) return rc;
    const unsigned fgrid = (unsigned)(ngp_div_up(n_out, TG_THREADS) < 148 * 16 ? ngp_div_up(n_out, TG_THREADS) : 148 * 16);
    k_tg_fixed_finish<<<fgrid, TG_THREADS, 0, st>>>((const long long *)acc, hdr, grad_embeddings, n_out);
    NGP_CHECK_LAUNCH("ngp_grid_encode_backward_table_fixed");
    return NGP_OK;
}

static int tg_bits(uint32_t v) {
    int b = 0;
    while ((1ull << b) <= v) ++b;
    return b;
}

extern "C" int64_t ngp_grid_backward_table_workspace_bytes(int64_t B, int D, int L) {
    const int64_t n = B * L * (1ll << D);
    // 2 key buffers + 2 payload buffers + radix-sort scratch (histograms / tile descriptors: well under n/4 + 16 MiB)
    // + per-256-element-block partials of the segmented reduction (16 B flags + 2 x 8 floats)
    return 4 * n * 4 + n / 4 + (16ll << 20) + (n / TG_THREADS + 2) * 80;
}

template <typename G, int D>
static int tg_launch_reduce(int C, const uint32_t *keys, const uint32_t *ids, int64_t n, uint32_t sentinel,
                            const void *grad, int grad_layout, const float *inputs, const int32_t *offsets, int64_t B,
                            int L, float S, uint32_t H, uint32_t gridtype, int align, float *gt, TgPartial *part,
                            float *lead, float *trail, cudaStream_t st) {
    const unsigned grid = (unsigned)ngp_div_up(n, TG_THREADS);
    const unsigned sgrid = (unsigned)ngp_div_up((int64_t)grid, TG_THREADS);
#define TG_RED(CC)                                                                                                       \
    k_tg_reduce<G, D, CC><<<grid, TG_THREADS, 0, st>>>(keys, ids, n, sentinel, (const G *)grad, grad_layout, inputs, \
                                                       offsets, B, L, S, H, gridtype, align, gt, part, lead, trail); \
    k_tg_stitch<CC><<<sgrid, TG_THREADS, 0, st>>>(part, lead, trail, (int64_t)grid, gt)
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
                                              const int32_t *offsets, int64_t total_rows_,
                                              float *grad_embeddings, int64_t B, int D, int C, int L, float S,
                                              uint32_t H, uint32_t gridtype, int align_corners, void *workspace,
                                              int64_t workspace_bytes, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(grad, "ngp_grid_encode_backward_table");
    NGP_REQUIRE(grad && inputs && offsets && total_rows_ > 0 && total_rows_ < (1ll << 32) && grad_embeddings, NGP_ERR_ARG,
                "ngp_grid_encode_backward_table: null pointer");
    NGP_REQUIRE(L > 0 && L <= TG_MAX_L, NGP_ERR_ARG, "ngp_grid_encode_backward_table: L out of range");
    NGP_REQUIRE(D == 2 || D == 3, NGP_ERR_UNSUPPORTED, "GridEncoder: D must be 2 or 3 (got %d)", D);
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t total_rows = (uint32_t)total_rows_;
    cudaError_t e = cudaMemsetAsync(grad_embeddings, 0, (size_t)total_rows * C * sizeof(float), st);
    NGP_REQUIRE(e == cudaSuccess, NGP_ERR_CUDA, "ngp_grid_encode_backward_table: memset: %s", cudaGetErrorString(e));
    if (B == 0) return NGP_OK;
    const int64_t n = B * L * (1ll << D);
    NGP_REQUIRE(n < (1ll << 31), NGP_ERR_UNSUPPORTED, "ngp_grid_encode_backward_table: B*L*2^D must be < 2^31");
    NGP_REQUIRE(workspace && workspace_bytes >= ngp_grid_backward_table_workspace_bytes(B, D, L), NGP_ERR_ARG,
                "ngp_grid_encode_backward_table: workspace too small");
    uint32_t *keys_a = (uint32_t *)workspace, *keys_b = keys_a + n, *ids_a = keys_b + n, *ids_b = ids_a + n;
    const int64_t nblk = n / TG_THREADS + 2;
    TgPartial *part = (TgPartial *)(ids_b + n);
    float *lead = (float *)(part + nblk), *trail = lead + nblk * 8;
    void *cub_tmp = (void *)(trail + nblk * 8);
    size_t cub_avail = (size_t)workspace_bytes - (size_t)(4 * n * 4) - (size_t)(nblk * 80);
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
                                                 offsets, B, L, S, H, gridtype, align_corners, grad_embeddings, part, lead, trail, st)
                    : tg_launch_reduce<float, 2>(C, dk.Current(), dv.Current(), n, sentinel, grad, grad_layout, inputs,
                                                 offsets, B, L, S, H, gridtype, align_corners, grad_embeddings, part, lead, trail, st);
    else if (grad_dtype == NGP_DTYPE_F16)
        rc = D == 3 ? tg_launch_reduce<__half, 3>(C, dk.Current(), dv.Current(), n, sentinel, grad, grad_layout, inputs,
                                                  offsets, B, L, S, H, gridtype, align_corners, grad_embeddings, part, lead, trail, st)
                    : tg_launch_reduce<__half, 2>(C, dk.Current(), dv.Current(), n, sentinel, grad, grad_layout, inputs,
                                                  offsets, B, L, S, H, gridtype, align_corners, grad_embeddings, part, lead, trail, st);
    else {
        ngp_set_error("ngp_grid_encode_backward_table: bad grad_dtype");
        return NGP_ERR_UNSUPPORTED;
    }
    if (rc != NGP_OK) return rc;
    NGP_CHECK_LAUNCH("ngp_grid_encode_backward_table(reduce)");
    return NGP_OK;
}
