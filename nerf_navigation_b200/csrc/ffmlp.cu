// ffmlp.cu — fully fused 64-wide fp16 MLP on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM).
// Replaces the reference's _ffmlp extension: ffmlp_forward / ffmlp_inference / ffmlp_backward
// (ffmlp/src/ffmlp.h:8-14, kernels ffmlp.cu:331-518, host ffmlp.cu:525-895; Python ffmlp/ffmlp.py:15-169).
//
// Semantics kept: weights are one flat fp16 vector [W_in (64 x in)] [W_h (64 x 64)] x (n-1) [W_out (16 x 64)], every
// block row-major [out, in] (ffmlp.cu:632, :89); hidden activation ReLU or None, output activation None
// (ffmlp.py:107-108); activations are rounded to fp16 after every layer; forward_buffer [n, B, 64] holds the post-
// activation hidden layers, outputs [B, 16].  Differences (documented in DESIGN.md): products accumulate in fp32 inside
// TMEM (the reference accumulates in fp16 wmma fragments), B needs no padding to 128, everything runs on the caller's
// stream (the reference forks n+1 side streams for split-K weight gradients, ffmlp.cu:711-730).
//
// Kernel shape: a CTA = 128 threads owns tiles of 128 rows.  Thread r stages row r of the tile in shared memory as the
// A operand (chunk-major no-swizzle layout, tc5.cuh), one elected thread issues the K/16 tcgen05.mma of a layer
// (M = 128 rows, N = 64 | 16, fp32 accumulators in TMEM), tcgen05.commit signals an mbarrier, then thread r reads
// ITS row of the accumulator back with tcgen05.ld (TMEM lane r = tile row r), applies the activation, rounds to fp16 and
// writes the row as the next layer's A operand — per-row activations never leave the SM except for the buffers the
// reference API returns.  Several CTAs are resident per SM so one tile's epilogue overlaps another tile's MMAs.
#include "common.cuh"
#include "tc5.cuh"

#define FF_ROWS 128
#define FF_HID 64
#define FF_OUT 16
#define FF_ACT_RELU 0
#define FF_ACT_NONE 6

struct FFShared {  // fixed part; the tiles follow in dynamic shared memory
    uint64_t bar;
    uint32_t tmem;
    uint32_t pad;
};

__device__ __forceinline__ uint4 ff_ldg16(const __half *p) { return __ldg(reinterpret_cast<const uint4 *>(p)); }

// stage a row-major [N, K] fp16 weight block as a K-major chunk-major tile (rows = N)
__device__ __forceinline__ void ff_stage_weights(unsigned char *dst, const __half *__restrict__ w, int N, int K) {
    const int kcs = K >> 3;
    for (int i = threadIdx.x; i < N * kcs; i += blockDim.x) {
        const int n = i / kcs, kc = i - n * kcs;
        *reinterpret_cast<uint4 *>(dst + tc5::chunk_off(N, n, kc)) = ff_ldg16(w + (size_t)n * K + kc * 8);
    }
}
// stage the TRANSPOSE of a row-major [N, K] block: tile rows = K index, tile K = N index (for the backward products)
__device__ __forceinline__ void ff_stage_weights_t(unsigned char *dst, const __half *__restrict__ w, int N, int K) {
    __half *d = reinterpret_cast<__half *>(dst);
    for (int i = threadIdx.x; i < N * K; i += blockDim.x) {
        const int n = i / K, k = i - n * K;  // element W[n][k] -> tile row k, tile column n
        d[(tc5::chunk_off(K, k, n >> 3) >> 1) + (n & 7)] = w[i];
    }
}

// issue the MMAs of one layer: D[128 x N] = A[128 x K] * W[N x K]^T
__device__ __forceinline__ void ff_issue_layer(uint32_t d_tmem, uint32_t a_addr, uint32_t w_addr, int N, int K, uint32_t bar) {
    const uint32_t idesc = tc5::idesc_f16(FF_ROWS, N, tc5::MAJOR_K, tc5::MAJOR_K);
    for (int ks = 0; ks < (K >> 4); ++ks) {
        const uint64_t ad = tc5::smem_desc(a_addr + ks * 2 * FF_ROWS * 16, FF_ROWS * 16, 128);
        const uint64_t bd = tc5::smem_desc(w_addr + ks * 2 * N * 16, N * 16, 128);
        tc5::mma_f16(d_tmem, ad, bd, idesc, ks > 0);
    }
    tc5::mma_commit(bar);
}

// ------------------------------------------------------------------------------------------------ forward / inference
__global__ void __launch_bounds__(FF_ROWS)
k_ffmlp_forward(const __half *__restrict__ x, const __half *__restrict__ w, __half *__restrict__ fwd_buf,
                __half *__restrict__ out, int64_t B, int in_dim, int n_layers, int act) {
    extern __shared__ __align__(128) unsigned char ff_smem[];
    FFShared *sh = reinterpret_cast<FFShared *>(ff_smem);
    unsigned char *Xs = ff_smem + 128;                          // [128 x in]
    unsigned char *Hs = Xs + FF_ROWS * in_dim * 2;              // [128 x 64]
    unsigned char *Win = Hs + FF_ROWS * FF_HID * 2;             // [64 x in]
    unsigned char *Wh = Win + FF_HID * in_dim * 2;              // (n-1) x [64 x 64]
    unsigned char *Wout = Wh + (n_layers - 1) * FF_HID * FF_HID * 2;  // [16 x 64]
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const uint32_t bar = tc5::smem_u32(&sh->bar);

    if (tid == 0) {
        tc5::mbar_init(bar, 1);
        tc5::fence_mbar_init();
    }
    if (warp == 0) tc5::tmem_alloc(tc5::smem_u32(&sh->tmem), 64);
    ff_stage_weights(Win, w, FF_HID, in_dim);
    for (int l = 0; l < n_layers - 1; ++l)
        ff_stage_weights(Wh + l * FF_HID * FF_HID * 2, w + FF_HID * in_dim + l * FF_HID * FF_HID, FF_HID, FF_HID);
    ff_stage_weights(Wout, w + FF_HID * in_dim + (n_layers - 1) * FF_HID * FF_HID, FF_OUT, FF_HID);
    tc5::fence_async_smem();
    tc5::fence_before_sync();
    __syncthreads();
    tc5::fence_after_sync();
    const uint32_t tmem = __shfl_sync(0xffffffffu, sh->tmem, 0);
    const uint32_t trow = tmem + ((uint32_t)(warp * 32) << 16);  // this warp's 32 TMEM lanes
    const uint32_t xs_a = tc5::smem_u32(Xs), hs_a = tc5::smem_u32(Hs), win_a = tc5::smem_u32(Win),
                   wh_a = tc5::smem_u32(Wh), wout_a = tc5::smem_u32(Wout);
    uint32_t phase = 0;

    const int64_t ntiles = (B + FF_ROWS - 1) / FF_ROWS;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t row = tile * FF_ROWS + tid;
        const bool inb = row < B;
        // ---- stage the input rows
        for (int kc = 0; kc < (in_dim >> 3); ++kc) {
            uint4 v = make_uint4(0u, 0u, 0u, 0u);
            if (inb) v = ff_ldg16(x + row * in_dim + kc * 8);
            *reinterpret_cast<uint4 *>(Xs + tc5::chunk_off(FF_ROWS, tid, kc)) = v;
        }
        tc5::fence_async_smem();
        tc5::fence_before_sync();
        __syncthreads();
        // ---- n hidden layers
        for (int l = 0; l < n_layers; ++l) {
            if (tid == 0) {
                tc5::fence_after_sync();
                if (l == 0) ff_issue_layer(tmem, xs_a, win_a, FF_HID, in_dim, bar);
                else ff_issue_layer(tmem, hs_a, wh_a + (l - 1) * FF_HID * FF_HID * 2, FF_HID, FF_HID, bar);
            }
            tc5::mbar_wait(bar, phase);
            phase ^= 1u;
            tc5::fence_after_sync();
#pragma unroll
            for (int c = 0; c < 4; ++c) {  // 16 accumulator columns at a time
                float v[16];
                tc5::tmem_ld16(trow + c * 16, v);
                if (act == FF_ACT_RELU) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.0f);
                }
                const uint4 lo = tc5::pack8(v, 0), hi = tc5::pack8(v, 8);
                *reinterpret_cast<uint4 *>(Hs + tc5::chunk_off(FF_ROWS, tid, 2 * c)) = lo;
                *reinterpret_cast<uint4 *>(Hs + tc5::chunk_off(FF_ROWS, tid, 2 * c + 1)) = hi;
                if (fwd_buf && inb) {
                    uint4 *dst = reinterpret_cast<uint4 *>(fwd_buf + ((size_t)l * B + row) * FF_HID + c * 16);
                    dst[0] = lo;
                    dst[1] = hi;
                }
            }
            tc5::fence_async_smem();
            tc5::fence_before_sync();
            __syncthreads();
        }
        // ---- output layer (N = 16, no activation)
        if (tid == 0) {
            tc5::fence_after_sync();
            ff_issue_layer(tmem, hs_a, wout_a, FF_OUT, FF_HID, bar);
        }
        tc5::mbar_wait(bar, phase);
        phase ^= 1u;
        tc5::fence_after_sync();
        {
            float v[16];
            tc5::tmem_ld16(trow, v);
            if (inb) {
                uint4 *dst = reinterpret_cast<uint4 *>(out + row * FF_OUT);
                dst[0] = tc5::pack8(v, 0);
                dst[1] = tc5::pack8(v, 8);
            }
        }
        tc5::fence_before_sync();
        __syncthreads();  // the next tile's first MMA overwrites the accumulator columns just read
    }
    if (warp == 0) tc5::tmem_dealloc(tmem, 64);
}

static size_t ff_forward_smem(int in_dim, int n_layers) {
    return 128 + (size_t)FF_ROWS * in_dim * 2 + FF_ROWS * FF_HID * 2 + (size_t)FF_HID * in_dim * 2 +
           (size_t)(n_layers - 1) * FF_HID * FF_HID * 2 + FF_OUT * FF_HID * 2;
}

static int ff_check(int64_t B, int input_dim, int output_dim, int hidden_dim, int num_layers, int activation,
                    int output_activation, const char *who) {
    NGP_REQUIRE(B >= 0, NGP_ERR_ARG, "%s: negative B", who);
    NGP_REQUIRE(hidden_dim == FF_HID, NGP_ERR_UNSUPPORTED, "%s: hidden_dim must be 64 (got %d)", who, hidden_dim);
    NGP_REQUIRE(input_dim > 0 && input_dim % 16 == 0 && input_dim <= 256, NGP_ERR_UNSUPPORTED,
                "%s: input_dim must be 16*m, m in 1..16 (got %d)", who, input_dim);
    NGP_REQUIRE(output_dim == FF_OUT, NGP_ERR_UNSUPPORTED, "%s: padded output_dim must be 16 (got %d)", who, output_dim);
    NGP_REQUIRE(num_layers >= 2 && num_layers <= 8, NGP_ERR_UNSUPPORTED, "%s: num_layers must be in 2..8 (got %d)", who, num_layers);
    NGP_REQUIRE(activation == FF_ACT_RELU || activation == FF_ACT_NONE, NGP_ERR_UNSUPPORTED,
                "%s: hidden activation must be relu (0) or none (6); got %d", who, activation);
    NGP_REQUIRE(output_activation == FF_ACT_NONE, NGP_ERR_UNSUPPORTED, "%s: output activation must be none (6)", who);
    return NGP_OK;
}

static int ff_sm_count() {
    int dev = 0, sms = NGP_SM_COUNT;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms;
}

// default forward: the warp-specialised multi-tile kernel of ffmlp5.cu (returns -1 if one slot does not fit in shared
// memory); NGP_B200_FFMLP_IMPL=v1 selects the one-tile-per-CTA kernel above (the cross-check)
int f5_ffmlp_forward(const void *inputs, const void *weights, int64_t B, int input_dim, int num_layers, int activation,
                     void *forward_buffer, void *outputs, cudaStream_t stream);
static int ff_impl_v1() {
    static int v = -1;
    if (v < 0) {
        const char *e = getenv("NGP_B200_FFMLP_IMPL");
        v = (e && strcmp(e, "v1") == 0) ? 1 : 0;
    }
    return v;
}

extern "C" int ngp_ffmlp_forward(const void *inputs, const void *weights, int64_t B, int input_dim, int output_dim,
                                 int hidden_dim, int num_layers, int activation, int output_activation,
                                 void *forward_buffer, void *outputs, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(inputs, "ngp_ffmlp_forward");
    if (B == 0) return NGP_OK;
    NGP_REQUIRE(inputs && weights && outputs, NGP_ERR_ARG, "ngp_ffmlp_forward: null pointer");
    int rc = ff_check(B, input_dim, output_dim, hidden_dim, num_layers, activation, output_activation, "ngp_ffmlp_forward");
    if (rc) return rc;
    if (!ff_impl_v1()) {
        rc = f5_ffmlp_forward(inputs, weights, B, input_dim, num_layers, activation, forward_buffer, outputs, (cudaStream_t)stream);
        if (rc != -1) return rc;
    }
    const size_t smem = ff_forward_smem(input_dim, num_layers);
    NGP_REQUIRE(smem <= 227 * 1024, NGP_ERR_UNSUPPORTED, "ngp_ffmlp_forward: configuration needs %zu B of shared memory", smem);
    if (cudaFuncSetAttribute(k_ffmlp_forward, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        ngp_set_error("ngp_ffmlp_forward: cannot reserve %zu B of shared memory", smem);
        return NGP_ERR_CUDA;
    }
    cudaFuncSetAttribute(k_ffmlp_forward, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    int per_sm = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_ffmlp_forward, FF_ROWS, smem);
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 8) per_sm = 8;  // 8 x 64 TMEM columns = 512
    const int64_t tiles = ngp_div_up(B, FF_ROWS), resident = (int64_t)ff_sm_count() * per_sm;
    const int grid = (int)(tiles < resident ? tiles : resident);
    k_ffmlp_forward<<<grid, FF_ROWS, smem, (cudaStream_t)stream>>>(
        (const __half *)inputs, (const __half *)weights, (__half *)forward_buffer, (__half *)outputs, B, input_dim,
        num_layers, activation);
    NGP_CHECK_LAUNCH("ngp_ffmlp_forward");
    return NGP_OK;
}

// ------------------------------------------------------------------------------------------------ backward
// One kernel produces the hidden-layer gradients (kernel_mlp_fused_backward, ffmlp.cu:410-518), dL/dinput for any
// input width (the reference fuses it only when input_dim == 64 and otherwise launches CUTLASS, ffmlp.cu:795-796,
// 880-887) AND the weight gradients (the reference's CUTLASS split-K products on side streams, ffmlp.cu:795-875).
// Weight gradients are transposed products over the batch rows, dW[j][i] = sum_m dH[m][j] * Src[m][i]; the SAME shared
// memory tiles that feed the row-wise chain products are re-read MN-major (tc5.cuh) as A = dH^T (M = 64) and
// B = Src^T, accumulating in TMEM columns that persist across all tiles of the CTA.  Each CTA finally writes its fp32
// partial sums to the workspace and k_ffmlp_wgrad_reduce adds the partials in CTA order (deterministic).
struct FFBwdLayout {  // TMEM column map
    int d_cols;       // working accumulator: max(64, in_dim)
    int wout_col;     // 16 columns: dW_out^T [j][o]
    int wh_col;       // (n-1) x 64 columns: dW_h[l] [j][i]
    int win_col;      // in_dim columns: dW_in [j][i]
    int total;
};
__host__ __device__ inline FFBwdLayout ff_bwd_layout(int in_dim, int n_layers) {
    FFBwdLayout L;
    L.d_cols = in_dim > FF_HID ? in_dim : FF_HID;
    L.wout_col = L.d_cols;
    L.wh_col = L.wout_col + FF_OUT;
    L.win_col = L.wh_col + (n_layers - 1) * FF_HID;
    L.total = L.win_col + in_dim;
    return L;
}
__host__ __device__ inline int ff_pow2_cols(int c) { return c <= 32 ? 32 : c <= 64 ? 64 : c <= 128 ? 128 : c <= 256 ? 256 : 512; }

// dW (+)= A^T B over the 128 rows of the tiles: A tile [128 x 64] read as M = 64, B tile [128 x N] read as N
__device__ __forceinline__ void ff_issue_wgrad(uint32_t d_tmem, uint32_t a_addr, uint32_t b_addr, int N, bool first) {
    const uint32_t idesc = tc5::idesc_f16(FF_HID, N, tc5::MAJOR_MN, tc5::MAJOR_MN);
    for (int ks = 0; ks < FF_ROWS / 16; ++ks) {
        const uint64_t ad = tc5::smem_desc(a_addr + ks * 256, 128, FF_ROWS * 16);
        const uint64_t bd = tc5::smem_desc(b_addr + ks * 256, 128, FF_ROWS * 16);
        tc5::mma_f16(d_tmem, ad, bd, idesc, !(first && ks == 0));
    }
}

// thread r copies row r of a row-major [*, width] fp16 matrix into a chunk-major tile (zero rows past B)
__device__ __forceinline__ void ff_stage_rows(unsigned char *dst, const __half *__restrict__ src, int64_t row, bool inb,
                                              int width) {
    for (int kc = 0; kc < (width >> 3); ++kc) {
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (inb) v = ff_ldg16(src + row * width + kc * 8);
        *reinterpret_cast<uint4 *>(dst + tc5::chunk_off(FF_ROWS, threadIdx.x, kc)) = v;
    }
}

__device__ __forceinline__ uint4 ff_relu_mask8(uint4 g, uint4 f) {  // g where f > 0 else 0, 8 halves
    const __half2 z = __float2half2_rn(0.0f);
    __half2 *gp = reinterpret_cast<__half2 *>(&g);
    const __half2 *fp = reinterpret_cast<const __half2 *>(&f);
#pragma unroll
    for (int i = 0; i < 4; ++i) gp[i] = __hmul2(gp[i], __hgt2(fp[i], z));
    return g;
}

__global__ void __launch_bounds__(FF_ROWS)
k_ffmlp_backward(const __half *__restrict__ grad, const __half *__restrict__ x, const __half *__restrict__ w,
                 const __half *__restrict__ fwd_buf, __half *__restrict__ bwd_buf, __half *__restrict__ grad_x,
                 float *__restrict__ partials, int64_t B, int in_dim, int n_layers, int act, int tmem_cols) {
    extern __shared__ __align__(128) unsigned char ff_smem[];
    FFShared *sh = reinterpret_cast<FFShared *>(ff_smem);
    const int fw = in_dim > FF_HID ? in_dim : FF_HID;  // width of the two source tiles
    unsigned char *Gs = ff_smem + 128;                              // [128 x 16]  dL/doutput
    unsigned char *Hs = Gs + FF_ROWS * FF_OUT * 2;                  // [128 x 64]  dL/dhidden of the current layer
    unsigned char *Fs0 = Hs + FF_ROWS * FF_HID * 2;                 // [128 x fw]  forward activations / inputs (ping)
    unsigned char *Fs1 = Fs0 + FF_ROWS * fw * 2;                    // (pong)
    unsigned char *WoutT = Fs1 + FF_ROWS * fw * 2;                  // [64 x 16]   W_out^T
    unsigned char *WhT = WoutT + FF_HID * FF_OUT * 2;               // (n-1) x [64 x 64]  W_h^T
    unsigned char *WinT = WhT + (n_layers - 1) * FF_HID * FF_HID * 2;  // [in x 64]  W_in^T
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    const uint32_t bar = tc5::smem_u32(&sh->bar);
    const FFBwdLayout TL = ff_bwd_layout(in_dim, n_layers);

    if (tid == 0) {
        tc5::mbar_init(bar, 1);
        tc5::fence_mbar_init();
    }
    if (warp == 0) tc5::tmem_alloc(tc5::smem_u32(&sh->tmem), tmem_cols);
    const __half *w_h = w + FF_HID * in_dim, *w_out = w_h + (n_layers - 1) * FF_HID * FF_HID;
    ff_stage_weights_t(WoutT, w_out, FF_OUT, FF_HID);
    for (int l = 0; l < n_layers - 1; ++l) ff_stage_weights_t(WhT + l * FF_HID * FF_HID * 2, w_h + l * FF_HID * FF_HID, FF_HID, FF_HID);
    if (grad_x) ff_stage_weights_t(WinT, w, FF_HID, in_dim);
    tc5::fence_async_smem();
    tc5::fence_before_sync();
    __syncthreads();
    tc5::fence_after_sync();
    const uint32_t tmem = __shfl_sync(0xffffffffu, sh->tmem, 0);
    const uint32_t trow = tmem + ((uint32_t)(warp * 32) << 16);
    const uint32_t gs_a = tc5::smem_u32(Gs), hs_a = tc5::smem_u32(Hs), f_a[2] = {tc5::smem_u32(Fs0), tc5::smem_u32(Fs1)},
                   woutT_a = tc5::smem_u32(WoutT), whT_a = tc5::smem_u32(WhT), winT_a = tc5::smem_u32(WinT);
    unsigned char *Fs[2] = {Fs0, Fs1};
    uint32_t phase = 0;
    bool first = true;

    const int64_t ntiles = (B + FF_ROWS - 1) / FF_ROWS;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t row = tile * FF_ROWS + tid;
        const bool inb = row < B;
        ff_stage_rows(Gs, grad, row, inb, FF_OUT);
        ff_stage_rows(Fs[0], fwd_buf + (size_t)(n_layers - 1) * B * FF_HID, row, inb, FF_HID);
        tc5::fence_async_smem();
        tc5::fence_before_sync();
        __syncthreads();
        if (tid == 0) {
            tc5::fence_after_sync();
            ff_issue_wgrad(tmem + TL.wout_col, f_a[0], gs_a, FF_OUT, first);   // dW_out^T[j][o] += F_{n-1}^T G
            ff_issue_layer(tmem, gs_a, woutT_a, FF_HID, FF_OUT, bar);          // D = G W_out
        }
        int cur = 0;  // Fs[cur] holds the forward activations of the layer whose gradient is in D
        for (int k = 0; k < n_layers; ++k) {
            const int layer = n_layers - 1 - k;
            // source rows of the weight gradient of `layer` (previous activations, or the network input): other buffer
            if (layer > 0) ff_stage_rows(Fs[cur ^ 1], fwd_buf + (size_t)(layer - 1) * B * FF_HID, row, inb, FF_HID);
            else ff_stage_rows(Fs[cur ^ 1], x, row, inb, in_dim);
            tc5::mbar_wait(bar, phase);
            phase ^= 1u;
            tc5::fence_after_sync();
            // ---- epilogue: dH_layer = D .* relu'(F_layer), fp16, -> Hs (+ backward_buffer[k])
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                float v[16];
                tc5::tmem_ld16(trow + c * 16, v);
                uint4 lo = tc5::pack8(v, 0), hi = tc5::pack8(v, 8);
                if (act == FF_ACT_RELU) {
                    lo = ff_relu_mask8(lo, *reinterpret_cast<const uint4 *>(Fs[cur] + tc5::chunk_off(FF_ROWS, tid, 2 * c)));
                    hi = ff_relu_mask8(hi, *reinterpret_cast<const uint4 *>(Fs[cur] + tc5::chunk_off(FF_ROWS, tid, 2 * c + 1)));
                }
                *reinterpret_cast<uint4 *>(Hs + tc5::chunk_off(FF_ROWS, tid, 2 * c)) = lo;
                *reinterpret_cast<uint4 *>(Hs + tc5::chunk_off(FF_ROWS, tid, 2 * c + 1)) = hi;
                if (bwd_buf && inb) {
                    uint4 *dst = reinterpret_cast<uint4 *>(bwd_buf + ((size_t)k * B + row) * FF_HID + c * 16);
                    dst[0] = lo;
                    dst[1] = hi;
                }
            }
            tc5::fence_async_smem();
            tc5::fence_before_sync();
            __syncthreads();
            if (tid == 0) {
                tc5::fence_after_sync();
                if (layer > 0) {
                    ff_issue_wgrad(tmem + TL.wh_col + (layer - 1) * FF_HID, hs_a, f_a[cur ^ 1], FF_HID, first);
                    ff_issue_layer(tmem, hs_a, whT_a + (layer - 1) * FF_HID * FF_HID * 2, FF_HID, FF_HID, bar);
                } else {
                    ff_issue_wgrad(tmem + TL.win_col, hs_a, f_a[cur ^ 1], in_dim, first);
                    if (grad_x) ff_issue_layer(tmem, hs_a, winT_a, in_dim, FF_HID, bar);
                    else tc5::mma_commit(bar);
                }
            }
            cur ^= 1;
        }
        // ---- dL/dinput (or just drain the weight-gradient MMAs of layer 0)
        tc5::mbar_wait(bar, phase);
        phase ^= 1u;
        tc5::fence_after_sync();
        if (grad_x) {
            for (int c = 0; c < (in_dim >> 4); ++c) {
                float v[16];
                tc5::tmem_ld16(trow + c * 16, v);
                if (inb) {
                    uint4 *dst = reinterpret_cast<uint4 *>(grad_x + row * in_dim + c * 16);
                    dst[0] = tc5::pack8(v, 0);
                    dst[1] = tc5::pack8(v, 8);
                }
            }
        }
        first = false;
        tc5::fence_before_sync();
        __syncthreads();
    }

    // ---- this CTA's weight-gradient partial sums: M = 64 accumulators live in lanes 0-15 of each warp's quadrant
    if (partials) {
        float *mine = partials + (size_t)blockIdx.x * (FF_HID * (in_dim + FF_HID * (n_layers - 1) + FF_OUT));
        const int j = warp * 16 + (lane & 15);
        const bool own = lane < 16;
        const bool have = !first;  // a CTA without tiles contributes zeros
        float v[16];
        // W_in [j][i]
        for (int c = 0; c < (in_dim >> 4); ++c) {
            if (have) tc5::tmem_ld16(trow + TL.win_col + c * 16, v);
            if (own)
#pragma unroll
                for (int i = 0; i < 16; ++i) mine[j * in_dim + c * 16 + i] = have ? v[i] : 0.0f;
        }
        float *ph = mine + FF_HID * in_dim;
        for (int l = 0; l < n_layers - 1; ++l)
            for (int c = 0; c < 4; ++c) {
                if (have) tc5::tmem_ld16(trow + TL.wh_col + l * FF_HID + c * 16, v);
                if (own)
#pragma unroll
                    for (int i = 0; i < 16; ++i) ph[l * FF_HID * FF_HID + j * FF_HID + c * 16 + i] = have ? v[i] : 0.0f;
            }
        float *po = ph + (n_layers - 1) * FF_HID * FF_HID;  // W_out [o][j] from dW_out^T [j][o]
        if (have) tc5::tmem_ld16(trow + TL.wout_col, v);
        if (own)
#pragma unroll
            for (int o = 0; o < 16; ++o) po[o * FF_HID + j] = have ? v[o] : 0.0f;
    }
    tc5::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc5::tmem_dealloc(tmem, tmem_cols);
}

__global__ void k_ffmlp_wgrad_reduce(const float *__restrict__ partials, int n_cta, int n_params, __half *__restrict__ gw) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_params) return;
    float s = 0.0f;
    for (int c = 0; c < n_cta; ++c) s += partials[(size_t)c * n_params + p];
    gw[p] = __float2half_rn(s);
}

static size_t ff_backward_smem(int in_dim, int n_layers) {
    const int fw = in_dim > FF_HID ? in_dim : FF_HID;
    return 128 + FF_ROWS * FF_OUT * 2 + FF_ROWS * FF_HID * 2 + 2 * (size_t)FF_ROWS * fw * 2 + FF_HID * FF_OUT * 2 +
           (size_t)(n_layers - 1) * FF_HID * FF_HID * 2 + (size_t)in_dim * FF_HID * 2;
}
static int ff_backward_grid_max() { return ff_sm_count() * 2; }

extern "C" int64_t ngp_ffmlp_backward_workspace_bytes(int64_t B, int input_dim, int num_layers) {
    (void)B;
    return (int64_t)ff_backward_grid_max() * FF_HID * (input_dim + FF_HID * (num_layers - 1) + FF_OUT) * 4;
}

extern "C" int ngp_ffmlp_backward(const void *grad, const void *inputs, const void *weights, const void *forward_buffer,
                                  int64_t B, int input_dim, int output_dim, int hidden_dim, int num_layers,
                                  int activation, int output_activation, int calc_grad_inputs, void *backward_buffer,
                                  void *grad_inputs, void *grad_weights, void *workspace, int64_t workspace_bytes,
                                  ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(grad, "ngp_ffmlp_backward");
    NGP_REQUIRE(grad && inputs && weights && forward_buffer, NGP_ERR_ARG, "ngp_ffmlp_backward: null pointer");
    NGP_REQUIRE(!calc_grad_inputs || grad_inputs, NGP_ERR_ARG, "ngp_ffmlp_backward: calc_grad_inputs needs grad_inputs");
    int rc = ff_check(B, input_dim, output_dim, hidden_dim, num_layers, activation, output_activation, "ngp_ffmlp_backward");
    if (rc) return rc;
    const int n_params = FF_HID * (input_dim + FF_HID * (num_layers - 1) + FF_OUT);
    const FFBwdLayout TL = ff_bwd_layout(input_dim, num_layers);
    NGP_REQUIRE(TL.total <= 512, NGP_ERR_UNSUPPORTED,
                "ngp_ffmlp_backward: input_dim %d with %d layers needs %d TMEM columns (max 512)", input_dim, num_layers, TL.total);
    const int cols = ff_pow2_cols(TL.total);
    cudaStream_t st = (cudaStream_t)stream;
    if (B == 0) {
        if (grad_weights) cudaMemsetAsync(grad_weights, 0, (size_t)n_params * 2, st);
        return NGP_OK;
    }
    const size_t smem = ff_backward_smem(input_dim, num_layers);
    NGP_REQUIRE(smem <= 227 * 1024, NGP_ERR_UNSUPPORTED, "ngp_ffmlp_backward: configuration needs %zu B of shared memory", smem);
    if (cudaFuncSetAttribute(k_ffmlp_backward, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        ngp_set_error("ngp_ffmlp_backward: cannot reserve %zu B of shared memory", smem);
        return NGP_ERR_CUDA;
    }
    // CTAs per SM from the resources themselves: cudaOccupancyMaxActiveBlocksPerMultiprocessor answered 1 for this kernel
    // (67.7 KB of dynamic shared memory, 40 registers, 128 threads) in most processes and 2 in some; with one CTA per SM
    // the launch takes 1.46 ms, with two 0.88 ms (ncu: warps_active 6.3 % vs 12.4 %).  The CTAs are independent and
    // stride over the tiles by gridDim, so a CTA that does not fit simply runs in a second wave.
    cudaFuncSetAttribute(k_ffmlp_backward, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    int per_sm = (int)((227 * 1024) / (smem + 1024));
    if (per_sm > 512 / cols) per_sm = 512 / cols;
    if (per_sm > 2) per_sm = 2;   // ngp_ffmlp_backward_workspace_bytes sizes the partials for 2 CTAs per SM
    if (per_sm < 1) per_sm = 1;
    const int64_t tiles = ngp_div_up(B, FF_ROWS), resident = (int64_t)ff_sm_count() * per_sm;
    const int grid = (int)(tiles < resident ? tiles : resident);
    float *partials = nullptr;
    if (grad_weights) {
        NGP_REQUIRE(workspace && workspace_bytes >= (int64_t)grid * n_params * 4, NGP_ERR_ARG,
                    "ngp_ffmlp_backward: workspace too small (see ngp_ffmlp_backward_workspace_bytes)");
        partials = (float *)workspace;
    }
    k_ffmlp_backward<<<grid, FF_ROWS, smem, st>>>((const __half *)grad, (const __half *)inputs, (const __half *)weights,
                                                  (const __half *)forward_buffer, (__half *)backward_buffer,
                                                  calc_grad_inputs ? (__half *)grad_inputs : nullptr, partials, B,
                                                  input_dim, num_layers, activation, cols);
    NGP_CHECK_LAUNCH("ngp_ffmlp_backward");
    if (grad_weights) {
        k_ffmlp_wgrad_reduce<<<(n_params + 255) / 256, 256, 0, st>>>(partials, grid, n_params, (__half *)grad_weights);
        NGP_CHECK_LAUNCH("ngp_ffmlp_backward (reduce)");
    }
    return NGP_OK;
}
