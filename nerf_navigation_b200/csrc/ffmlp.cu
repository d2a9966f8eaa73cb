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
    const int tid = threadIdx.x, warp = tid >> 5;
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
    const uint32_t tmem = sh->tmem;
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

extern "C" int ngp_ffmlp_forward(const void *inputs, const void *weights, int64_t B, int input_dim, int output_dim,
                                 int hidden_dim, int num_layers, int activation, int output_activation,
                                 void *forward_buffer, void *outputs, ngp_stream_t stream) {
    NGP_REQUIRE(inputs && weights && outputs, NGP_ERR_ARG, "ngp_ffmlp_forward: null pointer");
    int rc = ff_check(B, input_dim, output_dim, hidden_dim, num_layers, activation, output_activation, "ngp_ffmlp_forward");
    if (rc) return rc;
    if (B == 0) return NGP_OK;
    const size_t smem = ff_forward_smem(input_dim, num_layers);
    NGP_REQUIRE(smem <= 227 * 1024, NGP_ERR_UNSUPPORTED, "ngp_ffmlp_forward: configuration needs %zu B of shared memory", smem);
    if (cudaFuncSetAttribute(k_ffmlp_forward, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        ngp_set_error("ngp_ffmlp_forward: cannot reserve %zu B of shared memory", smem);
        return NGP_ERR_CUDA;
    }
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
