// wgrad.cu — weight gradient of a bias-free Linear layer of the field MLPs in the training configuration:
//     dW [N_out, N_in] = dY^T X,   dY [M, N_out], X [M, N_in],   M = samples of the step (~3e5), N_out, N_in <= 64.
// Reference path: nn.Linear backward of sigma_net / color_net (nerf/network.py:33-68) inside Trainer.train_step; torch
// hands this to a SIMT sgemm whose tile shape (128x32, K = M) leaves most of the chip idle (0.3 ms per layer, 1.5 ms
// of a 5.5 ms step).  Here the long sample axis is the MMA's K dimension: every CTA reduces a contiguous slice of the
// samples into a private 64 x 64 tile on the tensor cores (mma.sync m16n8k8, 3xTF32 split products, fp32 accumulate),
// the per-CTA partials are summed in CTA order by a second kernel — deterministic, no float atomics.
// Bound: HBM/L2 streaming of dY and X once (4 (N_out + N_in) bytes per sample).
#include "fieldcore.cuh"

#define WG_THREADS 128
#define WG_MAXN 64
#ifndef WG_TWO_LEVEL
#define WG_TWO_LEVEL 1
#endif

// One k8-step worth of operands in MMA fragment order.
struct WgFrag {
    float a[4];
    float b[8][2];
};

__device__ __forceinline__ void wg_load(WgFrag &f, const float *__restrict__ dY, const float *__restrict__ X, int64_t m0,
                                        int64_t m_end, int n_out, int n_in, int o0, int g, int t, int ntiles) {
    // A[o][k] = dY[m0 + k][o0 + o]: a0 (row g, k t), a1 (row g+8, k t), a2 (row g, k t+4), a3 (row g+8, k t+4)
    // B[k][n] = X[m0 + k][8 nt + n]: b0 (k t, n g), b1 (k t+4, n g)
    const int64_t ma = m0 + t, mb = m0 + t + 4;
    const bool va = ma < m_end, vb = mb < m_end;
    const bool oa = o0 + g < n_out, ob = o0 + g + 8 < n_out;
    f.a[0] = (va && oa) ? __ldg(dY + ma * n_out + o0 + g) : 0.0f;
    f.a[1] = (va && ob) ? __ldg(dY + ma * n_out + o0 + g + 8) : 0.0f;
    f.a[2] = (vb && oa) ? __ldg(dY + mb * n_out + o0 + g) : 0.0f;
    f.a[3] = (vb && ob) ? __ldg(dY + mb * n_out + o0 + g + 8) : 0.0f;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
        const int i = 8 * nt + g;
        const bool vi = nt < ntiles && i < n_in;
        f.b[nt][0] = (va && vi) ? __ldg(X + ma * n_in + i) : 0.0f;
        f.b[nt][1] = (vb && vi) ? __ldg(X + mb * n_in + i) : 0.0f;
    }
}

// partials: [nblocks][64][64].  The CTA's four warps are (output tile of 16 features) x (slice of the CTA's samples):
// layers with fewer than 64 outputs split the samples further instead of idling warps; slices of one tile are added in
// warp order through shared memory.
__global__ void __launch_bounds__(WG_THREADS)
k_wgrad_partial(const float *__restrict__ dY, const float *__restrict__ X, int64_t M, int n_out, int n_in,
                int64_t rows_per_block, float *__restrict__ partials) {
    __shared__ float red[WG_THREADS / 32][8][4][32];
    __shared__ float lvl2[WG_THREADS / 32][8][4][32];   // second accumulation level, private to its thread
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    const int mtiles = (n_out + 15) >> 4;                 // 1..4
    const int ksplit = mtiles == 1 ? 4 : (mtiles == 2 ? 2 : 1);
    const int mtile = warp / ksplit, kpart = warp - mtile * ksplit;
    const int o0 = mtile * 16;
    const int ntiles = (n_in + 7) >> 3;
    float acc[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.0f;
    const int64_t blk_begin = (int64_t)blockIdx.x * rows_per_block;
    const int64_t blk_end = min(M, blk_begin + rows_per_block);
    const int64_t per = ((rows_per_block / ksplit + 7) >> 3) << 3;
    const int64_t m_begin = blk_begin + kpart * per;
    const int64_t m_end = min(blk_end, m_begin + per);
    const bool active = mtile < mtiles && m_begin < m_end;
    if (active) {
        // two-level accumulation: the MMA accumulators take 256 samples, then go into a second level kept in shared
        // memory (no extra registers) — a running fp32 sum over the CTA's thousands of samples loses ~sqrt(n) ulp (seen as
        // 2.6e-5 of the largest entry between two partitions of the same 2.6x10^6-sample batch), 256-sample blocks summed
        // in a second level lose a hundredth of that (7e-8)
#pragma unroll
        for (int nt = 0; nt < 8; ++nt)
#pragma unroll
            for (int c = 0; c < 4; ++c) lvl2[warp][nt][c][lane] = 0.0f;
        int in_blk = 0;
        WgFrag cur, nxt;
        wg_load(cur, dY, X, m_begin, m_end, n_out, n_in, o0, g, t, ntiles);
        for (int64_t m0 = m_begin; m0 < m_end; m0 += 8) {
            if (m0 + 8 < m_end) wg_load(nxt, dY, X, m0 + 8, m_end, n_out, n_in, o0, g, t, ntiles);   // in flight during the MMAs
            uint32_t ahi[4], alo[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) fq_split(cur.a[c], ahi[c], alo[c]);
#pragma unroll
            for (int nt = 0; nt < 8; ++nt)
                if (nt < ntiles) fq_mma3(acc[nt], ahi, alo, fq_split_pair(cur.b[nt][0], cur.b[nt][1]));
            cur = nxt;
            if (WG_TWO_LEVEL && ++in_blk == 32) {
                in_blk = 0;
#pragma unroll
                for (int nt = 0; nt < 8; ++nt)
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        lvl2[warp][nt][c][lane] = __fadd_rn(lvl2[warp][nt][c][lane], acc[nt][c]);
                        acc[nt][c] = 0.0f;
                    }
            }
        }
#pragma unroll
        for (int nt = 0; nt < 8; ++nt)
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[nt][c] = __fadd_rn(lvl2[warp][nt][c][lane], acc[nt][c]);
    }
    if (ksplit > 1) {   // add the sample slices of a tile in warp order
#pragma unroll
        for (int nt = 0; nt < 8; ++nt)
#pragma unroll
            for (int c = 0; c < 4; ++c) red[warp][nt][c][lane] = acc[nt][c];
        __syncthreads();
        if (kpart == 0) {
            for (int k = 1; k < ksplit; ++k)
#pragma unroll
                for (int nt = 0; nt < 8; ++nt)
#pragma unroll
                    for (int c = 0; c < 4; ++c) acc[nt][c] = __fadd_rn(acc[nt][c], red[warp + k][nt][c][lane]);
        }
    }
    if (kpart != 0 || mtile >= 4) return;
    float *out = partials + (int64_t)blockIdx.x * WG_MAXN * WG_MAXN;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {   // c0,c1: (row g, cols 2t, 2t+1); c2,c3: (row g+8, ...)
        const int col = 8 * nt + 2 * t;
        out[(o0 + g) * WG_MAXN + col] = acc[nt][0];
        out[(o0 + g) * WG_MAXN + col + 1] = acc[nt][1];
        out[(o0 + g + 8) * WG_MAXN + col] = acc[nt][2];
        out[(o0 + g + 8) * WG_MAXN + col + 1] = acc[nt][3];
    }
}

// dW[e] = sum over the CTAs' partials, in double, in a fixed order: 8 lanes per output element take the partials
// b = j, j + 8, ... each, then the 8 sub-sums are added in lane order (deterministic; 8x shorter dependent chains than one
// thread per element: 20 -> ~5 us per layer)
__global__ void __launch_bounds__(256)
k_wgrad_reduce(const float *__restrict__ partials, int nblocks, int n_out, int n_in, float *__restrict__ dW) {
    const int e = blockIdx.x * 32 + (threadIdx.x >> 3), j = threadIdx.x & 7;
    const bool in = e < n_out * n_in;
    const int o = in ? e / n_in : 0, i = in ? e - o * n_in : 0;
    double s = 0.0;
    if (in)
        for (int b = j; b < nblocks; b += 8) s += (double)__ldg(partials + (int64_t)b * WG_MAXN * WG_MAXN + o * WG_MAXN + i);
    // ordered sum of the 8 sub-sums: lane 0 of the group adds them 0, 1, ..., 7
    double tot = s;
#pragma unroll
    for (int k = 1; k < 8; ++k) {
        const double v = __shfl_down_sync(0xffffffffu, s, k, 8);
        tot += v;
    }
    if (in && j == 0) dW[e] = (float)tot;
}

static int wg_blocks(int64_t M) {
    int dev = 0, sms = NGP_SM_COUNT;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t want = ngp_div_up(M, 256);   // at least 256 samples per CTA
    const int64_t cap = (int64_t)sms * 4;
    return (int)(want < 1 ? 1 : (want < cap ? want : cap));
}

extern "C" int64_t ngp_linear_wgrad_workspace_bytes(int64_t M) {
    return (int64_t)wg_blocks(M) * WG_MAXN * WG_MAXN * (int64_t)sizeof(float);
}

extern "C" int ngp_linear_wgrad(const float *dY, const float *X, float *dW, int64_t M, int n_out, int n_in, void *workspace,
                                int64_t workspace_bytes, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(dY, "ngp_linear_wgrad");
    NGP_REQUIRE(dY && X && dW, NGP_ERR_ARG, "ngp_linear_wgrad: null pointer");
    NGP_REQUIRE(n_out >= 1 && n_out <= WG_MAXN && n_in >= 1 && n_in <= WG_MAXN && M >= 0, NGP_ERR_UNSUPPORTED,
                "ngp_linear_wgrad: needs 1 <= n_out, n_in <= 64 (got %d, %d)", n_out, n_in);
    cudaStream_t st = (cudaStream_t)stream;
    if (M == 0) {
        cudaMemsetAsync(dW, 0, (size_t)n_out * n_in * sizeof(float), st);
        return NGP_OK;
    }
    const int blocks = wg_blocks(M);
    NGP_REQUIRE(workspace && workspace_bytes >= ngp_linear_wgrad_workspace_bytes(M), NGP_ERR_ARG,
                "ngp_linear_wgrad: workspace too small");
    int64_t rows = ngp_div_up(M, blocks);
    rows = ngp_div_up(rows, 8) * 8;
    k_wgrad_partial<<<blocks, WG_THREADS, 0, st>>>(dY, X, M, n_out, n_in, rows, (float *)workspace);
    NGP_CHECK_LAUNCH("ngp_linear_wgrad(partial)");
    k_wgrad_reduce<<<(n_out * n_in + 31) / 32, 256, 0, st>>>((const float *)workspace, blocks, n_out, n_in, dW);
    NGP_CHECK_LAUNCH("ngp_linear_wgrad(reduce)");
    return NGP_OK;
}

// ------------------------------------------------------------------------------------------------ allocation-free variant
// ngp_linear_wgrad_static: the same two kernels with the per-CTA partials in a file-scope device array instead of a
// caller-provided workspace, for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc).  No counterpart in the reference.
//   * limit: M <= NGP_LINEAR_WGRAD_STATIC_MAX_M = 65536 samples (256 CTAs x 64 x 64 floats = 4 MiB, held by every GPU that
//     loads the library until the process exits); larger M is refused with NGP_ERR_UNSUPPORTED;
//   * the array is shared by every call on a device and keeps the last call's contents.  Nothing has to start at zero:
//     every partial the reduction reads (rows < 16 ceil(n_out / 16), all 64 columns, of the call's own CTAs) is written by
//     the same call's first kernel.  Calls that overlap on the GPU (different streams, or graphs replayed concurrently)
//     race on it: serialise them;
//   * in host code `g_wgrad_static_scratch` is only a stand-in for the symbol, not a device address: the address is
//     looked up once per device with cudaGetSymbolAddress (that device current) and cached.  The first call on a device
//     does this lookup; make one call before starting a graph capture.
#define WG_STATIC_MAX_M 65536
#define WG_STATIC_BLOCKS (WG_STATIC_MAX_M / 256)
#define WG_STATIC_MAX_DEVICES 64
static __device__ __align__(256) float g_wgrad_static_scratch[(size_t)WG_STATIC_BLOCKS * WG_MAXN * WG_MAXN];

static float *wg_static_ptr(int dev) {
    static float *ptrs[WG_STATIC_MAX_DEVICES] = {};
    if (dev < 0 || dev >= WG_STATIC_MAX_DEVICES) return nullptr;
    float *p = __atomic_load_n(&ptrs[dev], __ATOMIC_ACQUIRE);
    if (p) return p;
    void *sym = nullptr;   // the caller made `dev` current: the lookup returns this device's copy of the array
    if (cudaGetSymbolAddress(&sym, g_wgrad_static_scratch) != cudaSuccess || !sym) {
        cudaGetLastError();
        return nullptr;
    }
    __atomic_store_n(&ptrs[dev], (float *)sym, __ATOMIC_RELEASE);   // racing first calls store the same value
    return (float *)sym;
}

extern "C" int ngp_linear_wgrad_static(const float *dY, const float *X, float *dW, int64_t M, int n_out, int n_in,
                                       ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(dY, "ngp_linear_wgrad_static");
    NGP_REQUIRE(dY && X && dW, NGP_ERR_ARG, "ngp_linear_wgrad_static: null pointer");
    NGP_REQUIRE(n_out >= 1 && n_out <= WG_MAXN && n_in >= 1 && n_in <= WG_MAXN && M >= 0, NGP_ERR_UNSUPPORTED,
                "ngp_linear_wgrad_static: needs 1 <= n_out, n_in <= 64 (got %d, %d)", n_out, n_in);
    NGP_REQUIRE(M <= WG_STATIC_MAX_M, NGP_ERR_UNSUPPORTED,
                "ngp_linear_wgrad_static: M = %lld exceeds the static scratch limit of %d samples; use ngp_linear_wgrad with a "
                "caller-provided workspace", (long long)M, WG_STATIC_MAX_M);
    cudaStream_t st = (cudaStream_t)stream;
    if (M == 0) {
        cudaMemsetAsync(dW, 0, (size_t)n_out * n_in * sizeof(float), st);
        return NGP_OK;
    }
    int dev = 0;
    NGP_REQUIRE(cudaGetDevice(&dev) == cudaSuccess, NGP_ERR_CUDA, "ngp_linear_wgrad_static: cudaGetDevice failed");
    float *scratch = wg_static_ptr(dev);
    NGP_REQUIRE(scratch != nullptr, NGP_ERR_CUDA, "ngp_linear_wgrad_static: cudaGetSymbolAddress failed on device %d", dev);
    int blocks = wg_blocks(M);
    if (blocks > WG_STATIC_BLOCKS) blocks = WG_STATIC_BLOCKS;   // cannot happen for M <= 65536 (>= 256 samples per CTA)
    int64_t rows = ngp_div_up(M, blocks);
    rows = ngp_div_up(rows, 8) * 8;
    k_wgrad_partial<<<blocks, WG_THREADS, 0, st>>>(dY, X, M, n_out, n_in, rows, scratch);
    NGP_CHECK_LAUNCH("ngp_linear_wgrad_static(partial)");
    k_wgrad_reduce<<<(n_out * n_in + 31) / 32, 256, 0, st>>>(scratch, blocks, n_out, n_in, dW);
    NGP_CHECK_LAUNCH("ngp_linear_wgrad_static(reduce)");
    return NGP_OK;
}
