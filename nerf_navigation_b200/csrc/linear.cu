// linear.cu — Y [M, n] = X [M, k] . B for the bias-free Linear layers of the field MLPs when they train:
//     forward   y  = x W^T     (B[k][n] = W[n][k],  W [n_out, n_in] = nn.Linear.weight)
//     backward  dx = dy W      (B[k][n] = W[k][n])
// Reference path: nn.Linear forward / input-gradient of sigma_net / color_net (nerf/network.py:33-68) inside
// Trainer.train_step.  M is the sample count of the step (~3e5), n and k are <= 64: torch hands these to SIMT sgemm
// tiles at ~20 TFLOP/s (0.6 ms for the ten products of a step).  Here every warp owns 16 rows, reads its A fragments
// straight from the row-major activations, multiplies on the tensor cores (mma.sync m16n8k8, 3xTF32 split products, fp32
// accumulate — the same arithmetic as the fused field kernels) against weight fragments staged pre-split in shared
// memory, and streams the result out: one pass over X and Y.
// Bound: HBM streaming, 4 (k + n) bytes per row.
#include "fieldcore.cuh"

#define LIN_THREADS 128
#define LIN_MAX 64

// VEC (k_dim % 16 == 0, X 16-byte aligned): the order of the K index inside an MMA is free as long as A and B agree, so
// lane t takes the four CONSECUTIVE inputs 16j + 4t .. 4t+3 of its rows with one 16-byte load and feeds them to two
// k-tiles (slots t / t+4 of tile 2j = inputs 4t, 4t+1; of tile 2j+1 = inputs 4t+2, 4t+3): a quarter of the load
// instructions and L1 wavefronts of the scalar fragment order (the scalar form is LSU-bound: 64-77 % of the data pipe).
template <int NT, bool VEC>
__global__ void __launch_bounds__(LIN_THREADS)
k_linear_small(const float *__restrict__ X, const float *__restrict__ W, float *__restrict__ Y, int64_t M, int k_dim,
               int n_dim, int ld_n, int ld_k) {
    __shared__ float4 frag[NT * (LIN_MAX / 8) * 32];   // [nt][kt][lane]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int ktiles = (k_dim + 7) >> 3;
    for (int i = tid; i < NT * ktiles * 32; i += LIN_THREADS) {
        const int ln = i & 31, kt = (i >> 5) % ktiles, nt = (i >> 5) / ktiles, gg = ln >> 2, tt = ln & 3;
        const int n = 8 * nt + gg;
        const int k0 = VEC ? 16 * (kt >> 1) + 4 * tt + 2 * (kt & 1) : 8 * kt + tt, k1 = VEC ? k0 + 1 : k0 + 4;
        const float b0 = (n < n_dim && k0 < k_dim) ? W[(int64_t)n * ld_n + (int64_t)k0 * ld_k] : 0.0f;
        const float b1 = (n < n_dim && k1 < k_dim) ? W[(int64_t)n * ld_n + (int64_t)k1 * ld_k] : 0.0f;
        frag[(nt * (LIN_MAX / 8) + kt) * 32 + ln] = fq_split_pair(b0, b1);
    }
    __syncthreads();
    const int64_t tiles = (M + 15) >> 4;
    for (int64_t tile = (int64_t)blockIdx.x * (LIN_THREADS / 32) + warp; tile < tiles;
         tile += (int64_t)gridDim.x * (LIN_THREADS / 32)) {
        const int64_t r0 = tile * 16 + g, r1 = r0 + 8;
        const bool v0 = r0 < M, v1 = r1 < M;
        const float *x0 = X + r0 * k_dim, *x1 = X + r1 * k_dim;
        float acc[NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.0f;
        if (VEC) {
            for (int j = 0; j < (ktiles >> 1); ++j) {
                const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
                const float4 p0 = v0 ? __ldg(reinterpret_cast<const float4 *>(x0 + 16 * j) + t) : z;
                const float4 p1 = v1 ? __ldg(reinterpret_cast<const float4 *>(x1 + 16 * j) + t) : z;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint32_t ahi[4], alo[4];
                    fq_split(h ? p0.z : p0.x, ahi[0], alo[0]);
                    fq_split(h ? p1.z : p1.x, ahi[1], alo[1]);
                    fq_split(h ? p0.w : p0.y, ahi[2], alo[2]);
                    fq_split(h ? p1.w : p1.y, ahi[3], alo[3]);
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt)
                        fq_mma3(acc[nt], ahi, alo, frag[(nt * (LIN_MAX / 8) + 2 * j + h) * 32 + lane]);
                }
            }
        } else {
            for (int kt = 0; kt < ktiles; ++kt) {
                const int ka = 8 * kt + t, kb = ka + 4;
                float a[4];
                a[0] = (v0 && ka < k_dim) ? __ldg(x0 + ka) : 0.0f;
                a[1] = (v1 && ka < k_dim) ? __ldg(x1 + ka) : 0.0f;
                a[2] = (v0 && kb < k_dim) ? __ldg(x0 + kb) : 0.0f;
                a[3] = (v1 && kb < k_dim) ? __ldg(x1 + kb) : 0.0f;
                uint32_t ahi[4], alo[4];
#pragma unroll
                for (int c = 0; c < 4; ++c) fq_split(a[c], ahi[c], alo[c]);
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) fq_mma3(acc[nt], ahi, alo, frag[(nt * (LIN_MAX / 8) + kt) * 32 + lane]);
            }
        }
        // c0,c1: (row g, cols 8nt+2t, +1); c2,c3: (row g+8, ...)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int col = 8 * nt + 2 * t;
            if (col + 1 < n_dim && (n_dim & 1) == 0) {
                if (v0) *reinterpret_cast<float2 *>(Y + r0 * n_dim + col) = make_float2(acc[nt][0], acc[nt][1]);
                if (v1) *reinterpret_cast<float2 *>(Y + r1 * n_dim + col) = make_float2(acc[nt][2], acc[nt][3]);
            } else {
                if (col < n_dim) {
                    if (v0) Y[r0 * n_dim + col] = acc[nt][0];
                    if (v1) Y[r1 * n_dim + col] = acc[nt][2];
                }
                if (col + 1 < n_dim) {
                    if (v0) Y[r0 * n_dim + col + 1] = acc[nt][1];
                    if (v1) Y[r1 * n_dim + col + 1] = acc[nt][3];
                }
            }
        }
    }
}

// transpose_w = 0: Y [M, n_out] = X [M, n_in] W^T (forward);  1: Y [M, n_in] = X [M, n_out] W (input gradient).
// W is always nn.Linear.weight [n_out, n_in], row-major.
extern "C" int ngp_linear_small(const float *X, const float *W, float *Y, int64_t M, int n_out, int n_in, int transpose_w,
                                ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(X, "ngp_linear_small");
    NGP_REQUIRE(n_out >= 1 && n_out <= LIN_MAX && n_in >= 1 && n_in <= LIN_MAX && M >= 0, NGP_ERR_UNSUPPORTED,
                "ngp_linear_small: needs 1 <= n_out, n_in <= 64 (got %d, %d)", n_out, n_in);
    if (M == 0) return NGP_OK;
    NGP_REQUIRE(X && W && Y, NGP_ERR_ARG, "ngp_linear_small: null pointer");
    NGP_REQUIRE(((uintptr_t)Y & 7u) == 0, NGP_ERR_ARG, "ngp_linear_small: Y must be 8-byte aligned");
    const int k_dim = transpose_w ? n_out : n_in, n_dim = transpose_w ? n_in : n_out;
    const int ld_n = transpose_w ? 1 : n_in, ld_k = transpose_w ? n_in : 1;   // B[k][n] = W[n * ld_n + k * ld_k]
    int dev = 0, sms = NGP_SM_COUNT;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t want = ngp_div_up(ngp_div_up(M, 16), LIN_THREADS / 32);
    const int64_t cap = (int64_t)sms * 8;
    const unsigned grid = (unsigned)(want < cap ? want : cap);
    cudaStream_t st = (cudaStream_t)stream;
    const int nt = (n_dim + 7) >> 3;
    const bool vec = (k_dim % 16) == 0 && ((uintptr_t)X & 15u) == 0;
#define LIN_GO(NT)                                                                                        \
    if (vec) k_linear_small<NT, true><<<grid, LIN_THREADS, 0, st>>>(X, W, Y, M, k_dim, n_dim, ld_n, ld_k); \
    else k_linear_small<NT, false><<<grid, LIN_THREADS, 0, st>>>(X, W, Y, M, k_dim, n_dim, ld_n, ld_k)
    switch (nt) {
        case 1: LIN_GO(1); break;
        case 2: LIN_GO(2); break;
        case 3: LIN_GO(3); break;
        case 4: LIN_GO(4); break;
        case 5: LIN_GO(5); break;
        case 6: LIN_GO(6); break;
        case 7: LIN_GO(7); break;
        default: LIN_GO(8); break;
    }
#undef LIN_GO
    NGP_CHECK_LAUNCH("ngp_linear_small");
    return NGP_OK;
}
