// sh.cu — SHEncoder forward / backward (degree <= 4).  Reference: shencoder/src/shencoder.cu:27-383.
// The backward recomputes the Jacobian from the directions instead of reading a stored dy_dx [B, 3*deg^2].
#include "common.cuh"
#include "sh.cuh"

__global__ void __launch_bounds__(256)
k_sh_forward(const float *__restrict__ inputs, float *__restrict__ outputs, int64_t B, int degree,
             float *__restrict__ dy_dx) {
    const int64_t b = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (b >= B) return;
    const float x = inputs[b * 3], y = inputs[b * 3 + 1], z = inputs[b * 3 + 2];
    const int n = degree * degree;
    float v[16];
    ngp_sh4(x, y, z, v);
    float *o = outputs + b * n;
#pragma unroll
    for (int i = 0; i < 16; ++i)
        if (i < n) o[i] = v[i];
    if (dy_dx) {
        float gx[16], gy[16], gz[16];
        ngp_sh4_jac(x, y, z, gx, gy, gz);
        float *g = dy_dx + b * 3 * n;
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if (i < n) { g[i] = gx[i]; g[n + i] = gy[i]; g[2 * n + i] = gz[i]; }
    }
}

__global__ void __launch_bounds__(256)
k_sh_backward(const float *__restrict__ grad, const float *__restrict__ inputs, int64_t B, int degree,
              float *__restrict__ grad_inputs) {
    const int64_t b = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (b >= B) return;
    const float x = inputs[b * 3], y = inputs[b * 3 + 1], z = inputs[b * 3 + 2];
    const int n = degree * degree;
    float gx[16], gy[16], gz[16];
    ngp_sh4_jac(x, y, z, gx, gy, gz);
    float ax = 0.0f, ay = 0.0f, az = 0.0f;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        if (i < n) {
            const float g = grad[b * n + i];
            ax = fmaf(g, gx[i], ax);
            ay = fmaf(g, gy[i], ay);
            az = fmaf(g, gz[i], az);
        }
    }
    // the reference accumulates into a pre-zeroed buffer (shencoder.cu:378-381); same result
    grad_inputs[b * 3] = ax; grad_inputs[b * 3 + 1] = ay; grad_inputs[b * 3 + 2] = az;
}

extern "C" int ngp_sh_encode_forward(const float *inputs, float *outputs, int64_t B, int D, int degree,
                                     int calc_grad_inputs, float *dy_dx, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(inputs, "ngp_sh_encode_forward");
    if (B == 0) return NGP_OK;
    NGP_REQUIRE(inputs && outputs, NGP_ERR_ARG, "ngp_sh_encode_forward: null pointer");
    NGP_REQUIRE(D == 3, NGP_ERR_UNSUPPORTED, "SHEncoder: input_dim must be 3 (got %d)", D);
    NGP_REQUIRE(degree >= 1 && degree <= 4, NGP_ERR_UNSUPPORTED, "SHEncoder: degree must be in [1,4] (got %d)", degree);
    k_sh_forward<<<(unsigned)ngp_div_up(B, 256), 256, 0, (cudaStream_t)stream>>>(inputs, outputs, B, degree,
                                                                               calc_grad_inputs ? dy_dx : nullptr);
    NGP_CHECK_LAUNCH("ngp_sh_encode_forward");
    return NGP_OK;
}

extern "C" int ngp_sh_encode_backward(const float *grad, const float *inputs, int64_t B, int D, int degree,
                                      float *grad_inputs, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(grad, "ngp_sh_encode_backward");
    if (B == 0) return NGP_OK;
    NGP_REQUIRE(grad && inputs && grad_inputs, NGP_ERR_ARG, "ngp_sh_encode_backward: null pointer");
    NGP_REQUIRE(D == 3, NGP_ERR_UNSUPPORTED, "SHEncoder: input_dim must be 3 (got %d)", D);
    NGP_REQUIRE(degree >= 1 && degree <= 4, NGP_ERR_UNSUPPORTED, "SHEncoder: degree must be in [1,4] (got %d)", degree);
    k_sh_backward<<<(unsigned)ngp_div_up(B, 256), 256, 0, (cudaStream_t)stream>>>(grad, inputs, B, degree, grad_inputs);
    NGP_CHECK_LAUNCH("ngp_sh_encode_backward");
    return NGP_OK;
}
