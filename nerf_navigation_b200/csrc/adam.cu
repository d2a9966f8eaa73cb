// adam.cu — fused Adam step for the training configuration (BASELINE config 5).
// Reference: torch.optim.Adam(model.get_params(lr), betas=(0.9, 0.99), eps=1e-15) (main_nerf.py:126) stepped by
// Trainer.train_one_epoch (nerf/utils.py:784-791), optionally followed by an EMA of the parameters (decay 0.95,
// nerf/utils.py:814-815).  torch's single-tensor Adam arithmetic is followed operation by operation:
//     m = lerp(m, g, 1 - b1);  v = b2 v + (1 - b2) g g;  p -= (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
// One pass over the hash table (12.66 M fp32 parameters): 16 B read + 12 B written per parameter (+ 8 B with the EMA),
// HBM-bound; the all-reduce's 1/world is folded into `grad_scale` so the mean needs no pass.  torch.amp.GradScaler
// semantics (nerf/utils.py:784-791 scaler.step / scaler.update): `amp_scale` (device, the loss scale: gradients are divided
// by it) and `found_inf` (device, non-zero = some gradient was inf / nan: the step is skipped, moments and EMA untouched)
// are read by the kernel, so a skipped step needs no host round trip.
#include "common.cuh"

#define ADAM_THREADS 256

struct AdamScalars {
    float lr_c1;      // lr / (1 - b1^t)
    float sqrt_c2;    // sqrt(1 - b2^t)
    float b1, b2, eps, grad_scale, ema_decay;
};

__device__ __forceinline__ void adam_one(float &p, float g, float &m, float &v, const AdamScalars &s) {
    g *= s.grad_scale;
    m = fmaf(1.0f - s.b1, g - m, m);                         // torch: exp_avg.lerp_(grad, 1 - beta1)
    v = fmaf(__fmul_rn(1.0f - s.b2, g), g, __fmul_rn(v, s.b2));  // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
    const float denom = __fadd_rn(__fdiv_rn(sqrtf(v), s.sqrt_c2), s.eps);
    p = __fsub_rn(p, __fmul_rn(s.lr_c1, __fdiv_rn(m, denom)));   // param.addcdiv_(exp_avg, denom, value=-step_size)
}

__global__ void __launch_bounds__(ADAM_THREADS)
k_adam_step(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m, float *__restrict__ v,
            float *__restrict__ ema, int64_t n, AdamScalars s, const float *__restrict__ amp_scale,
            const float *__restrict__ found_inf) {
    if (found_inf && *found_inf != 0.0f) return;           // GradScaler: skip the whole step
    if (amp_scale) s.grad_scale = __fdiv_rn(s.grad_scale, *amp_scale);
    const int64_t n4 = n >> 2;
    const int64_t stride = (int64_t)gridDim.x * ADAM_THREADS;
    for (int64_t i = (int64_t)blockIdx.x * ADAM_THREADS + threadIdx.x; i < n4; i += stride) {
        float4 pp = reinterpret_cast<float4 *>(p)[i];
        const float4 gg = __ldg(reinterpret_cast<const float4 *>(g) + i);
        float4 mm = reinterpret_cast<float4 *>(m)[i], vv = reinterpret_cast<float4 *>(v)[i];
        adam_one(pp.x, gg.x, mm.x, vv.x, s);
        adam_one(pp.y, gg.y, mm.y, vv.y, s);
        adam_one(pp.z, gg.z, mm.z, vv.z, s);
        adam_one(pp.w, gg.w, mm.w, vv.w, s);
        reinterpret_cast<float4 *>(p)[i] = pp;
        reinterpret_cast<float4 *>(m)[i] = mm;
        reinterpret_cast<float4 *>(v)[i] = vv;
        if (ema) {  // shadow = decay * shadow + (1 - decay) * param   (torch_ema semantics: shadow -= (1-decay)(shadow-param))
            float4 e = reinterpret_cast<float4 *>(ema)[i];
            e.x = fmaf(1.0f - s.ema_decay, pp.x - e.x, e.x);
            e.y = fmaf(1.0f - s.ema_decay, pp.y - e.y, e.y);
            e.z = fmaf(1.0f - s.ema_decay, pp.z - e.z, e.z);
            e.w = fmaf(1.0f - s.ema_decay, pp.w - e.w, e.w);
            reinterpret_cast<float4 *>(ema)[i] = e;
        }
    }
    // tail (n not a multiple of 4)
    const int64_t t = (n4 << 2) + (int64_t)blockIdx.x * ADAM_THREADS + threadIdx.x;
    if (t < n) {
        float pp = p[t], mm = m[t], vv = v[t];
        adam_one(pp, g[t], mm, vv, s);
        p[t] = pp; m[t] = mm; v[t] = vv;
        if (ema) ema[t] = fmaf(1.0f - s.ema_decay, pp - ema[t], ema[t]);
    }
}

extern "C" int ngp_adam_step(float *param, const float *grad, float *exp_avg, float *exp_avg_sq, float *ema, int64_t n,
                             float lr, float beta1, float beta2, float eps, int64_t step, float grad_scale,
                             float ema_decay, const float *amp_scale, const float *found_inf, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(param, "ngp_adam_step");
    if (n == 0) return NGP_OK;
    NGP_REQUIRE(param && grad && exp_avg && exp_avg_sq, NGP_ERR_ARG, "ngp_adam_step: null pointer");
    NGP_REQUIRE(n >= 0 && step >= 1, NGP_ERR_ARG, "ngp_adam_step: bad n / step (step counts from 1)");
    NGP_REQUIRE((((uintptr_t)param | (uintptr_t)grad | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq | (uintptr_t)ema) & 15u) == 0,
                NGP_ERR_ARG, "ngp_adam_step: pointers must be 16-byte aligned");
    AdamScalars s;
    // bias corrections in double like torch's python scalars (1 - beta ** step), then rounded once
    const double c1 = 1.0 - pow((double)beta1, (double)step), c2 = 1.0 - pow((double)beta2, (double)step);
    s.lr_c1 = (float)((double)lr / c1);
    s.sqrt_c2 = (float)sqrt(c2);
    s.b1 = beta1; s.b2 = beta2; s.eps = eps; s.grad_scale = grad_scale; s.ema_decay = ema_decay;
    int dev = 0, sms = NGP_SM_COUNT;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t want = ngp_div_up(ngp_div_up(n, 4) > 0 ? ngp_div_up(n, 4) : 1, ADAM_THREADS);
    const int64_t cap = (int64_t)sms * 8;
    const int blocks = (int)(want < cap ? want : cap);
    k_adam_step<<<blocks, ADAM_THREADS, 0, (cudaStream_t)stream>>>(param, grad, exp_avg, exp_avg_sq, ema, n, s, amp_scale,
                                                                   found_inf);
    NGP_CHECK_LAUNCH("ngp_adam_step");
    return NGP_OK;
}
