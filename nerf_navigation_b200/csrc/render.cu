// render.cu — fused inference renderer: AABB near/far -> occupancy-grid march -> hash-grid field + colour MLP ->
// front-to-back compositing with early termination -> background / depth normalisation, in ONE persistent kernel.
// Replaces the host loop of NeRFRenderer.run_cuda's eval branch (nerf/renderer.py:325-374):
//     while step < max_steps: march_rays (raymarching.cu:706-814) -> self(xyzs, dirs) (network.py:95-124)
//                             -> composite_rays (:828-913) -> rays_alive = rays_alive[rays_alive >= 0]   (host sync)
// Per-sample positions, sigma and rgb never reach HBM: HBM sees 24 B in and 20 B out per RAY.
//
// Organisation (same warp-cooperative field core as renderfield.cu): a warp owns 16 ray slots; lane (g, t) marches
// slot g + 8 (t & 1) (two lanes per slot compute identical values), the 16 pending samples form one mma.sync row tile,
// the field is evaluated, every slot composites its sample, and a slot whose ray ended (t >= far, T < 1e-4, sample
// cap) writes its pixel and takes the next ray from a global counter (one warp-aggregated atomic per refill).  So the
// row tile stays full while rays of very different lengths come and go — the role of the reference's alive-list
// compaction, without its per-iteration host round trip.
//
// Differences from the reference loop, all inside its fp tolerance: t is the marcher's own abscissa (the reference
// re-sums deltas in composite_rays, :876, a <= 1 ulp wobble that depends on how its n_step chunks fall); a ray takes at
// most max_steps samples (reference: max_steps rounded up to its last chunk of <= 8); perturb is not supported here.
#include <float.h>

#include "march.cuh"
#include "renderfield.cuh"

#define RI_SLOTS (RF_WARPS * 16)

struct __align__(16) RenderRaysSmem {
    RenderFieldSmem f;
    float ox[RI_SLOTS], oy[RI_SLOTS], oz[RI_SLOTS], dx[RI_SLOTS], dy[RI_SLOTS], dz[RI_SLOTS];
    float t[RI_SLOTS], near[RI_SLOTS], far[RI_SLOTS];
    float ws[RI_SLOTS], depth[RI_SLOTS], r[RI_SLOTS], g[RI_SLOTS], b[RI_SLOTS];
    uint32_t nsamp[RI_SLOTS];
};

struct RenderRaysParams {
    const float *rays_o, *rays_d, *aabb, *bg;
    int64_t N;
    int bg_stride;
    float min_near, density_scale, bound;
    uint32_t max_steps;
    const float *table;
    int align_corners;
    float *image, *depth, *weights_sum;
    unsigned long long *counter;
    unsigned long long *nsamples;  // optional statistics: total samples evaluated
};

// write the finished ray of `slot` (renderer.py:371-374: image + (1 - ws) * bg; clamp(depth - near, 0) / (far - near))
__device__ __forceinline__ void ri_finish(const RenderRaysSmem &S, const RenderRaysParams &P, int slot, int64_t id) {
    const float ws = S.ws[slot], om = 1.0f - ws;
    const float *bg = P.bg ? P.bg + (int64_t)P.bg_stride * id : nullptr;
    const float b0 = bg ? __ldg(bg) : 1.0f, b1 = bg ? __ldg(bg + 1) : 1.0f, b2 = bg ? __ldg(bg + 2) : 1.0f;
    P.image[id * 3] = __fadd_rn(S.r[slot], __fmul_rn(om, b0));
    P.image[id * 3 + 1] = __fadd_rn(S.g[slot], __fmul_rn(om, b1));
    P.image[id * 3 + 2] = __fadd_rn(S.b[slot], __fmul_rn(om, b2));
    const float near = S.near[slot], far = S.far[slot];
    P.depth[id] = __fdiv_rn(fmaxf(S.depth[slot] - near, 0.0f), far - near);
    if (P.weights_sum) P.weights_sum[id] = ws;
}

template <bool SAFE>
__device__ __forceinline__ void ri_run(RenderRaysSmem &S, const RenderRaysParams &P, const MarchCfg &cfg) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int myh = t & 1;                    // which row of the lane's pair (g, g + 8) this lane marches
    const int slot = warp * 16 + g + 8 * myh;
    const bool writer = t < 2;                // one of the slot's two lanes does the global writes / atomics
    const int primary = lane & ~2;
    int64_t id = -1;                          // -1: slot wants a ray; -2: no rays left
    bool have = false;
    float px = 0.0f, py = 0.0f, pz = 0.0f, dt = 0.0f;
    unsigned long long evaluated = 0;

    for (;;) {
        // ---- phase A: every slot ends up with a pending sample (or permanently empty)
        for (;;) {
            const unsigned need = __ballot_sync(0xffffffffu, id == -1 && writer);
            if (need) {
                unsigned long long base = 0;
                if (lane == 0) base = atomicAdd(P.counter, (unsigned long long)__popc(need));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (id == -1) {
                    const int64_t nid = (int64_t)base + __popc(need & ((1u << primary) - 1u));
                    if (nid < P.N) {
                        id = nid;
                        const float o[3] = {__ldg(P.rays_o + nid * 3), __ldg(P.rays_o + nid * 3 + 1), __ldg(P.rays_o + nid * 3 + 2)};
                        const float d[3] = {__ldg(P.rays_d + nid * 3), __ldg(P.rays_d + nid * 3 + 1), __ldg(P.rays_d + nid * 3 + 2)};
                        float near, far;
                        ngp_near_far_ray(o, d, P.aabb, P.min_near, near, far);
                        S.ox[slot] = o[0]; S.oy[slot] = o[1]; S.oz[slot] = o[2];
                        S.dx[slot] = d[0]; S.dy[slot] = d[1]; S.dz[slot] = d[2];
                        S.t[slot] = near; S.near[slot] = near; S.far[slot] = far;
                        S.ws[slot] = 0.0f; S.depth[slot] = 0.0f; S.r[slot] = 0.0f; S.g[slot] = 0.0f; S.b[slot] = 0.0f;
                        S.nsamp[slot] = 0u;
                    } else {
                        id = -2;
                    }
                }
            }
            __syncwarp();
            // both lanes of a slot read its state, then (after the barrier) both write back identical values
            const bool go = id >= 0 && !have;
            MarchRay r;
            r.ox = S.ox[slot]; r.oy = S.oy[slot]; r.oz = S.oz[slot];
            r.dx = S.dx[slot]; r.dy = S.dy[slot]; r.dz = S.dz[slot];
            r.rdx = 1.0f / r.dx; r.rdy = 1.0f / r.dy; r.rdz = 1.0f / r.dz;
            float tc = S.t[slot];
            const float far = S.far[slot];
            bool done = go && S.nsamp[slot] >= P.max_steps;
            __syncwarp();
            if (go) {
                while (!done) {
                    if (!(tc < far)) { done = true; break; }
                    if (ngp_march_step(r, cfg, tc, px, py, pz, dt)) { have = true; break; }
                }
                S.t[slot] = tc;
                if (done) {
                    if (writer) ri_finish(S, P, slot, id);
                    id = -1;
                }
            }
            __syncwarp();
            if (!__any_sync(0xffffffffu, id == -1)) break;
        }
        if (!__any_sync(0xffffffffu, id >= 0)) break;

        // ---- phase B: the field at the warp's 16 pending samples
        float x01[2][3], dir[2][3];
        bool valid[2];
        {
            const float qx = __shfl_xor_sync(0xffffffffu, px, 1), qy = __shfl_xor_sync(0xffffffffu, py, 1);
            const float qz = __shfl_xor_sync(0xffffffffu, pz, 1);
            const bool qhave = __shfl_xor_sync(0xffffffffu, (int)have, 1) != 0;
            const float den = 2.0f * P.bound;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const bool mine = h == myh;
                valid[h] = mine ? have : qhave;
                const float sx = mine ? px : qx, sy = mine ? py : qy, sz = mine ? pz : qz;
                x01[h][0] = valid[h] ? __fdiv_rn(sx + P.bound, den) : 0.5f;   // grid.py:144
                x01[h][1] = valid[h] ? __fdiv_rn(sy + P.bound, den) : 0.5f;
                x01[h][2] = valid[h] ? __fdiv_rn(sz + P.bound, den) : 0.5f;
                const int sl = warp * 16 + g + 8 * h;
                dir[h][0] = S.dx[sl]; dir[h][1] = S.dy[sl]; dir[h][2] = S.dz[sl];
            }
        }
        float o[2][4], o3[1][4];
        uint32_t m_sig[2], m_c1[2], m_c2[2];
        rf_eval16<SAFE>(S.f, P.table, P.align_corners, lane, x01, valid, dir, o, o3, m_sig, m_c1, m_c2);

        // results live on quad lanes 0 (sigma, r, g) and 1 (b) of each row pair: hand them to the slot's lanes
        const int q0 = lane & ~3;
        const float sg0 = __shfl_sync(0xffffffffu, o[0][0], q0), sg1 = __shfl_sync(0xffffffffu, o[0][2], q0);
        const float r0 = __shfl_sync(0xffffffffu, o3[0][0], q0), r1 = __shfl_sync(0xffffffffu, o3[0][2], q0);
        const float g0 = __shfl_sync(0xffffffffu, o3[0][1], q0), g1 = __shfl_sync(0xffffffffu, o3[0][3], q0);
        const float b0 = __shfl_sync(0xffffffffu, o3[0][0], q0 + 1), b1 = __shfl_sync(0xffffffffu, o3[0][2], q0 + 1);

        // ---- phase C: composite the slot's sample (raymarching.cu:860-893)
        const float ws = S.ws[slot], tprev = S.t[slot], dprev = S.depth[slot];
        const float rprev = S.r[slot], gprev = S.g[slot], bprev = S.b[slot];
        const uint32_t nprev = S.nsamp[slot];
        __syncwarp();
        bool ended = false;
        if (have) {
            const float sigma = expf(myh ? sg1 : sg0) * P.density_scale;   // trunc_exp fwd; renderer.py:355
            const float cr = 1.0f / (1.0f + expf(-(myh ? r1 : r0)));
            const float cg = 1.0f / (1.0f + expf(-(myh ? g1 : g0)));
            const float cb = 1.0f / (1.0f + expf(-(myh ? b1 : b0)));
            const float alpha = 1.0f - __expf(-sigma * dt);
            const float T = 1.0f - ws;
            const float w = alpha * T;
            const float tn = tprev + dt;
            S.ws[slot] = ws + w; S.t[slot] = tn; S.depth[slot] = fmaf(w, tn, dprev);
            S.r[slot] = fmaf(w, cr, rprev); S.g[slot] = fmaf(w, cg, gprev); S.b[slot] = fmaf(w, cb, bprev);
            S.nsamp[slot] = nprev + 1u;
            have = false;
            if (writer) ++evaluated;
            ended = (double)T < 1e-4;   // :880 (the sample that saw T < 1e-4 is still accumulated)
        }
        __syncwarp();
        if (ended) {
            if (writer) ri_finish(S, P, slot, id);
            id = -1;
        }
        __syncwarp();
    }
    if (P.nsamples) {
#pragma unroll
        for (int o2 = 16; o2 > 0; o2 >>= 1) evaluated += __shfl_xor_sync(0xffffffffu, evaluated, o2);
        if (lane == 0 && evaluated) atomicAdd(P.nsamples, evaluated);
    }
}

__global__ void __launch_bounds__(RF_THREADS, 1)
k_render_rays_infer(RenderRaysParams P, MarchCfg cfg, const int32_t *__restrict__ offsets, float S_, uint32_t H,
                    uint32_t gridtype, const float *__restrict__ Ws0, const float *__restrict__ Ws1,
                    const float *__restrict__ Wc0, const float *__restrict__ Wc1, const float *__restrict__ Wc2) {
    extern __shared__ __align__(16) unsigned char ri_smem_raw[];
    RenderRaysSmem &S = *reinterpret_cast<RenderRaysSmem *>(ri_smem_raw);
    fq_stage_levels(S.f.lv, &S.f.all_safe, offsets, S_, H, gridtype, P.align_corners);
    rf_stage_forward_weights(S.f, Ws0, Ws1, Wc0, Wc1, Wc2);
    __syncthreads();
    if (S.f.all_safe) ri_run<true>(S, P, cfg);
    else ri_run<false>(S, P, cfg);
}

extern "C" int64_t ngp_render_rays_infer_workspace_bytes(void) { return 16; }

extern "C" int ngp_render_rays_infer(const float *rays_o, const float *rays_d, int64_t N, const float *aabb, float min_near,
                                     const uint8_t *grid, float bound, float dt_gamma, uint32_t max_steps, uint32_t C,
                                     uint32_t Hgrid, const float *embeddings, const int32_t *offsets, float S, uint32_t H,
                                     uint32_t gridtype, int align_corners, const float *Ws0, const float *Ws1,
                                     const float *Wc0, const float *Wc1, const float *Wc2, float density_scale,
                                     const float *bg_color, int bg_stride, float *image, float *depth,
                                     float *weights_sum, void *workspace, ngp_stream_t stream) {
    NGP_REQUIRE(rays_o && rays_d && aabb && grid && embeddings && offsets && Ws0 && Ws1 && Wc0 && Wc1 && Wc2 && image &&
                    depth && workspace,
                NGP_ERR_ARG, "ngp_render_rays_infer: null pointer");
    NGP_REQUIRE(bound > 0.0f && N >= 0, NGP_ERR_ARG, "ngp_render_rays_infer: bad bound / N");
    NGP_REQUIRE(C >= 1 && C <= 8 && Hgrid >= 1 && Hgrid <= 1024, NGP_ERR_ARG, "ngp_render_rays_infer: bad cascade/grid size");
    NGP_REQUIRE(bg_stride == 0 || bg_stride == 3, NGP_ERR_ARG, "ngp_render_rays_infer: bg_stride must be 0 or 3");
    NGP_REQUIRE(((uintptr_t)embeddings & 15u) == 0 && ((uintptr_t)workspace & 7u) == 0, NGP_ERR_ARG,
                "ngp_render_rays_infer: embeddings must be 16-byte aligned, workspace 8-byte aligned");
    if (N == 0) return NGP_OK;
    const int smem = (int)sizeof(RenderRaysSmem);
    if (cudaFuncSetAttribute(k_render_rays_infer, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
        ngp_set_error("ngp_render_rays_infer: cannot reserve %d B of shared memory", smem);
        return NGP_ERR_CUDA;
    }
    int dev = 0, sms = NGP_SM_COUNT;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaMemsetAsync(workspace, 0, 16, (cudaStream_t)stream) != cudaSuccess) {
        ngp_set_error("ngp_render_rays_infer: workspace memset failed");
        return NGP_ERR_CUDA;
    }
    RenderRaysParams P;
    P.rays_o = rays_o; P.rays_d = rays_d; P.aabb = aabb; P.bg = bg_color; P.N = N; P.bg_stride = bg_stride;
    P.min_near = min_near; P.density_scale = density_scale; P.bound = bound; P.max_steps = max_steps;
    P.table = embeddings; P.align_corners = align_corners;
    P.image = image; P.depth = depth; P.weights_sum = weights_sum;
    P.counter = (unsigned long long *)workspace;
    P.nsamples = (unsigned long long *)workspace + 1;
    const MarchCfg cfg = ngp_march_cfg(grid, bound, dt_gamma, max_steps, C, Hgrid);
    const int64_t tiles = ngp_div_up(N, RI_SLOTS);
    const int nblocks = (int)(tiles < sms ? tiles : sms);
    k_render_rays_infer<<<nblocks, RF_THREADS, smem, (cudaStream_t)stream>>>(P, cfg, offsets, S, H, gridtype, Ws0, Ws1, Wc0,
                                                                             Wc1, Wc2);
    NGP_CHECK_LAUNCH("ngp_render_rays_infer");
    return NGP_OK;
}
