// render.cu — fused inference renderer: AABB near/far -> occupancy-grid march -> hash-grid field + colour MLP ->
// front-to-back compositing with early termination -> background / depth normalisation, in ONE persistent kernel.
// Replaces the host loop of NeRFRenderer.run_cuda's eval branch (nerf/renderer.py:325-374):
//     while step < max_steps: march_rays (raymarching.cu:706-814) -> self(xyzs, dirs) (network.py:95-124)
//                             -> composite_rays (:828-913) -> rays_alive = rays_alive[rays_alive >= 0]   (host sync)
// Per-sample positions, sigma and rgb never reach HBM: HBM sees 24 B in and 20 B out per RAY.
//
// Organisation (same warp-cooperative field core as renderfield.cu): a warp owns 32 ray slots, one per lane.  Each lane
// marches its own ray a bounded number of voxel steps at a time and queues the samples it finds (t, dt) in a small
// per-slot FIFO in shared memory, so a ray crossing a long empty stretch never stalls the others.  The heads of up to 16
// non-empty FIFOs form one mma.sync row tile, the field is evaluated, and each of those lanes composites its sample in
// ray order.  A slot whose ray ended (t >= far with an empty FIFO, T < 1e-4, sample cap) writes its pixel and takes the
// next ray from a global counter (one warp-aggregated atomic per refill).  The row tile stays full while rays of very
// different lengths come and go — the role of the reference's alive-list compaction, without its per-iteration host
// round trip.
//
// Differences from the reference loop, all inside its fp tolerance: t is the marcher's own abscissa (the reference
// re-sums deltas in composite_rays, :876, a <= 1 ulp wobble that depends on how its n_step chunks fall); a ray takes at
// most max_steps samples (reference: max_steps rounded up to its last chunk of <= 8); perturb is not supported here.
#include <float.h>

#include "march.cuh"
#include "renderfield.cuh"

#define RI_SLOTS (RF_WARPS * 32)
#ifndef RI_FIFO
#define RI_FIFO 8          // queued samples per slot
#endif
#ifndef RI_BUDGET
#define RI_BUDGET 16       // voxel steps a lane may take per marching phase
#endif
#ifndef RI_MARCH_BELOW
#define RI_MARCH_BELOW 20  // march when fewer slots than this have a queued sample
#endif

struct __align__(16) RenderRaysSmem {
    RenderFieldSmem f;
    float ox[RI_SLOTS], oy[RI_SLOTS], oz[RI_SLOTS], dx[RI_SLOTS], dy[RI_SLOTS], dz[RI_SLOTS];
    float t[RI_SLOTS], near[RI_SLOTS], far[RI_SLOTS];
    float ws[RI_SLOTS], depth[RI_SLOTS], r[RI_SLOTS], g[RI_SLOTS], b[RI_SLOTS];
    float qt[RI_FIFO][RI_SLOTS], qdt[RI_FIFO][RI_SLOTS];  // per-slot FIFO of pending samples
    float tin[RF_WARPS][6][16];                            // the warp's row tile: sample position, view direction
    float tout[RF_WARPS][4][16];                           // log sigma, colour logits of the tile rows
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
    int32_t *counts;               // optional: samples composited per ray
    unsigned long long *counter;
    unsigned long long *nsamples;  // optional statistics: total samples evaluated
};

// write the finished ray of `slot` (renderer.py:371-374: image + (1 - ws) * bg; clamp(depth - near, 0) / (far - near))
__device__ __forceinline__ void ri_finish(const RenderRaysSmem &S, const RenderRaysParams &P, int slot, int64_t id,
                                          uint32_t ncomp) {
    if (P.counts) P.counts[id] = (int32_t)ncomp;
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
    const int slot = tid;                     // lane-private ray slot
    int64_t id = -1;                          // -1: slot wants a ray; -2: no rays left
    int cnt = 0, head = 0;                    // FIFO fill / read position
    uint32_t marched = 0;                     // samples queued so far for this ray (<= max_steps)
    uint32_t ncomp = 0;                       // samples composited so far for this ray
    bool mend = false;                        // the march of this ray is over (t >= far or sample cap)
    unsigned rot = 0;
    unsigned long long evaluated = 0;
    float (*tin)[16] = S.tin[warp];
    float (*tout)[16] = S.tout[warp];

#ifndef RI_LOCKSTEP
#define RI_LOCKSTEP 0   // 1: one CTA barrier per round so the 12 warps walk the unrolled MLP code together (I-cache
                        // experiment; measured slower, 179 vs 118 ms: warps in different phases overlap gathers and MMAs)
#endif
    bool warp_done = false;
    for (;;) {
        unsigned avail = 0u;
        if (!warp_done) {
            // ---- A1: refill empty slots (one atomic per warp)
            const unsigned need = __ballot_sync(0xffffffffu, id == -1);
            if (need) {
                unsigned long long base = 0;
                if (lane == 0) base = atomicAdd(P.counter, (unsigned long long)__popc(need));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (id == -1) {
                    const int64_t nid = (int64_t)base + __popc(need & ((1u << lane) - 1u));
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
                        cnt = 0; head = 0; marched = 0; ncomp = 0; mend = false;
                    } else {
                        id = -2;
                    }
                }
            }
            // ---- A2: bounded marching, only when the queues run low
            if (__popc(__ballot_sync(0xffffffffu, cnt > 0)) < RI_MARCH_BELOW) {
                if (id >= 0 && !mend && cnt < RI_FIFO) {
                    MarchRay r;
                    r.ox = S.ox[slot]; r.oy = S.oy[slot]; r.oz = S.oz[slot];
                    r.dx = S.dx[slot]; r.dy = S.dy[slot]; r.dz = S.dz[slot];
                    r.rdx = 1.0f / r.dx; r.rdy = 1.0f / r.dy; r.rdz = 1.0f / r.dz;
                    float tc = S.t[slot];
                    const float far = S.far[slot];
                    float px, py, pz, dt;
                    for (int budget = RI_BUDGET; budget > 0 && cnt < RI_FIFO; --budget) {
                        if (!(tc < far) || marched >= P.max_steps) { mend = true; break; }
                        if (ngp_march_step(r, cfg, tc, px, py, pz, dt)) {
                            const int w = (head + cnt) % RI_FIFO;
                            S.qt[w][slot] = tc;
                            S.qdt[w][slot] = dt;
                            tc += dt;
                            ++cnt;
                            ++marched;
                        }
                    }
                    S.t[slot] = tc;
                }
            }
            // ---- A3: rays whose march is over and whose queue is drained are done
            if (id >= 0 && mend && cnt == 0) {
                ri_finish(S, P, slot, id, ncomp);
                id = -1;
            }

            avail = __ballot_sync(0xffffffffu, cnt > 0);
            if (avail == 0 && !__any_sync(0xffffffffu, id != -2)) warp_done = true;   // every slot has run out of rays
        }
#if RI_LOCKSTEP
        if (!__syncthreads_or(!warp_done)) break;
#else
        if (warp_done) break;
#endif
        if (avail == 0) continue;
        // ---- B: tile = heads of up to 16 non-empty queues (rotating priority), row = rank among them
        const unsigned ravail = __funnelshift_r(avail, avail, rot);
        const int pos = (lane - (int)rot) & 31;
        const int row = __popc(ravail & ((1u << pos) - 1u));
        const bool sel = cnt > 0 && row < 16;
        const int nrows = min(__popc(avail), 16);
        rot = (rot + 13u) & 31u;
        float st = 0.0f, sdt = 0.0f;
        if (sel) {
            st = S.qt[head][slot];
            sdt = S.qdt[head][slot];
            const float dxv = S.dx[slot], dyv = S.dy[slot], dzv = S.dz[slot];
            tin[0][row] = ngp_clampf(fmaf(st, dxv, S.ox[slot]), -P.bound, P.bound);   // = the marcher's sample position
            tin[1][row] = ngp_clampf(fmaf(st, dyv, S.oy[slot]), -P.bound, P.bound);
            tin[2][row] = ngp_clampf(fmaf(st, dzv, S.oz[slot]), -P.bound, P.bound);
            tin[3][row] = dxv; tin[4][row] = dyv; tin[5][row] = dzv;
        }
        __syncwarp();
        float x01[2][3], dir[2][3];
        bool valid[2];
        {
            const float den = 2.0f * P.bound;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int rr = g + 8 * h;
                valid[h] = rr < nrows;
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    x01[h][a] = valid[h] ? __fdiv_rn(tin[a][rr] + P.bound, den) : 0.5f;   // grid.py:144
                    dir[h][a] = valid[h] ? tin[3 + a][rr] : 0.0f;
                }
            }
        }
        float o[2][4], o3[1][4];
        uint32_t m_sig[2], m_c1[2], m_c2[2];
        rf_eval16<SAFE>(S.f, P.table, P.align_corners, lane, x01, valid, dir, o, o3, m_sig, m_c1, m_c2);
        if (t == 0) {
            tout[0][g] = o[0][0]; tout[0][g + 8] = o[0][2];
            tout[1][g] = o3[0][0]; tout[1][g + 8] = o3[0][2];
            tout[2][g] = o3[0][1]; tout[2][g + 8] = o3[0][3];
        } else if (t == 1) {
            tout[3][g] = o3[0][0]; tout[3][g + 8] = o3[0][2];
        }
        __syncwarp();
        // ---- C: composite the slot's sample (raymarching.cu:860-893)
        if (sel) {
            const float sigma = expf(tout[0][row]) * P.density_scale;   // trunc_exp fwd; renderer.py:355
            const float cr = 1.0f / (1.0f + expf(-tout[1][row]));
            const float cg = 1.0f / (1.0f + expf(-tout[2][row]));
            const float cb = 1.0f / (1.0f + expf(-tout[3][row]));
            const float ws = S.ws[slot];
            const float alpha = 1.0f - __expf(-sigma * sdt);
            const float T = 1.0f - ws;
            const float w = alpha * T;
            const float tn = st + sdt;
            S.ws[slot] = ws + w;
            S.depth[slot] = fmaf(w, tn, S.depth[slot]);
            S.r[slot] = fmaf(w, cr, S.r[slot]); S.g[slot] = fmaf(w, cg, S.g[slot]); S.b[slot] = fmaf(w, cb, S.b[slot]);
            head = (head + 1) % RI_FIFO;
            --cnt;
            ++evaluated;
            ++ncomp;
            if ((double)T < 1e-4) {   // :880 (the sample that saw T < 1e-4 is still accumulated); queued samples are dropped
                ri_finish(S, P, slot, id, ncomp);
                id = -1;
                cnt = 0;
            }
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

extern "C" int64_t ngp_render_rays_infer_workspace_bytes(void) { return 16 + 24 * 8; }   // counter, samples, 24 dev statistics

// default implementation: the warp-specialised tcgen05 kernel of render5.cu; NGP_B200_FIELD_IMPL=mma / ngp_set_field_impl(0)
// selects the mma.sync kernel above (kept as the independent cross-check)
int ngp_field_impl_tc5();
int r5_render_rays_infer(const float *rays_o, const float *rays_d, int64_t N, const float *aabb, float min_near,
                         const uint8_t *grid, float bound, float dt_gamma, uint32_t max_steps, uint32_t C, uint32_t Hgrid,
                         const float *embeddings, const int32_t *offsets, float S, uint32_t H, uint32_t gridtype,
                         int align_corners, const float *Ws0, const float *Ws1, const float *Wc0, const float *Wc1,
                         const float *Wc2, float density_scale, const float *bg_color, int bg_stride, float *image,
                         float *depth, float *weights_sum, int32_t *counts, void *workspace, cudaStream_t stream);

extern "C" int ngp_render_rays_infer(const float *rays_o, const float *rays_d, int64_t N, const float *aabb, float min_near,
                                     const uint8_t *grid, float bound, float dt_gamma, uint32_t max_steps, uint32_t C,
                                     uint32_t Hgrid, const float *embeddings, const int32_t *offsets, float S, uint32_t H,
                                     uint32_t gridtype, int align_corners, const float *Ws0, const float *Ws1,
                                     const float *Wc0, const float *Wc1, const float *Wc2, float density_scale,
                                     const float *bg_color, int bg_stride, float *image, float *depth,
                                     float *weights_sum, int32_t *counts, void *workspace, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(rays_o, "ngp_render_rays_infer");
    if (N == 0) return NGP_OK;
    NGP_REQUIRE(rays_o && rays_d && aabb && grid && embeddings && offsets && Ws0 && Ws1 && Wc0 && Wc1 && Wc2 && image &&
                    depth && workspace,
                NGP_ERR_ARG, "ngp_render_rays_infer: null pointer");
    NGP_REQUIRE(bound > 0.0f && N >= 0, NGP_ERR_ARG, "ngp_render_rays_infer: bad bound / N");
    NGP_REQUIRE(C >= 1 && C <= 8 && Hgrid >= 1 && Hgrid <= 1024, NGP_ERR_ARG, "ngp_render_rays_infer: bad cascade/grid size");
    NGP_REQUIRE(bg_stride == 0 || bg_stride == 3, NGP_ERR_ARG, "ngp_render_rays_infer: bg_stride must be 0 or 3");
    NGP_REQUIRE(((uintptr_t)embeddings & 15u) == 0 && ((uintptr_t)workspace & 7u) == 0, NGP_ERR_ARG,
                "ngp_render_rays_infer: embeddings must be 16-byte aligned, workspace 8-byte aligned");
    if (cudaMemsetAsync(workspace, 0, 16 + 24 * 8, (cudaStream_t)stream) != cudaSuccess) {
        ngp_set_error("ngp_render_rays_infer: workspace memset failed");
        return NGP_ERR_CUDA;
    }
    if (ngp_field_impl_tc5())
        return r5_render_rays_infer(rays_o, rays_d, N, aabb, min_near, grid, bound, dt_gamma, max_steps, C, Hgrid, embeddings,
                                    offsets, S, H, gridtype, align_corners, Ws0, Ws1, Wc0, Wc1, Wc2, density_scale, bg_color,
                                    bg_stride, image, depth, weights_sum, counts, workspace, (cudaStream_t)stream);
    const int smem = (int)sizeof(RenderRaysSmem);
    if (cudaFuncSetAttribute(k_render_rays_infer, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
        ngp_set_error("ngp_render_rays_infer: cannot reserve %d B of shared memory", smem);
        return NGP_ERR_CUDA;
    }
    int dev = 0, sms = NGP_SM_COUNT;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    RenderRaysParams P;
    P.rays_o = rays_o; P.rays_d = rays_d; P.aabb = aabb; P.bg = bg_color; P.N = N; P.bg_stride = bg_stride;
    P.min_near = min_near; P.density_scale = density_scale; P.bound = bound; P.max_steps = max_steps;
    P.table = embeddings; P.align_corners = align_corners;
    P.image = image; P.depth = depth; P.weights_sum = weights_sum; P.counts = counts;
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
