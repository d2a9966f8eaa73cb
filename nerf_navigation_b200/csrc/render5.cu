// render5.cu — fused inference renderer on the Blackwell tensor-core path: the whole eval branch of
// NeRFRenderer.run_cuda (nerf/renderer.py:325-374: near/far -> loop { march_rays (raymarching.cu:706-814) ->
// self(xyzs, dirs) (network.py:95-124) -> composite_rays (:828-913) -> compaction } -> bg / depth post) as ONE persistent,
// warp-specialised kernel.  Same tcgen05 field core as cfield5.cu (128-sample tiles, A operands and accumulators in TMEM,
// 3xTF32), fed by marcher warps instead of a sample array:
//
//   marcher warps (8)  own R5_NR = 768 ray slots in shared memory (3 per lane): they fetch rays (one warp-aggregated atomic
//                      per refill), compute near/far, march each ray into a small per-slot FIFO of (t, dt) samples (up to
//                      R5_BUDGET voxel steps per slot per pass, the slots' occupancy loads in flight together), build
//                      their 16 rows of the next tile from the FIFO heads of rays that have fewer than R5_INFLIGHT
//                      samples in flight (ballot compaction), and write finished pixels;
//   gather warps (8)   thread = tile row: hash-grid encode of the row's sample -> features straight into TMEM;
//   MMA threads (2)    the five layers of a tile as tcgen05.mma kind::tf32 (one thread per TMEM slot);
//   chain warps (4)    the epilogues (ReLU / SH / split to hi, lo) and, after the last layer, the compositing step of the
//                      row's ray (raymarching.cu:860-893) on the ray's accumulators in shared memory.
//
// A ray has at most R5_INFLIGHT = 2 samples in flight, always in consecutive tiles, and the chain threads composite tile
// after tile: compositing is in ray order, and a sample that arrives after its ray fell under T < 1e-4 (status bit R5_DEAD)
// is dropped uncounted, so early termination is exact (at most one speculative sample per ray is evaluated and thrown
// away).  Tiles stay full because there are more ray slots (768) than rows in flight (R5_TBUF = 8 row buffers x 128).
// Roles hand tiles over through mbarriers; the end of the work is a pair of zero-row tiles.  Per-sample positions, sigma
// and rgb never reach HBM: 24 B in and 20 B (+4 B count) out per RAY.  Differences from the reference loop are those of
// render.cu (documented there): t is the marcher's own abscissa, at most max_steps samples per ray, no perturb.
// Measured decisions (R5_STATS cycle counters, ncu source page): the marcher is instruction-fetch bound, not latency
// bound — rolled loops, out-of-line cold paths and the bit-exact fast cell/exit arithmetic of march.cuh bought more than
// larger budgets or deeper FIFOs, which made it slower.
#include <float.h>

#include <cub/cub.cuh>

#include "cfield5.cuh"
#include "march.cuh"

#define R5_MARCH_WARP0 (C5_MMA_WARP + 2)                       // warps 14 .. 21
#ifndef R5_MARCH_WARPS
#define R5_MARCH_WARPS 8
#endif
#define R5_THREADS ((R5_MARCH_WARP0 + R5_MARCH_WARPS) * 32)    // 704 (80 registers per thread)
#ifndef R5_PER_LANE
#define R5_PER_LANE (24 / R5_MARCH_WARPS)                      // ray slots per marcher lane (768 slots per CTA)
#endif
#define R5_PER_WARP (32 * R5_PER_LANE)
#define R5_NR (R5_MARCH_WARPS * R5_PER_WARP)                   // 768 ray slots per CTA
#define R5_PART (C5_ROWS / R5_MARCH_WARPS)                     // tile rows built by one marcher warp (32)
#define R5_FIFO 4                                              // queued samples per ray slot
#ifndef R5_TBUF_LOG2
#define R5_TBUF_LOG2 3
#endif
#define R5_TBUF (1 << R5_TBUF_LOG2)                             // tile row buffers (tile n: buffer n % R5_TBUF, TMEM slot n & 1)
#ifndef R5_INFLIGHT
#define R5_INFLIGHT 2                                          // samples of one ray that may be in flight (in consecutive tiles)
#endif
#ifndef R5_BUDGET
#define R5_BUDGET 2                                            // marcher steps a ray may take per pass
#endif
#define R5_INVALID 0xffffffffu
// ray-slot status (written by the marcher when idle, by the chain thread while the sample is in flight)
// ray-slot status word: bits 0-1 = samples in flight (the marcher adds one per issued sample, the chain thread takes it off
// after compositing: shared-memory atomics), bit 2 = the ray saw T < 1e-4 and is over (later samples are not composited)
#define R5_IDLE 0u
#define R5_DEAD 4u
#define R5_NFLY(st) ((st) & 3u)

// marcher-private state word: samples queued so far (20 bits) | FIFO fill (3) | FIFO head (2) | march ended (1)
#define R5_MS_MARCHED(m) ((m) & 0xfffffu)
#define R5_MS_CNT(m) (((m) >> 20) & 7u)
#define R5_MS_HEAD(m) (((m) >> 23) & 3u)
#define R5_MS_END(m) (((m) >> 25) & 1u)
#define R5_MS_PACK(marched, cnt, head, end) ((marched) | ((cnt) << 20) | ((head) << 23) | ((end) << 25))

// -DR5_STATS=1: per-role cycle accounting (dev aid, scripts/check_r5.py prints it): every wait is timed with clock64 by
// lane 0 of the warp and summed into workspace[2 + index]
#ifndef R5_STATS
#define R5_STATS 0
#endif
#if R5_STATS
#define R5_STAT_DECL long long stat_[24] = {0}
#define R5_WAIT(bar, par, idx)                       \
    do {                                             \
        const long long t0_ = clock64();             \
        tc5::mbar_wait_wd((bar), (par));             \
        stat_[idx] += clock64() - t0_;               \
    } while (0)
#define R5_STAT_ADD(idx, v) stat_[idx] += (long long)(v)
#define R5_STAT_FLUSH(P, lane)                                                                       \
    do {                                                                                             \
        if ((lane) == 0)                                                                             \
            for (int q_ = 0; q_ < 24; ++q_)                                                          \
                if (stat_[q_]) atomicAdd((P).counter + 2 + q_, (unsigned long long)stat_[q_]);       \
    } while (0)
#else
#define R5_STAT_DECL
#define R5_WAIT(bar, par, idx) tc5::mbar_wait_wd((bar), (par))
#define R5_STAT_ADD(idx, v)
#define R5_STAT_FLUSH(P, lane)
#endif

struct R5Tile {
    float x[C5_ROWS], y[C5_ROWS], z[C5_ROWS];        // sample position (the marcher's clamped o + t d)
    float dx[C5_ROWS], dy[C5_ROWS], dz[C5_ROWS];     // view direction
    float tn[C5_ROWS], dt[C5_ROWS];                  // t + dt and dt of the sample
    float lsig[C5_ROWS];                             // log sigma (chain step 2 -> step 5)
    uint32_t slot[C5_ROWS];                          // ray slot of the row, R5_INVALID for an unused row
    int nrows[R5_MARCH_WARPS];
};

struct R5Rays {
    float ox[R5_NR], oy[R5_NR], oz[R5_NR], dx[R5_NR], dy[R5_NR], dz[R5_NR];
    float rdx[R5_NR], rdy[R5_NR], rdz[R5_NR];        // 1 / d (raymarching.cu:338)
    float near[R5_NR], far[R5_NR], t[R5_NR];
    int id[R5_NR];                                   // ray index; -1: slot wants a ray; -2: no rays left
    uint32_t mstate[R5_NR];
    float qt[R5_FIFO][R5_NR], qdt[R5_FIFO][R5_NR];
    float ws[R5_NR], depth[R5_NR], r[R5_NR], g[R5_NR], b[R5_NR];
    uint32_t ncomp[R5_NR];                           // samples composited for the ray
    uint32_t status[R5_NR];
};

template <int TB>
struct R5BarsT {
    uint64_t rows_full[TB];   // 128 marcher threads: the tile's rows are in the buffer
    uint64_t rows_free[TB];   // 128 chain threads: the tile has been composited, the buffer may be rewritten
};

struct __align__(128) R5Smem {
    static constexpr int TBUF = R5_TBUF, TBUF_LOG2 = R5_TBUF_LOG2;
    C5FwdWeights w;
    FqLevel lv[FQ_L];
    C5Bars bars[2];
    R5BarsT<R5_TBUF> tb;
    uint32_t tmem;
    int all_safe;
    float mipb[8], miprb[8];       // min(2^level, bound) and its reciprocal (raymarching.cu:376-377)
    uint16_t mlist[R5_MARCH_WARPS][R5_PER_WARP];   // marcher warps: slots that want a step this pass
    int chain_two[2];              // chain warps: does the current group take two tiles (double-buffered by group parity)
    R5Tile tile[R5_TBUF];
    R5Rays rays;
};

struct R5Params {
    const float *rays_o, *rays_d, *aabb, *bg;
    int64_t N;
    int bg_stride;
    float min_near, density_scale, bound;
    uint32_t max_steps;
    const float *table;
    int align_corners;
    float *image, *depth, *weights_sum;
    int32_t *counts;
    unsigned long long *counter;
    unsigned long long *nsamples;
    // pre-marched sample lists (k_render5_rays_infer_lists): samples of ray i are tdt[ends[i] - counts[i] .. ends[i])
    const int32_t *plan_counts;
    const long long *plan_ends;
    const float2 *tdt;
    // pre-marched SEGMENTS (k_render5_rays_infer_lists<true>): ray i has nsegs[i] runs of consecutive samples, run k =
    // (first t, number of samples as int bits) at segs[i * seg_cap + k]; nsegs[i] < 0: more than seg_cap runs, the ray is
    // rendered by the marcher-inside kernel from ray_list instead
    const float2 *segs;
    const int32_t *nsegs;
    int seg_cap;
    float dt_min, dt_max, dt_gamma;
    int lvl_dt;
    // marcher-inside kernel on a subset: ray indices ray_list[0 .. *ray_list_count) instead of 0 .. N
    const int32_t *ray_list;
    const unsigned long long *ray_list_count;
};

__device__ __forceinline__ uint32_t r5_ld_status(const uint32_t *p) { return *(const volatile uint32_t *)p; }
__device__ __forceinline__ void r5_st_status(uint32_t *p, uint32_t v) { *(volatile uint32_t *)p = v; }

__device__ __forceinline__ bool r5_tile_end(const R5Tile &T) {
    int n = 0;
#pragma unroll
    for (int w = 0; w < R5_MARCH_WARPS; ++w) n += T.nrows[w];
    return n == 0;
}

// ------------------------------------------------------------------------------------------------ marcher warps
// renderer.py:371-374: image + (1 - ws) * bg; clamp(depth - near, 0) / (far - near)
template <class RAYS>
static __device__ __noinline__ void r5_finish(const RAYS &R, const R5Params &P, int slot, int64_t id) {
    const float ws = R.ws[slot], om = 1.0f - ws;
    const float *bg = P.bg ? P.bg + (int64_t)P.bg_stride * id : nullptr;
    const float b0 = bg ? __ldg(bg) : 1.0f, b1 = bg ? __ldg(bg + 1) : 1.0f, b2 = bg ? __ldg(bg + 2) : 1.0f;
    P.image[id * 3] = __fadd_rn(R.r[slot], __fmul_rn(om, b0));
    P.image[id * 3 + 1] = __fadd_rn(R.g[slot], __fmul_rn(om, b1));
    P.image[id * 3 + 2] = __fadd_rn(R.b[slot], __fmul_rn(om, b2));
    const float near = R.near[slot], far = R.far[slot];
    P.depth[id] = __fdiv_rn(fmaxf(R.depth[slot] - near, 0.0f), far - near);
    if (P.weights_sum) P.weights_sum[id] = ws;
    if (P.counts) P.counts[id] = (int32_t)R.ncomp[slot];
}

// initialise slot `slot` with ray `nid`
static __device__ __noinline__ void r5_load_ray(R5Rays &R, const R5Params &P, int slot, int64_t nid) {
    const float o[3] = {__ldg(P.rays_o + nid * 3), __ldg(P.rays_o + nid * 3 + 1), __ldg(P.rays_o + nid * 3 + 2)};
    const float d[3] = {__ldg(P.rays_d + nid * 3), __ldg(P.rays_d + nid * 3 + 1), __ldg(P.rays_d + nid * 3 + 2)};
    float near, far;
    ngp_near_far_ray(o, d, P.aabb, P.min_near, near, far);
    R.ox[slot] = o[0]; R.oy[slot] = o[1]; R.oz[slot] = o[2];
    R.dx[slot] = d[0]; R.dy[slot] = d[1]; R.dz[slot] = d[2];
    R.rdx[slot] = 1.0f / d[0]; R.rdy[slot] = 1.0f / d[1]; R.rdz[slot] = 1.0f / d[2];
    R.t[slot] = near; R.near[slot] = near; R.far[slot] = far;
    R.ws[slot] = 0.0f; R.depth[slot] = 0.0f; R.r[slot] = 0.0f; R.g[slot] = 0.0f; R.b[slot] = 0.0f;
    R.ncomp[slot] = 0u;
    R.mstate[slot] = 0u;
    R.id[slot] = (int)nid;
}

// One pass of a marcher warp over its R5_PER_WARP slots:
//   1. classify every slot from one read of (id, status, mstate): eligible for a tile / finished / wants marching;
//   2. retire finished rays and refill their slots (one atomic per warp);
//   3. if R5_PART slots are eligible (or nothing more can come) and the tile buffer is free: build the warp's part of tile n;
//   4. march: the slots that want a step are compacted into a list (ballots) and the lanes walk the list, so every lane
//      probes a cell no matter how the work is spread over the slots; R5_ILP list entries per lane are in flight at once
//      (a step is a dependent chain: position -> cell -> occupancy bit -> next t).
// All marcher state lives in shared memory, so any lane of the warp can step any of the warp's slots; the phases are
// separated by __syncwarp().
__device__ __forceinline__ void r5_marcher(R5Smem &S, const R5Params &P, const MarchCfg &cfg, int w, int lane) {
    R5Rays &R = S.rays;
    const int slot0 = w * R5_PER_WARP + lane;
    const unsigned lt_mask = (1u << lane) - 1u;
    uint16_t *list = S.mlist[w];
#pragma unroll 1
    for (int j = 0; j < R5_PER_LANE; ++j) {
        const int slot = slot0 + 32 * j;
        R.id[slot] = -1;
        R.mstate[slot] = 0u;
        r5_st_status(&R.status[slot], R5_IDLE);
    }
    bool rays_left = true;
    unsigned long long evaluated = 0;
    int64_t n = 0;                                  // next tile to build
    int rot = 0;
    unsigned needs = (1u << R5_PER_LANE) - 1u;      // bit j: slot j of this lane wants a ray
    bool ended = false;
    R5_STAT_DECL;
    R5_STAT_ADD(17, -clock64());
    for (;;) {
        R5_STAT_ADD(13, 1);
        // ---- 2a. refill the slots that want a ray: one atomic per warp
        if (!rays_left) needs = 0u;
        if (__any_sync(0xffffffffu, needs != 0u)) {
            const int mine = __popc(needs);
            int pre = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, pre, o);
                if (lane >= o) pre += v;
            }
            const int total = __shfl_sync(0xffffffffu, pre, 31);
            unsigned long long base = 0;
            if (lane == 31) base = atomicAdd(P.counter, (unsigned long long)total);
            base = __shfl_sync(0xffffffffu, base, 31);
            int64_t nid = (int64_t)base + pre - mine;
            const int64_t n_rays = P.ray_list ? (int64_t)*P.ray_list_count : P.N;
            while (needs) {
                const int j = __ffs(needs) - 1;
                needs &= needs - 1u;
                const int slot = slot0 + 32 * j;
                if (nid < n_rays) {
                    r5_load_ray(R, P, slot, P.ray_list ? (int64_t)__ldg(P.ray_list + nid) : nid);
                    ++nid;
                } else {
                    R.id[slot] = -2;
                    rays_left = false;
                }
            }
        }
        // ---- 1. classify
        unsigned elig = 0u;       // bit j: idle ray with a queued sample
        unsigned retire = 0u;     // bit j: the ray of slot j is over (dead, or marched to the end and drained)
        unsigned want = 0u;       // bit j: the ray wants a marcher step (queue has room, march not over)
        bool pending = false;     // some slot may still become eligible without a tile being built
        {
            int id[R5_PER_LANE];
            uint32_t st[R5_PER_LANE], ms[R5_PER_LANE];
#pragma unroll
            for (int j = 0; j < R5_PER_LANE; ++j) {
                const int slot = slot0 + 32 * j;
                id[j] = R.id[slot];
                st[j] = r5_ld_status(&R.status[slot]);
                ms[j] = R.mstate[slot];
            }
#pragma unroll
            for (int j = 0; j < R5_PER_LANE; ++j) {
                if (id[j] < 0) continue;
                const uint32_t cnt = R5_MS_CNT(ms[j]);
                const bool end = R5_MS_END(ms[j]) != 0u;
                const bool dead = (st[j] & R5_DEAD) != 0u;
                const uint32_t fly = R5_NFLY(st[j]);
                if (!dead && !end && cnt < R5_FIFO) want |= 1u << j;
                if (!dead && fly < R5_INFLIGHT && cnt > 0u) elig |= 1u << j;
                else if (fly == 0u && (dead || (end && cnt == 0u))) retire |= 1u << j;
                else pending = true;   // in flight (comes back) or still marching
            }
        }
        // ---- 2b. retire finished rays (their slots are refilled at the top of the next pass)
        if (retire) pending = true;
        while (retire) {
            const int j = __ffs(retire) - 1;
            retire &= retire - 1u;
            const int slot = slot0 + 32 * j;
            __threadfence_block();         // the chain thread wrote the accumulators before the status
            r5_finish(R, P, slot, R.id[slot]);
            evaluated += R.ncomp[slot];
            if (rays_left) {
                R.id[slot] = -1;
                needs |= 1u << j;
            } else {
                R.id[slot] = -2;
            }
            r5_st_status(&R.status[slot], R5_IDLE);
        }
        // ---- 3. build this warp's part of tile n: R5_PART eligible slots, or all of them if nothing else can come
        int E = __popc(elig);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) E += __shfl_xor_sync(0xffffffffu, E, o);
        const bool any_pending = __any_sync(0xffffffffu, pending);
        const bool any_want = __any_sync(0xffffffffu, want != 0u);
        bool build = E >= R5_PART || !any_pending;
        if (!build) R5_STAT_ADD(15, 1);
        R5_STAT_ADD(20, E);
        const int b = (int)(n & (R5_TBUF - 1));
        if (build && n >= R5_TBUF) {
            uint32_t ok = 0;
            if (lane == 0) ok = tc5::mbar_try(tc5::smem_u32(&S.tb.rows_free[b]), (uint32_t)(((n >> R5_TBUF_LOG2) - 1) & 1));
            build = __shfl_sync(0xffffffffu, ok, 0) != 0u;     // buffer still in use: keep marching
            if (!build) R5_STAT_ADD(16, 1);
        }
        if (build) {
            R5Tile &T = S.tile[b];
            int nrows = 0;
#pragma unroll 1
            for (int jj = 0; jj < R5_PER_LANE && nrows < R5_PART; ++jj) {
                int j = jj + rot;
                if (j >= R5_PER_LANE) j -= R5_PER_LANE;
                const bool e = (elig >> j) & 1u;
                const unsigned m = __ballot_sync(0xffffffffu, e);
                const int row = nrows + __popc(m & lt_mask);
                if (e && row < R5_PART) {
                    const int slot = slot0 + 32 * j;
                    const uint32_t ms = R.mstate[slot];
                    const uint32_t head = R5_MS_HEAD(ms);
                    const float st = R.qt[head][slot], sdt = R.qdt[head][slot];
                    R.mstate[slot] = R5_MS_PACK(R5_MS_MARCHED(ms), R5_MS_CNT(ms) - 1u, (head + 1u) % R5_FIFO, R5_MS_END(ms));
                    if (!R5_MS_END(ms)) want |= 1u << j;      // its queue has room again
                    const float dxv = R.dx[slot], dyv = R.dy[slot], dzv = R.dz[slot];
                    const int tr = w * R5_PART + row;
                    T.x[tr] = ngp_clampf(fmaf(st, dxv, R.ox[slot]), -P.bound, P.bound);   // = the marcher's sample position
                    T.y[tr] = ngp_clampf(fmaf(st, dyv, R.oy[slot]), -P.bound, P.bound);
                    T.z[tr] = ngp_clampf(fmaf(st, dzv, R.oz[slot]), -P.bound, P.bound);
                    T.dx[tr] = dxv; T.dy[tr] = dyv; T.dz[tr] = dzv;
                    T.tn[tr] = st + sdt;
                    T.dt[tr] = sdt;
                    T.slot[tr] = (uint32_t)slot;
                    atomicAdd(&R.status[slot], 1u);
                }
                nrows = min(R5_PART, nrows + __popc(m));
            }
            rot = rot + 1 >= R5_PER_LANE ? 0 : rot + 1;
            for (int row = nrows + lane; row < R5_PART; row += 32) T.slot[w * R5_PART + row] = R5_INVALID;
            if (lane == 0) T.nrows[w] = nrows;
            tc5::mbar_arrive(tc5::smem_u32(&S.tb.rows_full[b]));
            ++n;
            R5_STAT_ADD(14, 1);
            R5_STAT_ADD(21, nrows);
            if (nrows == 0) {
                // this part is empty for good (every slot is out of rays).  The work ends with the first tile whose other
                // parts are empty too; one more empty tile follows so that the roles of both TMEM slots see an end.
                R5_WAIT(tc5::smem_u32(&S.tb.rows_full[b]), (uint32_t)(((n - 1) >> R5_TBUF_LOG2) & 1), 18);
                if (r5_tile_end(T)) {
                    const int b2 = (int)(n & (R5_TBUF - 1));
                    if (n >= R5_TBUF) tc5::mbar_wait_wd(tc5::smem_u32(&S.tb.rows_free[b2]), (uint32_t)(((n >> R5_TBUF_LOG2) - 1) & 1));
                    R5Tile &T2 = S.tile[b2];
                    for (int row = lane; row < R5_PART; row += 32) T2.slot[w * R5_PART + row] = R5_INVALID;
                    if (lane == 0) T2.nrows[w] = 0;
                    tc5::mbar_arrive(tc5::smem_u32(&S.tb.rows_full[b2]));
                    ended = true;
                }
            }
        } else if (!any_want) {
            __nanosleep(200);   // nothing to march, nothing to build: rays are in flight — leave the issue slots to the others
        }
        if (ended) break;
        // ---- 4. march: compact the slots that want a step, then walk the list
        int nlist = 0;
#pragma unroll
        for (int j = 0; j < R5_PER_LANE; ++j) {
            const bool wj = (want >> j) & 1u;
            const unsigned m = __ballot_sync(0xffffffffu, wj);
            if (wj) list[nlist + __popc(m & lt_mask)] = (uint16_t)(slot0 + 32 * j);
            nlist += __popc(m);
        }
        R5_STAT_ADD(19, nlist);
        __syncwarp();
        // Each lane walks its share of the list (entries lane, lane + 32, ...) with TWO rays in hand, alternating between them
        // every iteration: while the occupancy load of one is in flight the other's step is finished and its next load
        // issued (rolled: one copy of the step in the instruction cache).  A ray keeps its chain for up to R5_BUDGET steps
        // or until its queue is full, so a pass moves a ray through several empty voxels and the per-pass bookkeeping
        // (classification, compaction, tile building) is paid once per several steps.
        {
            int i = lane;
            bool haveA = false, haveB = false;
            int slotA = 0, lvlA = 0, stepsA = 0, slotB = 0, lvlB = 0, stepsB = 0;
            float tA = 0.0f, tB = 0.0f;
            uint32_t cellA = 0u, nxyzA = 0u, msA = 0u, byteA = 0u, cellB = 0u, nxyzB = 0u, msB = 0u, byteB = 0u;
#pragma unroll 1
            for (;;) {
                if (haveA) {   // finish the step whose load was issued two iterations ago
                    if ((byteA >> (cellA & 7u)) & 1u) {
                        const float dt = cfg.lvl_dt >= 0 ? cfg.dt_min : ngp_clampf(tA * cfg.dt_gamma, cfg.dt_min, cfg.dt_max);
                        const uint32_t wpos = (R5_MS_HEAD(msA) + R5_MS_CNT(msA)) % R5_FIFO;
                        R.qt[wpos][slotA] = tA;
                        R.qdt[wpos][slotA] = dt;
                        tA += dt;
                        msA += (1u << 20) + 1u;   // one more queued, one more marched
                    } else {
                        MarchRay r;
                        r.ox = R.ox[slotA]; r.oy = R.oy[slotA]; r.oz = R.oz[slotA];
                        r.dx = R.dx[slotA]; r.dy = R.dy[slotA]; r.dz = R.dz[slotA];
                        r.rdx = R.rdx[slotA]; r.rdy = R.rdy[slotA]; r.rdz = R.rdz[slotA];
                        const float tt = ngp_march_exit_fast(r, cfg, S.mipb, tA, nxyzA, lvlA);
                        if (cfg.lvl_dt >= 0) {   // dt_gamma == 0: the recurrence of ngp_march_advance is t += dt_min
                            do {
                                tA += cfg.dt_min;
                            } while (tA < tt);
                        } else {
                            do {
                                tA = ngp_march_advance(tA, cfg);
                            } while (tA < tt);
                        }
                    }
                    --stepsA;
                    bool go = stepsA > 0 && R5_MS_CNT(msA) < R5_FIFO;
                    if (go && (!(tA < R.far[slotA]) || R5_MS_MARCHED(msA) >= P.max_steps)) {   // raymarching.cu:730
                        msA |= 1u << 25;
                        go = false;
                    }
                    if (!go) {
                        R.t[slotA] = tA;
                        R.mstate[slotA] = msA;
                        haveA = false;
                    }
                }
                if (!haveA && i < nlist) {   // take the next list entry
                    slotA = (int)list[i];
                    i += 32;
                    msA = R.mstate[slotA];
                    tA = R.t[slotA];
                    stepsA = R5_BUDGET;
                    if (!(tA < R.far[slotA]) || R5_MS_MARCHED(msA) >= P.max_steps) R.mstate[slotA] = msA | (1u << 25);
                    else haveA = true;
                }
                if (haveA) {                 // issue the occupancy load of the ray's current cell
                    MarchRay r;
                    r.ox = R.ox[slotA]; r.oy = R.oy[slotA]; r.oz = R.oz[slotA];
                    r.dx = R.dx[slotA]; r.dy = R.dy[slotA]; r.dz = R.dz[slotA];
                    float dt;
                    cellA = ngp_march_cell_fast(r, cfg, S.miprb, tA, dt, lvlA, nxyzA);
                    byteA = (uint32_t)__ldg(cfg.grid + (cellA >> 3));
                }
                if (!haveA && !haveB && i >= nlist) break;
                // swap the two chains — unless the other one is idle and there is nothing left to give it (then this lane
                // keeps stepping its one ray instead of spending every other iteration on an empty chain)
                if (haveB || i < nlist) {
                    { const bool h = haveA; haveA = haveB; haveB = h; }
                    { const int v = slotA; slotA = slotB; slotB = v; }
                    { const int v = lvlA; lvlA = lvlB; lvlB = v; }
                    { const int v = stepsA; stepsA = stepsB; stepsB = v; }
                    { const float v = tA; tA = tB; tB = v; }
                    { const uint32_t v = cellA; cellA = cellB; cellB = v; }
                    { const uint32_t v = nxyzA; nxyzA = nxyzB; nxyzB = v; }
                    { const uint32_t v = msA; msA = msB; msB = v; }
                    { const uint32_t v = byteA; byteA = byteB; byteB = v; }
                }
            }
        }
        __syncwarp();
    }
    if (P.nsamples) {
#pragma unroll
        for (int o2 = 16; o2 > 0; o2 >>= 1) evaluated += __shfl_xor_sync(0xffffffffu, evaluated, o2);
        if (lane == 0 && evaluated) atomicAdd(P.nsamples, evaluated);
    }
    R5_STAT_ADD(17, clock64());
    R5_STAT_FLUSH(P, lane);
}

// ------------------------------------------------------------------------------------------------ gather warps
template <bool SAFE, class SM>
__device__ __forceinline__ void r5_gather(SM &S, const R5Params &P, uint32_t tmem, int warp, int lane) {
    const int s = warp >> 2, q = warp & 3, row = 32 * q + lane;
    const uint32_t lane_addr = tmem + s * C5_SLOT_COLS + ((uint32_t)(32 * q) << 16);
    const uint32_t bar_full = tc5::smem_u32(&S.bars[s].f_full), bar_empty = tc5::smem_u32(&S.bars[s].f_empty);
    const float den = 2.0f * P.bound;
    R5_STAT_DECL;
    R5_STAT_ADD(2, -clock64());
    for (int64_t it = 0;; ++it) {
        const int64_t n = 2 * it + s;
        const int b = (int)(n & (SM::TBUF - 1));
        R5_WAIT(tc5::smem_u32(&S.tb.rows_full[b]), (uint32_t)((n >> SM::TBUF_LOG2) & 1), 0);
        const R5Tile &T = S.tile[b];
        if (r5_tile_end(T)) break;
        bool valid = T.slot[row] != R5_INVALID;
        float x01[3];
        x01[0] = __fdiv_rn((valid ? T.x[row] : 0.0f) + P.bound, den);   // grid.py:144
        x01[1] = __fdiv_rn((valid ? T.y[row] : 0.0f) + P.bound, den);
        x01[2] = __fdiv_rn((valid ? T.z[row] : 0.0f) + P.bound, den);
#pragma unroll
        for (int a = 0; a < 3; ++a) valid = valid && !(x01[a] < 0.0f || x01[a] > 1.0f);
        if (!valid) x01[0] = x01[1] = x01[2] = 0.5f;
#pragma unroll 1
        for (int lp = 0; lp < 8; ++lp) {
            float feat[4];
            c5_encode_pair<SAFE>(P.table, S.lv, lp, x01, valid, P.align_corners, feat);
            if (lp == 0) {
                if (it > 0) R5_WAIT(bar_empty, (uint32_t)((it - 1) & 1), 1);
                tc5::fence_after_sync();
            }
            c5_put_features(lane_addr, lp, feat);
        }
        tc5::tmem_st_wait();
        tc5::fence_before_sync();
        tc5::mbar_arrive(bar_full);
    }
    R5_STAT_ADD(2, clock64());
    R5_STAT_FLUSH(P, lane);
}

// ------------------------------------------------------------------------------------------------ MMA threads
template <class SM>
__device__ __forceinline__ void r5_mma(SM &S, const R5Params &P, uint32_t tmem, int s) {
    R5_STAT_DECL;
    R5_STAT_ADD(7, -clock64());
    const uint32_t t = tmem + s * C5_SLOT_COLS;
    const uint32_t bar_ff = tc5::smem_u32(&S.bars[s].f_full), bar_fe = tc5::smem_u32(&S.bars[s].f_empty),
                   bar_d = tc5::smem_u32(&S.bars[s].d_full), bar_a = tc5::smem_u32(&S.bars[s].a_ready),
                   bar_t = tc5::smem_u32(&S.bars[s].t_free);
    const uint32_t s0h = tc5::smem_u32(S.w.s0[0]), s0l = tc5::smem_u32(S.w.s0[1]), s1h = tc5::smem_u32(S.w.s1[0]),
                   s1l = tc5::smem_u32(S.w.s1[1]), c0h = tc5::smem_u32(S.w.c0[0]), c0l = tc5::smem_u32(S.w.c0[1]),
                   c1h = tc5::smem_u32(S.w.c1[0]), c1l = tc5::smem_u32(S.w.c1[1]), c2h = tc5::smem_u32(S.w.c2[0]),
                   c2l = tc5::smem_u32(S.w.c2[1]);
    for (int64_t it = 0;; ++it) {
        const int64_t n = 2 * it + s;
        const int b = (int)(n & (SM::TBUF - 1));
        R5_WAIT(tc5::smem_u32(&S.tb.rows_full[b]), (uint32_t)((n >> SM::TBUF_LOG2) & 1), 3);
        if (r5_tile_end(S.tile[b])) break;
        const uint32_t pa = (uint32_t)(4 * it);
        R5_WAIT(bar_ff, (uint32_t)(it & 1), 4);
        if (it > 0) R5_WAIT(bar_t, (uint32_t)((it - 1) & 1), 5);
        tc5::fence_after_sync();
        c5_issue_layer<64, 32>(t + C5_D, t + C5_FHI, t + C5_FLO, s0h, s0l);   // sigma L1: features -> 64
        tc5::mma_commit(bar_fe);
        tc5::mma_commit(bar_d);
        R5_WAIT(bar_a, (pa + 0u) & 1u, 6);
        tc5::fence_after_sync();
        c5_issue_layer<16, 64>(t + C5_D, t + C5_HHI, t + C5_HLO, s1h, s1l);   // sigma L2: 64 -> 16
        tc5::mma_commit(bar_d);
        R5_WAIT(bar_a, (pa + 1u) & 1u, 6);
        tc5::fence_after_sync();
        c5_issue_layer<64, 32>(t + C5_D, t + C5_HHI, t + C5_HLO, c0h, c0l);   // colour L1: [SH | o] (32) -> 64
        tc5::mma_commit(bar_d);
        R5_WAIT(bar_a, (pa + 2u) & 1u, 6);
        tc5::fence_after_sync();
        c5_issue_layer<64, 64>(t + C5_D, t + C5_HHI, t + C5_HLO, c1h, c1l);   // colour L2: 64 -> 64
        tc5::mma_commit(bar_d);
        R5_WAIT(bar_a, (pa + 3u) & 1u, 6);
        tc5::fence_after_sync();
        c5_issue_layer<16, 64>(t + C5_D, t + C5_HHI, t + C5_HLO, c2h, c2l);   // colour L3: 64 -> 3 (16)
        tc5::mma_commit(bar_d);
    }
    R5_STAT_ADD(7, clock64());
    R5_STAT_FLUSH(P, 0);
}

// ------------------------------------------------------------------------------------------------ chain warps
// Tiles are taken in order, two at a time when the second one's rows are already there (its epilogues then run while the
// first tile's next layer is on the tensor core, as in cfield5.cu), otherwise one — waiting for a tile that the marchers
// can only build once THIS tile's rays have been composited would deadlock.  d_full of a TMEM slot completes 5 times per
// tile: wait k of the slot's tile number `it` has parity (5 it + k) & 1.
template <class SM>
__device__ __forceinline__ void r5_chain(SM &S, const R5Params &P, uint32_t tmem, int warp, int lane) {
    const int q = warp & 3, row = 32 * q + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q) << 16);
    auto &R = S.rays;
    int64_t n0 = 0;
    uint32_t group = 0;
    R5_STAT_DECL;
    R5_STAT_ADD(10, -clock64());
    for (;;) {
        R5_WAIT(tc5::smem_u32(&S.tb.rows_full[n0 & (SM::TBUF - 1)]), (uint32_t)((n0 >> SM::TBUF_LOG2) & 1), 8);
        if (r5_tile_end(S.tile[n0 & (SM::TBUF - 1)])) break;
        // one decision for the four chain warps (they hand the same tiles to the MMA threads step by step, so taking a tile
        // in one warp and not in another would deadlock): the first chain warp tests, the others read its verdict
        if (warp == C5_CHAIN_WARP0 && lane == 0) {
            const int64_t n1 = n0 + 1;
            bool ok = tc5::mbar_try(tc5::smem_u32(&S.tb.rows_full[n1 & (SM::TBUF - 1)]), (uint32_t)((n1 >> SM::TBUF_LOG2) & 1));
            if (ok && r5_tile_end(S.tile[n1 & (SM::TBUF - 1)])) ok = false;
            S.chain_two[group & 1] = ok ? 1 : 0;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const int n_act = S.chain_two[group & 1] ? 2 : 1;
        ++group;
        R5_STAT_ADD(12, 1);
        R5_STAT_ADD(11, n_act - 1);
#define R5_TILE_VARS                                                                                \
        const int64_t n = n0 + i;                                                                   \
        const int s = (int)(n & 1);                                                                 \
        R5Tile &T = S.tile[n & (SM::TBUF - 1)];                                                      \
        const uint32_t par = (uint32_t)(5 * (n >> 1));                                              \
        const uint32_t lane_addr = lane_base + s * C5_SLOT_COLS;                                    \
        const uint32_t bar_d = tc5::smem_u32(&S.bars[s].d_full), bar_a = tc5::smem_u32(&S.bars[s].a_ready);
        // ---- sigma hidden
#pragma unroll 1
        for (int i = 0; i < n_act; ++i) {
            R5_TILE_VARS
            (void)T;
            R5_WAIT(bar_d, (par + 0u) & 1u, 9);
            tc5::fence_after_sync();
            c5_relu_to_a<false>(lane_addr);
            c5_chain_arrive(bar_a);
        }
        // ---- sigma-net output o (column 0 = log sigma, 1..15 = geo_feat) -> colour input [SH(d) | 0, geo_feat]
#pragma unroll 1
        for (int i = 0; i < n_act; ++i) {
            R5_TILE_VARS
            const float dir[3] = {T.dx[row], T.dy[row], T.dz[row]};
            R5_WAIT(bar_d, (par + 1u) & 1u, 9);
            tc5::fence_after_sync();
            float o[16];
            c5_get16(lane_addr, 0, o);
            T.lsig[row] = o[0];
            o[0] = 0.0f;
            c5_put16(lane_addr, 16, o);
            float sh[16];
            ngp_sh4(dir[0], dir[1], dir[2], sh);
            c5_put16(lane_addr, 0, sh);
            c5_chain_arrive(bar_a);
        }
        // ---- colour hidden 1, 2
#pragma unroll 1
        for (int k = 0; k < 2; ++k) {
#pragma unroll 1
            for (int i = 0; i < n_act; ++i) {
                R5_TILE_VARS
                (void)T;
                R5_WAIT(bar_d, (par + 2u + (uint32_t)k) & 1u, 9);
                tc5::fence_after_sync();
                c5_relu_to_a<false>(lane_addr);
                c5_chain_arrive(bar_a);
            }
        }
        // ---- colour logits -> sigmoid, and the compositing step of the row's ray (raymarching.cu:860-893)
#pragma unroll 1
        for (int i = 0; i < n_act; ++i) {
            R5_TILE_VARS
            (void)bar_a;
            const uint32_t slot = T.slot[row];
            const float tn = T.tn[row], sdt = T.dt[row], lsig = T.lsig[row];
            R5_WAIT(bar_d, (par + 4u) & 1u, 9);
            tc5::fence_after_sync();
            uint32_t lg[4];
            tc5::tmem_ld4_nowait(lane_addr + C5_D, lg);
            tc5::tmem_ld_wait();
            tc5::fence_before_sync();
            tc5::mbar_arrive(tc5::smem_u32(&S.bars[s].t_free));
            if (slot != R5_INVALID) {
                bool dead = (r5_ld_status(&R.status[slot]) & R5_DEAD) != 0u;   // an earlier sample of the ray ended it
                if (!dead) {
                    const float sigma = expf(lsig) * P.density_scale;   // trunc_exp fwd; renderer.py:355
                    const float cr = 1.0f / (1.0f + expf(-__uint_as_float(lg[0])));
                    const float cg = 1.0f / (1.0f + expf(-__uint_as_float(lg[1])));
                    const float cb = 1.0f / (1.0f + expf(-__uint_as_float(lg[2])));
                    const float ws = R.ws[slot];
                    const float alpha = 1.0f - __expf(-sigma * sdt);
                    const float Tr = 1.0f - ws;
                    const float wgt = alpha * Tr;
                    R.ws[slot] = ws + wgt;
                    R.depth[slot] = fmaf(wgt, tn, R.depth[slot]);
                    R.r[slot] = fmaf(wgt, cr, R.r[slot]);
                    R.g[slot] = fmaf(wgt, cg, R.g[slot]);
                    R.b[slot] = fmaf(wgt, cb, R.b[slot]);
                    R.ncomp[slot] += 1u;
                    dead = (double)Tr < 1e-4;   // :880 — the sample that saw T < 1e-4 is still accumulated, then the ray is over
                }
                __threadfence_block();
                if (dead) atomicOr(&R.status[slot], R5_DEAD);
                atomicAdd(&R.status[slot], 0xffffffffu);   // one sample fewer in flight (after the flag: the marcher retires at 0)
            }
            // the two samples a ray may have in flight sit in consecutive tiles and in the same 32-row band (its marcher
            // warp's rows), i.e. this warp composites both, one after the other: order and visibility between its lanes
            __syncwarp();
            tc5::mbar_arrive(tc5::smem_u32(&S.tb.rows_free[n & (SM::TBUF - 1)]));
        }
        n0 += n_act;
    }
    R5_STAT_ADD(10, clock64());
    R5_STAT_FLUSH(P, lane);
}

__global__ void __launch_bounds__(R5_THREADS, 1)
k_render5_rays_infer(R5Params P, MarchCfg cfg, const int32_t *__restrict__ offsets, float S_, uint32_t H, uint32_t gridtype,
                     const float *__restrict__ Ws0, const float *__restrict__ Ws1, const float *__restrict__ Wc0,
                     const float *__restrict__ Wc1, const float *__restrict__ Wc2) {
    extern __shared__ __align__(128) unsigned char r5_smem_raw[];
    R5Smem &S = *reinterpret_cast<R5Smem *>(r5_smem_raw);
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    fq_stage_levels(S.lv, &S.all_safe, offsets, S_, H, gridtype, P.align_corners);
    c5_stage_forward_weights(S.w, Ws0, Ws1, Wc0, Wc1, Wc2);
    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            tc5::mbar_init(tc5::smem_u32(&S.bars[s].f_full), 128);
            tc5::mbar_init(tc5::smem_u32(&S.bars[s].f_empty), 1);
            tc5::mbar_init(tc5::smem_u32(&S.bars[s].d_full), 1);
            tc5::mbar_init(tc5::smem_u32(&S.bars[s].a_ready), 128);
            tc5::mbar_init(tc5::smem_u32(&S.bars[s].t_free), 128);
            tc5::mbar_init(tc5::smem_u32(&S.bars[s].g_full), 1);
        }
        for (int b = 0; b < R5_TBUF; ++b) {
            tc5::mbar_init(tc5::smem_u32(&S.tb.rows_full[b]), R5_MARCH_WARPS * 32);
            tc5::mbar_init(tc5::smem_u32(&S.tb.rows_free[b]), 128);
        }
        tc5::fence_mbar_init();
    }
    if (tid < 8) {
        S.mipb[tid] = fminf((float)(1 << tid), cfg.bound);
        S.miprb[tid] = 1.0f / S.mipb[tid];
    }
    if (warp == C5_MMA_WARP) tc5::tmem_alloc(tc5::smem_u32(&S.tmem), 512);
    tc5::fence_async_smem();
    tc5::fence_before_sync();
    __syncthreads();
    tc5::fence_after_sync();
    const uint32_t tmem = __shfl_sync(0xffffffffu, S.tmem, 0);

    if (warp < C5_GATHER_WARPS) {
        if (S.all_safe) r5_gather<true, R5Smem>(S, P, tmem, warp, lane);
        else r5_gather<false, R5Smem>(S, P, tmem, warp, lane);
    } else if (warp < C5_MMA_WARP) {
        r5_chain<R5Smem>(S, P, tmem, warp, lane);
    } else if (warp < R5_MARCH_WARP0) {
        if (lane == 0) r5_mma<R5Smem>(S, P, tmem, warp - C5_MMA_WARP);
        __syncwarp();
    } else {
        r5_marcher(S, P, cfg, warp - R5_MARCH_WARP0, lane);
    }
    tc5::fence_before_sync();
    __syncthreads();
    if (warp == C5_MMA_WARP) tc5::tmem_dealloc(tmem, 512);
}

// ================================================================================================ pre-marched variant
// What bounds the kernel above is its marcher: occupancy marching is a chain of dependent steps (position -> cell -> bit ->
// next t) that only thread-level parallelism hides, and 8 warps per SM offer 1/8 of what a full-occupancy kernel does
// (R5_STATS: 2.7 marcher passes of 8,800 cycles per 15 tile rows; gather warps 37 %, chain warps 43 % of their time
// waiting).  Here the marching moves OUT of the persistent kernel into a thread-per-ray pre-pass at full occupancy
// (k_r5_premarch: the same cell / exit arithmetic, so the same samples bit for bit) that leaves every ray's (t, dt) list in
// global memory — a count pass, an inclusive scan, a write pass; or, in the single-call form further down, one pass that
// records each ray's RUNS of samples — and the persistent kernel's 8 "feeder" warps only hand list entries to tiles: classify the slots, retire / refill rays, ballot-compact the eligible ones into rows.  Early
// termination stays exact (samples after a ray's end are never evaluated; only their marching is wasted), per-ray counts
// stay the oracle's.  The ray store shrinks (no sample FIFOs, no reciprocal directions, 512 slots): 162 KB of shared memory
// instead of 219 KB, which gives the gathers L1 back (dfield5.cu: the carve-out is part of the roofline).  Measured on one
// 800x800 frame: 4 / 8 / 16 row buffers (142 / 162 / 223 KB) -> 22.2 / 20.7 / 29.2 ms.
#ifndef R5L_PER_LANE
#define R5L_PER_LANE 2
#endif
#define R5L_PER_WARP (32 * R5L_PER_LANE)
#define R5L_NR (R5_MARCH_WARPS * R5L_PER_WARP)     // 512 ray slots per CTA
#ifndef R5L_TBUF_LOG2
#define R5L_TBUF_LOG2 3
#endif

struct R5RaysL {
    float ox[R5L_NR], oy[R5L_NR], oz[R5L_NR], dx[R5L_NR], dy[R5L_NR], dz[R5L_NR];
    float near[R5L_NR], far[R5L_NR];
    int id[R5L_NR];                                  // ray index; -1: slot wants a ray; -2: no rays left
    uint32_t cursor[R5L_NR], count[R5L_NR];          // lists: samples handed to tiles so far / in the ray's list
                                                     // segments: runs opened so far / runs of the ray
    long long off[R5L_NR];                           // first list entry / first segment of the ray
    uint32_t rem[R5L_NR];                            // segments: samples left in the open run
    float tcur[R5L_NR];                              // segments: t of the run's next sample
    float ws[R5L_NR], depth[R5L_NR], r[R5L_NR], g[R5L_NR], b[R5L_NR];
    uint32_t ncomp[R5L_NR];
    uint32_t status[R5L_NR];
};

struct __align__(128) R5SmemL {
    static constexpr int TBUF = 1 << R5L_TBUF_LOG2, TBUF_LOG2 = R5L_TBUF_LOG2;
    C5FwdWeights w;
    FqLevel lv[FQ_L];
    C5Bars bars[2];
    R5BarsT<(1 << R5L_TBUF_LOG2)> tb;
    uint32_t tmem;
    int all_safe;
    int chain_two[2];
    R5Tile tile[1 << R5L_TBUF_LOG2];
    R5RaysL rays;
};

template <bool SEG>
static __device__ __noinline__ void r5l_load_ray(R5RaysL &R, const R5Params &P, int slot, int64_t nid) {
    const float o[3] = {__ldg(P.rays_o + nid * 3), __ldg(P.rays_o + nid * 3 + 1), __ldg(P.rays_o + nid * 3 + 2)};
    const float d[3] = {__ldg(P.rays_d + nid * 3), __ldg(P.rays_d + nid * 3 + 1), __ldg(P.rays_d + nid * 3 + 2)};
    float near, far;
    ngp_near_far_ray(o, d, P.aabb, P.min_near, near, far);
    R.ox[slot] = o[0]; R.oy[slot] = o[1]; R.oz[slot] = o[2];
    R.dx[slot] = d[0]; R.dy[slot] = d[1]; R.dz[slot] = d[2];
    R.near[slot] = near; R.far[slot] = far;
    R.ws[slot] = 0.0f; R.depth[slot] = 0.0f; R.r[slot] = 0.0f; R.g[slot] = 0.0f; R.b[slot] = 0.0f;
    R.ncomp[slot] = 0u;
    R.cursor[slot] = 0u;
    if (SEG) {
        const int ns = __ldg(P.nsegs + nid);
        R.count[slot] = ns < 0 ? 0xffffffffu : (uint32_t)ns;   // 0xffffffff: not this kernel's ray (see r5l_feeder)
        R.off[slot] = (long long)nid * P.seg_cap;
        R.rem[slot] = 0u;
    } else {
        const uint32_t cnt = (uint32_t)__ldg(P.plan_counts + nid);
        R.count[slot] = cnt;
        R.off[slot] = __ldg(P.plan_ends + nid) - (long long)cnt;
    }
    R.id[slot] = (int)nid;
}

// One pass of a feeder warp over its R5L_PER_WARP slots: classify (eligible / finished / waiting), retire and refill,
// build the warp's R5_PART rows of tile n from the eligible slots' next list entries.
template <bool SEG>
__device__ __forceinline__ void r5l_feeder(R5SmemL &S, const R5Params &P, int w, int lane) {
    R5RaysL &R = S.rays;
    const int slot0 = w * R5L_PER_WARP + lane;
    const unsigned lt_mask = (1u << lane) - 1u;
#pragma unroll 1
    for (int j = 0; j < R5L_PER_LANE; ++j) {
        const int slot = slot0 + 32 * j;
        R.id[slot] = -1;
        r5_st_status(&R.status[slot], R5_IDLE);
    }
    bool rays_left = true;
    unsigned long long evaluated = 0;
    int64_t n = 0;                                  // next tile to build
    int rot = 0;
    unsigned needs = (1u << R5L_PER_LANE) - 1u;     // bit j: slot j of this lane wants a ray
    bool ended = false;
    R5_STAT_DECL;
    R5_STAT_ADD(17, -clock64());
    for (;;) {
        R5_STAT_ADD(13, 1);
        // ---- refill the slots that want a ray: one atomic per warp
        if (!rays_left) needs = 0u;
        if (__any_sync(0xffffffffu, needs != 0u)) {
            const int mine = __popc(needs);
            int pre = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, pre, o);
                if (lane >= o) pre += v;
            }
            const int total = __shfl_sync(0xffffffffu, pre, 31);
            unsigned long long base = 0;
            if (lane == 31) base = atomicAdd(P.counter, (unsigned long long)total);
            base = __shfl_sync(0xffffffffu, base, 31);
            int64_t nid = (int64_t)base + pre - mine;
            while (needs) {
                const int j = __ffs(needs) - 1;
                needs &= needs - 1u;
                const int slot = slot0 + 32 * j;
                if (nid < P.N) {
                    r5l_load_ray<SEG>(R, P, slot, nid);
                    ++nid;
                } else {
                    R.id[slot] = -2;
                    rays_left = false;
                }
            }
        }
        // ---- classify
        unsigned elig = 0u, retire = 0u;
        bool pending = false;
#pragma unroll
        for (int j = 0; j < R5L_PER_LANE; ++j) {
            const int slot = slot0 + 32 * j;
            if (R.id[slot] < 0) continue;
            const uint32_t st = r5_ld_status(&R.status[slot]);
            const uint32_t rcount = R.count[slot];
            const bool foreign = SEG && rcount == 0xffffffffu;   // too many runs: the marcher-inside kernel renders this ray
            const bool dead = (st & R5_DEAD) != 0u;
            const bool more = !foreign && (R.cursor[slot] < rcount || (SEG && R.rem[slot] > 0u));
            const uint32_t fly = R5_NFLY(st);
            if (!dead && more && fly < R5_INFLIGHT) elig |= 1u << j;
            else if (fly == 0u && (dead || !more)) retire |= 1u << j;
            else pending = true;   // in flight: comes back
        }
        // ---- retire finished rays (their slots are refilled at the top of the next pass)
        if (retire) pending = true;
        while (retire) {
            const int j = __ffs(retire) - 1;
            retire &= retire - 1u;
            const int slot = slot0 + 32 * j;
            __threadfence_block();         // the chain thread wrote the accumulators before the status
            if (!(SEG && R.count[slot] == 0xffffffffu)) r5_finish(R, P, slot, R.id[slot]);
            evaluated += R.ncomp[slot];
            if (rays_left) {
                R.id[slot] = -1;
                needs |= 1u << j;
            } else {
                R.id[slot] = -2;
            }
            r5_st_status(&R.status[slot], R5_IDLE);
        }
        // ---- build this warp's part of tile n: R5_PART eligible slots, or all of them if nothing else can come
        int E = __popc(elig);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) E += __shfl_xor_sync(0xffffffffu, E, o);
        const bool any_pending = __any_sync(0xffffffffu, pending);
        bool build = E >= R5_PART || !any_pending;
        if (!build) R5_STAT_ADD(15, 1);
        R5_STAT_ADD(20, E);
        const int b = (int)(n & (R5SmemL::TBUF - 1));
        if (build && n >= R5SmemL::TBUF) {
            // the rows are ready, the buffer is what is missing: block on its barrier (the eligible slots stay eligible —
            // in-flight counts only fall meanwhile) instead of polling pass after pass in the other roles' issue slots
            if (lane == 0) R5_WAIT(tc5::smem_u32(&S.tb.rows_free[b]), (uint32_t)(((n >> R5L_TBUF_LOG2) - 1) & 1), 16);
            __syncwarp();
        }
        if (build) {
            R5Tile &T = S.tile[b];
            int nrows = 0;
#pragma unroll 1
            for (int jj = 0; jj < R5L_PER_LANE && nrows < R5_PART; ++jj) {
                int j = jj + rot;
                if (j >= R5L_PER_LANE) j -= R5L_PER_LANE;
                const bool e = (elig >> j) & 1u;
                const unsigned m = __ballot_sync(0xffffffffu, e);
                const int row = nrows + __popc(m & lt_mask);
                if (e && row < R5_PART) {
                    const int slot = slot0 + 32 * j;
                    float st, sdt;
                    if (SEG) {
                        uint32_t rem = R.rem[slot];
                        float tc = R.tcur[slot];
                        if (rem == 0u) {   // open the ray's next run
                            const uint32_t cur = R.cursor[slot];
                            const float2 sg = __ldg(P.segs + R.off[slot] + cur);
                            R.cursor[slot] = cur + 1u;
                            tc = sg.x;
                            rem = (uint32_t)__float_as_int(sg.y);
                        }
                        st = tc;   // the marcher's recurrence inside a run: t += dt(t)
                        sdt = P.lvl_dt >= 0 ? P.dt_min : ngp_clampf(tc * P.dt_gamma, P.dt_min, P.dt_max);
                        R.tcur[slot] = tc + sdt;
                        R.rem[slot] = rem - 1u;
                    } else {
                        const uint32_t cur = R.cursor[slot];
                        const float2 sm = __ldg(P.tdt + R.off[slot] + cur);
                        R.cursor[slot] = cur + 1u;
                        st = sm.x; sdt = sm.y;
                    }
                    const float dxv = R.dx[slot], dyv = R.dy[slot], dzv = R.dz[slot];
                    const int tr = w * R5_PART + row;
                    T.x[tr] = ngp_clampf(fmaf(st, dxv, R.ox[slot]), -P.bound, P.bound);   // = the marcher's sample position
                    T.y[tr] = ngp_clampf(fmaf(st, dyv, R.oy[slot]), -P.bound, P.bound);
                    T.z[tr] = ngp_clampf(fmaf(st, dzv, R.oz[slot]), -P.bound, P.bound);
                    T.dx[tr] = dxv; T.dy[tr] = dyv; T.dz[tr] = dzv;
                    T.tn[tr] = st + sdt;
                    T.dt[tr] = sdt;
                    T.slot[tr] = (uint32_t)slot;
                    atomicAdd(&R.status[slot], 1u);
                }
                nrows = min(R5_PART, nrows + __popc(m));
            }
            rot = rot + 1 >= R5L_PER_LANE ? 0 : rot + 1;
            for (int row = nrows + lane; row < R5_PART; row += 32) T.slot[w * R5_PART + row] = R5_INVALID;
            if (lane == 0) T.nrows[w] = nrows;
            tc5::mbar_arrive(tc5::smem_u32(&S.tb.rows_full[b]));
            ++n;
            R5_STAT_ADD(14, 1);
            R5_STAT_ADD(21, nrows);
            if (nrows == 0) {
                // this part is empty for good (every slot is out of rays): the end protocol of r5_marcher
                R5_WAIT(tc5::smem_u32(&S.tb.rows_full[b]), (uint32_t)(((n - 1) >> R5L_TBUF_LOG2) & 1), 18);
                if (r5_tile_end(T)) {
                    const int b2 = (int)(n & (R5SmemL::TBUF - 1));
                    if (n >= R5SmemL::TBUF)
                        tc5::mbar_wait_wd(tc5::smem_u32(&S.tb.rows_free[b2]), (uint32_t)(((n >> R5L_TBUF_LOG2) - 1) & 1));
                    R5Tile &T2 = S.tile[b2];
                    for (int row = lane; row < R5_PART; row += 32) T2.slot[w * R5_PART + row] = R5_INVALID;
                    if (lane == 0) T2.nrows[w] = 0;
                    tc5::mbar_arrive(tc5::smem_u32(&S.tb.rows_full[b2]));
                    ended = true;
                }
            }
        } else {
            __nanosleep(400);   // too few eligible rays: they come back a tile at a time (several us); leave the issue slots to the others
        }
        if (ended) break;
        __syncwarp();
    }
    if (P.nsamples) {
#pragma unroll
        for (int o2 = 16; o2 > 0; o2 >>= 1) evaluated += __shfl_xor_sync(0xffffffffu, evaluated, o2);
        if (lane == 0 && evaluated) atomicAdd(P.nsamples, evaluated);
    }
    R5_STAT_ADD(17, clock64());
    R5_STAT_FLUSH(P, lane);
}

// Thread-per-ray marcher of the pre-pass: r5_marcher's sequence exactly (same cell / exit functions, same end tests).
// MODE 0: counts[i] = number of samples.  MODE 1: the samples go to tdt[ends[i] - counts[i] ...].
// MODE 2: the RUNS of consecutive samples go to segs[i * seg_cap ...] as (t of the first sample, number of samples): inside
// a run the marcher's recurrence is t += dt(t), which the feeder replays, so a run is all it needs; a ray with more than
// seg_cap runs gets nsegs[i] = -1 and its index is appended to ovf_list (rendered by the marcher-inside kernel).
template <int MODE>
__global__ void __launch_bounds__(256)
k_r5_premarch(const float *__restrict__ rays_o, const float *__restrict__ rays_d, int64_t N, const float *__restrict__ aabb,
              float min_near, MarchCfg cfg, uint32_t max_steps, int32_t *__restrict__ counts, const long long *__restrict__ ends,
              float2 *__restrict__ tdt, int seg_cap, int32_t *__restrict__ ovf_list, unsigned long long *__restrict__ ovf_count) {
    __shared__ float mipb[8], miprb[8];
    if (threadIdx.x < 8) {
        mipb[threadIdx.x] = fminf((float)(1 << threadIdx.x), cfg.bound);
        miprb[threadIdx.x] = 1.0f / mipb[threadIdx.x];
    }
    __syncthreads();
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const float o[3] = {__ldg(rays_o + i * 3), __ldg(rays_o + i * 3 + 1), __ldg(rays_o + i * 3 + 2)};
    const float d[3] = {__ldg(rays_d + i * 3), __ldg(rays_d + i * 3 + 1), __ldg(rays_d + i * 3 + 2)};
    float near, far;
    ngp_near_far_ray(o, d, aabb, min_near, near, far);
    MarchRay r;
    r.ox = o[0]; r.oy = o[1]; r.oz = o[2]; r.dx = d[0]; r.dy = d[1]; r.dz = d[2];
    r.rdx = 1.0f / d[0]; r.rdy = 1.0f / d[1]; r.rdz = 1.0f / d[2];
    float2 *out = nullptr;
    if (MODE == 1) out = tdt + (ends[i] - (long long)counts[i]);
    if (MODE == 2) out = tdt + i * seg_cap;
    float t = near;
    uint32_t marched = 0;
    int nseg = 0;               // MODE 2: runs closed so far, the open run (run_n > 0) and whether the ray ran out of room
    float run_t0 = 0.0f;
    uint32_t run_n = 0;
    bool overflow = false;
#pragma unroll 1
    while (t < far && marched < max_steps) {   // raymarching.cu:730
        float dt;
        int lvl;
        uint32_t nxyz;
        const uint32_t cell = ngp_march_cell_fast(r, cfg, miprb, t, dt, lvl, nxyz);
        const uint32_t byte = (uint32_t)__ldg(cfg.grid + (cell >> 3));
        if ((byte >> (cell & 7u)) & 1u) {
            const float sdt = cfg.lvl_dt >= 0 ? cfg.dt_min : ngp_clampf(t * cfg.dt_gamma, cfg.dt_min, cfg.dt_max);
            if (MODE == 1) out[marched] = make_float2(t, sdt);
            if (MODE == 2) {
                if (run_n == 0) {
                    if (nseg == seg_cap) {
                        overflow = true;
                        break;
                    }
                    run_t0 = t;
                }
                ++run_n;
            }
            t += sdt;
            ++marched;
        } else {
            if (MODE == 2 && run_n > 0) {
                out[nseg++] = make_float2(run_t0, __int_as_float((int)run_n));
                run_n = 0;
            }
            const float tt = ngp_march_exit_fast(r, cfg, mipb, t, nxyz, lvl);
            if (cfg.lvl_dt >= 0) {
                do {
                    t += cfg.dt_min;
                } while (t < tt);
            } else {
                do {
                    t = ngp_march_advance(t, cfg);
                } while (t < tt);
            }
        }
    }
    if (MODE == 0) counts[i] = (int32_t)marched;
    if (MODE == 2) {
        if (overflow) {
            counts[i] = -1;
            ovf_list[atomicAdd(ovf_count, 1ull)] = (int32_t)i;
        } else {
            if (run_n > 0) out[nseg++] = make_float2(run_t0, __int_as_float((int)run_n));
            counts[i] = nseg;
        }
    }
}

template <bool SEG>
__global__ void __launch_bounds__(R5_THREADS, 1)
k_render5_rays_infer_lists(R5Params P, const int32_t *__restrict__ offsets, float S_, uint32_t H, uint32_t gridtype,
                           const float *__restrict__ Ws0, const float *__restrict__ Ws1, const float *__restrict__ Wc0,
                           const float *__restrict__ Wc1, const float *__restrict__ Wc2) {
    extern __shared__ __align__(128) unsigned char r5_smem_raw[];
    R5SmemL &S = *reinterpret_cast<R5SmemL *>(r5_smem_raw);
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    fq_stage_levels(S.lv, &S.all_safe, offsets, S_, H, gridtype, P.align_corners);
    c5_stage_forward_weights(S.w, Ws0, Ws1, Wc0, Wc1, Wc2);
    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            tc5::mbar_init(tc5::smem_u32(&S.bars[s].f_full), 128);
            tc5::mbar_init(tc5::smem_u32(&S.bars[s].f_empty), 1);
            tc5::mbar_init(tc5::smem_u32(&S.bars[s].d_full), 1);
            tc5::mbar_init(tc5::smem_u32(&S.bars[s].a_ready), 128);
            tc5::mbar_init(tc5::smem_u32(&S.bars[s].t_free), 128);
            tc5::mbar_init(tc5::smem_u32(&S.bars[s].g_full), 1);
        }
        for (int b = 0; b < R5SmemL::TBUF; ++b) {
            tc5::mbar_init(tc5::smem_u32(&S.tb.rows_full[b]), R5_MARCH_WARPS * 32);
            tc5::mbar_init(tc5::smem_u32(&S.tb.rows_free[b]), 128);
        }
        tc5::fence_mbar_init();
    }
    if (warp == C5_MMA_WARP) tc5::tmem_alloc(tc5::smem_u32(&S.tmem), 512);
    tc5::fence_async_smem();
    tc5::fence_before_sync();
    __syncthreads();
    tc5::fence_after_sync();
    const uint32_t tmem = __shfl_sync(0xffffffffu, S.tmem, 0);
    if (warp < C5_GATHER_WARPS) {
        if (S.all_safe) r5_gather<true, R5SmemL>(S, P, tmem, warp, lane);
        else r5_gather<false, R5SmemL>(S, P, tmem, warp, lane);
    } else if (warp < C5_MMA_WARP) {
        r5_chain<R5SmemL>(S, P, tmem, warp, lane);
    } else if (warp < R5_MARCH_WARP0) {
        if (lane == 0) r5_mma<R5SmemL>(S, P, tmem, warp - C5_MMA_WARP);
        __syncwarp();
    } else {
        r5l_feeder<SEG>(S, P, warp - R5_MARCH_WARP0, lane);
    }
    tc5::fence_before_sync();
    __syncthreads();
    if (warp == C5_MMA_WARP) tc5::tmem_dealloc(tmem, 512);
}

// ---- entry points of the pre-marched renderer (declared in include/ngp_b200.h)
struct R5ToI64 {
    __host__ __device__ long long operator()(const int32_t &v) const { return (long long)v; }
};

extern "C" int64_t ngp_render_plan_workspace_bytes(int64_t N) {
    size_t need = 0;
    cub::TransformInputIterator<long long, R5ToI64, const int32_t *> it(nullptr, R5ToI64());
    cub::DeviceScan::InclusiveSum(nullptr, need, it, (long long *)nullptr, (int)(N < 1 ? 1 : N));
    return (int64_t)need + 256;
}

extern "C" int ngp_render_plan(const float *rays_o, const float *rays_d, int64_t N, const float *aabb, float min_near,
                               const uint8_t *grid, float bound, float dt_gamma, uint32_t max_steps, uint32_t C, uint32_t Hgrid,
                               int32_t *counts, int64_t *ends, void *workspace, int64_t workspace_bytes, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(rays_o, "ngp_render_plan");
    if (N == 0) return NGP_OK;
    NGP_REQUIRE(rays_o && rays_d && aabb && grid && counts && ends && workspace, NGP_ERR_ARG, "ngp_render_plan: null pointer");
    NGP_REQUIRE(bound > 0.0f && N > 0 && N < (1ll << 31) && max_steps < (1u << 20), NGP_ERR_ARG, "ngp_render_plan: bad bound / N / max_steps");
    NGP_REQUIRE(C >= 1 && C <= 8 && Hgrid >= 1 && Hgrid <= 1024, NGP_ERR_ARG, "ngp_render_plan: bad cascade/grid size");
    NGP_REQUIRE(workspace_bytes >= ngp_render_plan_workspace_bytes(N), NGP_ERR_ARG, "ngp_render_plan: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    const MarchCfg cfg = ngp_march_cfg(grid, bound, dt_gamma, max_steps, C, Hgrid);
    k_r5_premarch<0><<<(unsigned)ngp_div_up(N, 256), 256, 0, st>>>(rays_o, rays_d, N, aabb, min_near, cfg, max_steps, counts, nullptr,
                                                                  nullptr, 0, nullptr, nullptr);
    NGP_CHECK_LAUNCH("ngp_render_plan(count)");
    size_t need = (size_t)workspace_bytes;
    cub::TransformInputIterator<long long, R5ToI64, const int32_t *> it(counts, R5ToI64());
    cudaError_t e = cub::DeviceScan::InclusiveSum(workspace, need, it, (long long *)ends, (int)N, st);
    NGP_REQUIRE(e == cudaSuccess, NGP_ERR_CUDA, "ngp_render_plan: scan: %s", cudaGetErrorString(e));
    return NGP_OK;
}

extern "C" int ngp_render_rays_infer_planned(const float *rays_o, const float *rays_d, int64_t N, const float *aabb, float min_near,
                                             const uint8_t *grid, float bound, float dt_gamma, uint32_t max_steps, uint32_t C,
                                             uint32_t Hgrid, const float *embeddings, const int32_t *offsets, float S, uint32_t H,
                                             uint32_t gridtype, int align_corners, const float *Ws0, const float *Ws1,
                                             const float *Wc0, const float *Wc1, const float *Wc2, float density_scale,
                                             const float *bg_color, int bg_stride, float *image, float *depth, float *weights_sum,
                                             int32_t *counts, const int32_t *plan_counts, const int64_t *plan_ends, float *samples,
                                             int64_t samples_capacity, void *workspace, ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(rays_o, "ngp_render_rays_infer_planned");
    if (N == 0) return NGP_OK;
    NGP_REQUIRE(rays_o && rays_d && aabb && grid && embeddings && offsets && Ws0 && Ws1 && Wc0 && Wc1 && Wc2 && image && depth &&
                    workspace && plan_counts && plan_ends && (samples || samples_capacity == 0),
                NGP_ERR_ARG, "ngp_render_rays_infer_planned: null pointer");
    NGP_REQUIRE(bound > 0.0f && N > 0 && N < (1ll << 31) && max_steps < (1u << 20), NGP_ERR_ARG,
                "ngp_render_rays_infer_planned: bad bound / N / max_steps");
    NGP_REQUIRE(C >= 1 && C <= 8 && Hgrid >= 1 && Hgrid <= 1024, NGP_ERR_ARG, "ngp_render_rays_infer_planned: bad cascade/grid size");
    NGP_REQUIRE(bg_stride == 0 || bg_stride == 3, NGP_ERR_ARG, "ngp_render_rays_infer_planned: bg_stride must be 0 or 3");
    NGP_REQUIRE(((uintptr_t)embeddings & 15u) == 0 && ((uintptr_t)workspace & 7u) == 0 && ((uintptr_t)samples & 7u) == 0, NGP_ERR_ARG,
                "ngp_render_rays_infer_planned: embeddings must be 16-byte, workspace and samples 8-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(workspace, 0, 16 + 24 * 8, st) != cudaSuccess) {
        ngp_set_error("ngp_render_rays_infer_planned: workspace memset failed");
        return NGP_ERR_CUDA;
    }
    (void)samples_capacity;   // the caller sized `samples` from plan_ends[N - 1] (ngp_render_plan); the lists cannot exceed it
    const MarchCfg cfg = ngp_march_cfg(grid, bound, dt_gamma, max_steps, C, Hgrid);
    k_r5_premarch<1><<<(unsigned)ngp_div_up(N, 256), 256, 0, st>>>(rays_o, rays_d, N, aabb, min_near, cfg, max_steps,
                                                                  const_cast<int32_t *>(plan_counts), (const long long *)plan_ends,
                                                                  (float2 *)samples, 0, nullptr, nullptr);
    NGP_CHECK_LAUNCH("ngp_render_rays_infer_planned(write)");
    const int smem = (int)sizeof(R5SmemL);
    if (cudaFuncSetAttribute(k_render5_rays_infer_lists<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
        ngp_set_error("ngp_render_rays_infer_planned: cannot reserve %d B of shared memory", smem);
        return NGP_ERR_CUDA;
    }
    int dev = 0, sms = NGP_SM_COUNT;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    R5Params P;
    P.rays_o = rays_o; P.rays_d = rays_d; P.aabb = aabb; P.bg = bg_color; P.N = N; P.bg_stride = bg_stride;
    P.min_near = min_near; P.density_scale = density_scale; P.bound = bound; P.max_steps = max_steps;
    P.table = embeddings; P.align_corners = align_corners;
    P.image = image; P.depth = depth; P.weights_sum = weights_sum; P.counts = counts;
    P.counter = (unsigned long long *)workspace;
    P.nsamples = (unsigned long long *)workspace + 1;
    P.plan_counts = plan_counts; P.plan_ends = (const long long *)plan_ends; P.tdt = (const float2 *)samples;
    P.segs = nullptr; P.nsegs = nullptr; P.seg_cap = 0; P.dt_min = cfg.dt_min; P.dt_max = cfg.dt_max; P.dt_gamma = cfg.dt_gamma;
    P.lvl_dt = cfg.lvl_dt; P.ray_list = nullptr; P.ray_list_count = nullptr;
    const int64_t want = ngp_div_up(N, R5L_NR / 2);
    const int nblocks = (int)(want < sms ? want : sms);
    k_render5_rays_infer_lists<false><<<nblocks, R5_THREADS, smem, st>>>(P, offsets, S, H, gridtype, Ws0, Ws1, Wc0, Wc1, Wc2);
    NGP_CHECK_LAUNCH("ngp_render_rays_infer_planned");
    return NGP_OK;
}

// ---- single-call form: runs instead of lists, no host read
// Layout of the workspace: 512-byte header (u64[0] ray counter of the segment kernel, [1] samples evaluated, [2..25] stats,
// [26] rays with more than seg_cap runs, [32] ray counter of the marcher-inside kernel, [34..57] its stats), nsegs i32 [N],
// overflow ray list i32 [N], segs float2 [N x seg_cap].
static int64_t r5_align8(int64_t v) { return (v + 7) & ~7ll; }
extern "C" int64_t ngp_render_rays_infer_segments_workspace_bytes(int64_t N, int seg_cap) {
    if (N < 0 || seg_cap < 1) return -1;
    return 512 + r5_align8(4 * N) + r5_align8(4 * N) + N * (int64_t)seg_cap * 8;
}

extern "C" int ngp_render_rays_infer_segments(const float *rays_o, const float *rays_d, int64_t N, const float *aabb, float min_near,
                                              const uint8_t *grid, float bound, float dt_gamma, uint32_t max_steps, uint32_t C,
                                              uint32_t Hgrid, const float *embeddings, const int32_t *offsets, float S, uint32_t H,
                                              uint32_t gridtype, int align_corners, const float *Ws0, const float *Ws1,
                                              const float *Wc0, const float *Wc1, const float *Wc2, float density_scale,
                                              const float *bg_color, int bg_stride, float *image, float *depth, float *weights_sum,
                                              int32_t *counts, int seg_cap, void *workspace, int64_t workspace_bytes,
                                              ngp_stream_t stream) {
    NGP_REQUIRE_DEVICE(rays_o, "ngp_render_rays_infer_segments");
    if (N == 0) return NGP_OK;
    NGP_REQUIRE(rays_o && rays_d && aabb && grid && embeddings && offsets && Ws0 && Ws1 && Wc0 && Wc1 && Wc2 && image && depth &&
                    workspace,
                NGP_ERR_ARG, "ngp_render_rays_infer_segments: null pointer");
    NGP_REQUIRE(bound > 0.0f && N > 0 && N < (1ll << 31) && max_steps < (1u << 20), NGP_ERR_ARG,
                "ngp_render_rays_infer_segments: bad bound / N / max_steps");
    NGP_REQUIRE(C >= 1 && C <= 8 && Hgrid >= 1 && Hgrid <= 1024, NGP_ERR_ARG, "ngp_render_rays_infer_segments: bad cascade/grid size");
    NGP_REQUIRE(bg_stride == 0 || bg_stride == 3, NGP_ERR_ARG, "ngp_render_rays_infer_segments: bg_stride must be 0 or 3");
    NGP_REQUIRE(seg_cap >= 1 && seg_cap <= 1024, NGP_ERR_ARG, "ngp_render_rays_infer_segments: seg_cap must be in 1..1024");
    NGP_REQUIRE(((uintptr_t)embeddings & 15u) == 0 && ((uintptr_t)workspace & 7u) == 0, NGP_ERR_ARG,
                "ngp_render_rays_infer_segments: embeddings must be 16-byte aligned, workspace 8-byte aligned");
    NGP_REQUIRE(workspace_bytes >= ngp_render_rays_infer_segments_workspace_bytes(N, seg_cap), NGP_ERR_ARG,
                "ngp_render_rays_infer_segments: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(workspace, 0, 512, st) != cudaSuccess) {
        ngp_set_error("ngp_render_rays_infer_segments: workspace memset failed");
        return NGP_ERR_CUDA;
    }
    unsigned long long *hdr = (unsigned long long *)workspace;
    int32_t *nsegs = (int32_t *)((char *)workspace + 512);
    int32_t *ovf = (int32_t *)((char *)nsegs + r5_align8(4 * N));
    float2 *segs = (float2 *)((char *)ovf + r5_align8(4 * N));
    const MarchCfg cfg = ngp_march_cfg(grid, bound, dt_gamma, max_steps, C, Hgrid);
    k_r5_premarch<2><<<(unsigned)ngp_div_up(N, 256), 256, 0, st>>>(rays_o, rays_d, N, aabb, min_near, cfg, max_steps, nsegs, nullptr,
                                                                  segs, seg_cap, ovf, hdr + 26);
    NGP_CHECK_LAUNCH("ngp_render_rays_infer_segments(pre-pass)");
    int dev = 0, sms = NGP_SM_COUNT;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    R5Params P;
    P.rays_o = rays_o; P.rays_d = rays_d; P.aabb = aabb; P.bg = bg_color; P.N = N; P.bg_stride = bg_stride;
    P.min_near = min_near; P.density_scale = density_scale; P.bound = bound; P.max_steps = max_steps;
    P.table = embeddings; P.align_corners = align_corners;
    P.image = image; P.depth = depth; P.weights_sum = weights_sum; P.counts = counts;
    P.counter = hdr; P.nsamples = hdr + 1;
    P.plan_counts = nullptr; P.plan_ends = nullptr; P.tdt = nullptr;
    P.segs = segs; P.nsegs = nsegs; P.seg_cap = seg_cap; P.dt_min = cfg.dt_min; P.dt_max = cfg.dt_max; P.dt_gamma = cfg.dt_gamma;
    P.lvl_dt = cfg.lvl_dt; P.ray_list = nullptr; P.ray_list_count = nullptr;
    {
        const int smem = (int)sizeof(R5SmemL);
        if (cudaFuncSetAttribute(k_render5_rays_infer_lists<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
            ngp_set_error("ngp_render_rays_infer_segments: cannot reserve %d B of shared memory", smem);
            return NGP_ERR_CUDA;
        }
        const int64_t want = ngp_div_up(N, R5L_NR / 2);
        k_render5_rays_infer_lists<true><<<(int)(want < sms ? want : sms), R5_THREADS, smem, st>>>(P, offsets, S, H, gridtype, Ws0, Ws1,
                                                                                                   Wc0, Wc1, Wc2);
        NGP_CHECK_LAUNCH("ngp_render_rays_infer_segments");
    }
    {   // rays with more than seg_cap runs (their number is only known on the device): the kernel with the marcher inside,
        // over the overflow list; CTAs that find no ray leave through the end protocol at once
        const int smem = (int)sizeof(R5Smem);
        if (cudaFuncSetAttribute(k_render5_rays_infer, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
            ngp_set_error("ngp_render_rays_infer_segments: cannot reserve %d B of shared memory", smem);
            return NGP_ERR_CUDA;
        }
        P.counter = hdr + 32;
        P.segs = nullptr; P.nsegs = nullptr;
        P.ray_list = ovf; P.ray_list_count = hdr + 26;
        const int64_t want = ngp_div_up(N, R5_NR / 2);
        k_render5_rays_infer<<<(int)(want < sms ? want : sms), R5_THREADS, smem, st>>>(P, cfg, offsets, S, H, gridtype, Ws0, Ws1, Wc0,
                                                                                      Wc1, Wc2);
        NGP_CHECK_LAUNCH("ngp_render_rays_infer_segments(overflow rays)");
    }
    return NGP_OK;
}

// ------------------------------------------------------------------------------------------------ launcher (render.cu)
int r5_render_rays_infer(const float *rays_o, const float *rays_d, int64_t N, const float *aabb, float min_near,
                         const uint8_t *grid, float bound, float dt_gamma, uint32_t max_steps, uint32_t C, uint32_t Hgrid,
                         const float *embeddings, const int32_t *offsets, float S, uint32_t H, uint32_t gridtype,
                         int align_corners, const float *Ws0, const float *Ws1, const float *Wc0, const float *Wc1,
                         const float *Wc2, float density_scale, const float *bg_color, int bg_stride, float *image,
                         float *depth, float *weights_sum, int32_t *counts, void *workspace, cudaStream_t stream) {
    NGP_REQUIRE(N < (1ll << 31) && max_steps < (1u << 20), NGP_ERR_UNSUPPORTED,
                "ngp_render_rays_infer (tcgen05): N must be < 2^31 and max_steps < 2^20");
    const int smem = (int)sizeof(R5Smem);
    if (cudaFuncSetAttribute(k_render5_rays_infer, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
        ngp_set_error("ngp_render_rays_infer: cannot reserve %d B of shared memory", smem);
        return NGP_ERR_CUDA;
    }
    int dev = 0, sms = NGP_SM_COUNT;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    R5Params P;
    P.rays_o = rays_o; P.rays_d = rays_d; P.aabb = aabb; P.bg = bg_color; P.N = N; P.bg_stride = bg_stride;
    P.min_near = min_near; P.density_scale = density_scale; P.bound = bound; P.max_steps = max_steps;
    P.table = embeddings; P.align_corners = align_corners;
    P.image = image; P.depth = depth; P.weights_sum = weights_sum; P.counts = counts;
    P.counter = (unsigned long long *)workspace;
    P.nsamples = (unsigned long long *)workspace + 1;
    P.plan_counts = nullptr; P.plan_ends = nullptr; P.tdt = nullptr;
    const MarchCfg cfg = ngp_march_cfg(grid, bound, dt_gamma, max_steps, C, Hgrid);
    P.segs = nullptr; P.nsegs = nullptr; P.seg_cap = 0; P.dt_min = cfg.dt_min; P.dt_max = cfg.dt_max; P.dt_gamma = cfg.dt_gamma;
    P.lvl_dt = cfg.lvl_dt; P.ray_list = nullptr; P.ray_list_count = nullptr;
    // one CTA per SM (a CTA owns all 512 TMEM columns); small batches use fewer CTAs so that tiles still fill
    const int64_t want = ngp_div_up(N, R5_NR / 2);
    const int nblocks = (int)(want < sms ? want : sms);
    k_render5_rays_infer<<<nblocks, R5_THREADS, smem, stream>>>(P, cfg, offsets, S, H, gridtype, Ws0, Ws1, Wc0, Wc1, Wc2);
    NGP_CHECK_LAUNCH("ngp_render_rays_infer (tcgen05)");
    return NGP_OK;
}
