/*
 * ngp_oracle.c — TEST INFRASTRUCTURE ONLY.  Parity status: see oracle/README.md.
 *
 * Plain-C CPU restatement of the reference's neural-field hot path (stein7/nerf-navigation,
 * a torch-ngp fork).  It is the checker for the CUDA product and the timed "port" CPU baseline.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (nerf_navigation_b200/) never does.
 *
 * Each function cites the reference file:line whose behaviour it restates.  The code is written
 * from the algorithm description, with loops over points/rays instead of CUDA threads, explicit
 * fmaf() where nvcc's default -fmad=true contracts the reference expression, and everything else
 * compiled with -ffp-contract=off so no other fusion happens.
 *
 * Build: see oracle/Makefile  (gcc -O2 -fopenmp -ffp-contract=off -shared -fPIC).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_MAX_D 3
#define ORC_MAX_C 8

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void orc_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* ------------------------------------------------------------------------------------------
 * Hash-grid encoder.  Reference: gridencoder/src/gridencoder.cu.
 * ---------------------------------------------------------------------------------------- */

/* Per-level scale, gridencoder.cu:126  (scale = exp2f(level*S)*H - 1).  Computed once and passed
 * to every other entry so CPU libm / GPU libdevice exp2f differences cannot flip cell indices:
 * tests hand the SAME scale table to the oracle and to the CUDA path. */
void orc_grid_level_scales(int L, float S, uint32_t H, float *scales) {
    for (int l = 0; l < L; ++l) scales[l] = fmaf(exp2f((float)l * S), (float)H, -1.0f);
}

/* Spatial hash, gridencoder.cu:35-50: xor of coordinate*prime with primes {1, 2654435761, 805459861}. */
static inline uint32_t orc_hash(const uint32_t *pg, int D) {
    static const uint32_t primes[ORC_MAX_D] = {1u, 2654435761u, 805459861u};
    uint32_t h = 0;
    for (int d = 0; d < D; ++d) h ^= pg[d] * primes[d];
    return h;
}

/* Row index of one grid corner, gridencoder.cu:53-72.  Dense stride walk while the running stride
 * still fits in the level; hash only if gridtype==0 and the dense extent overflowed the level. */
static inline uint32_t orc_row(uint32_t gridtype, int align_corners, uint32_t rows, uint32_t resolution,
                               const uint32_t *pg, int D) {
    uint32_t stride = 1, index = 0;
    for (int d = 0; d < D && stride <= rows; ++d) {
        index += pg[d] * stride;
        stride *= align_corners ? resolution : (resolution + 1);
    }
    if (gridtype == 0 && stride > rows) index = orc_hash(pg, D);
    return index % rows;
}

typedef struct {
    float frac[ORC_MAX_D];
    uint32_t cell[ORC_MAX_D];
    uint32_t rows, resolution;
    float scale;
    int oob;
} orc_cell_t;

/* Cell location of one point at one level, gridencoder.cu:99-138. */
static inline void orc_locate(const float *x, int D, const int32_t *offsets, const float *scales, int level,
                              int align_corners, orc_cell_t *c) {
    c->oob = 0;
    for (int d = 0; d < D; ++d)
        if (x[d] < 0.0f || x[d] > 1.0f) c->oob = 1;
    c->rows = (uint32_t)(offsets[level + 1] - offsets[level]);
    c->scale = scales[level];
    c->resolution = (uint32_t)ceilf(c->scale) + 1u;
    if (c->oob) return;
    for (int d = 0; d < D; ++d) {
        float p = fmaf(x[d], c->scale, align_corners ? 0.0f : 0.5f);
        float fl = floorf(p);
        c->cell[d] = (uint32_t)fl;
        c->frac[d] = p - (float)c->cell[d];
    }
}

/*
 * Forward, gridencoder.cu:75-224.
 *   x01     [B,D]  inputs already mapped to [0,1]            (grid.py:144)
 *   table   [rows_total, C] fp32
 *   out     [L,B,C]   (reference layout, gridencoder.cu:96; Python permutes to [B,L*C], grid.py:52)
 *   dy_dx   [B,L,D,C] or NULL (gridencoder.cu:113,180-223)
 *   indices [B,L,2^D] u32 row indices (level-local, before *C) or NULL — debug output for the bit-exact test
 */
static inline void orc_encode_point_level(const float *x, const float *table, const int32_t *offsets,
                                          const float *scales, int l, int D, int C, uint32_t gridtype,
                                          int align_corners, float *o, float *g, uint32_t *ix) {
    const int NC = 1 << D;
    orc_cell_t c;
    orc_locate(x, D, offsets, scales, l, align_corners, &c);
    if (c.oob) {
        for (int ch = 0; ch < C; ++ch) o[ch] = 0.0f;
        if (g) for (int k = 0; k < D * C; ++k) g[k] = 0.0f;
        if (ix) for (int k = 0; k < NC; ++k) ix[k] = 0xFFFFFFFFu;
        return;
    }
    const float *lvl = table + (int64_t)(uint32_t)offsets[l] * C;
    float acc[ORC_MAX_C] = {0};
    for (int corner = 0; corner < NC; ++corner) {
        float w = 1.0f;
        uint32_t pg[ORC_MAX_D];
        for (int d = 0; d < D; ++d) {
            if (corner & (1 << d)) { w *= c.frac[d]; pg[d] = c.cell[d] + 1; }
            else { w *= 1.0f - c.frac[d]; pg[d] = c.cell[d]; }
        }
        uint32_t row = orc_row(gridtype, align_corners, c.rows, c.resolution, pg, D);
        if (ix) ix[corner] = row;
        for (int ch = 0; ch < C; ++ch) acc[ch] = fmaf(w, lvl[(int64_t)row * C + ch], acc[ch]);
    }
    for (int ch = 0; ch < C; ++ch) o[ch] = acc[ch];
    if (!g) return;
    for (int gd = 0; gd < D; ++gd) {
        float ga[ORC_MAX_C] = {0};
        for (int k = 0; k < (1 << (D - 1)); ++k) {
            float w = c.scale;
            uint32_t pg[ORC_MAX_D];
            for (int nd = 0; nd < D - 1; ++nd) {
                int d = nd >= gd ? nd + 1 : nd;
                if (k & (1 << nd)) { w *= c.frac[d]; pg[d] = c.cell[d] + 1; }
                else { w *= 1.0f - c.frac[d]; pg[d] = c.cell[d]; }
            }
            pg[gd] = c.cell[gd];
            uint32_t left = orc_row(gridtype, align_corners, c.rows, c.resolution, pg, D);
            pg[gd] = c.cell[gd] + 1;
            uint32_t right = orc_row(gridtype, align_corners, c.rows, c.resolution, pg, D);
            for (int ch = 0; ch < C; ++ch)
                ga[ch] = fmaf(w, lvl[(int64_t)right * C + ch] - lvl[(int64_t)left * C + ch], ga[ch]);
        }
        for (int ch = 0; ch < C; ++ch) g[gd * C + ch] = ga[ch];
    }
}

void orc_grid_encode_forward(const float *x01, const float *table, const int32_t *offsets, const float *scales,
                             int64_t B, int D, int C, int L, uint32_t gridtype, int align_corners,
                             float *out, float *dy_dx, uint32_t *indices) {
    const int NC = 1 << D;
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < B; ++b)
        for (int l = 0; l < L; ++l)
            orc_encode_point_level(x01 + b * D, table, offsets, scales, l, D, C, gridtype, align_corners,
                                   out + ((int64_t)l * B + b) * C,
                                   dy_dx ? dy_dx + ((b * L + l) * D) * C : NULL,
                                   indices ? indices + (b * L + l) * NC : NULL);
}

/*
 * Table gradient, gridencoder.cu:227-314.  The reference scatters w*grad with float atomics in
 * arbitrary order; here contributions are accumulated in float64 in point order and rounded once,
 * i.e. the order-independent value every fp32 summation order approximates.
 *   grad [L,B,C] ; grad_table [rows_total, C] (overwritten)
 */
void orc_grid_backward_table(const float *grad, const float *x01, const int32_t *offsets, const float *scales,
                             int64_t B, int D, int C, int L, uint32_t gridtype, int align_corners,
                             float *grad_table) {
    const int NC = 1 << D;
    const int64_t total = (int64_t)offsets[L] * C;
    double *acc = (double *)calloc((size_t)total, sizeof(double));
#pragma omp parallel for schedule(static)
    for (int l = 0; l < L; ++l) { /* levels own disjoint table ranges: race-free */
        double *lvl = acc + (int64_t)(uint32_t)offsets[l] * C;
        for (int64_t b = 0; b < B; ++b) {
            orc_cell_t c;
            orc_locate(x01 + b * D, D, offsets, scales, l, align_corners, &c);
            if (c.oob) continue; /* gridencoder.cu:254-259 */
            const float *g = grad + ((int64_t)l * B + b) * C;
            for (int corner = 0; corner < NC; ++corner) {
                float w = 1.0f;
                uint32_t pg[ORC_MAX_D];
                for (int d = 0; d < D; ++d) {
                    if (corner & (1 << d)) { w *= c.frac[d]; pg[d] = c.cell[d] + 1; }
                    else { w *= 1.0f - c.frac[d]; pg[d] = c.cell[d]; }
                }
                uint32_t row = orc_row(gridtype, align_corners, c.rows, c.resolution, pg, D);
                for (int ch = 0; ch < C; ++ch) lvl[(int64_t)row * C + ch] += (double)(w * g[ch]);
            }
        }
    }
    for (int64_t i = 0; i < total; ++i) grad_table[i] = (float)acc[i];
    free(acc);
}

/* Input gradient, gridencoder.cu:317-343:  dL/dx01[b,d] = sum_{l,c} grad[l,b,c] * dy_dx[b,l,d,c]. */
void orc_grid_backward_input(const float *grad, const float *dy_dx, int64_t B, int D, int C, int L,
                             float *grad_x01) {
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < B; ++b) {
        for (int d = 0; d < D; ++d) {
            float r = 0.0f;
            for (int l = 0; l < L; ++l)
                for (int ch = 0; ch < C; ++ch)
                    r = fmaf(grad[((int64_t)l * B + b) * C + ch], dy_dx[((b * L + l) * D + d) * C + ch], r);
            grad_x01[b * D + d] = r;
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * Field query: GridEncoder.forward -> sigma MLP -> trunc_exp, and its backward to x.
 * Reference: nerf/network.py:126-143 (density), gridencoder/grid.py:140-156, activation.py:5-17.
 * sigma net (torch variant, network.py:33-47): h = relu(feat @ W0^T), o = h @ W1^T, no bias,
 * W0 [hidden, L*C], W1 [n_out, hidden] row-major like nn.Linear.weight.
 * ---------------------------------------------------------------------------------------- */

#define ORC_MAX_FEAT 64
#define ORC_MAX_HID 128
#define ORC_MAX_OUT 16

/*
 * x [B,3] world coords in [-bound,bound].  Outputs: sigma [B], geo [B,n_out-1] (may be NULL).
 * If grad_sigma != NULL also runs the backward: upstream dL/dsigma [B] (+ optional dL/dgeo [B,n_out-1])
 * -> grad_x [B,3] (world coordinates, i.e. including the 1/(2*bound) of grid.py:144).
 * inv2b = 1/(2*bound) as float (torch CUDA divides by a scalar as multiply-by-reciprocal).
 */
void orc_field_query(const float *x, float bound, const float *table, const int32_t *offsets, const float *scales,
                     int64_t B, int C, int L, uint32_t gridtype, int align_corners,
                     const float *W0, const float *W1, int hidden, int n_out,
                     float *sigma, float *geo,
                     const float *grad_sigma, const float *grad_geo, float *grad_x) {
    const int D = 3, F = L * C;
    const float inv2b = 1.0f / (2.0f * bound);
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < B; ++b) {
        float x01[3], feat[ORC_MAX_FEAT], jac[ORC_MAX_FEAT * 3], h[ORC_MAX_HID], o[ORC_MAX_OUT];
        for (int d = 0; d < D; ++d) x01[d] = (x[b * 3 + d] + bound) * inv2b;
        float lo[ORC_MAX_C * 1], lj[3 * ORC_MAX_C];
        for (int l = 0; l < L; ++l) {
            orc_encode_point_level(x01, table, offsets, scales, l, D, C, gridtype, align_corners,
                                   lo, grad_sigma ? lj : NULL, NULL);
            for (int ch = 0; ch < C; ++ch) feat[l * C + ch] = lo[ch];
            if (grad_sigma)
                for (int d = 0; d < D; ++d)
                    for (int ch = 0; ch < C; ++ch) jac[(l * C + ch) * 3 + d] = lj[d * C + ch];
        }
        for (int j = 0; j < hidden; ++j) {
            float a = 0.0f;
            for (int k = 0; k < F; ++k) a = fmaf(feat[k], W0[j * F + k], a);
            h[j] = a > 0.0f ? a : 0.0f;
        }
        for (int i = 0; i < n_out; ++i) {
            float a = 0.0f;
            for (int j = 0; j < hidden; ++j) a = fmaf(h[j], W1[i * hidden + j], a);
            o[i] = a;
        }
        sigma[b] = expf(o[0]);
        if (geo) for (int i = 1; i < n_out; ++i) geo[b * (n_out - 1) + i - 1] = o[i];
        if (!grad_sigma) continue;
        /* backward: activation.py:13-17 -> g*exp(clamp(o0,-15,15)); Linear^T; relu mask; encoder input-backward */
        float go[ORC_MAX_OUT], gh[ORC_MAX_HID], gf[ORC_MAX_FEAT];
        float oc = o[0] < -15.0f ? -15.0f : (o[0] > 15.0f ? 15.0f : o[0]);
        go[0] = grad_sigma[b] * expf(oc);
        for (int i = 1; i < n_out; ++i) go[i] = grad_geo ? grad_geo[b * (n_out - 1) + i - 1] : 0.0f;
        for (int j = 0; j < hidden; ++j) {
            float a = 0.0f;
            for (int i = 0; i < n_out; ++i) a = fmaf(go[i], W1[i * hidden + j], a);
            gh[j] = h[j] > 0.0f ? a : 0.0f;
        }
        for (int k = 0; k < F; ++k) {
            float a = 0.0f;
            for (int j = 0; j < hidden; ++j) a = fmaf(gh[j], W0[j * F + k], a);
            gf[k] = a;
        }
        for (int d = 0; d < D; ++d) {
            float a = 0.0f;
            for (int k = 0; k < F; ++k) a = fmaf(gf[k], jac[k * 3 + d], a);
            grad_x[b * 3 + d] = a * inv2b;
        }
    }
}

/* Generic bias-free MLP forward, fp32 (nn.Linear stack with ReLU between layers, network.py:104-122).
 * dims[n_layers+1]; weights concatenated row-major [out,in] per layer.  out_act: 0 none, 1 sigmoid. */
void orc_mlp_forward(const float *x, const float *weights, const int *dims, int n_layers, int out_act,
                     int64_t B, float *y) {
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < B; ++b) {
        float cur[ORC_MAX_HID], nxt[ORC_MAX_HID];
        for (int k = 0; k < dims[0]; ++k) cur[k] = x[b * dims[0] + k];
        const float *W = weights;
        for (int l = 0; l < n_layers; ++l) {
            int in = dims[l], out = dims[l + 1];
            for (int j = 0; j < out; ++j) {
                float a = 0.0f;
                for (int k = 0; k < in; ++k) a = fmaf(cur[k], W[j * in + k], a);
                if (l != n_layers - 1) a = a > 0.0f ? a : 0.0f;
                else if (out_act == 1) a = 1.0f / (1.0f + expf(-a));
                nxt[j] = a;
            }
            memcpy(cur, nxt, sizeof(float) * out);
            W += (int64_t)in * out;
        }
        for (int j = 0; j < dims[n_layers]; ++j) y[b * dims[n_layers] + j] = cur[j];
    }
}

/* ------------------------------------------------------------------------------------------
 * Spherical harmonics, degree <= 4.  Reference: shencoder/src/shencoder.cu:44-70 (values),
 * :131-150, :205-224, :279-298 (Jacobian).  Standard real-SH polynomials in (x,y,z).
 * out [B,deg^2]; dy_dx [B,3,deg^2] or NULL.
 * ---------------------------------------------------------------------------------------- */
void orc_sh_encode(const float *dirs, int64_t B, int degree, float *out, float *dy_dx) {
    const int n = degree * degree;
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < B; ++b) {
        const float x = dirs[b * 3], y = dirs[b * 3 + 1], z = dirs[b * 3 + 2];
        const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
        float v[16], gx[16] = {0}, gy[16] = {0}, gz[16] = {0};
        const float c1 = 0.48860251190291987f, c1d = 0.48860251190291992f, c2 = 1.0925484305920792f;
        v[0] = 0.28209479177387814f;
        v[1] = -c1 * y; v[2] = c1 * z; v[3] = -c1 * x;
        gy[1] = -c1d; gz[2] = c1d; gx[3] = -c1d;
        v[4] = c2 * xy; v[5] = -c2 * yz;
        v[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
        v[7] = -c2 * xz;
        v[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
        gx[4] = c2 * y; gy[4] = c2 * x;
        gy[5] = -c2 * z; gz[5] = -c2 * y;
        gz[6] = 1.8923493915151202f * z;
        gx[7] = -c2 * z; gz[7] = -c2 * x;
        gx[8] = c2 * x; gy[8] = -c2 * y;
        v[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2);
        v[10] = 2.8906114426405538f * xy * z;
        v[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2);
        v[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
        v[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2);
        v[14] = 1.4453057213202769f * z * (x2 - y2);
        v[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
        gx[9] = -3.5402615395598609f * xy;
        gy[9] = -1.7701307697799304f * x2 + 1.7701307697799304f * y2;
        gx[10] = 2.8906114426405538f * yz; gy[10] = 2.8906114426405538f * xz; gz[10] = 2.8906114426405538f * xy;
        gy[11] = 0.45704579946446572f - 2.2852289973223288f * z2; gz[11] = -4.5704579946446566f * yz;
        gz[12] = 5.597644988851731f * z2 - 1.1195289977703462f;
        gx[13] = 0.45704579946446572f - 2.2852289973223288f * z2; gz[13] = -4.5704579946446566f * xz;
        gx[14] = 2.8906114426405538f * xz; gy[14] = -2.8906114426405538f * yz;
        gz[14] = 1.4453057213202769f * x2 - 1.4453057213202769f * y2;
        gx[15] = -1.7701307697799304f * x2 + 1.7701307697799304f * y2; gy[15] = 3.5402615395598609f * xy;
        for (int i = 0; i < n; ++i) out[b * n + i] = v[i];
        if (dy_dx)
            for (int i = 0; i < n; ++i) {
                dy_dx[(b * 3 + 0) * n + i] = gx[i];
                dy_dx[(b * 3 + 1) * n + i] = gy[i];
                dy_dx[(b * 3 + 2) * n + i] = gz[i];
            }
    }
}

/* ------------------------------------------------------------------------------------------
 * Ray ops.  Reference: raymarching/src/raymarching.cu (+ pcg32.h for perturbation).
 * ---------------------------------------------------------------------------------------- */

static inline float orc_clampf(float v, float lo, float hi) { return fminf(hi, fmaxf(lo, v)); }

/* Ray / AABB slab test, raymarching.cu:93-147.  Miss => near = far = FLT_MAX. */
void orc_near_far_from_aabb(const float *rays_o, const float *rays_d, const float *aabb, int64_t N,
                            float min_near, float *nears, float *fars) {
#pragma omp parallel for schedule(static)
    for (int64_t n = 0; n < N; ++n) {
        const float *o = rays_o + n * 3, *d = rays_d + n * 3;
        float tn = 0.0f, tf = 0.0f;
        int miss = 0;
        for (int a = 0; a < 3 && !miss; ++a) {
            float r = 1.0f / d[a];
            float lo = (aabb[a] - o[a]) * r, hi = (aabb[a + 3] - o[a]) * r;
            if (lo > hi) { float t = lo; lo = hi; hi = t; }
            if (a == 0) { tn = lo; tf = hi; continue; }
            if (tn > hi || lo > tf) { miss = 1; break; }
            if (lo > tn) tn = lo;
            if (hi < tf) tf = hi;
        }
        if (miss) { nears[n] = fars[n] = FLT_MAX; continue; }
        if (tn < min_near) tn = min_near;
        nears[n] = tn;
        fars[n] = tf;
    }
}

/* 10-bit-per-axis Morton code, raymarching.cu:58-83 (magic-multiply bit spreading). */
static inline uint32_t orc_spread3(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
static inline uint32_t orc_morton(uint32_t x, uint32_t y, uint32_t z) {
    return orc_spread3(x) | (orc_spread3(y) << 1) | (orc_spread3(z) << 2);
}
static inline uint32_t orc_compact3(uint32_t x) {
    x &= 0x49249249u;
    x = (x | (x >> 2)) & 0xc30c30c3u;
    x = (x | (x >> 4)) & 0x0f00f00fu;
    x = (x | (x >> 8)) & 0xff0000ffu;
    x = (x | (x >> 16)) & 0x0000ffffu;
    return x;
}
/* raymarching.cu:216-262 */
void orc_morton3D(const int32_t *coords, int64_t N, int32_t *indices) {
    for (int64_t n = 0; n < N; ++n)
        indices[n] = (int32_t)orc_morton((uint32_t)coords[n * 3], (uint32_t)coords[n * 3 + 1], (uint32_t)coords[n * 3 + 2]);
}
void orc_morton3D_invert(const int32_t *indices, int64_t N, int32_t *coords) {
    for (int64_t n = 0; n < N; ++n) {
        uint32_t v = (uint32_t)indices[n];
        coords[n * 3] = (int32_t)orc_compact3(v);
        coords[n * 3 + 1] = (int32_t)orc_compact3(v >> 1);
        coords[n * 3 + 2] = (int32_t)orc_compact3(v >> 2);
    }
}

/* Occupancy bit packing, raymarching.cu:269-291: bit i of byte n = grid[8n+i] > thresh. */
void orc_packbits(const float *grid, int64_t N, float thresh, uint8_t *bitfield) {
#pragma omp parallel for schedule(static)
    for (int64_t n = 0; n < N; ++n) {
        uint8_t bits = 0;
        for (int i = 0; i < 8; ++i)
            if (grid[n * 8 + i] > thresh) bits |= (uint8_t)(1u << i);
        bitfield[n] = bits;
    }
}

/* PCG32 (O'Neill), as used by raymarching/src/pcg32.h:57-72 (seed), :107-116 (next_float), :149-170 (advance). */
typedef struct { uint64_t state, inc; } orc_pcg32;
#define ORC_PCG_MULT 0x5851f42d4c957f2dULL
static inline uint32_t orc_pcg_next(orc_pcg32 *r) {
    uint64_t old = r->state;
    r->state = old * ORC_PCG_MULT + r->inc;
    uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    uint32_t rot = (uint32_t)(old >> 59u);
    return (xs >> rot) | (xs << ((~rot + 1u) & 31));
}
static inline void orc_pcg_seed(orc_pcg32 *r, uint64_t initstate, uint64_t initseq) {
    r->state = 0;
    r->inc = (initseq << 1u) | 1u;
    orc_pcg_next(r);
    r->state += initstate;
    orc_pcg_next(r);
}
static inline void orc_pcg_advance(orc_pcg32 *r, uint64_t delta) {
    uint64_t cur_mult = ORC_PCG_MULT, cur_plus = r->inc, acc_mult = 1, acc_plus = 0;
    while (delta > 0) {
        if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
        cur_plus = (cur_mult + 1) * cur_plus;
        cur_mult *= cur_mult;
        delta >>= 1;
    }
    r->state = acc_mult * r->state + acc_plus;
}
static inline float orc_pcg_float(orc_pcg32 *r) {
    union { uint32_t u; float f; } x;
    x.u = (orc_pcg_next(r) >> 9) | 0x3f800000u;
    return x.f - 1.0f;
}
/* test hook: n-th ray's jitter for a given seed */
float orc_pcg32_jitter(uint64_t seed, uint64_t n) {
    orc_pcg32 r;
    orc_pcg_seed(&r, seed, 1);
    orc_pcg_advance(&r, n);
    return orc_pcg_float(&r);
}

typedef struct {
    float ox, oy, oz, dx, dy, dz, rdx, rdy, rdz;
    float bound, dt_gamma, dt_min, dt_max, rH;
    uint32_t C, H;
    const uint8_t *grid;
} orc_march_t;

static inline void orc_march_setup(orc_march_t *m, const float *o, const float *d, const uint8_t *grid, float bound,
                                   float dt_gamma, uint32_t max_steps, uint32_t C, uint32_t H) {
    m->ox = o[0]; m->oy = o[1]; m->oz = o[2];
    m->dx = d[0]; m->dy = d[1]; m->dz = d[2];
    m->rdx = 1.0f / d[0]; m->rdy = 1.0f / d[1]; m->rdz = 1.0f / d[2];
    m->bound = bound; m->dt_gamma = dt_gamma;
    m->dt_min = 2.0f * 1.7320508075688772f / (float)max_steps;               /* raymarching.cu:347 */
    m->dt_max = 2.0f * 1.7320508075688772f * (float)(1 << (C - 1)) / (float)H; /* :348 */
    m->rH = 1.0f / (float)H;
    m->C = C; m->H = H; m->grid = grid;
}

static inline int orc_frexp_level(float v, uint32_t C) { /* raymarching.cu:44-56 */
    int e;
    frexpf(v, &e);
    return (int)fminf((float)C - 1.0f, fmaxf(0.0f, (float)e));
}

/*
 * One marcher iteration at parameter t (raymarching.cu:361-403 == :430-482 == :748-790):
 * returns 1 if the cell is occupied (sample position in p[3], step in *dt, t unchanged),
 * else 0 after advancing *t past the empty voxel.
 */
static inline int orc_march_step(const orc_march_t *m, float *t, float *p, float *dt_out) {
    const float tt0 = *t;
    const float x = orc_clampf(fmaf(tt0, m->dx, m->ox), -m->bound, m->bound);
    const float y = orc_clampf(fmaf(tt0, m->dy, m->oy), -m->bound, m->bound);
    const float z = orc_clampf(fmaf(tt0, m->dz, m->oz), -m->bound, m->bound);
    const float dt = orc_clampf(tt0 * m->dt_gamma, m->dt_min, m->dt_max);
    const float mx = fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z)));
    int lp = orc_frexp_level(mx, m->C);
    int ld = orc_frexp_level(dt * (float)m->H * 0.5f, m->C);
    const int level = lp > ld ? lp : ld;
    const float mip_bound = fminf((float)(1 << level), m->bound);
    const float mip_rbound = 1.0f / mip_bound;
    /* fp64 product (the literal 0.5 is a double, :378-380), rounded to float by the clamp(float,...) call */
    const float Hm1 = (float)(m->H - 1);
    const int nx = (int)orc_clampf((float)(0.5 * (double)fmaf(x, mip_rbound, 1.0f) * (double)m->H), 0.0f, Hm1);
    const int ny = (int)orc_clampf((float)(0.5 * (double)fmaf(y, mip_rbound, 1.0f) * (double)m->H), 0.0f, Hm1);
    const int nz = (int)orc_clampf((float)(0.5 * (double)fmaf(z, mip_rbound, 1.0f) * (double)m->H), 0.0f, Hm1);
    const uint32_t index = (uint32_t)level * m->H * m->H * m->H + orc_morton((uint32_t)nx, (uint32_t)ny, (uint32_t)nz);
    const int occ = (m->grid[index >> 3] >> (index & 7)) & 1;
    if (occ) {
        p[0] = x; p[1] = y; p[2] = z;
        *dt_out = dt;
        return 1;
    }
    /* distance to the voxel's exit face, :393-402 */
    const float sx = copysignf(1.0f, m->dx), sy = copysignf(1.0f, m->dy), sz = copysignf(1.0f, m->dz);
    const float tx = (fmaf(fmaf(((float)nx + 0.5f + 0.5f * sx) * m->rH, 2.0f, -1.0f), mip_bound, -x)) * m->rdx;
    const float ty = (fmaf(fmaf(((float)ny + 0.5f + 0.5f * sy) * m->rH, 2.0f, -1.0f), mip_bound, -y)) * m->rdy;
    const float tz = (fmaf(fmaf(((float)nz + 0.5f + 0.5f * sz) * m->rH, 2.0f, -1.0f), mip_bound, -z)) * m->rdz;
    const float tt = tt0 + fmaxf(0.0f, fminf(tx, fminf(ty, tz)));
    float tc = tt0;
    do { tc += orc_clampf(tc * m->dt_gamma, m->dt_min, m->dt_max); } while (tc < tt);
    *t = tc;
    return 0;
}

/*
 * Training marcher, raymarching.cu:313-484.  Two passes per ray (count, then write).  The reference
 * reserves output slots with atomicAdd (:409-410) so its row order / offsets are nondeterministic;
 * this restatement uses the canonical order (ray id ascending, offsets = exclusive scan of counts),
 * which is what the reference produces up to that permutation.  rays[n] = (ray id, offset, count).
 * A ray whose offset+count >= M is listed but writes no samples (:420).  counter[0] += total samples,
 * counter[1] += N.  perturb: jitter t0 by dt_min * pcg32(seed 42).advance(n).next_float() (:352-355, :489).
 */
void orc_march_rays_train(const float *rays_o, const float *rays_d, const uint8_t *grid, float bound, float dt_gamma,
                          uint32_t max_steps, int64_t N, uint32_t C, uint32_t H, int64_t M,
                          const float *nears, const float *fars,
                          float *xyzs, float *dirs, float *deltas, int32_t *rays, int32_t *counter, uint32_t perturb) {
    int32_t *counts = (int32_t *)malloc(sizeof(int32_t) * (size_t)(N > 0 ? N : 1));
    float *t0s = (float *)malloc(sizeof(float) * (size_t)(N > 0 ? N : 1));
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t n = 0; n < N; ++n) {
        orc_march_t m;
        orc_march_setup(&m, rays_o + n * 3, rays_d + n * 3, grid, bound, dt_gamma, max_steps, C, H);
        float t = nears[n];
        if (perturb) {
            orc_pcg32 r;
            orc_pcg_seed(&r, 42, 1);
            orc_pcg_advance(&r, (uint64_t)n);
            t = fmaf(m.dt_min, orc_pcg_float(&r), t); /* nvcc contracts near + dt_min*u to one fma */
        }
        t0s[n] = t;
        const float far = fars[n];
        uint32_t steps = 0;
        float p[3], dt;
        while (t < far && steps < max_steps)
            if (orc_march_step(&m, &t, p, &dt)) { ++steps; t += dt; }
        counts[n] = (int32_t)steps;
    }
    int64_t base = counter[0];
    for (int64_t n = 0; n < N; ++n) {
        rays[n * 3] = (int32_t)n;
        rays[n * 3 + 1] = (int32_t)base;
        rays[n * 3 + 2] = counts[n];
        base += counts[n];
    }
    counter[0] = (int32_t)base;
    counter[1] += (int32_t)N;
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t n = 0; n < N; ++n) {
        const int64_t off = rays[n * 3 + 1];
        const uint32_t cnt = (uint32_t)counts[n];
        if (cnt == 0 || off + cnt >= M) continue;
        orc_march_t m;
        orc_march_setup(&m, rays_o + n * 3, rays_d + n * 3, grid, bound, dt_gamma, max_steps, C, H);
        float t = t0s[n], last_t = t0s[n];
        const float far = fars[n];
        uint32_t step = 0;
        float p[3], dt;
        while (t < far && step < cnt) {
            if (!orc_march_step(&m, &t, p, &dt)) continue;
            const int64_t k = off + step;
            xyzs[k * 3] = p[0]; xyzs[k * 3 + 1] = p[1]; xyzs[k * 3 + 2] = p[2];
            dirs[k * 3] = m.dx; dirs[k * 3 + 1] = m.dy; dirs[k * 3 + 2] = m.dz;
            t += dt;
            deltas[k * 2] = dt;
            deltas[k * 2 + 1] = t - last_t;
            last_t = t;
            ++step;
        }
    }
    free(counts);
    free(t0s);
}

/* Compositing forward, raymarching.cu:505-582.  expf here; the CUDA reference uses __expf (fast intrinsic). */
void orc_composite_rays_train_forward(const float *sigmas, const float *rgbs, const float *deltas, const int32_t *rays,
                                      int64_t M, int64_t N, float *weights_sum, float *depth, float *image) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t n = 0; n < N; ++n) {
        const uint32_t id = (uint32_t)rays[n * 3], off = (uint32_t)rays[n * 3 + 1], cnt = (uint32_t)rays[n * 3 + 2];
        float T = 1.0f, r = 0, g = 0, b = 0, ws = 0, t = 0, d = 0;
        if (cnt != 0 && (int64_t)off + cnt < M) {
            for (uint32_t s = 0; s < cnt; ++s) {
                const int64_t k = (int64_t)off + s;
                const float alpha = 1.0f - expf(-sigmas[k] * deltas[k * 2]);
                const float w = alpha * T;
                r = fmaf(w, rgbs[k * 3], r);
                g = fmaf(w, rgbs[k * 3 + 1], g);
                b = fmaf(w, rgbs[k * 3 + 2], b);
                t += deltas[k * 2 + 1];
                d = fmaf(w, t, d);
                ws += w;
                T *= 1.0f - alpha;
                if (T < 1e-4f) break;
            }
        }
        weights_sum[id] = ws; depth[id] = d;
        image[id * 3] = r; image[id * 3 + 1] = g; image[id * 3 + 2] = b;
    }
}

/* Compositing backward, raymarching.cu:606-688.  grad buffers must be pre-zeroed by the caller
 * (raymarching.py:270-275); the terminating sample (T<1e-4) gets no gradient (:658-661). */
void orc_composite_rays_train_backward(const float *grad_weights_sum, const float *grad_image, const float *sigmas,
                                       const float *rgbs, const float *deltas, const int32_t *rays,
                                       const float *weights_sum, const float *image, int64_t M, int64_t N,
                                       float *grad_sigmas, float *grad_rgbs) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t n = 0; n < N; ++n) {
        const uint32_t id = (uint32_t)rays[n * 3], off = (uint32_t)rays[n * 3 + 1], cnt = (uint32_t)rays[n * 3 + 2];
        if (cnt == 0 || (int64_t)off + cnt >= M) continue;
        const float gr = grad_image[id * 3], gg = grad_image[id * 3 + 1], gb = grad_image[id * 3 + 2];
        const float gws = grad_weights_sum[id];
        const float rf = image[id * 3], gf = image[id * 3 + 1], bf = image[id * 3 + 2], wsf = weights_sum[id];
        float T = 1.0f, r = 0, g = 0, b = 0;
        for (uint32_t s = 0; s < cnt; ++s) {
            const int64_t k = (int64_t)off + s;
            const float alpha = 1.0f - expf(-sigmas[k] * deltas[k * 2]);
            const float w = alpha * T;
            r = fmaf(w, rgbs[k * 3], r);
            g = fmaf(w, rgbs[k * 3 + 1], g);
            b = fmaf(w, rgbs[k * 3 + 2], b);
            T *= 1.0f - alpha;
            if (T < 1e-4f) break;
            grad_rgbs[k * 3] = gr * w; grad_rgbs[k * 3 + 1] = gg * w; grad_rgbs[k * 3 + 2] = gb * w;
            grad_sigmas[k] = deltas[k * 2] * (gr * (T * rgbs[k * 3] - (rf - r)) + gg * (T * rgbs[k * 3 + 1] - (gf - g)) +
                                              gb * (T * rgbs[k * 3 + 2] - (bf - b)) + gws * (1.0f - wsf));
        }
    }
}

/* Inference marcher, raymarching.cu:706-814: up to n_step samples for each alive ray, from rays_t[id];
 * outputs must be pre-zeroed (delta==0 marks an unused slot).  perturb seeds pcg32(perturb), advance(n). */
void orc_march_rays(int64_t n_alive, uint32_t n_step, const int32_t *rays_alive, const float *rays_t,
                    const float *rays_o, const float *rays_d, float bound, float dt_gamma, uint32_t max_steps,
                    uint32_t C, uint32_t H, const uint8_t *grid, const float *nears, const float *fars,
                    float *xyzs, float *dirs, float *deltas, uint32_t perturb) {
    (void)nears;
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t n = 0; n < n_alive; ++n) {
        const int32_t id = rays_alive[n];
        orc_march_t m;
        orc_march_setup(&m, rays_o + (int64_t)id * 3, rays_d + (int64_t)id * 3, grid, bound, dt_gamma, max_steps, C, H);
        float t = rays_t[id];
        if (perturb) {
            orc_pcg32 r;
            orc_pcg_seed(&r, perturb, 1);
            orc_pcg_advance(&r, (uint64_t)n);
            t = fmaf(m.dt_min, orc_pcg_float(&r), t); /* nvcc contracts near + dt_min*u to one fma */
        }
        float last_t = t;
        const float far = fars[id];
        uint32_t step = 0;
        float p[3], dt;
        while (t < far && step < n_step) {
            if (!orc_march_step(&m, &t, p, &dt)) continue;
            const int64_t k = n * n_step + step;
            xyzs[k * 3] = p[0]; xyzs[k * 3 + 1] = p[1]; xyzs[k * 3 + 2] = p[2];
            dirs[k * 3] = m.dx; dirs[k * 3 + 1] = m.dy; dirs[k * 3 + 2] = m.dz;
            t += dt;
            deltas[k * 2] = dt;
            deltas[k * 2 + 1] = t - last_t;
            last_t = t;
            ++step;
        }
    }
}

/* Inference compositing, raymarching.cu:828-913: in-place accumulation with T = 1 - weight_sum;
 * a ray that ends early (delta==0 or T<1e-4) is marked dead (rays_alive[n] = -1), else rays_t advances. */
void orc_composite_rays(int64_t n_alive, uint32_t n_step, int32_t *rays_alive, float *rays_t, const float *sigmas,
                        const float *rgbs, const float *deltas, float *weights_sum, float *depth, float *image) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t n = 0; n < n_alive; ++n) {
        const int32_t id = rays_alive[n];
        float t = rays_t[id], ws = weights_sum[id], d = depth[id];
        float r = image[id * 3], g = image[id * 3 + 1], b = image[id * 3 + 2];
        uint32_t step = 0;
        while (step < n_step) {
            const int64_t k = n * n_step + step;
            if (deltas[k * 2] == 0.0f) break;
            const float alpha = 1.0f - expf(-sigmas[k] * deltas[k * 2]);
            const float T = 1.0f - ws;
            const float w = alpha * T;
            ws += w;
            t += deltas[k * 2 + 1];
            d = fmaf(w, t, d);
            r = fmaf(w, rgbs[k * 3], r);
            g = fmaf(w, rgbs[k * 3 + 1], g);
            b = fmaf(w, rgbs[k * 3 + 2], b);
            if (T < 1e-4f) break; /* the literal is a double in the source (:880); T<1e-4 in float vs double differs only at T==fl(1e-4) */
            ++step;
        }
        if (step < n_step) rays_alive[n] = -1;
        else rays_t[id] = t;
        weights_sum[id] = ws; depth[id] = d;
        image[id * 3] = r; image[id * 3 + 1] = g; image[id * 3 + 2] = b;
    }
}
