/*
 * TEST INFRASTRUCTURE — CPU restatement (plain C99) of the reference rasterizer hot path.
 * Not part of the product: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * reference legs may build, load or call this file.
 *
 * Restates, function by function, the reference at /root/reference/submodules/:
 *   DGR = diff-gaussian-rasterization/cuda_rasterizer, R2 = rasterizer2/cuda_rasterizer
 *   oc_preprocess      DGR/forward.cu:155-256 (+ :20-71 SH, :74-113 cov2D, :118-152 cov3D),
 *                      DGR/auxiliary.h:41-76,139-164 (ndc2Pix, getRect, transforms, in_frustum)
 *   oc_bin             DGR/rasterizer_impl.cu:35-50 (getHigherMsb), :70-111 (duplicateWithKeys),
 *                      :116-138 (identifyTileRanges), :277 (InclusiveSum), :303-308 (SortPairs: any
 *                      stable sort on bits [0,32+bit) is bit-identical)
 *   oc_blend_forward   DGR/forward.cu:261-374
 *   oc_blend_backward  DGR/backward.cu:399-557
 *   oc_preprocess_backward  DGR/backward.cu:144-274 (computeCov2DCUDA), :278-341 (computeCov3D),
 *                      :346-396 (preprocessCUDA), :20-139 (SH backward), auxiliary.h:107-117 (dnormvdv)
 *   oc_mark_visible    DGR/rasterizer_impl.cu:54-66
 *   oc_idxmap_*        R2/raster.cu:9-62, R2/rasterizer_impl.cu:26-42,578-615,680-847
 *
 * Pinning: the reference ships no golden vectors for this path. This
 * restatement is pinned against OUTPUTS OF THE REFERENCE ITSELF: the tests/golden .npz files were
 * produced by the unmodified reference kernels compiled for sm_100 (oracle/build_ref.sh) on a
 * B200 by tests/golden/make_golden.py, and tests/test_oracle_cpu.py checks this file against them.
 *
 * Rounding: the reference's integer outputs (radii, tiles_touched, depth bits in the sort keys,
 * ranges) depend on which multiplications nvcc contracted into FMAs. The geometry chain below
 * uses explicit fmaf() in the order the reference compiles to for sm_100 (verified bit-exact on
 * the device against the reference on 3.4 M Gaussians); everything else is plain C float
 * arithmetic built with -ffp-contract=off. expf/sinf/cosf come from libm, not from the GPU's
 * SFU, so n_contrib can differ from the reference on isolated pixels whose alpha sits within an
 * ulp of a threshold; tests allow for that explicitly.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define OC_BLOCK 16
#define OC_NFEAT 64

static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

/* a0*b0 + a1*b1 + a2*b2 as GLM's mat3 products compile: fma(a2,b2, fma(a0,b0, a1*b1)) */
static inline float dot3c(float a0, float b0, float a1, float b1, float a2, float b2) {
    return fmaf(a2, b2, fmaf(a0, b0, a1 * b1));
}
/* transformPoint4x3 / 4x4 row r (auxiliary.h:58-76): fma(z,m8, fma(x,m0, y*m4)) + m12 */
static inline float affine_row(const float* m, int r, float x, float y, float z) {
    return fmaf(z, m[8 + r], fmaf(x, m[r], y * m[4 + r])) + m[12 + r];
}
/* ndc2Pix (auxiliary.h:41-44) evaluates in double: ((v + 1.0) * S - 1.0) * 0.5 */
static inline float ndc2pix(float v, int S) {
    double t = (double)v + 1.0;
    t = fma(t, (double)S, -1.0);
    return (float)(t * 0.5);
}

void oc_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
int oc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ------------------------------------------------------------------------------------------
 * SH -> RGB (forward.cu:20-71).
 * ---------------------------------------------------------------------------------------- */
static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                               0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                               -0.5900435899266435f};

static void sh_color(int deg, const float* pos, const float* campos, const float* sh, float* rgb,
                     uint8_t* clamped) {
    /* FMA placement as the reference compiles for sm_100: one fma per coefficient, in order */
    float dx = pos[0] - campos[0], dy = pos[1] - campos[1], dz = pos[2] - campos[2];
    float len = sqrtf(fmaf(dz, dz, fmaf(dx, dx, dy * dy)));
    float x = dx / len, y = dy / len, z = dz / len;
    for (int c = 0; c < 3; ++c) {
#define S(k) sh[3 * (k) + c]
        float r = S(0) * SH_C0;
        if (deg > 0) {
            r = fmaf(-(y * SH_C1), S(1), r);
            r = fmaf(z * SH_C1, S(2), r);
            r = fmaf(-(x * SH_C1), S(3), r);
            if (deg > 1) {
                const float xx = x * x, yy = y * y, zz = z * z, xy = y * x, yz = z * y, xz = z * x;
                const float zz2 = zz + zz, xx_yy = xx - yy;
                r = fmaf(xy * SH_C2[0], S(4), r);
                r = fmaf(yz * SH_C2[1], S(5), r);
                r = fmaf(((zz2 - xx) - yy) * SH_C2[2], S(6), r);
                r = fmaf(xz * SH_C2[3], S(7), r);
                r = fmaf(xx_yy * SH_C2[4], S(8), r);
                if (deg > 2) {
                    const float q = fmaf(zz, 4.0f, -xx) - yy;
                    r = fmaf((y * SH_C3[0]) * fmaf(xx, 3.0f, -yy), S(9), r);
                    r = fmaf((xy * SH_C3[1]) * z, S(10), r);
                    r = fmaf((y * SH_C3[2]) * q, S(11), r);
                    r = fmaf((z * SH_C3[3]) * fmaf(yy, -3.0f, fmaf(xx, -3.0f, zz2)), S(12), r);
                    r = fmaf(q * (x * SH_C3[4]), S(13), r);
                    r = fmaf(xx_yy * (z * SH_C3[5]), S(14), r);
                    r = fmaf((x * SH_C3[6]) * fmaf(yy, -3.0f, xx), S(15), r);
                }
            }
        }
#undef S
        r += 0.5f;
        clamped[c] = r < 0.0f;
        rgb[c] = r > 0.0f ? r : 0.0f;
    }
}

/* computeCov3D (forward.cu:118-152): Sigma = (S R)^T (S R), un-normalised quaternion. */
static void cov3d(const float* s, float mod, const float* q, float* cov) {
    const float r = q[0], x = q[1], y = q[2], z = q[3];
    const float S0 = s[0] * mod, S1 = s[1] * mod, S2 = s[2] * mod;
    const float xz = x * z, rx = r * x, rz = r * z, yy = y * y, zz = z * z;
    const float xz_p_ry = fmaf(r, y, xz), xz_m_ry = fmaf(-r, y, xz);
    const float yz_m_rx = fmaf(y, z, -rx), yz_p_rx = fmaf(y, z, rx);
    const float xy_m_rz = fmaf(x, y, -rz), xy_p_rz = fmaf(x, y, rz);
    const float xx_p_yy = fmaf(x, x, yy), yy_p_zz = yy + zz, xx_p_zz = fmaf(x, x, zz);
    const float R00 = 1.0f - (yy_p_zz + yy_p_zz), R01 = xy_m_rz + xy_m_rz, R02 = xz_p_ry + xz_p_ry;
    const float R10 = xy_p_rz + xy_p_rz, R11 = 1.0f - (xx_p_zz + xx_p_zz), R12 = yz_m_rx + yz_m_rx;
    const float R20 = xz_m_ry + xz_m_ry, R21 = yz_p_rx + yz_p_rx, R22 = 1.0f - (xx_p_yy + xx_p_yy);
    const float M00 = S0 * R00, M01 = S1 * R01, M02 = S2 * R02;
    const float M10 = S0 * R10, M11 = S1 * R11, M12 = S2 * R12;
    const float M20 = S0 * R20, M21 = S1 * R21, M22 = S2 * R22;
    cov[0] = dot3c(M00, M00, M01, M01, M02, M02);
    cov[1] = dot3c(M00, M10, M01, M11, M02, M12);
    cov[2] = dot3c(M00, M20, M01, M21, M02, M22);
    cov[3] = dot3c(M10, M10, M11, M11, M12, M12);
    cov[4] = dot3c(M10, M20, M11, M21, M12, M22);
    cov[5] = dot3c(M20, M20, M21, M21, M22, M22);
}

typedef struct { float a, b, c, T[2][3], tx, ty, tz, txtz, tytz; } Cov2D;

/* computeCov2D (forward.cu:74-113): T = W*J, cov = T^T Vrk^T T, +0.3 on the diagonal. */
static Cov2D cov2d(const float* mean, float fx, float fy, float tanx, float tany, const float* c3,
                   const float* v) {
    Cov2D o;
    const float tx = affine_row(v, 0, mean[0], mean[1], mean[2]);
    const float ty = affine_row(v, 1, mean[0], mean[1], mean[2]);
    const float tz = affine_row(v, 2, mean[0], mean[1], mean[2]);
    const float limx = tanx * 1.3f, limy = tany * 1.3f;
    o.txtz = tx / tz; o.tytz = ty / tz;
    o.tx = fminf(fmaxf(o.txtz, -limx), limx) * tz;
    o.ty = fminf(fmaxf(o.tytz, -limy), limy) * tz;
    o.tz = tz;
    const float tz2 = tz * tz;
    const float J00 = fx / tz, J02 = (-o.tx * fx) / tz2, J11 = fy / tz, J12 = (-o.ty * fy) / tz2;
    o.T[0][0] = fmaf(v[2], J02, v[0] * J00);
    o.T[0][1] = fmaf(v[6], J02, v[4] * J00);
    o.T[0][2] = fmaf(v[10], J02, v[8] * J00);
    o.T[1][0] = fmaf(v[2], J12, v[1] * J11);
    o.T[1][1] = fmaf(v[6], J12, v[5] * J11);
    o.T[1][2] = fmaf(v[10], J12, v[9] * J11);
    const float (*T)[3] = o.T;
    const float X00 = dot3c(T[0][0], c3[0], T[0][1], c3[1], T[0][2], c3[2]);
    const float X01 = dot3c(T[1][0], c3[0], T[1][1], c3[1], T[1][2], c3[2]);
    const float X10 = dot3c(T[0][0], c3[1], T[0][1], c3[3], T[0][2], c3[4]);
    const float X11 = dot3c(T[1][0], c3[1], T[1][1], c3[3], T[1][2], c3[4]);
    const float X20 = dot3c(T[0][0], c3[2], T[0][1], c3[4], T[0][2], c3[5]);
    const float X21 = dot3c(T[1][0], c3[2], T[1][1], c3[4], T[1][2], c3[5]);
    o.a = dot3c(T[0][0], X00, T[0][1], X10, T[0][2], X20) + 0.3f;
    o.b = dot3c(T[0][0], X01, T[0][1], X11, T[0][2], X21);
    o.c = dot3c(T[1][0], X01, T[1][1], X11, T[1][2], X21) + 0.3f;
    return o;
}

/* getRect (auxiliary.h:46-56) */
static void get_rect(float px, float py, int radius, int gx, int gy, int* r) {
    const float rf = (float)radius;
    int a = (int)((px - rf) * 0.0625f), b = (int)((py - rf) * 0.0625f);
    int c = (int)((((px + rf) + 16.0f) + -1.0f) * 0.0625f), d = (int)((((py + rf) + 16.0f) + -1.0f) * 0.0625f);
    if (a < 0) a = 0; if (b < 0) b = 0; if (c < 0) c = 0; if (d < 0) d = 0;
    r[0] = a < gx ? a : gx; r[1] = b < gy ? b : gy; r[2] = c < gx ? c : gx; r[3] = d < gy ? d : gy;
}

/* K1 preprocessCUDA (forward.cu:155-256). Rows of culled Gaussians are written as zeros (the
 * reference leaves them uninitialised, so comparisons use radii > 0 rows only). */
void oc_preprocess(int P, int D, int M, int W, int H, float tanx, float tany, float scale_modifier,
                   const float* means3D, const float* shs, const float* colors_precomp,
                   const float* opacities, const float* scales, const float* rotations,
                   const float* cov3D_precomp, const float* view, const float* proj, const float* campos,
                   int32_t* radii, float* means2D, float* depths, float* cov3Ds, float* conic_opacity,
                   float* rgb, uint8_t* clamped, uint32_t* tiles_touched, int32_t* rects) {
    const int gx = (W + OC_BLOCK - 1) / OC_BLOCK, gy = (H + OC_BLOCK - 1) / OC_BLOCK;
    const float focal_y = H / (2.0f * tany), focal_x = W / (2.0f * tanx); /* rasterizer_impl.cu:222-223 */
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; ++i) {
        radii[i] = 0; tiles_touched[i] = 0; depths[i] = 0.0f;
        means2D[2 * i] = means2D[2 * i + 1] = 0.0f;
        for (int k = 0; k < 6; ++k) cov3Ds[6 * i + k] = 0.0f;
        for (int k = 0; k < 4; ++k) { conic_opacity[4 * i + k] = 0.0f; rects[4 * i + k] = 0; }
        for (int k = 0; k < 3; ++k) { rgb[3 * i + k] = 0.0f; clamped[3 * i + k] = 0; }
        const float* m = means3D + 3 * i;
        const float depth = affine_row(view, 2, m[0], m[1], m[2]);
        if (depth <= 0.2f) continue;                                     /* in_frustum */
        const float hx = affine_row(proj, 0, m[0], m[1], m[2]);
        const float hy = affine_row(proj, 1, m[0], m[1], m[2]);
        const float hw = affine_row(proj, 3, m[0], m[1], m[2]);
        const float p_w = 1.0f / (hw + 0.0000001f);
        const float projx = hx * p_w, projy = hy * p_w;
        float c3[6];
        if (cov3D_precomp) memcpy(c3, cov3D_precomp + 6 * i, sizeof c3);
        else cov3d(scales + 3 * i, scale_modifier, rotations + 4 * i, c3);
        const Cov2D c2 = cov2d(m, focal_x, focal_y, tanx, tany, c3, view);
        const float det = fmaf(c2.a, c2.c, -(c2.b * c2.b));
        if (det == 0.0f) continue;
        const float det_inv = 1.0f / det;
        const float mid = (c2.a + c2.c) * 0.5f;
        const float disc = sqrtf(fmaxf(fmaf(mid, mid, -det), 0.1f));
        const float lam = fmaxf(mid + disc, mid - disc);
        const int rad = (int)ceilf(sqrtf(lam) * 3.0f);
        const float ix = ndc2pix(projx, W), iy = ndc2pix(projy, H);
        int r[4];
        get_rect(ix, iy, rad, gx, gy, r);
        const int tiles = (r[2] - r[0]) * (r[3] - r[1]);
        if (tiles == 0) continue;
        if (colors_precomp) {
            for (int k = 0; k < 3; ++k) rgb[3 * i + k] = colors_precomp[3 * i + k];
        } else {
            sh_color(D, m, campos, shs + (size_t)i * M * 3, rgb + 3 * i, clamped + 3 * i);
        }
        for (int k = 0; k < 6; ++k) cov3Ds[6 * i + k] = c3[k];
        depths[i] = depth; radii[i] = rad;
        means2D[2 * i] = ix; means2D[2 * i + 1] = iy;
        conic_opacity[4 * i] = c2.c * det_inv;
        conic_opacity[4 * i + 1] = c2.b * -det_inv;
        conic_opacity[4 * i + 2] = c2.a * det_inv;
        conic_opacity[4 * i + 3] = opacities[i];
        tiles_touched[i] = (uint32_t)tiles;
        memcpy(rects + 4 * i, r, sizeof r);
    }
}

/* K10 checkFrustum (rasterizer_impl.cu:54-66) */
void oc_mark_visible(int P, const float* means3D, const float* view, uint8_t* present) {
    for (int i = 0; i < P; ++i)
        present[i] = !(affine_row(view, 2, means3D[3 * i], means3D[3 * i + 1], means3D[3 * i + 2]) <= 0.2f);
}

/* ------------------------------------------------------------------------------------------
 * Binning: InclusiveSum, duplicateWithKeys, stable sort on bits [0, 32+bit), identifyTileRanges.
 * ---------------------------------------------------------------------------------------- */
/* getHigherMsb (rasterizer_impl.cu:35-50) */
uint32_t oc_higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

/* Inclusive prefix sum; returns num_rendered (rasterizer_impl.cu:277-281). */
int64_t oc_scan(int P, const uint32_t* tiles_touched, uint32_t* point_offsets) {
    uint32_t s = 0;
    for (int i = 0; i < P; ++i) { s += tiles_touched[i]; point_offsets[i] = s; }
    return (int64_t)s;
}

/* Stable LSD radix sort of (key, value) pairs on key bits [0, end_bit). */
static void radix_sort_pairs(int64_t n, uint64_t* keys, uint32_t* vals, uint64_t* ktmp, uint32_t* vtmp,
                             int end_bit) {
    uint64_t *ki = keys, *ko = ktmp;
    uint32_t *vi = vals, *vo = vtmp;
    for (int shift = 0; shift < end_bit; shift += 8) {
        const int bits = end_bit - shift < 8 ? end_bit - shift : 8;
        const uint64_t mask = (1ull << bits) - 1;
        int64_t cnt[257];
        memset(cnt, 0, sizeof cnt);
        for (int64_t i = 0; i < n; ++i) cnt[((ki[i] >> shift) & mask) + 1]++;
        for (int d = 0; d < 256; ++d) cnt[d + 1] += cnt[d];
        for (int64_t i = 0; i < n; ++i) {
            const int64_t dst = cnt[(ki[i] >> shift) & mask]++;
            ko[dst] = ki[i]; vo[dst] = vi[i];
        }
        uint64_t* tk = ki; ki = ko; ko = tk;
        uint32_t* tv = vi; vi = vo; vo = tv;
    }
    if (ki != keys) { memcpy(keys, ki, 8 * (size_t)n); memcpy(vals, vi, 4 * (size_t)n); }
}

/* duplicateWithKeys + SortPairs + identifyTileRanges (rasterizer_impl.cu:70-138,289-317).
 * keys/vals: [num_rendered] sorted output; ranges: [tiles][2], zero for empty tiles. */
int oc_bin(int P, int W, int H, const int32_t* radii, const float* depths, const int32_t* rects,
           const uint32_t* point_offsets, int64_t num_rendered, uint64_t* keys, uint32_t* vals,
           uint32_t* ranges) {
    const int gx = (W + OC_BLOCK - 1) / OC_BLOCK, gy = (H + OC_BLOCK - 1) / OC_BLOCK;
    for (int i = 0; i < P; ++i) {
        if (radii[i] <= 0) continue;
        uint32_t off = i == 0 ? 0u : point_offsets[i - 1];
        const int32_t* r = rects + 4 * i;
        for (int y = r[1]; y < r[3]; ++y)
            for (int x = r[0]; x < r[2]; ++x) {
                uint64_t key = (uint64_t)(y * gx + x);
                key <<= 32;
                key |= f2u(depths[i]);
                keys[off] = key; vals[off] = (uint32_t)i; ++off;
            }
    }
    uint64_t* kt = (uint64_t*)malloc(8 * (size_t)(num_rendered > 0 ? num_rendered : 1));
    uint32_t* vt = (uint32_t*)malloc(4 * (size_t)(num_rendered > 0 ? num_rendered : 1));
    if (!kt || !vt) { free(kt); free(vt); return -1; }
    radix_sort_pairs(num_rendered, keys, vals, kt, vt, 32 + (int)oc_higher_msb((uint32_t)(gx * gy)));
    free(kt); free(vt);
    memset(ranges, 0, 8 * (size_t)gx * gy);
    for (int64_t i = 0; i < num_rendered; ++i) {
        const uint32_t cur = (uint32_t)(keys[i] >> 32);
        if (i == 0) ranges[2 * cur] = 0;
        else {
            const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
            if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)i; ranges[2 * cur] = (uint32_t)i; }
        }
        if (i == num_rendered - 1) ranges[2 * cur + 1] = (uint32_t)num_rendered;
    }
    return 0;
}

/* power exactly as renderCUDA compiles (forward.cu:332-335):
 * fma(fma(dx, dx*a, dy*(dy*c)), -0.5, -(dy*(dx*b))) */
static inline float blend_power(float dx, float dy, float a, float b, float c) {
    const float s = fmaf(dx, dx * a, dy * (dy * c));
    return fmaf(s, -0.5f, -(dy * (dx * b)));
}

/* K6 renderCUDA forward (forward.cu:261-374), one pixel at a time (pixels are independent; the
 * block-level early exit of the reference changes no output). counts[0] += traversed pairs,
 * counts[1] += contributing pairs (roofline accounting; may be NULL). */
void oc_blend_forward(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const float* means2D,
                      const float* conic_opacity, const float* rgb, const float* bg, float* final_T,
                      uint32_t* n_contrib, float* out_color, uint64_t* counts) {
    const int gx = (W + OC_BLOCK - 1) / OC_BLOCK;
    const size_t plane = (size_t)W * H;
    uint64_t trav = 0, contrib = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : trav, contrib)
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const int tile = (y / OC_BLOCK) * gx + x / OC_BLOCK;
            const uint32_t lo = ranges[2 * tile], hi = ranges[2 * tile + 1];
            const float pxf = (float)x, pyf = (float)y;
            float T = 1.0f, C[3] = {0, 0, 0};
            uint32_t contributor = 0, last = 0;
            for (uint32_t k = lo; k < hi; ++k) {
                ++contributor;
                const uint32_t g = point_list[k];
                const float dx = means2D[2 * g] - pxf, dy = means2D[2 * g + 1] - pyf;
                const float* co = conic_opacity + 4 * g;
                const float power = blend_power(dx, dy, co[0], co[1], co[2]);
                if (power > 0.0f) continue;
                const float alpha = fminf(0.99f, co[3] * expf(power));
                if (alpha < 1.0f / 255.0f) continue;
                const float test_T = T * (1.0f - alpha);
                if (test_T < 0.0001f) break;                       /* done = true */
                for (int ch = 0; ch < 3; ++ch) C[ch] = fmaf(T, alpha * rgb[3 * g + ch], C[ch]);
                T = test_T; last = contributor; ++contrib;
            }
            trav += contributor;
            const size_t pix = (size_t)y * W + x;
            final_T[pix] = T; n_contrib[pix] = last;
            for (int ch = 0; ch < 3; ++ch) out_color[ch * plane + pix] = fmaf(bg[ch], T, C[ch]);
        }
    if (counts) { counts[0] += trav; counts[1] += contrib; }
}

/* K7 renderCUDA backward (backward.cu:399-557). Per-Gaussian sums are accumulated in double
 * (the reference uses float atomics in arbitrary order, so it is itself only reproducible to
 * rounding); outputs are float, zero-initialised by the caller:
 * dL_dmean2D[P][3], dL_dconic[P][4] (a, b, -, c), dL_dopacity[P], dL_dcolor[P][3]. */
void oc_blend_backward(int P, int W, int H, const uint32_t* ranges, const uint32_t* point_list,
                       const float* means2D, const float* conic_opacity, const float* rgb, const float* bg,
                       const float* final_T, const uint32_t* n_contrib, const float* dL_dpix,
                       float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor) {
    const int gx = (W + OC_BLOCK - 1) / OC_BLOCK;
    const size_t plane = (size_t)W * H;
    double* acc = (double*)calloc((size_t)(P > 0 ? P : 1) * 9, sizeof(double));
    if (!acc) return;
    const float ddelx_dx = (float)(0.5 * W), ddely_dy = (float)(0.5 * H);     /* backward.cu:460-461 */
#pragma omp parallel for schedule(dynamic, 4)
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const int tile = (y / OC_BLOCK) * gx + x / OC_BLOCK;
            const uint32_t lo = ranges[2 * tile];
            const size_t pix = (size_t)y * W + x;
            const float pxf = (float)x, pyf = (float)y;
            const float T_final = final_T[pix];
            float T = T_final;
            const uint32_t last_contributor = n_contrib[pix];
            float accum_rec[3] = {0, 0, 0}, last_color[3] = {0, 0, 0}, dLp[3];
            for (int ch = 0; ch < 3; ++ch) dLp[ch] = dL_dpix[ch * plane + pix];
            float last_alpha = 0.0f;
            for (uint32_t k = last_contributor; k-- > 0;) {
                const uint32_t g = point_list[lo + k];
                const float dx = means2D[2 * g] - pxf, dy = means2D[2 * g + 1] - pyf;
                const float* co = conic_opacity + 4 * g;
                const float power = blend_power(dx, dy, co[0], co[1], co[2]);
                if (power > 0.0f) continue;
                const float G = expf(power);
                const float alpha = fminf(0.99f, co[3] * G);
                if (alpha < 1.0f / 255.0f) continue;
                T = T / (1.0f - alpha);
                const float dchannel_dcolor = alpha * T;
                float dL_dalpha = 0.0f;
                double* a = acc + (size_t)g * 9;
                for (int ch = 0; ch < 3; ++ch) {
                    const float c = rgb[3 * g + ch];
                    accum_rec[ch] = last_alpha * last_color[ch] + (1.0f - last_alpha) * accum_rec[ch];
                    last_color[ch] = c;
                    dL_dalpha += (c - accum_rec[ch]) * dLp[ch];
                    const double v = (double)(dchannel_dcolor * dLp[ch]);
#pragma omp atomic
                    a[ch] += v;
                }
                dL_dalpha *= T;
                last_alpha = alpha;
                float bg_dot = 0.0f;
                for (int ch = 0; ch < 3; ++ch) bg_dot += bg[ch] * dLp[ch];
                dL_dalpha += (-T_final / (1.0f - alpha)) * bg_dot;
                const float dL_dG = co[3] * dL_dalpha;
                const float gdx = G * dx, gdy = G * dy;
                const float dG_ddelx = -gdx * co[0] - gdy * co[1];
                const float dG_ddely = -gdy * co[2] - gdx * co[1];
                const double v3 = (double)(dL_dG * dG_ddelx * ddelx_dx), v4 = (double)(dL_dG * dG_ddely * ddely_dy);
                const double v5 = (double)(-0.5f * gdx * dx * dL_dG), v6 = (double)(-0.5f * gdx * dy * dL_dG);
                const double v7 = (double)(-0.5f * gdy * dy * dL_dG), v8 = (double)(G * dL_dalpha);
#pragma omp atomic
                a[3] += v3;
#pragma omp atomic
                a[4] += v4;
#pragma omp atomic
                a[5] += v5;
#pragma omp atomic
                a[6] += v6;
#pragma omp atomic
                a[7] += v7;
#pragma omp atomic
                a[8] += v8;
            }
        }
    for (int g = 0; g < P; ++g) {
        const double* a = acc + (size_t)g * 9;
        for (int ch = 0; ch < 3; ++ch) dL_dcolor[3 * g + ch] = (float)a[ch];
        dL_dmean2D[3 * g] = (float)a[3]; dL_dmean2D[3 * g + 1] = (float)a[4];
        dL_dconic[4 * g] = (float)a[5]; dL_dconic[4 * g + 1] = (float)a[6]; dL_dconic[4 * g + 3] = (float)a[7];
        dL_dopacity[g] = (float)a[8];
    }
    free(acc);
}

/* ------------------------------------------------------------------------------------------
 * Preprocess backward: computeCov2DCUDA (backward.cu:144-274), preprocessCUDA backward
 * (:346-396), SH backward (:20-139), computeCov3D backward (:278-341), dnormvdv
 * (auxiliary.h:107-117). Plain float arithmetic in source order (tolerance oracle).
 * ---------------------------------------------------------------------------------------- */
static void sh_backward(int deg, int M, const float* pos, const float* campos, const float* sh,
                        const uint8_t* clamped, const float* dL_dcolor, float* dL_dmean, float* dL_dsh) {
    (void)M;
    const float ox = pos[0] - campos[0], oy = pos[1] - campos[1], oz = pos[2] - campos[2];
    const float len = sqrtf(ox * ox + oy * oy + oz * oz);
    const float x = ox / len, y = oy / len, z = oz / len;
    float dRGB[3];
    for (int c = 0; c < 3; ++c) dRGB[c] = dL_dcolor[c] * (clamped[c] ? 0.0f : 1.0f);
    float dRGBdx[3] = {0, 0, 0}, dRGBdy[3] = {0, 0, 0}, dRGBdz[3] = {0, 0, 0};
#define S(k, c) sh[3 * (k) + (c)]
#define D(k, c) dL_dsh[3 * (k) + (c)]
    for (int c = 0; c < 3; ++c) {
        D(0, c) = SH_C0 * dRGB[c];
        if (deg > 0) {
            D(1, c) = -SH_C1 * y * dRGB[c];
            D(2, c) = SH_C1 * z * dRGB[c];
            D(3, c) = -SH_C1 * x * dRGB[c];
            dRGBdx[c] = -SH_C1 * S(3, c);
            dRGBdy[c] = -SH_C1 * S(1, c);
            dRGBdz[c] = SH_C1 * S(2, c);
            if (deg > 1) {
                const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                D(4, c) = SH_C2[0] * xy * dRGB[c];
                D(5, c) = SH_C2[1] * yz * dRGB[c];
                D(6, c) = SH_C2[2] * (2.f * zz - xx - yy) * dRGB[c];
                D(7, c) = SH_C2[3] * xz * dRGB[c];
                D(8, c) = SH_C2[4] * (xx - yy) * dRGB[c];
                dRGBdx[c] += SH_C2[0] * y * S(4, c) + SH_C2[2] * 2.f * -x * S(6, c) + SH_C2[3] * z * S(7, c) +
                             SH_C2[4] * 2.f * x * S(8, c);
                dRGBdy[c] += SH_C2[0] * x * S(4, c) + SH_C2[1] * z * S(5, c) + SH_C2[2] * 2.f * -y * S(6, c) +
                             SH_C2[4] * 2.f * -y * S(8, c);
                dRGBdz[c] += SH_C2[1] * y * S(5, c) + SH_C2[2] * 2.f * 2.f * z * S(6, c) + SH_C2[3] * x * S(7, c);
                if (deg > 2) {
                    D(9, c) = SH_C3[0] * y * (3.f * xx - yy) * dRGB[c];
                    D(10, c) = SH_C3[1] * xy * z * dRGB[c];
                    D(11, c) = SH_C3[2] * y * (4.f * zz - xx - yy) * dRGB[c];
                    D(12, c) = SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy) * dRGB[c];
                    D(13, c) = SH_C3[4] * x * (4.f * zz - xx - yy) * dRGB[c];
                    D(14, c) = SH_C3[5] * z * (xx - yy) * dRGB[c];
                    D(15, c) = SH_C3[6] * x * (xx - 3.f * yy) * dRGB[c];
                    dRGBdx[c] += SH_C3[0] * S(9, c) * 3.f * 2.f * xy + SH_C3[1] * S(10, c) * yz +
                                 SH_C3[2] * S(11, c) * -2.f * xy + SH_C3[3] * S(12, c) * -3.f * 2.f * xz +
                                 SH_C3[4] * S(13, c) * (-3.f * xx + 4.f * zz - yy) + SH_C3[5] * S(14, c) * 2.f * xz +
                                 SH_C3[6] * S(15, c) * 3.f * (xx - yy);
                    dRGBdy[c] += SH_C3[0] * S(9, c) * 3.f * (xx - yy) + SH_C3[1] * S(10, c) * xz +
                                 SH_C3[2] * S(11, c) * (-3.f * yy + 4.f * zz - xx) +
                                 SH_C3[3] * S(12, c) * -3.f * 2.f * yz + SH_C3[4] * S(13, c) * -2.f * xy +
                                 SH_C3[5] * S(14, c) * -2.f * yz + SH_C3[6] * S(15, c) * -3.f * 2.f * xy;
                    dRGBdz[c] += SH_C3[1] * S(10, c) * xy + SH_C3[2] * S(11, c) * 4.f * 2.f * yz +
                                 SH_C3[3] * S(12, c) * 3.f * (2.f * zz - xx - yy) +
                                 SH_C3[4] * S(13, c) * 4.f * 2.f * xz + SH_C3[5] * S(14, c) * (xx - yy);
                }
            }
        }
    }
#undef S
#undef D
    float dd[3] = {0, 0, 0};
    for (int c = 0; c < 3; ++c) { dd[0] += dRGBdx[c] * dRGB[c]; dd[1] += dRGBdy[c] * dRGB[c]; dd[2] += dRGBdz[c] * dRGB[c]; }
    /* dnormvdv (auxiliary.h:107-117) */
    const float sum2 = ox * ox + oy * oy + oz * oz;
    const float inv = 1.0f / sqrtf(sum2 * sum2 * sum2);
    dL_dmean[0] += ((sum2 - ox * ox) * dd[0] - oy * ox * dd[1] - oz * ox * dd[2]) * inv;
    dL_dmean[1] += (-ox * oy * dd[0] + (sum2 - oy * oy) * dd[1] - oz * oy * dd[2]) * inv;
    dL_dmean[2] += (-ox * oz * dd[0] - oy * oz * dd[1] + (sum2 - oz * oz) * dd[2]) * inv;
}

/* All outputs zero-initialised by the caller: dL_dmeans3D[P][3], dL_dcov3D[P][6], dL_dsh[P][M][3],
 * dL_dscales[P][3], dL_drots[P][4]. Inputs dL_dmean2D / dL_dconic / dL_dcolor come from
 * oc_blend_backward. */
void oc_preprocess_backward(int P, int D, int M, int W, int H, float tanx, float tany, float scale_modifier,
                            const float* means3D, const int32_t* radii, const float* shs,
                            const uint8_t* clamped, const float* scales, const float* rotations,
                            const float* cov3Ds, const float* view, const float* proj, const float* campos,
                            const float* dL_dmean2D, const float* dL_dconic, const float* dL_dcolor,
                            float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscales,
                            float* dL_drots) {
    const float h_y = H / (2.0f * tany), h_x = W / (2.0f * tanx);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; ++i) {
        if (!(radii[i] > 0)) continue;
        const float* m = means3D + 3 * i;
        const float* c3 = cov3Ds + 6 * i;
        const Cov2D c2 = cov2d(m, h_x, h_y, tanx, tany, c3, view);
        const float limx = 1.3f * tanx, limy = 1.3f * tany;
        const float xg = (c2.txtz < -limx || c2.txtz > limx) ? 0.f : 1.f;
        const float yg = (c2.tytz < -limy || c2.tytz > limy) ? 0.f : 1.f;
        const float gcx = dL_dconic[4 * i], gcy = dL_dconic[4 * i + 1], gcz = dL_dconic[4 * i + 3];
        const float a = c2.a, b = c2.b, c = c2.c;
        const float (*T)[3] = c2.T;
        const float denom = a * c - b * b;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        float* gc = dL_dcov3D + 6 * i;
        if (denom2inv != 0) {
            dL_da = denom2inv * (-c * c * gcx + 2 * b * c * gcy + (denom - a * c) * gcz);
            dL_dc = denom2inv * (-a * a * gcz + 2 * a * b * gcy + (denom - a * c) * gcx);
            dL_db = denom2inv * 2 * (b * c * gcx - (denom + 2 * b * b) * gcy + a * b * gcz);
            gc[0] = T[0][0] * T[0][0] * dL_da + T[0][0] * T[1][0] * dL_db + T[1][0] * T[1][0] * dL_dc;
            gc[3] = T[0][1] * T[0][1] * dL_da + T[0][1] * T[1][1] * dL_db + T[1][1] * T[1][1] * dL_dc;
            gc[5] = T[0][2] * T[0][2] * dL_da + T[0][2] * T[1][2] * dL_db + T[1][2] * T[1][2] * dL_dc;
            gc[1] = 2 * T[0][0] * T[0][1] * dL_da + (T[0][0] * T[1][1] + T[0][1] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][1] * dL_dc;
            gc[2] = 2 * T[0][0] * T[0][2] * dL_da + (T[0][0] * T[1][2] + T[0][2] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][2] * dL_dc;
            gc[4] = 2 * T[0][2] * T[0][1] * dL_da + (T[0][1] * T[1][2] + T[0][2] * T[1][1]) * dL_db + 2 * T[1][1] * T[1][2] * dL_dc;
        }
        const float V[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
        float dT0[3], dT1[3];
        for (int k = 0; k < 3; ++k) {
            const float r0 = T[0][0] * V[k][0] + T[0][1] * V[k][1] + T[0][2] * V[k][2];
            const float r1 = T[1][0] * V[k][0] + T[1][1] * V[k][1] + T[1][2] * V[k][2];
            dT0[k] = 2 * r0 * dL_da + r1 * dL_db;
            dT1[k] = 2 * r1 * dL_dc + r0 * dL_db;
        }
        /* W[c][r] with W[0] = (v0,v4,v8), W[1] = (v1,v5,v9), W[2] = (v2,v6,v10) */
        const float dJ00 = view[0] * dT0[0] + view[4] * dT0[1] + view[8] * dT0[2];
        const float dJ02 = view[2] * dT0[0] + view[6] * dT0[1] + view[10] * dT0[2];
        const float dJ11 = view[1] * dT1[0] + view[5] * dT1[1] + view[9] * dT1[2];
        const float dJ12 = view[2] * dT1[0] + view[6] * dT1[1] + view[10] * dT1[2];
        const float tz = 1.f / c2.tz, tz2 = tz * tz, tz3 = tz2 * tz;
        const float dtx = xg * -h_x * tz2 * dJ02;
        const float dty = yg * -h_y * tz2 * dJ12;
        const float dtz = -h_x * tz2 * dJ00 - h_y * tz2 * dJ11 + (2 * h_x * c2.tx) * tz3 * dJ02 + (2 * h_y * c2.ty) * tz3 * dJ12;
        float* gm = dL_dmeans3D + 3 * i;
        /* transformVec4x3Transpose */
        gm[0] = view[0] * dtx + view[1] * dty + view[2] * dtz;
        gm[1] = view[4] * dtx + view[5] * dty + view[6] * dtz;
        gm[2] = view[8] * dtx + view[9] * dty + view[10] * dtz;

        /* preprocessCUDA backward (backward.cu:370-387) */
        const float m_w = 1.0f / ((proj[3] * m[0] + proj[7] * m[1] + proj[11] * m[2] + proj[15]) + 0.0000001f);
        const float mul1 = (proj[0] * m[0] + proj[4] * m[1] + proj[8] * m[2] + proj[12]) * m_w * m_w;
        const float mul2 = (proj[1] * m[0] + proj[5] * m[1] + proj[9] * m[2] + proj[13]) * m_w * m_w;
        const float g2x = dL_dmean2D[3 * i], g2y = dL_dmean2D[3 * i + 1];
        gm[0] += (proj[0] * m_w - proj[3] * mul1) * g2x + (proj[1] * m_w - proj[3] * mul2) * g2y;
        gm[1] += (proj[4] * m_w - proj[7] * mul1) * g2x + (proj[5] * m_w - proj[7] * mul2) * g2y;
        gm[2] += (proj[8] * m_w - proj[11] * mul1) * g2x + (proj[9] * m_w - proj[11] * mul2) * g2y;
        if (shs)
            sh_backward(D, M, m, campos, shs + (size_t)i * M * 3, clamped + 3 * i, dL_dcolor + 3 * i, gm,
                        dL_dsh + (size_t)i * M * 3);
        if (scales) {
            /* computeCov3D backward (backward.cu:278-341); GLM is column-major: R[c][r] */
            const float* q = rotations + 4 * i;
            const float r = q[0], x = q[1], y = q[2], z = q[3];
            const float R[3][3] = {
                {1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
            const float s[3] = {scale_modifier * scales[3 * i], scale_modifier * scales[3 * i + 1],
                                scale_modifier * scales[3 * i + 2]};
            float Mm[3][3];                       /* M = S * R  ->  M[c][r] = s[r] * R[c][r] */
            for (int cc = 0; cc < 3; ++cc) for (int rr = 0; rr < 3; ++rr) Mm[cc][rr] = s[rr] * R[cc][rr];
            const float Dg[3][3] = {{gc[0], 0.5f * gc[1], 0.5f * gc[2]},
                                    {0.5f * gc[1], gc[3], 0.5f * gc[4]},
                                    {0.5f * gc[2], 0.5f * gc[4], gc[5]}};
            float dM[3][3];                       /* dL_dM = 2 * M * dL_dSigma: (A*B)[c][r] = sum_k A[k][r] B[c][k] */
            for (int cc = 0; cc < 3; ++cc)
                for (int rr = 0; rr < 3; ++rr)
                    dM[cc][rr] = 2.0f * Mm[0][rr] * Dg[cc][0] + 2.0f * Mm[1][rr] * Dg[cc][1] + 2.0f * Mm[2][rr] * Dg[cc][2];
            /* Rt[i][j] = R[j][i]; dL_dMt[i][j] = dM[j][i] */
            float t[3][3];
            float* gs = dL_dscales + 3 * i;
            for (int k = 0; k < 3; ++k) {
                gs[k] = R[0][k] * dM[0][k] + R[1][k] * dM[1][k] + R[2][k] * dM[2][k];
                for (int j = 0; j < 3; ++j) t[k][j] = dM[j][k] * s[k];
            }
            float* gq = dL_drots + 4 * i;
            gq[0] = 2 * z * (t[0][1] - t[1][0]) + 2 * y * (t[2][0] - t[0][2]) + 2 * x * (t[1][2] - t[2][1]);
            gq[1] = 2 * y * (t[1][0] + t[0][1]) + 2 * z * (t[2][0] + t[0][2]) + 2 * r * (t[1][2] - t[2][1]) - 4 * x * (t[2][2] + t[1][1]);
            gq[2] = 2 * x * (t[1][0] + t[0][1]) + 2 * r * (t[2][0] - t[0][2]) + 2 * z * (t[1][2] + t[2][1]) - 4 * y * (t[2][2] + t[0][0]);
            gq[3] = 2 * r * (t[0][1] - t[1][0]) + 2 * x * (t[2][0] + t[0][2]) + 2 * y * (t[1][2] + t[2][1]) - 4 * z * (t[1][1] + t[0][0]);
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * rasterizer2 "idxmap" extra channels (R2/raster.cu:9-62; R2/rasterizer_impl.cu:26-42,
 * 578-615,680-847). The reference sorts every (pixel, point) instance on the 63-bit key
 * x<<48 | y<<32 | depth_bits with a stable sort and takes the first entry of each pixel run;
 * with insertion order = ascending point index that entry is the instance with the smallest
 * (depth_bits, point index), which is what is computed here directly.
 * Outputs fully written: out_idx (-1 = miss), out_color[H][W][3], out_depth[H][W],
 * out_feature[H][W][64], radii[P]; returns the number of instances (the reference's return).
 * ---------------------------------------------------------------------------------------- */
int64_t oc_idxmap_forward(int P, int W, int H, const float* means3D, const float* view, const float* proj,
                          const float* campos, const float* feature_vector, int32_t* out_idx, float* out_color,
                          float* out_depth, float* out_feature, int32_t* radii) {
    const size_t npix = (size_t)W * H;
    uint64_t* zbuf = (uint64_t*)malloc(8 * npix);
    if (!zbuf) return -1;
    memset(zbuf, 0xff, 8 * npix);
    int64_t instances = 0;
    for (int i = 0; i < P; ++i) {
        radii[i] = 0;
        const float* m = means3D + 3 * i;
        const float hx = affine_row(proj, 0, m[0], m[1], m[2]);
        const float hy = affine_row(proj, 1, m[0], m[1], m[2]);
        const float hw = affine_row(proj, 3, m[0], m[1], m[2]);
        const float p_w = 1.0f / (hw + 0.0000001f);
        const float depth = affine_row(view, 2, m[0], m[1], m[2]);
        if (depth <= 0.2f) continue;
        const float fx = ndc2pix(hx * p_w, W), fy = ndc2pix(hy * p_w, H);
        const int cx = (int)fx, cy = (int)fy;
        const float rf = 3.0f / depth;                                  /* S / p_view.z, S = 3 */
        const int x0 = (int)fmaxf(0.0f, fx - rf), y0 = (int)fmaxf(0.0f, fy - rf);
        const int x1 = (int)fminf((float)W, (fx + rf) + 1.0f), y1 = (int)fminf((float)H, (fy + rf) + 1.0f);
        if (cx < 0 || cx >= W || cy < 0 || cy >= H) continue;
        radii[i] = 1;
        instances += (int64_t)(x1 - x0) * (y1 - y0);
        const uint64_t key = ((uint64_t)f2u(depth) << 32) | (uint32_t)i;
        for (int x = x0; x < x1; ++x)
            for (int y = y0; y < y1; ++y) {
                uint64_t* z = zbuf + (size_t)y * W + x;
                if (key < *z) *z = key;
            }
    }
    for (size_t pix = 0; pix < npix; ++pix) {
        float* f = out_feature + pix * OC_NFEAT;
        memset(f, 0, sizeof(float) * OC_NFEAT);
        out_color[3 * pix] = out_color[3 * pix + 1] = out_color[3 * pix + 2] = 0.0f;
        out_depth[pix] = 0.0f;
        if (zbuf[pix] == ~0ull) { out_idx[pix] = -1; continue; }
        const int g = (int)(uint32_t)(zbuf[pix] & 0xffffffffu);
        const float depth = u2f((uint32_t)(zbuf[pix] >> 32));
        float dx = means3D[3 * g] - campos[0], dy = means3D[3 * g + 1] - campos[1], dz = means3D[3 * g + 2] - campos[2];
        const float len = sqrtf(dx * dx + dy * dy + dz * dz);
        const float dir[3] = {dx / len, dy / len, dz / len};
        out_idx[pix] = g;
        f[0] = depth;
        for (int dim = 0; dim < 3; ++dim)                               /* positional_encoding_3d */
            for (int fr = 0; fr < 4; ++fr) {
                const float frequency = powf(2.0f, (float)fr);
                f[1 + dim * 8 + fr * 2] = sinf(frequency * 3.14159265358979323846f * dir[dim]);
                f[1 + dim * 8 + fr * 2 + 1] = cosf(frequency * 3.14159265358979323846f * dir[dim]);
            }
        for (int k = 0; k < 3; ++k) out_color[3 * pix + k] = dir[k];
        for (int k = 25; k < OC_NFEAT; ++k) f[k] = feature_vector[(size_t)g * OC_NFEAT + k];
        out_depth[pix] = depth;
    }
    free(zbuf);
    return instances;
}

/* CopyFeature (R2/rasterizer_impl.cu:820-847) with the race-free sum the reference intends:
 * dL_dfeature_vector[g][25..63] += grad_feature[pix][25..63] for every pixel that hit g.
 * dL_dfeature_vector is fully written. */
void oc_idxmap_backward(int P, int W, int H, const int32_t* out_idx, const float* grad_feature,
                        float* dL_dfeature_vector) {
    const size_t npix = (size_t)W * H;
    double* acc = (double*)calloc((size_t)(P > 0 ? P : 1) * OC_NFEAT, sizeof(double));
    if (!acc) return;
    for (size_t pix = 0; pix < npix; ++pix) {
        const int g = out_idx[pix];
        if (g < 0 || g >= P) continue;
        for (int k = 25; k < OC_NFEAT; ++k) acc[(size_t)g * OC_NFEAT + k] += (double)grad_feature[pix * OC_NFEAT + k];
    }
    for (size_t i = 0; i < (size_t)P * OC_NFEAT; ++i) dL_dfeature_vector[i] = (float)acc[i];
    free(acc);
}
