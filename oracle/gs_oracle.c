/*
 * gs_oracle.c -- CPU restatement of the reference differentiable Gaussian rasterizer.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (the CUDA library, the torch
 * binding, the Python package) may include, link or call this file.  It is the checker
 * used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg.
 *
 * Every function cites the reference file:line it restates.  Paths are relative to
 * /root/reference/submodules/diff-gaussian-rasterization (DGR).
 *
 * Parity pinning: the reference ships no tests or golden vectors for this path
 * (SURVEY.md section 4).  This oracle is pinned against (1) fixtures produced by the
 * reference's own CUDA build on a B200 (tests/golden/, generator committed next to them)
 * and (2) the reference's torch CPU paths eval_sh / build_covariance (tests/golden/).
 *
 * Floating-point contract.  The reference is compiled by nvcc with -fmad=true, so the
 * device code contracts some a*b+c into FMAs.  Which ones was read off the sm_100 SASS
 * of the reference kernels (tools/sass_expr.py); the forward per-Gaussian math below
 * uses fmaf() exactly where the reference binary has an FFMA and plain (uncontracted,
 * -ffp-contract=off) float ops elsewhere, so depth, means2D, cov3D, conic, radii and
 * tile rectangles are reproduced bit-for-bit on the host (div.rn/rcp.rn/sqrt.rn are
 * IEEE-correct on both sides).  exp() goes through MUFU.EX2 on the device, which the
 * host cannot reproduce bit-for-bit: blending and the backward pass are tolerance-level.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <limits.h>

#define BLOCK_X 16 /* DGR/cuda_rasterizer/config.h:16 */
#define BLOCK_Y 16 /* DGR/cuda_rasterizer/config.h:17 */

/* DGR/cuda_rasterizer/auxiliary.h:22-39 */
static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                               0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                               -0.5900435899266435f};

#define FMA(a, b, c) fmaf((a), (b), (c))

/* cvt.rzi.s32.f32: truncate, saturate, NaN -> 0 (what a C-style (int) cast compiles to on the GPU) */
static int f2i_rz(float x) {
    if (x != x) return 0;
    if (x >= 2147483648.0f) return INT_MAX;
    if (x <= -2147483648.0f) return INT_MIN;
    return (int)x;
}

static inline float minf_(float a, float b) { return fminf(a, b); }
static inline float maxf_(float a, float b) { return fmaxf(a, b); }

/* m0*x + m4*y + m8*z + m12 as the reference binary evaluates it (auxiliary.h:60-79) */
static inline float dot4(float m0, float m4, float m8, float m12, float x, float y, float z) {
    return FMA(z, m8, FMA(x, m0, y * m4)) + m12;
}
/* a0*b0 + a1*b1 + a2*b2 as the reference binary evaluates every GLM mat3 product term */
static inline float dot3(float a0, float b0, float a1, float b1, float a2, float b2) {
    return FMA(a2, b2, FMA(a0, b0, a1 * b1));
}

/* DGR/cuda_rasterizer/auxiliary.h:42-45 -- evaluated in double, narrowed once */
static inline float ndc2Pix(float v, int S) {
    return (float)(fma((double)v + 1.0, (double)S, -1.0) * 0.5);
}

/* DGR/cuda_rasterizer/auxiliary.h:48-58 */
static void getRect(float px, float py, int max_radius, uint32_t gx, uint32_t gy,
                    uint32_t* minx, uint32_t* miny, uint32_t* maxx, uint32_t* maxy) {
    float r = (float)max_radius;
    int a = f2i_rz((px - r) * 0.0625f);
    int b = f2i_rz((py - r) * 0.0625f);
    int c = f2i_rz((((px + r) + 16.0f) + -1.0f) * 0.0625f);
    int d = f2i_rz((((py + r) + 16.0f) + -1.0f) * 0.0625f);
    uint32_t ua = (uint32_t)(a > 0 ? a : 0), ub = (uint32_t)(b > 0 ? b : 0);
    uint32_t uc = (uint32_t)(c > 0 ? c : 0), ud = (uint32_t)(d > 0 ? d : 0);
    *minx = ua < gx ? ua : gx;
    *miny = ub < gy ? ub : gy;
    *maxx = uc < gx ? uc : gx;
    *maxy = ud < gy ? ud : gy;
}

/* DGR/cuda_rasterizer/forward.cu:146-180 (computeCov3D); GLM column-major products expanded,
 * including the zero terms of the diagonal S, in the order the binary keeps them. */
static void computeCov3D(const float* scale, float mod, const float* rot, float* cov3D) {
    float r = rot[0], x = rot[1], y = rot[2], z = rot[3];
    float sx = scale[0] * mod, sy = scale[1] * mod, sz = scale[2] * mod;
    float xz = x * z, rx = r * x, rz = r * z, yy = y * y, zz = z * z;
    float R02 = FMA(r, y, xz);   /* x*z + r*y */
    float R20 = FMA(-r, y, xz);  /* x*z - r*y */
    float R12 = FMA(y, z, -rx);  /* y*z - r*x */
    float R21 = FMA(y, z, rx);   /* y*z + r*x */
    float R01 = FMA(x, y, -rz);  /* x*y - r*z */
    float R10 = FMA(x, y, rz);   /* x*y + r*z */
    float d22 = FMA(x, x, yy);   /* x*x + y*y */
    float d00 = yy + zz;         /* y*y + z*z */
    float d11 = FMA(x, x, zz);   /* x*x + z*z */
    /* R as GLM columns: c0 = (a00,a01,a02), c1 = (a10,a11,a12), c2 = (a20,a21,a22) */
    float a00 = 1.0f - (d00 + d00), a01 = R01 + R01, a02 = R02 + R02;
    float a10 = R10 + R10, a11 = 1.0f - (d11 + d11), a12 = R12 + R12;
    float a20 = R20 + R20, a21 = R21 + R21, a22 = 1.0f - (d22 + d22);
    /* M = S * R, M[c][r] = S[0][r]*R[c][0] + S[1][r]*R[c][1] + S[2][r]*R[c][2] */
    float z00 = 0.0f * a00, z11 = 0.0f * a11, z21 = 0.0f * a21;
    float M00 = FMA(0.0f, a02, FMA(0.0f, a01, sx * a00));
    float M01 = FMA(0.0f, a02, FMA(sy, a01, z00));
    float M02 = FMA(sz, a02, FMA(0.0f, a01, z00));
    float M10 = FMA(0.0f, a12, FMA(sx, a10, z11));
    float M11 = FMA(0.0f, a12, FMA(0.0f, a10, sy * a11));
    float M12 = FMA(sz, a12, FMA(0.0f, a10, z11));
    float M20 = FMA(0.0f, a22, FMA(sx, a20, z21));
    float M21 = FMA(0.0f, a22, FMA(0.0f, a20, sy * a21));
    float M22 = FMA(sz, a22, FMA(0.0f, a20, z21));
    /* Sigma = transpose(M) * M, upper triangle */
    cov3D[0] = dot3(M00, M00, M01, M01, M02, M02);
    cov3D[1] = dot3(M00, M10, M01, M11, M02, M12);
    cov3D[2] = dot3(M00, M20, M01, M21, M02, M22);
    cov3D[3] = dot3(M10, M10, M11, M11, M12, M12);
    cov3D[4] = dot3(M10, M20, M11, M21, M12, M22);
    cov3D[5] = dot3(M20, M20, M21, M21, M22, M22);
}

/* DGR/cuda_rasterizer/forward.cu:84-140 (computeCov2D) */
static void computeCov2D(const float* mean, float focal_x, float focal_y, float tan_fovx,
                         float tan_fovy, const float* c, const float* V, float* cov) {
    float tx = dot4(V[0], V[4], V[8], V[12], mean[0], mean[1], mean[2]);
    float ty = dot4(V[1], V[5], V[9], V[13], mean[0], mean[1], mean[2]);
    float tz = dot4(V[2], V[6], V[10], V[14], mean[0], mean[1], mean[2]);
    float limx = tan_fovx * 1.3f, limy = tan_fovy * 1.3f;
    float cx = minf_(maxf_(tx / tz, -limx), limx);
    float cy = minf_(maxf_(ty / tz, -limy), limy);
    float tz2 = tz * tz;
    float J00 = focal_x / tz;
    float J02 = ((tz * -cx) * focal_x) / tz2;   /* -(focal_x * t.x) / (t.z*t.z), t.x = cx*tz */
    float J11 = focal_y / tz;
    float J12 = ((tz * -cy) * focal_y) / tz2;
    /* T = W * J (GLM): T[j][i] = W[0][i]*J[j][0] + W[1][i]*J[j][1] + W[2][i]*J[j][2] */
    float T00 = FMA(V[2], J02, FMA(V[0], J00, 0.0f * V[1]));
    float T01 = FMA(V[6], J02, FMA(V[4], J00, 0.0f * V[5]));
    float T02 = FMA(V[10], J02, FMA(V[8], J00, 0.0f * V[9]));
    float T10 = FMA(V[2], J12, FMA(0.0f, V[0], V[1] * J11));
    float T11 = FMA(V[6], J12, FMA(0.0f, V[4], V[5] * J11));
    float T12 = FMA(V[10], J12, FMA(0.0f, V[8], V[9] * J11));
    /* A = transpose(T) * transpose(Vrk); cov = A * T */
    float A00 = dot3(T00, c[0], T01, c[1], T02, c[2]);
    float A10 = dot3(T00, c[1], T01, c[3], T02, c[4]);
    float A20 = dot3(T00, c[2], T01, c[4], T02, c[5]);
    float A01 = dot3(T10, c[0], T11, c[1], T12, c[2]);
    float A11 = dot3(T10, c[1], T11, c[3], T12, c[4]);
    float A21 = dot3(T10, c[2], T11, c[4], T12, c[5]);
    float cov00 = FMA(T02, A20, FMA(T00, A00, T01 * A10));
    float cov01 = FMA(T02, A21, FMA(T00, A01, T01 * A11));
    float cov11 = FMA(T12, A21, FMA(T10, A01, T11 * A11));
    cov[0] = cov00 + 0.3f;
    cov[1] = cov01;
    cov[2] = cov11 + 0.3f;
}

/* DGR/cuda_rasterizer/forward.cu:21-81 (computeColorFromSH) */
static void computeColorFromSH(int deg, const float* mean, const float* campos, const float* sh,
                               float* rgb, uint8_t* clamped) {
    float dx = mean[0] - campos[0], dy = mean[1] - campos[1], dz = mean[2] - campos[2];
    float len = sqrtf(FMA(dz, dz, FMA(dx, dx, dy * dy)));
    float x = dx / len, y = dy / len, z = dz / len;
    float res[3];
    for (int c = 0; c < 3; c++) res[c] = sh[c] * SH_C0;
    if (deg > 0) {
        float ky = y * SH_C1, kz = z * SH_C1, kx = x * SH_C1;
        for (int c = 0; c < 3; c++) {
            float v = FMA(-ky, sh[3 + c], res[c]);
            v = FMA(kz, sh[6 + c], v);
            res[c] = FMA(-kx, sh[9 + c], v);
        }
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = y * x, yz = z * y, xz = z * x;
            float k4 = xy * SH_C2[0];
            float k5 = yz * SH_C2[1];
            float zz2 = zz + zz;
            float k6 = ((zz2 - xx) - yy) * SH_C2[2];
            float k7 = xz * SH_C2[3];
            float xxyy = xx - yy;
            float k8 = xxyy * SH_C2[4];
            for (int c = 0; c < 3; c++) {
                float v = FMA(k4, sh[12 + c], res[c]);
                v = FMA(k5, sh[15 + c], v);
                v = FMA(k6, sh[18 + c], v);
                v = FMA(k7, sh[21 + c], v);
                res[c] = FMA(k8, sh[24 + c], v);
            }
            if (deg > 2) {
                float k9 = (y * SH_C3[0]) * FMA(xx, 3.0f, -yy);
                float k10 = (xy * SH_C3[1]) * z;
                float q = FMA(zz, 4.0f, -xx) - yy; /* 4zz - xx - yy */
                float k11 = (y * SH_C3[2]) * q;
                float k12 = (z * SH_C3[3]) * FMA(yy, -3.0f, FMA(xx, -3.0f, zz2));
                float k13 = q * (x * SH_C3[4]);
                float k14 = xxyy * (z * SH_C3[5]);
                float k15 = (x * SH_C3[6]) * FMA(yy, -3.0f, xx);
                for (int c = 0; c < 3; c++) {
                    float v = FMA(k9, sh[27 + c], res[c]);
                    v = FMA(k10, sh[30 + c], v);
                    v = FMA(k11, sh[33 + c], v);
                    v = FMA(k12, sh[36 + c], v);
                    v = FMA(k13, sh[39 + c], v);
                    v = FMA(k14, sh[42 + c], v);
                    res[c] = FMA(k15, sh[45 + c], v);
                }
            }
        }
    }
    for (int c = 0; c < 3; c++) {
        float v = res[c] + 0.5f;
        clamped[c] = (v < 0.0f);
        rgb[c] = (v < 0.0f) ? 0.0f : v;
    }
}

/* DGR/cuda_rasterizer/auxiliary.h:144-175 (in_frustum: only the near-plane test is live) */
void gso_mark_visible(int P, const float* means, const float* V, uint8_t* present) {
    for (int i = 0; i < P; i++) {
        const float* m = means + 3 * i;
        float z = dot4(V[2], V[6], V[10], V[14], m[0], m[1], m[2]);
        present[i] = !(z <= 0.2f);
    }
}

/*
 * DGR/cuda_rasterizer/forward.cu:192-324 (preprocessCUDA) with focal computed as in
 * DGR/cuda_rasterizer/rasterizer_impl.cu:251-252.  All outputs are per-Gaussian arrays in the
 * reference's GeometryState layout (rasterizer_impl.cu:179-197).  Arrays are written only where
 * the reference writes them (callers zero them first).
 */
void gso_preprocess(int P, int D, int M, const float* means, const float* scales, float smod,
                    const float* rots, const float* opac, const float* shs,
                    const float* cov3D_pre, const float* colors_pre, const float* V,
                    const float* Pm, const float* campos, int W, int H, float tan_fovx,
                    float tan_fovy, int32_t* radii, float* xy, float* depths, float* cov3Ds,
                    float* rgb, float* conic_opacity, uint32_t* tiles_touched, uint8_t* clamped) {
    const float focal_y = H / (2.0f * tan_fovy);
    const float focal_x = W / (2.0f * tan_fovx);
    const uint32_t gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        radii[i] = 0;
        tiles_touched[i] = 0;
        const float* m = means + 3 * i;
        float depth = dot4(V[2], V[6], V[10], V[14], m[0], m[1], m[2]);
        if (depth <= 0.2f) continue;
        float hx = dot4(Pm[0], Pm[4], Pm[8], Pm[12], m[0], m[1], m[2]);
        float hy = dot4(Pm[1], Pm[5], Pm[9], Pm[13], m[0], m[1], m[2]);
        float hw = dot4(Pm[3], Pm[7], Pm[11], Pm[15], m[0], m[1], m[2]);
        float p_w = 1.0f / (hw + 0.0000001f);
        float projx = hx * p_w, projy = hy * p_w;
        const float* c3;
        float c3buf[6];
        if (cov3D_pre) {
            c3 = cov3D_pre + 6 * i;
        } else {
            computeCov3D(scales + 3 * i, smod, rots + 4 * i, c3buf);
            if (cov3Ds) memcpy(cov3Ds + 6 * i, c3buf, sizeof c3buf);
            c3 = c3buf;
        }
        float cov[3];
        computeCov2D(m, focal_x, focal_y, tan_fovx, tan_fovy, c3, V, cov);
        float det = FMA(cov[0], cov[2], -(cov[1] * cov[1]));
        if (det == 0.0f) continue;
        float det_inv = 1.0f / det;
        float conic[3] = {cov[2] * det_inv, cov[1] * -det_inv, cov[0] * det_inv};
        float mid = (cov[0] + cov[2]) * 0.5f;
        float s = sqrtf(maxf_(FMA(mid, mid, -det), 0.1f));
        float lambda1 = mid + s, lambda2 = mid - s;
        float my_radius = ceilf(sqrtf(maxf_(lambda1, lambda2)) * 3.0f);
        int irad = f2i_rz(my_radius);
        float px = ndc2Pix(projx, W), py = ndc2Pix(projy, H);
        uint32_t x0, y0, x1, y1;
        getRect(px, py, irad, gx, gy, &x0, &y0, &x1, &y1);
        if ((x1 - x0) * (y1 - y0) == 0) continue;
        if (!colors_pre) computeColorFromSH(D, m, campos, shs + (size_t)i * M * 3, rgb + 3 * i, clamped + 3 * i);
        depths[i] = depth;
        radii[i] = irad;
        xy[2 * i] = px;
        xy[2 * i + 1] = py;
        conic_opacity[4 * i] = conic[0];
        conic_opacity[4 * i + 1] = conic[1];
        conic_opacity[4 * i + 2] = conic[2];
        conic_opacity[4 * i + 3] = opac[i];
        tiles_touched[i] = (y1 - y0) * (x1 - x0);
    }
}

/* DGR/cuda_rasterizer/rasterizer_impl.cu:37-52 (getHigherMsb) */
uint32_t gso_higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

/*
 * Binning: inclusive scan (rasterizer_impl.cu:323-325), duplicateWithKeys (:78-126), a stable
 * LSD radix sort on key bits [0, 32+bit) (:357-374; cub::DeviceRadixSort is a stable LSD sort,
 * CUDA toolkit CUB, not vendored), identifyTileRanges (:133-159, with the memset at :377).
 * Returns num_rendered; point_list/keys must hold sum(tiles_touched) entries (pass NULL to count).
 */
int64_t gso_bin(int P, int W, int H, const float* xy, const float* depths, const int32_t* radii,
                const uint32_t* tiles_touched, uint32_t* point_list, uint64_t* keys_sorted,
                uint32_t* ranges /* 2 per tile */) {
    const uint32_t gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
    int64_t R = 0;
    for (int i = 0; i < P; i++) R += tiles_touched[i];
    if (!point_list) return R;
    uint64_t* k0 = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)(R ? R : 1));
    uint32_t* v0 = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(R ? R : 1));
    uint64_t* k1 = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)(R ? R : 1));
    uint32_t* v1 = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(R ? R : 1));
    size_t off = 0;
    for (int i = 0; i < P; i++) {
        if (radii[i] > 0) {
            uint32_t x0, y0, x1, y1;
            getRect(xy[2 * i], xy[2 * i + 1], radii[i], gx, gy, &x0, &y0, &x1, &y1);
            uint32_t dbits;
            memcpy(&dbits, depths + i, 4);
            for (uint32_t y = y0; y < y1; y++)
                for (uint32_t x = x0; x < x1; x++) {
                    uint64_t key = (uint64_t)(y * gx + x);
                    key <<= 32;
                    key |= dbits;
                    k0[off] = key;
                    v0[off] = (uint32_t)i;
                    off++;
                }
        }
    }
    int end_bit = 32 + (int)gso_higher_msb(gx * gy);
    size_t* cnt = (size_t*)malloc(sizeof(size_t) * 65536);
    for (int shift = 0; shift < end_bit; shift += 16) {
        int nb = end_bit - shift < 16 ? end_bit - shift : 16;
        uint64_t mask = ((uint64_t)1 << nb) - 1;
        memset(cnt, 0, sizeof(size_t) * 65536);
        for (size_t j = 0; j < (size_t)R; j++) cnt[(k0[j] >> shift) & mask]++;
        size_t acc = 0;
        for (size_t d = 0; d <= mask; d++) { size_t c = cnt[d]; cnt[d] = acc; acc += c; }
        for (size_t j = 0; j < (size_t)R; j++) {
            size_t dst = cnt[(k0[j] >> shift) & mask]++;
            k1[dst] = k0[j];
            v1[dst] = v0[j];
        }
        uint64_t* tk = k0; k0 = k1; k1 = tk;
        uint32_t* tv = v0; v0 = v1; v1 = tv;
    }
    memcpy(point_list, v0, sizeof(uint32_t) * (size_t)R);
    if (keys_sorted) memcpy(keys_sorted, k0, sizeof(uint64_t) * (size_t)R);
    memset(ranges, 0, sizeof(uint32_t) * 2 * gx * gy);
    for (int64_t j = 0; j < R; j++) {
        uint32_t cur = (uint32_t)(k0[j] >> 32);
        if (j == 0) ranges[2 * cur] = 0;
        else {
            uint32_t prev = (uint32_t)(k0[j - 1] >> 32);
            if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)j; ranges[2 * cur] = (uint32_t)j; }
        }
        if (j == R - 1) ranges[2 * cur + 1] = (uint32_t)R;
    }
    free(k0); free(v0); free(k1); free(v1); free(cnt);
    return R;
}

/* DGR/cuda_rasterizer/forward.cu:331-485 (renderCUDA, forward).  One pixel at a time; the block
 * vote / shared-memory batching of the reference changes nothing about per-pixel results. */
void gso_render(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const float* xy,
                const float* features, const float* conic_opacity, const float* bg,
                float* out_color, float* final_T, uint32_t* n_contrib) {
    const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
#pragma omp parallel for schedule(dynamic, 1) collapse(2)
    for (int ty = 0; ty < gy; ty++)
        for (int tx = 0; tx < gx; tx++) {
            uint32_t r0 = ranges[2 * (ty * gx + tx)], r1 = ranges[2 * (ty * gx + tx) + 1];
            for (int ly = 0; ly < BLOCK_Y; ly++)
                for (int lx = 0; lx < BLOCK_X; lx++) {
                    int pxi = tx * BLOCK_X + lx, pyi = ty * BLOCK_Y + ly;
                    if (pxi >= W || pyi >= H) continue;
                    float pixx = (float)pxi, pixy = (float)pyi;
                    float T = 1.0f, C[3] = {0, 0, 0};
                    uint32_t contributor = 0, last = 0;
                    for (uint32_t k = r0; k < r1; k++) {
                        contributor++;
                        uint32_t id = point_list[k];
                        float dx = xy[2 * id] - pixx, dy = xy[2 * id + 1] - pixy;
                        const float* co = conic_opacity + 4 * id;
                        /* -0.5f*(A dx dx + C dy dy) - B dx dy, contracted as in the binary */
                        float s = FMA(dx, dx * co[0], dy * (dy * co[2]));
                        float power = FMA(s, -0.5f, -(dy * (dx * co[1])));
                        if (power > 0.0f) continue;
                        float alpha = minf_(0.99f, co[3] * expf(power));
                        if (alpha < 1.0f / 255.0f) continue;
                        float test_T = T * (1.0f - alpha);
                        if (test_T < 0.0001f) break; /* done = true */
                        for (int ch = 0; ch < 3; ch++)
                            C[ch] = FMA(T, alpha * features[3 * id + ch], C[ch]);
                        T = test_T;
                        last = contributor;
                    }
                    size_t pid = (size_t)pyi * W + pxi;
                    final_T[pid] = T;
                    n_contrib[pid] = last;
                    for (int ch = 0; ch < 3; ch++)
                        out_color[(size_t)ch * H * W + pid] = FMA(bg[ch], T, C[ch]);
                }
        }
}

/* DGR/cuda_rasterizer/backward.cu:408-601 (renderCUDA, backward).  Per-pixel replay back to
 * front; per-Gaussian sums are accumulated in double (the deterministic reference sum; the GPU
 * uses float atomics in arbitrary order).  Outputs: dL_dmean2D (P,3), dL_dconic (P,4: x,y,-,w),
 * dL_dopacity (P), dL_dcolors (P,3), all float, zero-initialised here. */
void gso_render_backward(int P, int W, int H, const uint32_t* ranges, const uint32_t* point_list,
                         const float* bg, const float* xy, const float* conic_opacity,
                         const float* colors, const float* final_T, const uint32_t* n_contrib,
                         const float* dL_dpix, float* dL_dmean2D, float* dL_dconic,
                         float* dL_dopacity, float* dL_dcolors) {
    const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
    double* acc = (double*)calloc((size_t)P * 9, sizeof(double));
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
#pragma omp parallel for schedule(dynamic, 1) collapse(2)
    for (int ty = 0; ty < gy; ty++)
        for (int tx = 0; tx < gx; tx++) {
            uint32_t r0 = ranges[2 * (ty * gx + tx)];
            for (int ly = 0; ly < BLOCK_Y; ly++)
                for (int lx = 0; lx < BLOCK_X; lx++) {
                    int pxi = tx * BLOCK_X + lx, pyi = ty * BLOCK_Y + ly;
                    if (pxi >= W || pyi >= H) continue;
                    size_t pid = (size_t)pyi * W + pxi;
                    float pixx = (float)pxi, pixy = (float)pyi;
                    const float T_final = final_T[pid];
                    float T = T_final;
                    float dpix[3], accum_rec[3] = {0, 0, 0}, last_color[3] = {0, 0, 0};
                    for (int ch = 0; ch < 3; ch++) dpix[ch] = dL_dpix[(size_t)ch * H * W + pid];
                    float last_alpha = 0.0f;
                    float bg_dot_dpixel = 0.0f;
                    for (int ch = 0; ch < 3; ch++) bg_dot_dpixel += bg[ch] * dpix[ch];
                    for (int64_t c = (int64_t)n_contrib[pid] - 1; c >= 0; c--) {
                        uint32_t id = point_list[r0 + c];
                        float dx = xy[2 * id] - pixx, dy = xy[2 * id + 1] - pixy;
                        const float* co = conic_opacity + 4 * id;
                        float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
                        if (power > 0.0f) continue;
                        float G = expf(power);
                        float alpha = minf_(0.99f, co[3] * G);
                        if (alpha < 1.0f / 255.0f) continue;
                        T = T / (1.0f - alpha);
                        float dchannel_dcolor = alpha * T;
                        float dL_dalpha = 0.0f;
                        for (int ch = 0; ch < 3; ch++) {
                            float col = colors[3 * id + ch];
                            accum_rec[ch] = last_alpha * last_color[ch] + (1.0f - last_alpha) * accum_rec[ch];
                            last_color[ch] = col;
                            dL_dalpha += (col - accum_rec[ch]) * dpix[ch];
                            double v = (double)(dchannel_dcolor * dpix[ch]);
#pragma omp atomic
                            acc[(size_t)id * 9 + 6 + ch] += v;
                        }
                        dL_dalpha *= T;
                        last_alpha = alpha;
                        dL_dalpha += (-T_final / (1.0f - alpha)) * bg_dot_dpixel;
                        float dL_dG = co[3] * dL_dalpha;
                        float gdx = G * dx, gdy = G * dy;
                        float dG_ddelx = -gdx * co[0] - gdy * co[1];
                        float dG_ddely = -gdy * co[2] - gdx * co[1];
                        double g[6] = {(double)(dL_dG * dG_ddelx * ddelx_dx),
                                       (double)(dL_dG * dG_ddely * ddely_dy),
                                       (double)(-0.5f * gdx * dx * dL_dG),
                                       (double)(-0.5f * gdx * dy * dL_dG),
                                       (double)(-0.5f * gdy * dy * dL_dG),
                                       (double)(G * dL_dalpha)};
                        for (int q = 0; q < 6; q++) {
#pragma omp atomic
                            acc[(size_t)id * 9 + q] += g[q];
                        }
                    }
                }
        }
    for (int i = 0; i < P; i++) {
        const double* a = acc + (size_t)i * 9;
        dL_dmean2D[3 * i] = (float)a[0];
        dL_dmean2D[3 * i + 1] = (float)a[1];
        dL_dmean2D[3 * i + 2] = 0.0f;
        dL_dconic[4 * i] = (float)a[2];
        dL_dconic[4 * i + 1] = (float)a[3];
        dL_dconic[4 * i + 2] = 0.0f;
        dL_dconic[4 * i + 3] = (float)a[4];
        dL_dopacity[i] = (float)a[5];
        dL_dcolors[3 * i] = (float)a[6];
        dL_dcolors[3 * i + 1] = (float)a[7];
        dL_dcolors[3 * i + 2] = (float)a[8];
    }
    free(acc);
}

/* DGR/cuda_rasterizer/auxiliary.h:109-120 (dnormvdv, float3) */
static void dnormvdv3(const float* v, const float* dv, float* out) {
    float sum2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
    float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    out[0] = ((+sum2 - v[0] * v[0]) * dv[0] - v[1] * v[0] * dv[1] - v[2] * v[0] * dv[2]) * invsum32;
    out[1] = (-v[0] * v[1] * dv[0] + (sum2 - v[1] * v[1]) * dv[1] - v[2] * v[1] * dv[2]) * invsum32;
    out[2] = (-v[0] * v[2] * dv[0] - v[1] * v[2] * dv[1] + (sum2 - v[2] * v[2]) * dv[2]) * invsum32;
}

/* DGR/cuda_rasterizer/backward.cu:20-139 (computeColorFromSH, backward); dL_dmean is += */
static void sh_backward(int deg, int M, const float* mean, const float* campos, const float* sh,
                        const uint8_t* clamped, const float* dL_dcolor, float* dL_dmean,
                        float* dL_dsh) {
    float dir_orig[3] = {mean[0] - campos[0], mean[1] - campos[1], mean[2] - campos[2]};
    float len = sqrtf(dir_orig[0] * dir_orig[0] + dir_orig[1] * dir_orig[1] + dir_orig[2] * dir_orig[2]);
    float x = dir_orig[0] / len, y = dir_orig[1] / len, z = dir_orig[2] / len;
    float dRGB[3];
    for (int c = 0; c < 3; c++) dRGB[c] = dL_dcolor[c] * (clamped[c] ? 0.0f : 1.0f);
    float dx_[3] = {0, 0, 0}, dy_[3] = {0, 0, 0}, dz_[3] = {0, 0, 0};
    (void)M;
#define SHV(k, c) sh[3 * (k) + (c)]
#define DSH(k, w) for (int c = 0; c < 3; c++) dL_dsh[3 * (k) + c] = (w) * dRGB[c]
    DSH(0, SH_C0);
    if (deg > 0) {
        DSH(1, -SH_C1 * y);
        DSH(2, SH_C1 * z);
        DSH(3, -SH_C1 * x);
        for (int c = 0; c < 3; c++) {
            dx_[c] = -SH_C1 * SHV(3, c);
            dy_[c] = -SH_C1 * SHV(1, c);
            dz_[c] = SH_C1 * SHV(2, c);
        }
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            DSH(4, SH_C2[0] * xy);
            DSH(5, SH_C2[1] * yz);
            DSH(6, SH_C2[2] * (2.f * zz - xx - yy));
            DSH(7, SH_C2[3] * xz);
            DSH(8, SH_C2[4] * (xx - yy));
            for (int c = 0; c < 3; c++) {
                dx_[c] += SH_C2[0] * y * SHV(4, c) + SH_C2[2] * 2.f * -x * SHV(6, c) + SH_C2[3] * z * SHV(7, c) + SH_C2[4] * 2.f * x * SHV(8, c);
                dy_[c] += SH_C2[0] * x * SHV(4, c) + SH_C2[1] * z * SHV(5, c) + SH_C2[2] * 2.f * -y * SHV(6, c) + SH_C2[4] * 2.f * -y * SHV(8, c);
                dz_[c] += SH_C2[1] * y * SHV(5, c) + SH_C2[2] * 2.f * 2.f * z * SHV(6, c) + SH_C2[3] * x * SHV(7, c);
            }
            if (deg > 2) {
                DSH(9, SH_C3[0] * y * (3.f * xx - yy));
                DSH(10, SH_C3[1] * xy * z);
                DSH(11, SH_C3[2] * y * (4.f * zz - xx - yy));
                DSH(12, SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
                DSH(13, SH_C3[4] * x * (4.f * zz - xx - yy));
                DSH(14, SH_C3[5] * z * (xx - yy));
                DSH(15, SH_C3[6] * x * (xx - 3.f * yy));
                for (int c = 0; c < 3; c++) {
                    dx_[c] += (SH_C3[0] * SHV(9, c) * 3.f * 2.f * xy + SH_C3[1] * SHV(10, c) * yz +
                               SH_C3[2] * SHV(11, c) * -2.f * xy + SH_C3[3] * SHV(12, c) * -3.f * 2.f * xz +
                               SH_C3[4] * SHV(13, c) * (-3.f * xx + 4.f * zz - yy) +
                               SH_C3[5] * SHV(14, c) * 2.f * xz + SH_C3[6] * SHV(15, c) * 3.f * (xx - yy));
                    dy_[c] += (SH_C3[0] * SHV(9, c) * 3.f * (xx - yy) + SH_C3[1] * SHV(10, c) * xz +
                               SH_C3[2] * SHV(11, c) * (-3.f * yy + 4.f * zz - xx) +
                               SH_C3[3] * SHV(12, c) * -3.f * 2.f * yz + SH_C3[4] * SHV(13, c) * -2.f * xy +
                               SH_C3[5] * SHV(14, c) * -2.f * yz + SH_C3[6] * SHV(15, c) * -3.f * 2.f * xy);
                    dz_[c] += (SH_C3[1] * SHV(10, c) * xy + SH_C3[2] * SHV(11, c) * 4.f * 2.f * yz +
                               SH_C3[3] * SHV(12, c) * 3.f * (2.f * zz - xx - yy) +
                               SH_C3[4] * SHV(13, c) * 4.f * 2.f * xz + SH_C3[5] * SHV(14, c) * (xx - yy));
                }
            }
        }
    }
#undef SHV
#undef DSH
    float dL_ddir[3] = {dx_[0] * dRGB[0] + dx_[1] * dRGB[1] + dx_[2] * dRGB[2],
                        dy_[0] * dRGB[0] + dy_[1] * dRGB[1] + dy_[2] * dRGB[2],
                        dz_[0] * dRGB[0] + dz_[1] * dRGB[1] + dz_[2] * dRGB[2]};
    float dm[3];
    dnormvdv3(dir_orig, dL_ddir, dm);
    dL_dmean[0] += dm[0];
    dL_dmean[1] += dm[1];
    dL_dmean[2] += dm[2];
}

/* GLM helpers for the backward restatement: mat3 stored column-major, m[c][r] */
typedef struct { float m[3][3]; } mat3;
static mat3 mat3_mul(mat3 a, mat3 b) { /* glm operator*: r[j][i] = a[0][i]*b[j][0] + a[1][i]*b[j][1] + a[2][i]*b[j][2] */
    mat3 r;
    for (int j = 0; j < 3; j++)
        for (int i = 0; i < 3; i++)
            r.m[j][i] = a.m[0][i] * b.m[j][0] + a.m[1][i] * b.m[j][1] + a.m[2][i] * b.m[j][2];
    return r;
}
static mat3 mat3_T(mat3 a) {
    mat3 r;
    for (int j = 0; j < 3; j++)
        for (int i = 0; i < 3; i++) r.m[j][i] = a.m[i][j];
    return r;
}

/* DGR/cuda_rasterizer/backward.cu:144-277 (computeCov2DCUDA) for one Gaussian */
static void cov2d_backward(const float* mean, const float* cov3D, float h_x, float h_y,
                           float tan_fovx, float tan_fovy, const float* V, const float* dL_dconic4,
                           float* dL_dmean /* assigned */, float* dL_dcov /* 6 */) {
    float dL_dconic[3] = {dL_dconic4[0], dL_dconic4[1], dL_dconic4[3]};
    float t[3] = {V[0] * mean[0] + V[4] * mean[1] + V[8] * mean[2] + V[12],
                  V[1] * mean[0] + V[5] * mean[1] + V[9] * mean[2] + V[13],
                  V[2] * mean[0] + V[6] * mean[1] + V[10] * mean[2] + V[14]};
    const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
    const float txtz = t[0] / t[2], tytz = t[1] / t[2];
    t[0] = minf_(limx, maxf_(-limx, txtz)) * t[2];
    t[1] = minf_(limy, maxf_(-limy, tytz)) * t[2];
    const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.0f : 1.0f;
    const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.0f : 1.0f;
    mat3 J = {{{h_x / t[2], 0.0f, -(h_x * t[0]) / (t[2] * t[2])},
               {0.0f, h_y / t[2], -(h_y * t[1]) / (t[2] * t[2])},
               {0, 0, 0}}};
    mat3 Wm = {{{V[0], V[4], V[8]}, {V[1], V[5], V[9]}, {V[2], V[6], V[10]}}};
    mat3 Vrk = {{{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}}};
    mat3 T = mat3_mul(Wm, J);
    mat3 cov2D = mat3_mul(mat3_mul(mat3_T(T), mat3_T(Vrk)), T);
    float a = cov2D.m[0][0] += 0.3f;
    float b = cov2D.m[0][1];
    float c = cov2D.m[1][1] += 0.3f;
    float denom = a * c - b * b;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
#define TT(i, j) T.m[i][j]
    if (denom2inv != 0) {
        dL_da = denom2inv * (-c * c * dL_dconic[0] + 2 * b * c * dL_dconic[1] + (denom - a * c) * dL_dconic[2]);
        dL_dc = denom2inv * (-a * a * dL_dconic[2] + 2 * a * b * dL_dconic[1] + (denom - a * c) * dL_dconic[0]);
        dL_db = denom2inv * 2 * (b * c * dL_dconic[0] - (denom + 2 * b * b) * dL_dconic[1] + a * b * dL_dconic[2]);
        dL_dcov[0] = (TT(0, 0) * TT(0, 0) * dL_da + TT(0, 0) * TT(1, 0) * dL_db + TT(1, 0) * TT(1, 0) * dL_dc);
        dL_dcov[3] = (TT(0, 1) * TT(0, 1) * dL_da + TT(0, 1) * TT(1, 1) * dL_db + TT(1, 1) * TT(1, 1) * dL_dc);
        dL_dcov[5] = (TT(0, 2) * TT(0, 2) * dL_da + TT(0, 2) * TT(1, 2) * dL_db + TT(1, 2) * TT(1, 2) * dL_dc);
        dL_dcov[1] = 2 * TT(0, 0) * TT(0, 1) * dL_da + (TT(0, 0) * TT(1, 1) + TT(0, 1) * TT(1, 0)) * dL_db + 2 * TT(1, 0) * TT(1, 1) * dL_dc;
        dL_dcov[2] = 2 * TT(0, 0) * TT(0, 2) * dL_da + (TT(0, 0) * TT(1, 2) + TT(0, 2) * TT(1, 0)) * dL_db + 2 * TT(1, 0) * TT(1, 2) * dL_dc;
        dL_dcov[4] = 2 * TT(0, 2) * TT(0, 1) * dL_da + (TT(0, 1) * TT(1, 2) + TT(0, 2) * TT(1, 1)) * dL_db + 2 * TT(1, 1) * TT(1, 2) * dL_dc;
    } else {
        for (int i = 0; i < 6; i++) dL_dcov[i] = 0;
    }
#define VR(i, j) Vrk.m[i][j]
    float dL_dT00 = 2 * (TT(0, 0) * VR(0, 0) + TT(0, 1) * VR(0, 1) + TT(0, 2) * VR(0, 2)) * dL_da + (TT(1, 0) * VR(0, 0) + TT(1, 1) * VR(0, 1) + TT(1, 2) * VR(0, 2)) * dL_db;
    float dL_dT01 = 2 * (TT(0, 0) * VR(1, 0) + TT(0, 1) * VR(1, 1) + TT(0, 2) * VR(1, 2)) * dL_da + (TT(1, 0) * VR(1, 0) + TT(1, 1) * VR(1, 1) + TT(1, 2) * VR(1, 2)) * dL_db;
    float dL_dT02 = 2 * (TT(0, 0) * VR(2, 0) + TT(0, 1) * VR(2, 1) + TT(0, 2) * VR(2, 2)) * dL_da + (TT(1, 0) * VR(2, 0) + TT(1, 1) * VR(2, 1) + TT(1, 2) * VR(2, 2)) * dL_db;
    float dL_dT10 = 2 * (TT(1, 0) * VR(0, 0) + TT(1, 1) * VR(0, 1) + TT(1, 2) * VR(0, 2)) * dL_dc + (TT(0, 0) * VR(0, 0) + TT(0, 1) * VR(0, 1) + TT(0, 2) * VR(0, 2)) * dL_db;
    float dL_dT11 = 2 * (TT(1, 0) * VR(1, 0) + TT(1, 1) * VR(1, 1) + TT(1, 2) * VR(1, 2)) * dL_dc + (TT(0, 0) * VR(1, 0) + TT(0, 1) * VR(1, 1) + TT(0, 2) * VR(1, 2)) * dL_db;
    float dL_dT12 = 2 * (TT(1, 0) * VR(2, 0) + TT(1, 1) * VR(2, 1) + TT(1, 2) * VR(2, 2)) * dL_dc + (TT(0, 0) * VR(2, 0) + TT(0, 1) * VR(2, 1) + TT(0, 2) * VR(2, 2)) * dL_db;
#define WW(i, j) Wm.m[i][j]
    float dL_dJ00 = WW(0, 0) * dL_dT00 + WW(0, 1) * dL_dT01 + WW(0, 2) * dL_dT02;
    float dL_dJ02 = WW(2, 0) * dL_dT00 + WW(2, 1) * dL_dT01 + WW(2, 2) * dL_dT02;
    float dL_dJ11 = WW(1, 0) * dL_dT10 + WW(1, 1) * dL_dT11 + WW(1, 2) * dL_dT12;
    float dL_dJ12 = WW(2, 0) * dL_dT10 + WW(2, 1) * dL_dT11 + WW(2, 2) * dL_dT12;
#undef TT
#undef VR
#undef WW
    float tz = 1.f / t[2], tz2 = tz * tz, tz3 = tz2 * tz;
    float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
    float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
    float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t[0]) * tz3 * dL_dJ02 + (2 * h_y * t[1]) * tz3 * dL_dJ12;
    /* transformVec4x3Transpose, auxiliary.h:91-99 */
    dL_dmean[0] = V[0] * dL_dtx + V[1] * dL_dty + V[2] * dL_dtz;
    dL_dmean[1] = V[4] * dL_dtx + V[5] * dL_dty + V[6] * dL_dtz;
    dL_dmean[2] = V[8] * dL_dtx + V[9] * dL_dty + V[10] * dL_dtz;
}

/* DGR/cuda_rasterizer/backward.cu:281-344 (computeCov3D, backward) */
static void cov3d_backward(const float* scale, float mod, const float* rot, const float* dL_dcov3D,
                           float* dL_dscale, float* dL_drot) {
    float r = rot[0], x = rot[1], y = rot[2], z = rot[3];
    mat3 R = {{{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
               {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
               {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}}};
    mat3 S = {{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}};
    float s[3] = {mod * scale[0], mod * scale[1], mod * scale[2]};
    S.m[0][0] = s[0]; S.m[1][1] = s[1]; S.m[2][2] = s[2];
    mat3 M = mat3_mul(S, R);
    mat3 dL_dSigma = {{{dL_dcov3D[0], 0.5f * dL_dcov3D[1], 0.5f * dL_dcov3D[2]},
                       {0.5f * dL_dcov3D[1], dL_dcov3D[3], 0.5f * dL_dcov3D[4]},
                       {0.5f * dL_dcov3D[2], 0.5f * dL_dcov3D[4], dL_dcov3D[5]}}};
    mat3 M2;
    for (int j = 0; j < 3; j++) for (int i = 0; i < 3; i++) M2.m[j][i] = 2.0f * M.m[j][i];
    mat3 dL_dM = mat3_mul(M2, dL_dSigma);
    mat3 Rt = mat3_T(R);
    mat3 dL_dMt = mat3_T(dL_dM);
    for (int k = 0; k < 3; k++)
        dL_dscale[k] = Rt.m[k][0] * dL_dMt.m[k][0] + Rt.m[k][1] * dL_dMt.m[k][1] + Rt.m[k][2] * dL_dMt.m[k][2];
    for (int k = 0; k < 3; k++) for (int i = 0; i < 3; i++) dL_dMt.m[k][i] *= s[k];
#define D(i, j) dL_dMt.m[i][j]
    dL_drot[0] = 2 * z * (D(0, 1) - D(1, 0)) + 2 * y * (D(2, 0) - D(0, 2)) + 2 * x * (D(1, 2) - D(2, 1));
    dL_drot[1] = 2 * y * (D(1, 0) + D(0, 1)) + 2 * z * (D(2, 0) + D(0, 2)) + 2 * r * (D(1, 2) - D(2, 1)) - 4 * x * (D(2, 2) + D(1, 1));
    dL_drot[2] = 2 * x * (D(1, 0) + D(0, 1)) + 2 * r * (D(2, 0) - D(0, 2)) + 2 * z * (D(1, 2) + D(2, 1)) - 4 * y * (D(2, 2) + D(0, 0));
    dL_drot[3] = 2 * r * (D(0, 1) - D(1, 0)) + 2 * x * (D(2, 0) + D(0, 2)) + 2 * y * (D(1, 2) + D(2, 1)) - 4 * z * (D(1, 1) + D(0, 0));
#undef D
}

/*
 * DGR/cuda_rasterizer/backward.cu:610-679 (BACKWARD::preprocess = computeCov2DCUDA then
 * preprocessCUDA<3>), focal as rasterizer_impl.cu:459-460.  All output arrays must be
 * zero-initialised by the caller (the reference's torch::zeros, rasterize_points.cu:168-178).
 */
void gso_preprocess_backward(int P, int D, int M, const float* means, const int32_t* radii,
                             const float* shs, const uint8_t* clamped, const float* scales,
                             const float* rots, float smod, const float* cov3Ds, const float* V,
                             const float* Pm, int W, int H, float tan_fovx, float tan_fovy,
                             const float* campos, const float* dL_dmean2D, const float* dL_dconic,
                             float* dL_dmean3D, float* dL_dcolor, float* dL_dcov3D, float* dL_dsh,
                             float* dL_dscale, float* dL_drot) {
    const float focal_y = H / (2.0f * tan_fovy);
    const float focal_x = W / (2.0f * tan_fovx);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        if (!(radii[i] > 0)) continue;
        const float* m = means + 3 * i;
        cov2d_backward(m, cov3Ds + 6 * i, focal_x, focal_y, tan_fovx, tan_fovy, V, dL_dconic + 4 * i,
                       dL_dmean3D + 3 * i, dL_dcov3D + 6 * i);
        /* backward.cu:373-390 */
        float m_w = 1.0f / ((Pm[3] * m[0] + Pm[7] * m[1] + Pm[11] * m[2] + Pm[15]) + 0.0000001f);
        float mul1 = (Pm[0] * m[0] + Pm[4] * m[1] + Pm[8] * m[2] + Pm[12]) * m_w * m_w;
        float mul2 = (Pm[1] * m[0] + Pm[5] * m[1] + Pm[9] * m[2] + Pm[13]) * m_w * m_w;
        const float* g2 = dL_dmean2D + 3 * i;
        dL_dmean3D[3 * i + 0] += (Pm[0] * m_w - Pm[3] * mul1) * g2[0] + (Pm[1] * m_w - Pm[3] * mul2) * g2[1];
        dL_dmean3D[3 * i + 1] += (Pm[4] * m_w - Pm[7] * mul1) * g2[0] + (Pm[5] * m_w - Pm[7] * mul2) * g2[1];
        dL_dmean3D[3 * i + 2] += (Pm[8] * m_w - Pm[11] * mul1) * g2[0] + (Pm[9] * m_w - Pm[11] * mul2) * g2[1];
        if (shs)
            sh_backward(D, M, m, campos, shs + (size_t)i * M * 3, clamped + 3 * i, dL_dcolor + 3 * i,
                        dL_dmean3D + 3 * i, dL_dsh + (size_t)i * M * 3);
        if (scales)
            cov3d_backward(scales + 3 * i, smod, rots + 4 * i, dL_dcov3D + 6 * i, dL_dscale + 3 * i, dL_drot + 4 * i);
    }
}

/* Reference CPU-side torch path restated for timing/cross-checks:
 * utils/sh_utils.py:57-112 (eval_sh) + gaussian_renderer/__init__.py:93-99 are covered by
 * computeColorFromSH above; exposed here for the golden-vector test. */
void gso_sh_to_rgb(int P, int D, int M, const float* means, const float* campos, const float* shs,
                   float* rgb, uint8_t* clamped) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++)
        computeColorFromSH(D, means + 3 * i, campos, shs + (size_t)i * M * 3, rgb + 3 * i, clamped + 3 * i);
}

/* scene/gaussian_model.py:32-37 + utils/general_utils.py:78-150 (build_covariance_from_scaling_rotation) */
void gso_cov3d(int P, const float* scales, float smod, const float* rots, float* cov3D) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) computeCov3D(scales + 3 * i, smod, rots + 4 * i, cov3D + 6 * i);
}
