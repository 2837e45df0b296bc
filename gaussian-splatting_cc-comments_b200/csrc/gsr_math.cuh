// gsr_math.cuh -- per-Gaussian forward math with pinned rounding.
//
// radii, tile rectangles and depth bits must match the reference's nvcc-built kernels bit for
// bit (north_star), so every operation here is an explicit round-to-nearest intrinsic
// (__fmaf_rn / __fmul_rn / __fadd_rn ...) which neither cicc nor ptxas may re-associate or
// contract.  WHICH products are fused was read off the sm_100 SASS of the reference
// (tools/sass_expr.py): nvcc -fmad=true turns  a0*b0 + a1*b1 + a2*b2  into
// fma(a2,b2, fma(a0,b0, a1*b1)), and ptxas additionally fuses a few mul/sub pairs
// (det, mid^2-det, the quaternion products, some SH polynomials).
// Semantics restated: DGR/cuda_rasterizer/auxiliary.h:42-79, forward.cu:21-180,271-301.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gsr {

#define GSR_DI __device__ __forceinline__

GSR_DI float mulr(float a, float b) { return __fmul_rn(a, b); }
GSR_DI float addr(float a, float b) { return __fadd_rn(a, b); }
GSR_DI float subr(float a, float b) { return __fsub_rn(a, b); }
GSR_DI float fmar(float a, float b, float c) { return __fmaf_rn(a, b, c); }

// m0*x + m4*y + m8*z + m12  (transformPoint4x3/4x4 rows, auxiliary.h:60-79)
GSR_DI float dot4(float m0, float m4, float m8, float m12, float x, float y, float z) {
    return addr(fmar(z, m8, fmar(x, m0, mulr(y, m4))), m12);
}
// a0*b0 + a1*b1 + a2*b2 as every GLM mat3 product term is evaluated in the reference binary
GSR_DI float dot3(float a0, float b0, float a1, float b1, float a2, float b2) {
    return fmar(a2, b2, fmar(a0, b0, mulr(a1, b1)));
}

// auxiliary.h:42-45: ((v + 1.0) * S - 1.0) * 0.5 in double (one DFMA), narrowed once
GSR_DI float ndc2pix(float v, int S) {
    double t = __fma_rn(__dadd_rn((double)v, 1.0), (double)S, -1.0);
    return __double2float_rn(__dmul_rn(t, 0.5));
}

// auxiliary.h:48-58 (getRect); (int) casts are cvt.rzi.s32.f32 (saturating, NaN -> 0)
GSR_DI void get_rect(float px, float py, int max_radius, uint32_t gx, uint32_t gy,
                     uint32_t& x0, uint32_t& y0, uint32_t& x1, uint32_t& y1) {
    float r = __int2float_rn(max_radius);
    int a = __float2int_rz(mulr(subr(px, r), 0.0625f));
    int b = __float2int_rz(mulr(subr(py, r), 0.0625f));
    int c = __float2int_rz(mulr(addr(addr(addr(px, r), 16.0f), -1.0f), 0.0625f));
    int d = __float2int_rz(mulr(addr(addr(addr(py, r), 16.0f), -1.0f), 0.0625f));
    x0 = min(gx, (uint32_t)max(0, a));
    y0 = min(gy, (uint32_t)max(0, b));
    x1 = min(gx, (uint32_t)max(0, c));
    y1 = min(gy, (uint32_t)max(0, d));
}

// forward.cu:146-180 (computeCov3D).  rot = (r,x,y,z) un-normalised, S = diag(mod*scale).
// The zero terms of the diagonal S are kept: x*0 is not foldable without fast-math and the
// reference binary keeps them too (they only matter for the sign of zero / non-finite inputs).
GSR_DI void cov3d_from_scale_rot(float sx0, float sy0, float sz0, float mod, float4 q, float* c) {
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    const float sx = mulr(sx0, mod), sy = mulr(sy0, mod), sz = mulr(sz0, mod);
    const float xz = mulr(x, z), rx = mulr(r, x), rz = mulr(r, z), yy = mulr(y, y), zz = mulr(z, z);
    const float R02 = fmar(r, y, xz);
    const float R20 = fmar(-r, y, xz);
    const float R12 = fmar(y, z, -rx);
    const float R21 = fmar(y, z, rx);
    const float R01 = fmar(x, y, -rz);
    const float R10 = fmar(x, y, rz);
    const float d22 = fmar(x, x, yy);
    const float d00 = addr(yy, zz);
    const float d11 = fmar(x, x, zz);
    const float a00 = subr(1.0f, addr(d00, d00)), a01 = addr(R01, R01), a02 = addr(R02, R02);
    const float a10 = addr(R10, R10), a11 = subr(1.0f, addr(d11, d11)), a12 = addr(R12, R12);
    const float a20 = addr(R20, R20), a21 = addr(R21, R21), a22 = subr(1.0f, addr(d22, d22));
    const float z00 = mulr(0.0f, a00), z11 = mulr(0.0f, a11), z21 = mulr(0.0f, a21);
    const float M00 = fmar(0.0f, a02, fmar(0.0f, a01, mulr(sx, a00)));
    const float M01 = fmar(0.0f, a02, fmar(sy, a01, z00));
    const float M02 = fmar(sz, a02, fmar(0.0f, a01, z00));
    const float M10 = fmar(0.0f, a12, fmar(sx, a10, z11));
    const float M11 = fmar(0.0f, a12, fmar(0.0f, a10, mulr(sy, a11)));
    const float M12 = fmar(sz, a12, fmar(0.0f, a10, z11));
    const float M20 = fmar(0.0f, a22, fmar(sx, a20, z21));
    const float M21 = fmar(0.0f, a22, fmar(0.0f, a20, mulr(sy, a21)));
    const float M22 = fmar(sz, a22, fmar(0.0f, a20, z21));
    c[0] = dot3(M00, M00, M01, M01, M02, M02);
    c[1] = dot3(M00, M10, M01, M11, M02, M12);
    c[2] = dot3(M00, M20, M01, M21, M02, M22);
    c[3] = dot3(M10, M10, M11, M11, M12, M12);
    c[4] = dot3(M10, M20, M11, M21, M12, M22);
    c[5] = dot3(M20, M20, M21, M21, M22, M22);
}

// Intermediate T = W*J of the EWA projection (forward.cu:100-123); shared by forward and backward.
struct EwaT { float T00, T01, T02, T10, T11, T12; float tx, ty, tz; float cx, cy; };

GSR_DI EwaT ewa_T(float mx, float my, float mz, float focal_x, float focal_y, float tan_fovx,
                  float tan_fovy, const float* V) {
    EwaT e;
    e.tx = dot4(V[0], V[4], V[8], V[12], mx, my, mz);
    e.ty = dot4(V[1], V[5], V[9], V[13], mx, my, mz);
    e.tz = dot4(V[2], V[6], V[10], V[14], mx, my, mz);
    const float limx = mulr(tan_fovx, 1.3f), limy = mulr(tan_fovy, 1.3f);
    e.cx = fminf(fmaxf(__fdiv_rn(e.tx, e.tz), -limx), limx);
    e.cy = fminf(fmaxf(__fdiv_rn(e.ty, e.tz), -limy), limy);
    const float tz2 = mulr(e.tz, e.tz);
    const float J00 = __fdiv_rn(focal_x, e.tz);
    const float J02 = __fdiv_rn(mulr(mulr(e.tz, -e.cx), focal_x), tz2);
    const float J11 = __fdiv_rn(focal_y, e.tz);
    const float J12 = __fdiv_rn(mulr(mulr(e.tz, -e.cy), focal_y), tz2);
    e.T00 = fmar(V[2], J02, fmar(V[0], J00, mulr(0.0f, V[1])));
    e.T01 = fmar(V[6], J02, fmar(V[4], J00, mulr(0.0f, V[5])));
    e.T02 = fmar(V[10], J02, fmar(V[8], J00, mulr(0.0f, V[9])));
    e.T10 = fmar(V[2], J12, fmar(0.0f, V[0], mulr(V[1], J11)));
    e.T11 = fmar(V[6], J12, fmar(0.0f, V[4], mulr(V[5], J11)));
    e.T12 = fmar(V[10], J12, fmar(0.0f, V[8], mulr(V[9], J11)));
    return e;
}

// forward.cu:125-139: cov = T^T * Vrk^T * T, + 0.3 on the diagonal.  Returns (cov00, cov01, cov11).
GSR_DI float3 ewa_cov2d(const EwaT& e, const float* c) {
    const float A00 = dot3(e.T00, c[0], e.T01, c[1], e.T02, c[2]);
    const float A10 = dot3(e.T00, c[1], e.T01, c[3], e.T02, c[4]);
    const float A20 = dot3(e.T00, c[2], e.T01, c[4], e.T02, c[5]);
    const float A01 = dot3(e.T10, c[0], e.T11, c[1], e.T12, c[2]);
    const float A11 = dot3(e.T10, c[1], e.T11, c[3], e.T12, c[4]);
    const float A21 = dot3(e.T10, c[2], e.T11, c[4], e.T12, c[5]);
    const float cov00 = fmar(e.T02, A20, fmar(e.T00, A00, mulr(e.T01, A10)));
    const float cov01 = fmar(e.T02, A21, fmar(e.T00, A01, mulr(e.T01, A11)));
    const float cov11 = fmar(e.T12, A21, fmar(e.T10, A01, mulr(e.T11, A11)));
    return make_float3(addr(cov00, 0.3f), cov01, addr(cov11, 0.3f));
}

// SH constants (auxiliary.h:22-39)
#define GSR_SH_C0 0.28209479177387814f
#define GSR_SH_C1 0.4886025119029199f
#define GSR_SH_C2_0 1.0925484305920792f
#define GSR_SH_C2_1 -1.0925484305920792f
#define GSR_SH_C2_2 0.31539156525252005f
#define GSR_SH_C2_3 -1.0925484305920792f
#define GSR_SH_C2_4 0.5462742152960396f
#define GSR_SH_C3_0 -0.5900435899266435f
#define GSR_SH_C3_1 2.890611442640554f
#define GSR_SH_C3_2 -0.4570457994644658f
#define GSR_SH_C3_3 0.3731763325901154f
#define GSR_SH_C3_4 -0.4570457994644658f
#define GSR_SH_C3_5 1.445305721320277f
#define GSR_SH_C3_6 -0.5900435899266435f

// Basis weights k[0..K-1] of forward.cu:40-69 for the normalised direction (x,y,z),
// in the association order of the reference binary.  result = sum_k k[k]*sh[k] (chained FMAs).
GSR_DI void sh_basis(int deg, float x, float y, float z, float* k) {
    k[0] = GSR_SH_C0;
    if (deg > 0) {
        k[1] = -mulr(y, GSR_SH_C1);
        k[2] = mulr(z, GSR_SH_C1);
        k[3] = -mulr(x, GSR_SH_C1);
        if (deg > 1) {
            const float xx = mulr(x, x), yy = mulr(y, y), zz = mulr(z, z);
            const float xy = mulr(y, x), yz = mulr(z, y), xz = mulr(z, x);
            const float zz2 = addr(zz, zz);
            const float xxyy = subr(xx, yy);
            k[4] = mulr(xy, GSR_SH_C2_0);
            k[5] = mulr(yz, GSR_SH_C2_1);
            k[6] = mulr(subr(subr(zz2, xx), yy), GSR_SH_C2_2);
            k[7] = mulr(xz, GSR_SH_C2_3);
            k[8] = mulr(xxyy, GSR_SH_C2_4);
            if (deg > 2) {
                const float q = subr(fmar(zz, 4.0f, -xx), yy);
                k[9] = mulr(mulr(y, GSR_SH_C3_0), fmar(xx, 3.0f, -yy));
                k[10] = mulr(mulr(xy, GSR_SH_C3_1), z);
                k[11] = mulr(mulr(y, GSR_SH_C3_2), q);
                k[12] = mulr(mulr(z, GSR_SH_C3_3), fmar(yy, -3.0f, fmar(xx, -3.0f, zz2)));
                k[13] = mulr(q, mulr(x, GSR_SH_C3_4));
                k[14] = mulr(xxyy, mulr(z, GSR_SH_C3_5));
                k[15] = mulr(mulr(x, GSR_SH_C3_6), fmar(yy, -3.0f, xx));
            }
        }
    }
}

}  // namespace gsr
