// gsr_preprocess_bwd.cu -- backward of the per-Gaussian stage for sm_100a.
//
// Fuses computeCov2DCUDA (DGR/cuda_rasterizer/backward.cu:144-277) and preprocessCUDA<3>
// (backward) (:349-399, with computeColorFromSH :20-139 and computeCov3D :281-344) into one
// kernel, one thread per Gaussian.  It reads the (P,12) gradient accumulators the blending
// backward reduced into and WRITES every element of every output (zeros for Gaussians with
// radii <= 0), so callers need no 300 B/Gaussian zero-fill (rasterize_points.cu:168-178).
// cov3D is recomputed from scale/rotation with the forward's exact arithmetic instead of being
// stored in the forward and re-read here.  HBM-bound.
#include <cstdlib>
#include "gsr_internal.cuh"
#include "gsr_math.cuh"
#include "gsr_fused_params.cuh"
#include "gsr_smem.cuh"

namespace gsr {

struct F3 { float x, y, z; };
__device__ __forceinline__ F3 f3(float x, float y, float z) { F3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ F3 operator*(float s, F3 a) { return f3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ F3 operator+(F3 a, F3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float dot(F3 a, F3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// dnormvdv (auxiliary.h:109-120)
__device__ __forceinline__ F3 dnormvdv(F3 v, F3 dv) {
    const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
    const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    F3 r;
    r.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
    r.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
    r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
    return r;
}

// SH backward (backward.cu:20-139).  Writes dL_dsh[0..K) and returns dL/d(mean) through the
// normalised view direction.
// `sh` and `dL_dsh` may point to the SAME row (shared-memory staging): every read of sh happens before
// the first write of dL_dsh.  LD abstracts the load (read-only global path or plain shared-memory load).
template <bool SMEM>
__device__ __forceinline__ F3 sh_backward(int deg, const float* sh, F3 dir_orig, F3 dRGB, float* dL_dsh) {
    const float len = __fsqrt_rn(fmar(dir_orig.z, dir_orig.z, fmar(dir_orig.x, dir_orig.x, mulr(dir_orig.y, dir_orig.y))));
    const float x = __fdiv_rn(dir_orig.x, len), y = __fdiv_rn(dir_orig.y, len), z = __fdiv_rn(dir_orig.z, len);
    float k[16];
    sh_basis(deg, x, y, z, k);          // dRGB/dsh_k is the forward basis weight
    const int K = (deg + 1) * (deg + 1);
    F3 dx = f3(0, 0, 0), dy = f3(0, 0, 0), dz = f3(0, 0, 0);
#define SHV(j) (SMEM ? sh[(j)] : __ldg(sh + (j)))
#define SHC(i) f3(SHV(3 * (i)), SHV(3 * (i) + 1), SHV(3 * (i) + 2))
    if (deg > 0) {
        const F3 s1 = SHC(1), s2 = SHC(2), s3 = SHC(3);
        dx = -GSR_SH_C1 * s3;
        dy = -GSR_SH_C1 * s1;
        dz = GSR_SH_C1 * s2;
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            const F3 s4 = SHC(4), s5 = SHC(5), s6 = SHC(6), s7 = SHC(7), s8 = SHC(8);
            dx = dx + (GSR_SH_C2_0 * y * s4 + GSR_SH_C2_2 * 2.f * -x * s6 + GSR_SH_C2_3 * z * s7 + GSR_SH_C2_4 * 2.f * x * s8);
            dy = dy + (GSR_SH_C2_0 * x * s4 + GSR_SH_C2_1 * z * s5 + GSR_SH_C2_2 * 2.f * -y * s6 + GSR_SH_C2_4 * 2.f * -y * s8);
            dz = dz + (GSR_SH_C2_1 * y * s5 + GSR_SH_C2_2 * 2.f * 2.f * z * s6 + GSR_SH_C2_3 * x * s7);
            if (deg > 2) {
                const F3 s9 = SHC(9), s10 = SHC(10), s11 = SHC(11), s12 = SHC(12), s13 = SHC(13), s14 = SHC(14), s15 = SHC(15);
                dx = dx + (GSR_SH_C3_0 * 3.f * 2.f * xy * s9 + GSR_SH_C3_1 * yz * s10 + GSR_SH_C3_2 * -2.f * xy * s11 +
                           GSR_SH_C3_3 * -3.f * 2.f * xz * s12 + GSR_SH_C3_4 * (-3.f * xx + 4.f * zz - yy) * s13 +
                           GSR_SH_C3_5 * 2.f * xz * s14 + GSR_SH_C3_6 * 3.f * (xx - yy) * s15);
                dy = dy + (GSR_SH_C3_0 * 3.f * (xx - yy) * s9 + GSR_SH_C3_1 * xz * s10 + GSR_SH_C3_2 * (-3.f * yy + 4.f * zz - xx) * s11 +
                           GSR_SH_C3_3 * -3.f * 2.f * yz * s12 + GSR_SH_C3_4 * -2.f * xy * s13 + GSR_SH_C3_5 * -2.f * yz * s14 +
                           GSR_SH_C3_6 * -3.f * 2.f * xy * s15);
                dz = dz + (GSR_SH_C3_1 * xy * s10 + GSR_SH_C3_2 * 4.f * 2.f * yz * s11 + GSR_SH_C3_3 * 3.f * (2.f * zz - xx - yy) * s12 +
                           GSR_SH_C3_4 * 4.f * 2.f * xz * s13 + GSR_SH_C3_5 * (xx - yy) * s14);
            }
        }
    }
#undef SHC
#undef SHV
    for (int i = 0; i < K; i++) {
        dL_dsh[3 * i] = k[i] * dRGB.x;
        dL_dsh[3 * i + 1] = k[i] * dRGB.y;
        dL_dsh[3 * i + 2] = k[i] * dRGB.z;
    }
    const F3 dL_ddir = f3(dot(dx, dRGB), dot(dy, dRGB), dot(dz, dRGB));
    return dnormvdv(dir_orig, dL_ddir);
}

// SH backward on a shared-memory row, in place: the row holds sh[0..3K) on entry and dL_dsh[0..3K) (zero
// padded to a multiple of 4 floats) on exit.  Same derivative as sh_backward, written as a weighted sum
// dRGB/d{x,y,z} = sum_i w{x,y,z}_i * sh_i so that the row is consumed and overwritten one 16-byte chunk at
// a time (no 48-float register copy, no run-time indexed arrays).  Weights are d(basis_i)/d{x,y,z} of
// backward.cu:60-128; structurally zero ones are skipped at compile time.
template <int DEG>
__device__ __forceinline__ F3 sh_backward_row(uint32_t row, F3 dir_orig, F3 dRGB) {
    constexpr int K = (DEG + 1) * (DEG + 1);
    constexpr int NF = 3 * K, NV = (NF + 3) / 4;
    const float len = __fsqrt_rn(fmar(dir_orig.z, dir_orig.z, fmar(dir_orig.x, dir_orig.x, mulr(dir_orig.y, dir_orig.y))));
    const float x = __fdiv_rn(dir_orig.x, len), y = __fdiv_rn(dir_orig.y, len), z = __fdiv_rn(dir_orig.z, len);
    float k[16], wx[16], wy[16], wz[16];
    sh_basis(DEG, x, y, z, k);
#pragma unroll
    for (int i = 0; i < 16; i++) { wx[i] = 0.f; wy[i] = 0.f; wz[i] = 0.f; }
    if (DEG > 0) { wy[1] = -GSR_SH_C1; wz[2] = GSR_SH_C1; wx[3] = -GSR_SH_C1; }
    if (DEG > 1) {
        wx[4] = GSR_SH_C2_0 * y; wy[4] = GSR_SH_C2_0 * x;
        wy[5] = GSR_SH_C2_1 * z; wz[5] = GSR_SH_C2_1 * y;
        wx[6] = GSR_SH_C2_2 * 2.f * -x; wy[6] = GSR_SH_C2_2 * 2.f * -y; wz[6] = GSR_SH_C2_2 * 2.f * 2.f * z;
        wx[7] = GSR_SH_C2_3 * z; wz[7] = GSR_SH_C2_3 * x;
        wx[8] = GSR_SH_C2_4 * 2.f * x; wy[8] = GSR_SH_C2_4 * 2.f * -y;
    }
    if (DEG > 2) {
        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        wx[9] = GSR_SH_C3_0 * 3.f * 2.f * xy; wy[9] = GSR_SH_C3_0 * 3.f * (xx - yy);
        wx[10] = GSR_SH_C3_1 * yz; wy[10] = GSR_SH_C3_1 * xz; wz[10] = GSR_SH_C3_1 * xy;
        wx[11] = GSR_SH_C3_2 * -2.f * xy; wy[11] = GSR_SH_C3_2 * (-3.f * yy + 4.f * zz - xx); wz[11] = GSR_SH_C3_2 * 4.f * 2.f * yz;
        wx[12] = GSR_SH_C3_3 * -3.f * 2.f * xz; wy[12] = GSR_SH_C3_3 * -3.f * 2.f * yz; wz[12] = GSR_SH_C3_3 * 3.f * (2.f * zz - xx - yy);
        wx[13] = GSR_SH_C3_4 * (-3.f * xx + 4.f * zz - yy); wy[13] = GSR_SH_C3_4 * -2.f * xy; wz[13] = GSR_SH_C3_4 * 4.f * 2.f * xz;
        wx[14] = GSR_SH_C3_5 * 2.f * xz; wy[14] = GSR_SH_C3_5 * -2.f * yz; wz[14] = GSR_SH_C3_5 * (xx - yy);
        wx[15] = GSR_SH_C3_6 * 3.f * (xx - yy); wy[15] = GSR_SH_C3_6 * -3.f * 2.f * xy;
    }
    // structural zeros (compile-time): which weights exist at all
    constexpr unsigned HAS_X = 0xFFD8u /* i = 3,4,6,7,8,9..15 */, HAS_Y = 0xFF72u /* 1,4,5,6,8,9..15 */, HAS_Z = 0x7CE4u /* 2,5,6,7,10..14 */;
    float dxs[3] = {0.f, 0.f, 0.f}, dys[3] = {0.f, 0.f, 0.f}, dzs[3] = {0.f, 0.f, 0.f};
    const float dr[3] = {dRGB.x, dRGB.y, dRGB.z};
#pragma unroll
    for (int c4 = 0; c4 < NV; c4++) {
        const float4 v = lds128<0>(row + c4 * 16);
        const float e[4] = {v.x, v.y, v.z, v.w};
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int f = 4 * c4 + j;
            const int i = f / 3, ch = f % 3;
            if (f < NF) {
                if (i > 0) {
                    if ((HAS_X >> i) & 1u) dxs[ch] = fmaf(wx[i], e[j], dxs[ch]);
                    if ((HAS_Y >> i) & 1u) dys[ch] = fmaf(wy[i], e[j], dys[ch]);
                    if ((HAS_Z >> i) & 1u) dzs[ch] = fmaf(wz[i], e[j], dzs[ch]);
                }
                o[j] = k[i] * dr[ch];
            } else {
                o[j] = 0.f;
            }
        }
        sts128<0>(row + c4 * 16, make_float4(o[0], o[1], o[2], o[3]));
    }
    const F3 dL_ddir = f3(dxs[0] * dr[0] + dxs[1] * dr[1] + dxs[2] * dr[2], dys[0] * dr[0] + dys[1] * dr[1] + dys[2] * dr[2],
                          dzs[0] * dr[0] + dzs[1] * dr[1] + dzs[2] * dr[2]);
    return dnormvdv(dir_orig, dL_ddir);
}

// computeCov3D backward (backward.cu:281-344)
__device__ __forceinline__ void cov3d_backward(float sx0, float sy0, float sz0, float mod, float4 q, const float* g /* dL_dcov3D[6] */,
                                               float* dL_dscale, float* dL_drot) {
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    // R[c][r] as GLM columns
    const float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                           {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                           {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
    const float s[3] = {mod * sx0, mod * sy0, mod * sz0};
    const float dS[3][3] = {{g[0], 0.5f * g[1], 0.5f * g[2]}, {0.5f * g[1], g[3], 0.5f * g[4]}, {0.5f * g[2], 0.5f * g[4], g[5]}};
    // M[j][i] = s_i * R[j][i];  dL_dM[j][i] = sum_k 2*M[k][i]*dS[j][k]
    float dM[3][3];
#pragma unroll
    for (int j = 0; j < 3; j++)
#pragma unroll
        for (int i = 0; i < 3; i++)
            dM[j][i] = 2.0f * (s[i] * R[0][i]) * dS[j][0] + 2.0f * (s[i] * R[1][i]) * dS[j][1] + 2.0f * (s[i] * R[2][i]) * dS[j][2];
    // dL_dscale_k = dot(Rt[k], dL_dMt[k]) = sum_m R[m][k] * dM[m][k]
#pragma unroll
    for (int k = 0; k < 3; k++) dL_dscale[k] = R[0][k] * dM[0][k] + R[1][k] * dM[1][k] + R[2][k] * dM[2][k];
    // D(i,j) = dL_dMt[i][j] * s_i = dM[j][i] * s_i
#define D(i, j) (dM[j][i] * s[i])
    dL_drot[0] = 2 * z * (D(0, 1) - D(1, 0)) + 2 * y * (D(2, 0) - D(0, 2)) + 2 * x * (D(1, 2) - D(2, 1));
    dL_drot[1] = 2 * y * (D(1, 0) + D(0, 1)) + 2 * z * (D(2, 0) + D(0, 2)) + 2 * r * (D(1, 2) - D(2, 1)) - 4 * x * (D(2, 2) + D(1, 1));
    dL_drot[2] = 2 * x * (D(1, 0) + D(0, 1)) + 2 * r * (D(2, 0) - D(0, 2)) + 2 * z * (D(1, 2) + D(2, 1)) - 4 * y * (D(2, 2) + D(0, 0));
    dL_drot[3] = 2 * r * (D(0, 1) - D(1, 0)) + 2 * x * (D(2, 0) + D(0, 2)) + 2 * y * (D(1, 2) + D(2, 1)) - 4 * z * (D(1, 1) + D(0, 0));
#undef D
}

// Everything one Gaussian produces except dL_dsh (which goes straight to its row pointer).
struct BwdOut { float g[9]; float dmean[3]; float dcov[6]; float dscale[3]; float drot[4]; float drgb[3]; };

// The per-Gaussian backward math.  `sh_row` / `dsh_row` are this Gaussian's SH coefficients and SH-gradient
// row: global pointers in the generic kernel, the same shared-memory row in the staged kernel.
// Per-Gaussian inputs, fetched up front so their latency overlaps the cooperative SH-row load.
struct BwdIn {
    float4 g0, g1; float g8;
    float mx, my, mz;
    float4 q; float sx, sy, sz;
    uint32_t cl;
    bool vis;
    // raw-parameter path only (RAW instantiations): what the activations' derivatives need
    QuatNorm qn;          // normalisation of the raw quaternion (q holds the normalised one)
    float op;             // activated opacity, from the forward's record
};

template <bool VEC, bool RAW, bool SPLIT>
__device__ __forceinline__ void load_inputs(const PreprocessBwdArgs& a, int idx, bool vis, BwdIn& in) {
    in.vis = vis;
    in.g0 = in.g1 = in.q = make_float4(0, 0, 0, 0);
    in.g8 = in.mx = in.my = in.mz = in.sx = in.sy = in.sz = 0.f; in.cl = 0;
    if (!vis) return;
    const size_t i3 = 3 * (size_t)idx;
    const float4* ga = reinterpret_cast<const float4*>(a.grad_acc + 12 * (size_t)idx);
    in.g0 = ga[0]; in.g1 = ga[1];
    in.g8 = a.grad_acc[12 * (size_t)idx + 8];
    in.mx = __ldg(a.means3D + i3); in.my = __ldg(a.means3D + i3 + 1); in.mz = __ldg(a.means3D + i3 + 2);
    if (a.cov3D_precomp == nullptr) {
        if (VEC) in.q = __ldg(reinterpret_cast<const float4*>(a.rotations) + idx);
        else in.q = make_float4(__ldg(a.rotations + 4 * (size_t)idx), __ldg(a.rotations + 4 * (size_t)idx + 1),
                                __ldg(a.rotations + 4 * (size_t)idx + 2), __ldg(a.rotations + 4 * (size_t)idx + 3));
        in.sx = __ldg(a.scales + i3); in.sy = __ldg(a.scales + i3 + 1); in.sz = __ldg(a.scales + i3 + 2);
        if (RAW) {      // the forward's activations, same expressions (gsr_preprocess.cu)
            in.sx = act_exp(in.sx); in.sy = act_exp(in.sy); in.sz = act_exp(in.sz);
            in.qn = act_normalize(in.q);
            in.q = in.qn.q;
        }
    }
    if (RAW) in.op = reinterpret_cast<const float*>(a.packed + 3 * (size_t)idx + 1)[2];    // record b = (conic.z, thr, opacity, r)
    if (SPLIT || a.shs != nullptr) in.cl = a.clamped[idx];
}

// Returns the number of SH coefficients written to dsh_row (0 when not visible or no SHs).
template <bool SMEM, bool RAW, bool SPLIT>
__device__ __forceinline__ int bwd_one(const PreprocessBwdArgs& a, int idx, const BwdIn& in, const float* V, const float* Pm,
                                       const float* sh_row, float* dsh_row, BwdOut& o) {
#pragma unroll
    for (int i = 0; i < 9; i++) o.g[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 3; i++) { o.dmean[i] = 0.f; o.dscale[i] = 0.f; o.drgb[i] = 0.f; }
#pragma unroll
    for (int i = 0; i < 6; i++) o.dcov[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 4; i++) o.drot[i] = 0.f;
    if (!in.vis) return 0;
    float* g = o.g; float* dmean = o.dmean; float* dcov = o.dcov;
    const float4 g0 = in.g0, g1 = in.g1;
    g[5] = g1.y; g[6] = g1.z; g[7] = g1.w;
    g[8] = in.g8;
    const float mx = in.mx, my = in.my, mz = in.mz;
    float c3[6];
    const float4 q = in.q;
    const float sx = in.sx, sy = in.sy, sz = in.sz;
    if (a.cov3D_precomp != nullptr) {
#pragma unroll
        for (int i = 0; i < 6; i++) c3[i] = __ldg(a.cov3D_precomp + 6 * (size_t)idx + i);
    } else {
        cov3d_from_scale_rot(sx, sy, sz, a.scale_modifier, q, c3);
    }
    // ---- computeCov2DCUDA (backward.cu:144-277) ----
    const EwaT e = ewa_T(mx, my, mz, a.focal_x, a.focal_y, a.tan_fovx, a.tan_fovy, V);
    const float3 cov = ewa_cov2d(e, c3);
    {   // The blending backward reduced moments (gsr_render_bwd.cu); combine them with this Gaussian's conic, rebuilt
        // with the forward's exact arithmetic (gsr_preprocess.cu) instead of re-read from the 48-byte record:
        // dL_dmean2D = -(conic * [sum w dx, sum w dy]) * (W/2, H/2), dL_dconic = -1/2 [sum w dx dx, sum w dx dy, sum w dy dy]
        const float det = fmar(cov.x, cov.z, -mulr(cov.y, cov.y));
        const float det_inv = __frcp_rn(det);
        const float conx = mulr(cov.z, det_inv), cony = mulr(cov.y, -det_inv), conz = mulr(cov.x, det_inv);
        g[0] = (-g0.x * conx - g0.y * cony) * (0.5f * a.W);           // backward.cu:587
        g[1] = (-g0.y * conz - g0.x * cony) * (0.5f * a.H);           // backward.cu:588
        g[2] = -0.5f * g0.z; g[3] = -0.5f * g0.w; g[4] = -0.5f * g1.x;
    }
    const float limx = 1.3f * a.tan_fovx, limy = 1.3f * a.tan_fovy;
    const float txtz = e.tx / e.tz, tytz = e.ty / e.tz;
    const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    const float tcx = e.cx * e.tz, tcy = e.cy * e.tz;       // clamped t.x, t.y
    const float ca = cov.x, cb = cov.y, cc = cov.z;
    const float denom = ca * cc - cb * cb;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    const float dcx = g[2], dcy = g[3], dcz = g[4];           // dL_dconic .x .y .w
    if (denom2inv != 0) {
        dL_da = denom2inv * (-cc * cc * dcx + 2 * cb * cc * dcy + (denom - ca * cc) * dcz);
        dL_dc = denom2inv * (-ca * ca * dcz + 2 * ca * cb * dcy + (denom - ca * cc) * dcx);
        dL_db = denom2inv * 2 * (cb * cc * dcx - (denom + 2 * cb * cb) * dcy + ca * cb * dcz);
        dcov[0] = (e.T00 * e.T00 * dL_da + e.T00 * e.T10 * dL_db + e.T10 * e.T10 * dL_dc);
        dcov[3] = (e.T01 * e.T01 * dL_da + e.T01 * e.T11 * dL_db + e.T11 * e.T11 * dL_dc);
        dcov[5] = (e.T02 * e.T02 * dL_da + e.T02 * e.T12 * dL_db + e.T12 * e.T12 * dL_dc);
        dcov[1] = 2 * e.T00 * e.T01 * dL_da + (e.T00 * e.T11 + e.T01 * e.T10) * dL_db + 2 * e.T10 * e.T11 * dL_dc;
        dcov[2] = 2 * e.T00 * e.T02 * dL_da + (e.T00 * e.T12 + e.T02 * e.T10) * dL_db + 2 * e.T10 * e.T12 * dL_dc;
        dcov[4] = 2 * e.T02 * e.T01 * dL_da + (e.T01 * e.T12 + e.T02 * e.T11) * dL_db + 2 * e.T11 * e.T12 * dL_dc;
    }
    const float p0 = e.T00 * c3[0] + e.T01 * c3[1] + e.T02 * c3[2];
    const float p1 = e.T00 * c3[1] + e.T01 * c3[3] + e.T02 * c3[4];
    const float p2 = e.T00 * c3[2] + e.T01 * c3[4] + e.T02 * c3[5];
    const float q0 = e.T10 * c3[0] + e.T11 * c3[1] + e.T12 * c3[2];
    const float q1 = e.T10 * c3[1] + e.T11 * c3[3] + e.T12 * c3[4];
    const float q2 = e.T10 * c3[2] + e.T11 * c3[4] + e.T12 * c3[5];
    const float dT00 = 2 * p0 * dL_da + q0 * dL_db, dT01 = 2 * p1 * dL_da + q1 * dL_db, dT02 = 2 * p2 * dL_da + q2 * dL_db;
    const float dT10 = 2 * q0 * dL_dc + p0 * dL_db, dT11 = 2 * q1 * dL_dc + p1 * dL_db, dT12 = 2 * q2 * dL_dc + p2 * dL_db;
    const float dJ00 = V[0] * dT00 + V[4] * dT01 + V[8] * dT02;
    const float dJ02 = V[2] * dT00 + V[6] * dT01 + V[10] * dT02;
    const float dJ11 = V[1] * dT10 + V[5] * dT11 + V[9] * dT12;
    const float dJ12 = V[2] * dT10 + V[6] * dT11 + V[10] * dT12;
    const float tz = 1.f / e.tz, tz2 = tz * tz, tz3 = tz2 * tz;
    const float h_x = a.focal_x, h_y = a.focal_y;
    const float dtx = x_grad_mul * -h_x * tz2 * dJ02;
    const float dty = y_grad_mul * -h_y * tz2 * dJ12;
    const float dtz = -h_x * tz2 * dJ00 - h_y * tz2 * dJ11 + (2 * h_x * tcx) * tz3 * dJ02 + (2 * h_y * tcy) * tz3 * dJ12;
    dmean[0] = V[0] * dtx + V[1] * dty + V[2] * dtz;          // transformVec4x3Transpose
    dmean[1] = V[4] * dtx + V[5] * dty + V[6] * dtz;
    dmean[2] = V[8] * dtx + V[9] * dty + V[10] * dtz;
    // ---- preprocessCUDA backward: projection part (backward.cu:373-390) ----
    const float m_w = 1.0f / ((Pm[3] * mx + Pm[7] * my + Pm[11] * mz + Pm[15]) + 0.0000001f);
    const float mul1 = (Pm[0] * mx + Pm[4] * my + Pm[8] * mz + Pm[12]) * m_w * m_w;
    const float mul2 = (Pm[1] * mx + Pm[5] * my + Pm[9] * mz + Pm[13]) * m_w * m_w;
    dmean[0] += (Pm[0] * m_w - Pm[3] * mul1) * g[0] + (Pm[1] * m_w - Pm[3] * mul2) * g[1];
    dmean[1] += (Pm[4] * m_w - Pm[7] * mul1) * g[0] + (Pm[5] * m_w - Pm[7] * mul2) * g[1];
    dmean[2] += (Pm[8] * m_w - Pm[11] * mul1) * g[0] + (Pm[9] * m_w - Pm[11] * mul2) * g[1];
    // ---- SH (backward.cu:392-394) ----
    int sh_written = 0;
    if (SPLIT || a.shs != nullptr) {
        const uint32_t cl = in.cl;
        const F3 dRGB = f3((cl & 1) ? 0.f : g[6], (cl & 2) ? 0.f : g[7], (cl & 4) ? 0.f : g[8]);
        const F3 dir = f3(subr(mx, V[32]), subr(my, V[33]), subr(mz, V[34]));
        o.drgb[0] = dRGB.x; o.drgb[1] = dRGB.y; o.drgb[2] = dRGB.z;
        F3 dm;
        if (SMEM) {
            const uint32_t row = smem_addr(sh_row);
            switch (a.D) {
                case 0: dm = sh_backward_row<0>(row, dir, dRGB); break;
                case 1: dm = sh_backward_row<1>(row, dir, dRGB); break;
                case 2: dm = sh_backward_row<2>(row, dir, dRGB); break;
                default: dm = sh_backward_row<3>(row, dir, dRGB); break;
            }
        } else {
            dm = sh_backward<false>(a.D, sh_row, dir, dRGB, dsh_row);
        }
        sh_written = (a.D + 1) * (a.D + 1);
        dmean[0] += dm.x; dmean[1] += dm.y; dmean[2] += dm.z;
    }
    // ---- scale / rotation (backward.cu:396-398) ----
    if (a.scales != nullptr) cov3d_backward(sx, sy, sz, a.scale_modifier, q, dcov, o.dscale, o.drot);
    if (RAW) {      // chain rule through sigmoid / exp / normalize: the gradients leave w.r.t. the raw parameters
        g[5] = act_sigmoid_bwd(in.op, g[5]);
        if (a.scales != nullptr) {
            o.dscale[0] = act_exp_bwd(sx, o.dscale[0]); o.dscale[1] = act_exp_bwd(sy, o.dscale[1]); o.dscale[2] = act_exp_bwd(sz, o.dscale[2]);
            const float4 dr = act_normalize_bwd(in.qn, make_float4(o.drot[0], o.drot[1], o.drot[2], o.drot[3]));
            o.drot[0] = dr.x; o.drot[1] = dr.y; o.drot[2] = dr.z; o.drot[3] = dr.w;
        }
    }
    return sh_written;
}

__device__ __forceinline__ void load_camera(const PreprocessBwdArgs& a, float* s_cam) {
    if (threadIdx.x < 16) s_cam[threadIdx.x] = a.cam.view[threadIdx.x];
    else if (threadIdx.x < 32) s_cam[threadIdx.x] = a.cam.proj[threadIdx.x - 16];
    else if (threadIdx.x < 35) s_cam[threadIdx.x] = a.cam.campos[threadIdx.x - 32];
}

// ---- generic kernel: any M, any alignment; per-thread strided stores ------------------------------------
template <bool RAW>
__global__ void __launch_bounds__(256) preprocess_bwd_kernel_generic(PreprocessBwdArgs a) {
    __shared__ float s_cam[36];
    load_camera(a, s_cam);
    __syncthreads();
    const int idx = a.begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= a.P) return;
    const size_t i3 = 3 * (size_t)idx;
    float* dsh = a.dL_dsh ? a.dL_dsh + (size_t)idx * a.M * 3 : nullptr;
    BwdOut o;
    BwdIn in;
    load_inputs<false, RAW, false>(a, idx, a.radii[idx] > 0, in);
    const int sh_written = bwd_one<false, RAW, false>(a, idx, in, s_cam, s_cam + 16, a.shs ? a.shs + (size_t)idx * a.M * 3 : nullptr, dsh, o);
    // ---- write everything (zeros where the reference's pre-zeroed tensors stay untouched) ----
    if (a.dL_dconic) reinterpret_cast<float4*>(a.dL_dconic)[idx] = make_float4(o.g[2], o.g[3], 0.f, o.g[4]);
    if (a.dL_dcolor) { a.dL_dcolor[i3] = o.g[6]; a.dL_dcolor[i3 + 1] = o.g[7]; a.dL_dcolor[i3 + 2] = o.g[8]; }
    if (a.dL_dcov3D) {
#pragma unroll
        for (int i = 0; i < 6; i++) a.dL_dcov3D[6 * (size_t)idx + i] = o.dcov[i];
    }
    if (a.small_pack) {
        float* row = a.small_pack + 14 * (size_t)idx;
        row[0] = o.dmean[0]; row[1] = o.dmean[1]; row[2] = o.dmean[2];
        row[3] = o.g[0]; row[4] = o.g[1]; row[5] = 0.f;
        row[6] = o.g[5];
        row[7] = o.dscale[0]; row[8] = o.dscale[1]; row[9] = o.dscale[2];
#pragma unroll
        for (int i = 0; i < 4; i++) row[10 + i] = o.drot[i];
    } else {
        a.dL_dmean2D[i3] = o.g[0]; a.dL_dmean2D[i3 + 1] = o.g[1]; a.dL_dmean2D[i3 + 2] = 0.f;
        a.dL_dopacity[idx] = o.g[5];
        a.dL_dmean3D[i3] = o.dmean[0]; a.dL_dmean3D[i3 + 1] = o.dmean[1]; a.dL_dmean3D[i3 + 2] = o.dmean[2];
        a.dL_dscale[i3] = o.dscale[0]; a.dL_dscale[i3 + 1] = o.dscale[1]; a.dL_dscale[i3 + 2] = o.dscale[2];
#pragma unroll
        for (int i = 0; i < 4; i++) a.dL_drot[4 * (size_t)idx + i] = o.drot[i];
    }
    if (dsh) for (int i = 3 * sh_written; i < 3 * a.M; i++) dsh[i] = 0.f;
    if (a.dL_drgb) { a.dL_drgb[i3] = o.drgb[0]; a.dL_drgb[i3 + 1] = o.drgb[1]; a.dL_drgb[i3 + 2] = o.drgb[2]; }
}

// ---- staged kernel: the common case (3*M a multiple of 4, 16-byte aligned tensors) ------------------------
// 128 Gaussians per block.  The SH coefficient rows of the block's visible Gaussians are brought in with
// coalesced 16-byte loads into shared memory (odd row stride: conflict-free per-thread row walks), each
// thread overwrites its row in place with dL_dsh, and the rows leave as coalesced 16-byte stores (zeros for
// culled Gaussians and for coefficients beyond the active degree).  The 27 other floats per Gaussian are
// staged the same way, so every global store instruction of a warp covers one contiguous 512-byte span.
constexpr int PB = 128;
// Resident blocks per SM the staged kernels are compiled for.  6 -> 80 registers, no spills; at 8 (64 registers) ptxas spills
// ~100 bytes per thread and the kernel waits on its own spill stores: measured 0.319 ms at 8, 0.318 at 7, 0.284 at 6, 0.293 at
// 5, 0.322 at 4, 0.381 at 10 (C2, P = 3 M).
constexpr int PB_MINB = 6;
// floats of output staging per block; the last 14 per Gaussian are the packed small-gradient rows (multi-view exchange)
constexpr int PB_OUT = 44 * PB;     // mean2D 3 | conic 4 | opacity 1 | colour 3 | mean3D 3 | cov3D 6 | scale 3 | rot 4 | masked rgb 3

__device__ __forceinline__ void copy_out(float* __restrict__ dst, const float* __restrict__ src, int n) {
    int n4 = ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) ? (n >> 2) : 0;
    for (int i = threadIdx.x; i < n4; i += PB) reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(src)[i];
    for (int i = 4 * n4 + threadIdx.x; i < n; i += PB) dst[i] = src[i];
}

// RAW / SPLIT: the raw-parameter path (gsr_fused_params.cuh); <.., false, false> are the reference-shaped kernels.
template <int MINB, bool PACK, bool RAW, bool SPLIT>
__global__ void __launch_bounds__(PB, MINB) preprocess_bwd_kernel_staged(PreprocessBwdArgs a, int S /* smem row stride, floats */) {
    __shared__ float s_cam[36];
    __shared__ uint8_t s_vis[PB];
    GSR_DYN_SMEM(s_dyn);                               // SH rows (PB x S), later re-used for the 27 other outputs
    float* s_sh = s_dyn;
    float* s_out = s_dyn;
    load_camera(a, s_cam);
    const int row0 = a.begin + blockIdx.x * PB;
    const int nrows = min(PB, a.P - row0);
    const int t = threadIdx.x;
    const int idx = row0 + t;
    const bool valid = t < nrows;
    const bool vis = valid && (a.radii[idx] > 0);
    s_vis[t] = vis;
    __syncthreads();
    BwdIn in;
    load_inputs<true, RAW, SPLIT>(a, idx, vis, in);
    const int L = 3 * a.M;
    const int Kf = 3 * (a.D + 1) * (a.D + 1);               // live floats per row
    const int Kf4 = (Kf + 3) & ~3;
    const bool has_sh = SPLIT || a.shs != nullptr;
    const int qstep = L ? (4 * PB) / L : 0, rstep = L ? (4 * PB) % L : 0;
    if (SPLIT) {
        // rows [dc(3) | rest]: only the rest part is read by the backward (the DC coefficient has no direction derivative)
        stage_rest_rows<PB>(s_sh, S, a.f_rest + (size_t)row0 * (L - 3), nrows, L - 3, Kf - 3, s_vis, t);
        cp_async_commit_wait_all();
        __syncthreads();
    } else if (has_sh) {
        const float4* src = reinterpret_cast<const float4*>(a.shs + (size_t)row0 * L);
        const int n4 = nrows * L / 4;
        int f = 4 * t, row = f / L, col = f - row * L;
        // asynchronous 16-byte copies (LDGSTS): every thread has all of its row pieces in flight at once, next to
        // the per-Gaussian loads issued above
        const uint32_t sh_base = smem_addr(s_sh);
        for (int e4 = t; e4 < n4; e4 += PB) {
            if (s_vis[row] && col < Kf4)
                cp_async16(sh_base + 4u * (uint32_t)(row * S + col), src + e4);
            row += qstep; col += rstep;
            if (col >= L) { col -= L; row++; }
        }
        cp_async_commit_wait_all();
        __syncthreads();
    }
    BwdOut o;
    if (valid) {
        float* myrow = s_sh + t * S;
        bwd_one<true, RAW, SPLIT>(a, idx, in, s_cam, s_cam + 16, myrow, myrow, o);
    }
    __syncthreads();
    if (SPLIT) {
        store_split_rows<PB>(s_sh, S, a.dL_df_dc + (size_t)row0 * 3, a.dL_df_rest + (size_t)row0 * (L - 3), nrows, L - 3, Kf, s_vis, t);
    } else if (a.dL_dsh != nullptr && L > 0 && !a.skip_dsh) {
        float4* dst = reinterpret_cast<float4*>(a.dL_dsh + (size_t)row0 * L);
        const int n4 = nrows * L / 4;
        int f = 4 * t, row = f / L, col = f - row * L;
        for (int e4 = t; e4 < n4; e4 += PB) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (has_sh && s_vis[row] && col < Kf4) v = *reinterpret_cast<const float4*>(s_sh + row * S + col);
            dst[e4] = v;
            row += qstep; col += rstep;
            if (col >= L) { col -= L; row++; }
        }
    }
    __syncthreads();                                    // SH rows are out: the same memory now stages the rest
    constexpr bool pack = PACK;        // compile-time: the single-GPU instantiation is untouched by the exchange layout
    if (valid) {
        reinterpret_cast<float4*>(s_out + 3 * PB)[t] = make_float4(o.g[2], o.g[3], 0.f, o.g[4]);
        s_out[8 * PB + 3 * t] = o.g[6]; s_out[8 * PB + 3 * t + 1] = o.g[7]; s_out[8 * PB + 3 * t + 2] = o.g[8];
#pragma unroll
        for (int i = 0; i < 6; i++) s_out[14 * PB + 6 * t + i] = o.dcov[i];
        s_out[27 * PB + 3 * t] = o.drgb[0]; s_out[27 * PB + 3 * t + 1] = o.drgb[1]; s_out[27 * PB + 3 * t + 2] = o.drgb[2];
        if (pack) {        // one 56-byte row per Gaussian; the rows use the 14*PB floats the five tensors' stages would
            float* row = s_out + 30 * PB + 14 * t;
            row[0] = o.dmean[0]; row[1] = o.dmean[1]; row[2] = o.dmean[2];
            row[3] = o.g[0]; row[4] = o.g[1]; row[5] = 0.f;
            row[6] = o.g[5];
            row[7] = o.dscale[0]; row[8] = o.dscale[1]; row[9] = o.dscale[2];
            row[10] = o.drot[0]; row[11] = o.drot[1]; row[12] = o.drot[2]; row[13] = o.drot[3];
        } else {
            s_out[3 * t] = o.g[0]; s_out[3 * t + 1] = o.g[1]; s_out[3 * t + 2] = 0.f;
            s_out[7 * PB + t] = o.g[5];
            s_out[11 * PB + 3 * t] = o.dmean[0]; s_out[11 * PB + 3 * t + 1] = o.dmean[1]; s_out[11 * PB + 3 * t + 2] = o.dmean[2];
            s_out[20 * PB + 3 * t] = o.dscale[0]; s_out[20 * PB + 3 * t + 1] = o.dscale[1]; s_out[20 * PB + 3 * t + 2] = o.dscale[2];
            reinterpret_cast<float4*>(s_out + 23 * PB)[t] = make_float4(o.drot[0], o.drot[1], o.drot[2], o.drot[3]);
        }
    }
    __syncthreads();
    if (a.dL_dconic) copy_out(a.dL_dconic + (size_t)row0 * 4, s_out + 3 * PB, nrows * 4);
    if (a.dL_dcolor) copy_out(a.dL_dcolor + (size_t)row0 * 3, s_out + 8 * PB, nrows * 3);
    if (a.dL_dcov3D) copy_out(a.dL_dcov3D + (size_t)row0 * 6, s_out + 14 * PB, nrows * 6);
    if (a.dL_drgb) copy_out(a.dL_drgb + (size_t)row0 * 3, s_out + 27 * PB, nrows * 3);
    if (pack) {
        copy_out(a.small_pack + (size_t)row0 * 14, s_out + 30 * PB, nrows * 14);
    } else {
        copy_out(a.dL_dmean2D + (size_t)row0 * 3, s_out, nrows * 3);
        copy_out(a.dL_dopacity + row0, s_out + 7 * PB, nrows);
        copy_out(a.dL_dmean3D + (size_t)row0 * 3, s_out + 11 * PB, nrows * 3);
        copy_out(a.dL_dscale + (size_t)row0 * 3, s_out + 20 * PB, nrows * 3);
        copy_out(a.dL_drot + (size_t)row0 * 4, s_out + 23 * PB, nrows * 4);
    }
}

void launch_preprocess_backward(const PreprocessBwdArgs& a, cudaStream_t st) {
    if (a.P <= 0) return;
    const int L = 3 * a.M;
    const int S = ((L / 4) % 2 == 1 || L == 0) ? L : L + 4;      // 16-byte rows, odd number of 16-byte words: conflict-free
    size_t dyn = (size_t)PB_OUT * sizeof(float);
    if (a.shs && (size_t)PB * S * sizeof(float) > dyn) dyn = (size_t)PB * S * sizeof(float);
    const bool aligned = (reinterpret_cast<uintptr_t>(a.shs) & 15) == 0 && (reinterpret_cast<uintptr_t>(a.dL_dsh) & 15) == 0 &&
                         (reinterpret_cast<uintptr_t>(a.rotations) & 15) == 0;
    const bool staged = (L % 4 == 0) && aligned && dyn <= 40 * 1024 && (a.M == 0 || a.dL_dsh != nullptr);
    if (a.f_dc != nullptr) {
        // split SH storage: always staged (rows are filled with 4-byte copies, so any M works); gsr_backward has checked
        // M <= 16, the alignment of `rotations` and that no exchange option is combined with it
        const int S4 = (L + 3) / 4, Ss = 4 * (S4 % 2 == 1 ? S4 : S4 + 1);        // odd number of 16-byte words per row
        size_t dyns = (size_t)PB_OUT * sizeof(float);
        if ((size_t)PB * Ss * sizeof(float) > dyns) dyns = (size_t)PB * Ss * sizeof(float);
        if (a.raw_act) preprocess_bwd_kernel_staged<PB_MINB, false, true, true><<<(a.P + PB - 1) / PB, PB, dyns, st>>>(a, Ss);
        else preprocess_bwd_kernel_staged<PB_MINB, false, false, true><<<(a.P + PB - 1) / PB, PB, dyns, st>>>(a, Ss);
        count_launch(1);
        return;
    }
    // beside a collective (multi-view exchange): chunked, see chunk_plan
    int bounds[5];
    const int nchunk = a.chunked ? chunk_plan(a.P, 256, bounds) : 1;
    if (nchunk == 1) { bounds[0] = 0; bounds[1] = a.P; }
    for (int c = 0; c < nchunk; c++) {
        PreprocessBwdArgs ac = a;
        ac.begin = bounds[c]; ac.P = bounds[c + 1];
        const int n = ac.P - ac.begin;
        if (n <= 0) continue;
        if (staged && a.small_pack) preprocess_bwd_kernel_staged<PB_MINB, true, false, false><<<(n + PB - 1) / PB, PB, dyn, st>>>(ac, S);
        else if (staged && a.raw_act) preprocess_bwd_kernel_staged<PB_MINB, false, true, false><<<(n + PB - 1) / PB, PB, dyn, st>>>(ac, S);
        else if (staged) preprocess_bwd_kernel_staged<PB_MINB, false, false, false><<<(n + PB - 1) / PB, PB, dyn, st>>>(ac, S);
        else if (a.raw_act) preprocess_bwd_kernel_generic<true><<<(n + 255) / 256, 256, 0, st>>>(ac);
        else preprocess_bwd_kernel_generic<false><<<(n + 255) / 256, 256, 0, st>>>(ac);
        count_launch(1);
        if (a.chunk_events && c < a.n_chunk_events && c < nchunk - 1) cudaEventRecord((cudaEvent_t)a.chunk_events[c], st);
    }
    for (int c = nchunk - 1; a.chunk_events && c < a.n_chunk_events; c++) cudaEventRecord((cudaEvent_t)a.chunk_events[c], st);
}

}  // namespace gsr

// ------------------------------------------------------------------------------------------------
// Multi-view SH-gradient exchange fused with its producer math (include/gsr.h: gsr_multiview_sh_grad).
// Each rank's backward leaves 12 bytes per Gaussian (clamp-masked dL/dRGB) and its camera position in a
// peer-mapped block.  One kernel per rank pulls every rank's block over NVLink with plain coalesced 16-byte
// peer loads, rebuilds the SH basis for each view's direction and writes sum_v basis(dir_v) (x) dRGB_v as
// coalesced 16-byte rows -- 12 B per Gaussian per view cross the links instead of the 12*M B an all-reduce of
// dL_dsh moves twice.  128 Gaussians per block, staged through shared memory like the backward above.
namespace gsr {

constexpr int MV = 128;

template <int DEG>
__device__ __forceinline__ void mv_accumulate(float* acc, float x, float y, float z, float r, float g, float b) {
    constexpr int K = (DEG + 1) * (DEG + 1);
    float k[16];
    sh_basis(DEG, x, y, z, k);
#pragma unroll
    for (int i = 0; i < K; i++) {
        acc[3 * i] = fmaf(k[i], r, acc[3 * i]);
        acc[3 * i + 1] = fmaf(k[i], g, acc[3 * i + 1]);
        acc[3 * i + 2] = fmaf(k[i], b, acc[3 * i + 2]);
    }
}

template <int DEG>
__global__ void __launch_bounds__(MV) multiview_sh_grad_kernel(MultiviewShArgs a, int S) {
    constexpr int K = (DEG + 1) * (DEG + 1);
    constexpr int NF = 3 * K, NV = (NF + 3) / 4;
    GSR_DYN_SMEM(s_mv);                                  // [n_views][MV*3] colour gradients, then MV rows of S floats
    float* s_rgb = s_mv;
    float* s_rows = s_mv + a.n_views * MV * 3;
    __shared__ float s_campos[16 * 4];
    const int row0 = a.begin + blockIdx.x * MV;
    const int nrows = min(MV, a.P - row0);
    const int t = threadIdx.x;
    if (t < a.n_views * 4) s_campos[t] = a.campos ? a.campos[t] : a.blocks[t >> 2][t & 3];
    // ---- peer loads: each view's (nrows,3) slice is one contiguous span ----
    const int n = nrows * 3, n4 = n >> 2;
    for (int v = 0; v < a.n_views; v++) {
        const float* src = a.blocks[v] + 4 + (size_t)row0 * 3;      // 16-byte aligned: row0 is a multiple of 128
        float* dst = s_rgb + v * MV * 3;
        for (int i = t; i < n4; i += MV) reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(src)[i];
        for (int i = 4 * n4 + t; i < n; i += MV) dst[i] = src[i];
    }
    __syncthreads();
    if (t < nrows) {
        float acc[NF];
#pragma unroll
        for (int i = 0; i < NF; i++) acc[i] = 0.f;
        const size_t i3 = 3 * (size_t)(row0 + t);
        const float mx = __ldg(a.means3D + i3), my = __ldg(a.means3D + i3 + 1), mz = __ldg(a.means3D + i3 + 2);
        for (int v = 0; v < a.n_views; v++) {
            const float r = s_rgb[v * MV * 3 + 3 * t], g = s_rgb[v * MV * 3 + 3 * t + 1], b = s_rgb[v * MV * 3 + 3 * t + 2];
            if (r == 0.f && g == 0.f && b == 0.f) continue;         // culled / fully clamped in this view
            // direction exactly as the forward and the single-view backward compute it (forward.cu:28-31)
            const float dx = subr(mx, s_campos[4 * v]), dy = subr(my, s_campos[4 * v + 1]), dz = subr(mz, s_campos[4 * v + 2]);
            const float len = __fsqrt_rn(fmar(dz, dz, fmar(dx, dx, mulr(dy, dy))));
            mv_accumulate<DEG>(acc, __fdiv_rn(dx, len), __fdiv_rn(dy, len), __fdiv_rn(dz, len), r, g, b);
        }
        float* row = s_rows + t * S;
#pragma unroll
        for (int c = 0; c < NV; c++) {
            float4 v4;
            v4.x = acc[4 * c];
            v4.y = (4 * c + 1 < NF) ? acc[4 * c + 1] : 0.f;
            v4.z = (4 * c + 2 < NF) ? acc[4 * c + 2] : 0.f;
            v4.w = (4 * c + 3 < NF) ? acc[4 * c + 3] : 0.f;
            reinterpret_cast<float4*>(row)[c] = v4;
        }
    }
    __syncthreads();
    const int L = 3 * a.M;
    float4* dst = reinterpret_cast<float4*>(a.dL_dsh + (size_t)row0 * L);
    const int m4 = nrows * L / 4;
    const int qstep = (4 * MV) / L, rstep = (4 * MV) % L;
    int f = 4 * t, row = f / L, col = f - row * L;
    for (int e4 = t; e4 < m4; e4 += MV) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (col < 4 * NV) v = *reinterpret_cast<const float4*>(s_rows + row * S + col);
        dst[e4] = v;
        row += qstep; col += rstep;
        if (col >= L) { col -= L; row++; }
    }
}

// fallback for shapes the staged kernel does not take (3*M not a multiple of 4): one thread per Gaussian
__global__ void __launch_bounds__(256) multiview_sh_grad_kernel_generic(MultiviewShArgs a) {
    const int idx = a.begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= a.P) return;
    float acc[48];
    for (int i = 0; i < 48; i++) acc[i] = 0.f;
    const size_t i3 = 3 * (size_t)idx;
    const float mx = __ldg(a.means3D + i3), my = __ldg(a.means3D + i3 + 1), mz = __ldg(a.means3D + i3 + 2);
    const int K = (a.D + 1) * (a.D + 1);
    for (int v = 0; v < a.n_views; v++) {
        const float* blk = a.blocks[v];
        const float r = blk[4 + i3], g = blk[4 + i3 + 1], b = blk[4 + i3 + 2];
        if (r == 0.f && g == 0.f && b == 0.f) continue;
        const float* cp = a.campos ? a.campos + 4 * v : blk;
        const float dx = subr(mx, cp[0]), dy = subr(my, cp[1]), dz = subr(mz, cp[2]);
        const float len = __fsqrt_rn(fmar(dz, dz, fmar(dx, dx, mulr(dy, dy))));
        float k[16];
        sh_basis(a.D, __fdiv_rn(dx, len), __fdiv_rn(dy, len), __fdiv_rn(dz, len), k);
        for (int i = 0; i < K; i++) { acc[3 * i] += k[i] * r; acc[3 * i + 1] += k[i] * g; acc[3 * i + 2] += k[i] * b; }
    }
    float* out = a.dL_dsh + (size_t)idx * a.M * 3;
    for (int i = 0; i < 3 * a.M; i++) out[i] = (i < 3 * K) ? acc[i] : 0.f;
}

__global__ void __launch_bounds__(256) publish_drgb_kernel(int P, const float* __restrict__ grad_acc, const int* __restrict__ radii,
                                                           const uint8_t* __restrict__ clamped, float* __restrict__ out) {
    __shared__ float s[256 * 3];
    const int idx = blockIdx.x * 256 + threadIdx.x;
    float r = 0.f, g = 0.f, b = 0.f;
    if (idx < P && radii[idx] > 0) {
        const float4 q = reinterpret_cast<const float4*>(grad_acc + 12 * (size_t)idx)[1];    // acc[4..7]
        const uint32_t cl = clamped ? clamped[idx] : 0u;
        r = (cl & 1) ? 0.f : q.z; g = (cl & 2) ? 0.f : q.w; b = (cl & 4) ? 0.f : grad_acc[12 * (size_t)idx + 8];
    }
    s[3 * threadIdx.x] = r; s[3 * threadIdx.x + 1] = g; s[3 * threadIdx.x + 2] = b;
    __syncthreads();
    const size_t base = (size_t)blockIdx.x * 768;
    const size_t n = (size_t)P * 3;
    for (int i = threadIdx.x; i < 768; i += 256) if (base + i < n) out[base + i] = s[i];
}

// rows [begin, end) of the (P,14) packed block -> the five gradient tensors, staged through shared memory so that both
// sides are coalesced
__global__ void __launch_bounds__(256) unpack_small_kernel(int begin, int end, const float* __restrict__ pack, float* __restrict__ dmean3D,
                                                           float* __restrict__ dmean2D, float* __restrict__ dop, float* __restrict__ dscale,
                                                           float* __restrict__ drot) {
    __shared__ float s[256 * 14];
    const int row0 = begin + blockIdx.x * 256;
    const int nrows = min(256, end - row0);
    const float* src = pack + (size_t)row0 * 14;
    for (int i = threadIdx.x; i < nrows * 14; i += 256) s[i] = src[i];
    __syncthreads();
    for (int i = threadIdx.x; i < nrows * 3; i += 256) {
        const int r = i / 3, c = i - 3 * r;
        dmean3D[(size_t)row0 * 3 + i] = s[r * 14 + c];
        dmean2D[(size_t)row0 * 3 + i] = s[r * 14 + 3 + c];
        dscale[(size_t)row0 * 3 + i] = s[r * 14 + 7 + c];
    }
    for (int i = threadIdx.x; i < nrows; i += 256) dop[row0 + i] = s[i * 14 + 6];
    for (int i = threadIdx.x; i < nrows * 4; i += 256) drot[(size_t)row0 * 4 + i] = s[(i >> 2) * 14 + 10 + (i & 3)];
}

void launch_unpack_small(int begin, int end, const float* pack, float* dmean3D, float* dmean2D, float* dop, float* dscale,
                         float* drot, cudaStream_t st) {
    if (end <= begin) return;
    unpack_small_kernel<<<(end - begin + 255) / 256, 256, 0, st>>>(begin, end, pack, dmean3D, dmean2D, dop, dscale, drot);
    count_launch(1);
}

void launch_publish_drgb(int P, const float* grad_acc, const int* radii, const uint8_t* clamped, float* out, cudaStream_t st) {
    if (P <= 0) return;
    publish_drgb_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, grad_acc, radii, clamped, out);
    count_launch(1);
}

void launch_multiview_sh_grad(const MultiviewShArgs& a, cudaStream_t st) {
    if (a.P <= 0) return;
    const int L = 3 * a.M;
    const int S = ((L / 4) % 2 == 1) ? L : L + 4;
    const size_t dyn = ((size_t)a.n_views * MV * 3 + (size_t)MV * S) * sizeof(float);
    // runs beside the all-reduce of the small gradients: chunked, see chunk_plan
    int bounds[5];
    int nchunk = chunk_plan(a.P, 256, bounds);
    if (a.no_split) { nchunk = 1; bounds[0] = a.begin; bounds[1] = a.P; }
    for (int c = 0; c < nchunk; c++) {
        MultiviewShArgs ac = a;
        ac.begin = bounds[c]; ac.P = bounds[c + 1];
        const int n = ac.P - ac.begin;
        if (n <= 0) continue;
        if (L % 4 == 0 && dyn <= 44 * 1024) {
            const unsigned grid = (n + MV - 1) / MV;
            switch (a.D) {
                case 0: multiview_sh_grad_kernel<0><<<grid, MV, dyn, st>>>(ac, S); break;
                case 1: multiview_sh_grad_kernel<1><<<grid, MV, dyn, st>>>(ac, S); break;
                case 2: multiview_sh_grad_kernel<2><<<grid, MV, dyn, st>>>(ac, S); break;
                default: multiview_sh_grad_kernel<3><<<grid, MV, dyn, st>>>(ac, S); break;
            }
        } else {
            multiview_sh_grad_kernel_generic<<<(n + 255) / 256, 256, 0, st>>>(ac);
        }
        count_launch(1);
    }
}

}  // namespace gsr
