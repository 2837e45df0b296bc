// gsr_preprocess_bwd.cu -- backward of the per-Gaussian stage for sm_100a.
//
// Fuses computeCov2DCUDA (DGR/cuda_rasterizer/backward.cu:144-277) and preprocessCUDA<3>
// (backward) (:349-399, with computeColorFromSH :20-139 and computeCov3D :281-344) into one
// kernel, one thread per Gaussian.  It reads the (P,12) gradient accumulators the blending
// backward reduced into and WRITES every element of every output (zeros for Gaussians with
// radii <= 0), so callers need no 300 B/Gaussian zero-fill (rasterize_points.cu:168-178).
// cov3D is recomputed from scale/rotation with the forward's exact arithmetic instead of being
// stored in the forward and re-read here.  HBM-bound.
#include "gsr_internal.cuh"
#include "gsr_math.cuh"

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
struct BwdOut { float g[9]; float dmean[3]; float dcov[6]; float dscale[3]; float drot[4]; };

// The per-Gaussian backward math.  `sh_row` / `dsh_row` are this Gaussian's SH coefficients and SH-gradient
// row: global pointers in the generic kernel, the same shared-memory row in the staged kernel.
// Returns the number of SH coefficients written to dsh_row (0 when not visible or no SHs).
template <bool SMEM>
__device__ __forceinline__ int bwd_one(const PreprocessBwdArgs& a, int idx, const float* V, const float* Pm,
                                       const float* sh_row, float* dsh_row, BwdOut& o) {
    const size_t i3 = 3 * (size_t)idx;
#pragma unroll
    for (int i = 0; i < 9; i++) o.g[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 3; i++) { o.dmean[i] = 0.f; o.dscale[i] = 0.f; }
#pragma unroll
    for (int i = 0; i < 6; i++) o.dcov[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 4; i++) o.drot[i] = 0.f;
    if (!(a.radii[idx] > 0)) return 0;
    float* g = o.g; float* dmean = o.dmean; float* dcov = o.dcov;
    const float4* ga = reinterpret_cast<const float4*>(a.grad_acc + 12 * (size_t)idx);
    const float4 g0 = ga[0], g1 = ga[1];
    g[0] = g0.x; g[1] = g0.y; g[2] = g0.z; g[3] = g0.w; g[4] = g1.x; g[5] = g1.y; g[6] = g1.z; g[7] = g1.w;
    g[8] = a.grad_acc[12 * (size_t)idx + 8];
    const float mx = __ldg(a.means3D + i3), my = __ldg(a.means3D + i3 + 1), mz = __ldg(a.means3D + i3 + 2);
    float c3[6];
    float4 q = make_float4(0, 0, 0, 0);
    float sx = 0, sy = 0, sz = 0;
    if (a.cov3D_precomp != nullptr) {
#pragma unroll
        for (int i = 0; i < 6; i++) c3[i] = __ldg(a.cov3D_precomp + 6 * (size_t)idx + i);
    } else {
        q = make_float4(__ldg(a.rotations + 4 * (size_t)idx), __ldg(a.rotations + 4 * (size_t)idx + 1),
                        __ldg(a.rotations + 4 * (size_t)idx + 2), __ldg(a.rotations + 4 * (size_t)idx + 3));
        sx = __ldg(a.scales + i3); sy = __ldg(a.scales + i3 + 1); sz = __ldg(a.scales + i3 + 2);
        cov3d_from_scale_rot(sx, sy, sz, a.scale_modifier, q, c3);
    }
    // ---- computeCov2DCUDA (backward.cu:144-277) ----
    const EwaT e = ewa_T(mx, my, mz, a.focal_x, a.focal_y, a.tan_fovx, a.tan_fovy, V);
    const float3 cov = ewa_cov2d(e, c3);
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
    if (a.shs != nullptr) {
        const uint32_t cl = a.clamped[idx];
        const F3 dRGB = f3((cl & 1) ? 0.f : g[6], (cl & 2) ? 0.f : g[7], (cl & 4) ? 0.f : g[8]);
        const F3 dir = f3(subr(mx, V[32]), subr(my, V[33]), subr(mz, V[34]));
        const F3 dm = sh_backward<SMEM>(a.D, sh_row, dir, dRGB, dsh_row);
        sh_written = (a.D + 1) * (a.D + 1);
        dmean[0] += dm.x; dmean[1] += dm.y; dmean[2] += dm.z;
    }
    // ---- scale / rotation (backward.cu:396-398) ----
    if (a.scales != nullptr) cov3d_backward(sx, sy, sz, a.scale_modifier, q, dcov, o.dscale, o.drot);
    return sh_written;
}

__device__ __forceinline__ void load_camera(const PreprocessBwdArgs& a, float* s_cam) {
    if (threadIdx.x < 16) s_cam[threadIdx.x] = a.cam.view[threadIdx.x];
    else if (threadIdx.x < 32) s_cam[threadIdx.x] = a.cam.proj[threadIdx.x - 16];
    else if (threadIdx.x < 35) s_cam[threadIdx.x] = a.cam.campos[threadIdx.x - 32];
}

// ---- generic kernel: any M, any alignment; per-thread strided stores ------------------------------------
__global__ void __launch_bounds__(256) preprocess_bwd_kernel_generic(PreprocessBwdArgs a) {
    __shared__ float s_cam[36];
    load_camera(a, s_cam);
    __syncthreads();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= a.P) return;
    const size_t i3 = 3 * (size_t)idx;
    float* dsh = a.dL_dsh ? a.dL_dsh + (size_t)idx * a.M * 3 : nullptr;
    BwdOut o;
    const int sh_written = bwd_one<false>(a, idx, s_cam, s_cam + 16, a.shs ? a.shs + (size_t)idx * a.M * 3 : nullptr, dsh, o);
    // ---- write everything (zeros where the reference's pre-zeroed tensors stay untouched) ----
    a.dL_dmean2D[i3] = o.g[0]; a.dL_dmean2D[i3 + 1] = o.g[1]; a.dL_dmean2D[i3 + 2] = 0.f;
    if (a.dL_dconic) reinterpret_cast<float4*>(a.dL_dconic)[idx] = make_float4(o.g[2], o.g[3], 0.f, o.g[4]);
    a.dL_dopacity[idx] = o.g[5];
    a.dL_dcolor[i3] = o.g[6]; a.dL_dcolor[i3 + 1] = o.g[7]; a.dL_dcolor[i3 + 2] = o.g[8];
    a.dL_dmean3D[i3] = o.dmean[0]; a.dL_dmean3D[i3 + 1] = o.dmean[1]; a.dL_dmean3D[i3 + 2] = o.dmean[2];
#pragma unroll
    for (int i = 0; i < 6; i++) a.dL_dcov3D[6 * (size_t)idx + i] = o.dcov[i];
    a.dL_dscale[i3] = o.dscale[0]; a.dL_dscale[i3 + 1] = o.dscale[1]; a.dL_dscale[i3 + 2] = o.dscale[2];
#pragma unroll
    for (int i = 0; i < 4; i++) a.dL_drot[4 * (size_t)idx + i] = o.drot[i];
    if (dsh) for (int i = 3 * sh_written; i < 3 * a.M; i++) dsh[i] = 0.f;
}

// ---- staged kernel: the common case (3*M a multiple of 4, 16-byte aligned tensors) ------------------------
// 128 Gaussians per block.  The SH coefficient rows of the block's visible Gaussians are brought in with
// coalesced 16-byte loads into shared memory (odd row stride: conflict-free per-thread row walks), each
// thread overwrites its row in place with dL_dsh, and the rows leave as coalesced 16-byte stores (zeros for
// culled Gaussians and for coefficients beyond the active degree).  The 27 other floats per Gaussian are
// staged the same way, so every global store instruction of a warp covers one contiguous 512-byte span.
constexpr int PB = 128;
constexpr int PB_OUT = 27 * PB;     // mean2D 3 | conic 4 | opacity 1 | colour 3 | mean3D 3 | cov3D 6 | scale 3 | rot 4

__device__ __forceinline__ void copy_out(float* __restrict__ dst, const float* __restrict__ src, int n) {
    int n4 = ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) ? (n >> 2) : 0;
    for (int i = threadIdx.x; i < n4; i += PB) reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(src)[i];
    for (int i = 4 * n4 + threadIdx.x; i < n; i += PB) dst[i] = src[i];
}

__global__ void __launch_bounds__(PB) preprocess_bwd_kernel_staged(PreprocessBwdArgs a, int S /* smem row stride, floats */) {
    __shared__ float s_cam[36];
    __shared__ uint8_t s_vis[PB];
    __shared__ __align__(16) float s_out[PB_OUT];
    extern __shared__ __align__(16) float s_sh[];
    load_camera(a, s_cam);
    const int row0 = blockIdx.x * PB;
    const int nrows = min(PB, a.P - row0);
    const int t = threadIdx.x;
    const int idx = row0 + t;
    const bool valid = t < nrows;
    s_vis[t] = valid && (a.radii[idx] > 0);
    __syncthreads();
    const int L = 3 * a.M;
    const int Kf = 3 * (a.D + 1) * (a.D + 1);               // live floats per row
    const bool has_sh = a.shs != nullptr;
    const int qstep = L ? (4 * PB) / L : 0, rstep = L ? (4 * PB) % L : 0;
    if (has_sh) {
        const float4* src = reinterpret_cast<const float4*>(a.shs + (size_t)row0 * L);
        const int n4 = nrows * L / 4;
        int f = 4 * t, row = f / L, col = f - row * L;
        for (int e4 = t; e4 < n4; e4 += PB) {
            if (s_vis[row] && col < Kf) {
                const float4 v = __ldg(src + e4);
                float* d = s_sh + row * S + col;
                d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
            }
            row += qstep; col += rstep;
            if (col >= L) { col -= L; row++; }
        }
        __syncthreads();
    }
    if (valid) {
        BwdOut o;
        float* myrow = s_sh + t * S;
        bwd_one<true>(a, idx, s_cam, s_cam + 16, myrow, myrow, o);
        s_out[3 * t] = o.g[0]; s_out[3 * t + 1] = o.g[1]; s_out[3 * t + 2] = 0.f;
        reinterpret_cast<float4*>(s_out + 3 * PB)[t] = make_float4(o.g[2], o.g[3], 0.f, o.g[4]);
        s_out[7 * PB + t] = o.g[5];
        s_out[8 * PB + 3 * t] = o.g[6]; s_out[8 * PB + 3 * t + 1] = o.g[7]; s_out[8 * PB + 3 * t + 2] = o.g[8];
        s_out[11 * PB + 3 * t] = o.dmean[0]; s_out[11 * PB + 3 * t + 1] = o.dmean[1]; s_out[11 * PB + 3 * t + 2] = o.dmean[2];
#pragma unroll
        for (int i = 0; i < 6; i++) s_out[14 * PB + 6 * t + i] = o.dcov[i];
        s_out[20 * PB + 3 * t] = o.dscale[0]; s_out[20 * PB + 3 * t + 1] = o.dscale[1]; s_out[20 * PB + 3 * t + 2] = o.dscale[2];
        reinterpret_cast<float4*>(s_out + 23 * PB)[t] = make_float4(o.drot[0], o.drot[1], o.drot[2], o.drot[3]);
    }
    __syncthreads();
    copy_out(a.dL_dmean2D + (size_t)row0 * 3, s_out, nrows * 3);
    if (a.dL_dconic) copy_out(a.dL_dconic + (size_t)row0 * 4, s_out + 3 * PB, nrows * 4);
    copy_out(a.dL_dopacity + row0, s_out + 7 * PB, nrows);
    copy_out(a.dL_dcolor + (size_t)row0 * 3, s_out + 8 * PB, nrows * 3);
    copy_out(a.dL_dmean3D + (size_t)row0 * 3, s_out + 11 * PB, nrows * 3);
    copy_out(a.dL_dcov3D + (size_t)row0 * 6, s_out + 14 * PB, nrows * 6);
    copy_out(a.dL_dscale + (size_t)row0 * 3, s_out + 20 * PB, nrows * 3);
    copy_out(a.dL_drot + (size_t)row0 * 4, s_out + 23 * PB, nrows * 4);
    if (a.dL_dsh != nullptr && L > 0) {
        float4* dst = reinterpret_cast<float4*>(a.dL_dsh + (size_t)row0 * L);
        const int n4 = nrows * L / 4;
        const bool live = has_sh;
        int f = 4 * t, row = f / L, col = f - row * L;
        for (int e4 = t; e4 < n4; e4 += PB) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (live && s_vis[row] && col < Kf) {
                const float* sp = s_sh + row * S + col;
                v.x = sp[0];
                v.y = (col + 1 < Kf) ? sp[1] : 0.f;
                v.z = (col + 2 < Kf) ? sp[2] : 0.f;
                v.w = (col + 3 < Kf) ? sp[3] : 0.f;
            }
            dst[e4] = v;
            row += qstep; col += rstep;
            if (col >= L) { col -= L; row++; }
        }
    }
}

void launch_preprocess_backward(const PreprocessBwdArgs& a, cudaStream_t st) {
    if (a.P <= 0) return;
    const int L = 3 * a.M;
    const int S = (L % 2 == 0) ? L + 1 : L;
    const size_t dyn = a.shs ? (size_t)PB * S * sizeof(float) : 0;
    const bool aligned = (reinterpret_cast<uintptr_t>(a.shs) & 15) == 0 && (reinterpret_cast<uintptr_t>(a.dL_dsh) & 15) == 0;
    const bool staged = (L % 4 == 0) && aligned && dyn <= 32 * 1024 && (a.M == 0 || a.dL_dsh != nullptr);
    if (staged) preprocess_bwd_kernel_staged<<<(a.P + PB - 1) / PB, PB, dyn, st>>>(a, S);
    else preprocess_bwd_kernel_generic<<<(a.P + 255) / 256, 256, 0, st>>>(a);
    count_launch(1);
}

}  // namespace gsr
