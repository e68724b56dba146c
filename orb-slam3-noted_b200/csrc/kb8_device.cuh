// kb8_device.cuh — Kannala-Brandt (KB8) camera model on the device: project (KannalaBrandt8.cpp:24-41),
// unproject (:96-123, Newton iteration), Triangulate_ (:528-537, here a double-precision Jacobi solve for the
// null vector of the 4x4 DLT system instead of cv::SVD) and TriangulateMatches_ (:409-480).  Shared by the
// fisheye stereo kernel (k_kb8.cu) and the KB8 mode of the triangulation search (k_bow.cu).
#pragma once
#include "orbfe_internal.h"

static __device__ __forceinline__ void kb8_project(const float* c, const float p[3], float& u, float& v) {
    const float x2y2 = __fadd_rn(__fmul_rn(p[0], p[0]), __fmul_rn(p[1], p[1]));
    const float theta = atan2f(sqrtf(x2y2), p[2]);
    const float psi = atan2f(p[1], p[0]);
    const float t2 = __fmul_rn(theta, theta), t3 = __fmul_rn(theta, t2), t5 = __fmul_rn(t3, t2), t7 = __fmul_rn(t5, t2),
                t9 = __fmul_rn(t7, t2);
    const float r = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(theta, __fmul_rn(c[4], t3)), __fmul_rn(c[5], t5)), __fmul_rn(c[6], t7)),
                              __fmul_rn(c[7], t9));
    u = (float)((double)__fmul_rn(c[0], r) * cos((double)psi) + (double)c[2]);
    v = (float)((double)__fmul_rn(c[1], r) * sin((double)psi) + (double)c[3]);
}

static __device__ __forceinline__ void kb8_unproject(const float* c, float u, float v, float ray[3]) {
    const float pwx = __fdiv_rn(__fsub_rn(u, c[2]), c[0]), pwy = __fdiv_rn(__fsub_rn(v, c[3]), c[1]);
    float scale = 1.f;
    float theta_d = sqrtf(__fadd_rn(__fmul_rn(pwx, pwx), __fmul_rn(pwy, pwy)));
    theta_d = fminf(fmaxf(-1.57079632679489661923f, theta_d), 1.57079632679489661923f);
    if (theta_d > 1e-8f) {
        float theta = theta_d;
        for (int j = 0; j < 10; j++) {
            const float t2 = __fmul_rn(theta, theta), t4 = __fmul_rn(t2, t2), t6 = __fmul_rn(t4, t2), t8 = __fmul_rn(t4, t4);
            const float a = __fmul_rn(c[4], t2), b = __fmul_rn(c[5], t4), cc = __fmul_rn(c[6], t6), d = __fmul_rn(c[7], t8);
            const float num = __fsub_rn(__fmul_rn(theta, __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(1.f, a), b), cc), d)), theta_d);
            const float den = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(1.f, __fmul_rn(3.f, a)), __fmul_rn(5.f, b)), __fmul_rn(7.f, cc)),
                                        __fmul_rn(9.f, d));
            const float fix = __fdiv_rn(num, den);
            theta = __fsub_rn(theta, fix);
            if (fabsf(fix) < 1e-6f) break;
        }
        scale = __fdiv_rn((float)tan((double)theta), theta_d);
    }
    ray[0] = __fmul_rn(pwx, scale);
    ray[1] = __fmul_rn(pwy, scale);
    ray[2] = 1.f;
}

// Eigenvector of the smallest eigenvalue of A^T A (A row-major 4x4), cyclic Jacobi in double.
static __device__ void smallest_right_singular_vector(const float A[16], double x[4]) {
    double M[4][4], V[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            double s = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) s += (double)A[k * 4 + i] * (double)A[k * 4 + j];
            M[i][j] = s;
            V[i][j] = i == j ? 1.0 : 0.0;
        }
    for (int sweep = 0; sweep < 30; ++sweep) {
        double off = 0;
#pragma unroll
        for (int p = 0; p < 4; ++p)
#pragma unroll
            for (int q = p + 1; q < 4; ++q) off += M[p][q] * M[p][q];
        if (off < 1e-300) break;
#pragma unroll
        for (int p = 0; p < 4; ++p)
#pragma unroll
            for (int q = p + 1; q < 4; ++q) {
                if (M[p][q] == 0) continue;
                const double th = (M[q][q] - M[p][p]) / (2 * M[p][q]);
                const double t = (th >= 0 ? 1.0 : -1.0) / (fabs(th) + sqrt(th * th + 1));
                const double cs = 1 / sqrt(t * t + 1), sn = t * cs;
#pragma unroll
                for (int k = 0; k < 4; ++k) { const double a = M[k][p], b = M[k][q]; M[k][p] = cs * a - sn * b; M[k][q] = sn * a + cs * b; }
#pragma unroll
                for (int k = 0; k < 4; ++k) { const double a = M[p][k], b = M[q][k]; M[p][k] = cs * a - sn * b; M[q][k] = sn * a + cs * b; }
#pragma unroll
                for (int k = 0; k < 4; ++k) { const double a = V[k][p], b = V[k][q]; V[k][p] = cs * a - sn * b; V[k][q] = sn * a + cs * b; }
            }
    }
    int m = 0;
#pragma unroll
    for (int i = 1; i < 4; ++i)
        if (M[i][i] < M[m][m]) m = i;
#pragma unroll
    for (int k = 0; k < 4; ++k) x[k] = V[k][m];
}

// Returns z1 or -1 (KannalaBrandt8.cpp:338-407).
static __device__ __noinline__ float kb8_triangulate_matches(const float* cam1, const float* cam2, const float* R12, const float* t12, float u1, float v1, float u2, float v2, float sigmaLevel, float unc,
                                         float p3D[3]) {
    float r1[3], r2[3], r21[3];
    kb8_unproject(cam1, u1, v1, r1);
    kb8_unproject(cam2, u2, v2, r2);
#pragma unroll
    for (int i = 0; i < 3; ++i)
        r21[i] = __fadd_rn(__fadd_rn(__fmul_rn(R12[3 * i], r2[0]), __fmul_rn(R12[3 * i + 1], r2[1])), __fmul_rn(R12[3 * i + 2], r2[2]));
    double dot = 0, n1 = 0, n2 = 0;
#pragma unroll
    for (int i = 0; i < 3; ++i) { dot += (double)r1[i] * r21[i]; n1 += (double)r1[i] * r1[i]; n2 += (double)r21[i] * r21[i]; }
    const float cosParallaxRays = (float)(dot / (sqrt(n1) * sqrt(n2)));
    if ((double)cosParallaxRays > 0.9998) return -1.f;
    float R21[9], t21[3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) R21[3 * i + j] = R12[3 * j + i];
#pragma unroll
    for (int i = 0; i < 3; ++i)
        t21[i] = -__fadd_rn(__fadd_rn(__fmul_rn(R21[3 * i], t12[0]), __fmul_rn(R21[3 * i + 1], t12[1])), __fmul_rn(R21[3 * i + 2], t12[2]));
    const float p1x = r1[0], p1y = r1[1], p2x = r2[0], p2y = r2[1];
    const float A[16] = {-1.f, 0.f, p1x, 0.f,
                         0.f, -1.f, p1y, 0.f,
                         __fsub_rn(__fmul_rn(p2x, R21[6]), R21[0]), __fsub_rn(__fmul_rn(p2x, R21[7]), R21[1]),
                         __fsub_rn(__fmul_rn(p2x, R21[8]), R21[2]), __fsub_rn(__fmul_rn(p2x, t21[2]), t21[0]),
                         __fsub_rn(__fmul_rn(p2y, R21[6]), R21[3]), __fsub_rn(__fmul_rn(p2y, R21[7]), R21[4]),
                         __fsub_rn(__fmul_rn(p2y, R21[8]), R21[5]), __fsub_rn(__fmul_rn(p2y, t21[2]), t21[1])};
    double xh[4];
    smallest_right_singular_vector(A, xh);
    float x3D[3] = {(float)(xh[0] / xh[3]), (float)(xh[1] / xh[3]), (float)(xh[2] / xh[3])};
    const float z1 = x3D[2];
    if (!(z1 > 0.f)) return -1.f;
    const float z2 = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(R21[6], x3D[0]), __fmul_rn(R21[7], x3D[1])), __fmul_rn(R21[8], x3D[2])), t21[2]);
    if (!(z2 > 0.f)) return -1.f;
    float pu, pv;
    kb8_project(cam1, x3D, pu, pv);
    const float ex1 = __fsub_rn(pu, u1), ey1 = __fsub_rn(pv, v1);
    if ((double)__fadd_rn(__fmul_rn(ex1, ex1), __fmul_rn(ey1, ey1)) > 5.991 * (double)sigmaLevel) return -1.f;
    float x2[3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
        x2[i] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(R21[3 * i], x3D[0]), __fmul_rn(R21[3 * i + 1], x3D[1])), __fmul_rn(R21[3 * i + 2], x3D[2])), t21[i]);
    kb8_project(cam2, x2, pu, pv);
    const float ex2 = __fsub_rn(pu, u2), ey2 = __fsub_rn(pv, v2);
    if ((double)__fadd_rn(__fmul_rn(ex2, ex2), __fmul_rn(ey2, ey2)) > 5.991 * (double)unc) return -1.f;
    p3D[0] = x3D[0]; p3D[1] = x3D[1]; p3D[2] = x3D[2];
    return z1;
}

