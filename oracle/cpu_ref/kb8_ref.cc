// kb8_ref.cc — Kannala-Brandt fisheye model and stereo triangulation check (see match_ref.h).
// TEST INFRASTRUCTURE: oracle + timed CPU baseline only.
#include <algorithm>
#include <cmath>
#include <cstring>

#include "match_ref.h"

namespace orbref {

// KannalaBrandt8.cpp:24-40
void kb8Project(const Kb8& c, const float p[3], float uv[2]) {
    const float x2_plus_y2 = p[0] * p[0] + p[1] * p[1];
    const float theta = atan2f(sqrtf(x2_plus_y2), p[2]);
    const float psi = atan2f(p[1], p[0]);
    const float theta2 = theta * theta, theta3 = theta * theta2, theta5 = theta3 * theta2, theta7 = theta5 * theta2,
                theta9 = theta7 * theta2;
    const float r = theta + c.k0 * theta3 + c.k1 * theta5 + c.k2 * theta7 + c.k3 * theta9;
    uv[0] = (float)(c.fx * r * cos(psi) + c.cx);   // cos(float) promotes to double in the reference, too
    uv[1] = (float)(c.fy * r * sin(psi) + c.cy);
}

// KannalaBrandt8.cpp:96-123 (precision = 1e-6, KannalaBrandt8.h:39)
void kb8Unproject(const Kb8& c, float u, float v, float ray[3]) {
    const float precision = 1e-6f;
    const float pwx = (u - c.cx) / c.fx, pwy = (v - c.cy) / c.fy;
    float scale = 1.f;
    float theta_d = sqrtf(pwx * pwx + pwy * pwy);
    theta_d = fminf(fmaxf(-3.1415926535897932384626433832795 / 2.f, theta_d), 3.1415926535897932384626433832795 / 2.f);
    if (theta_d > 1e-8) {
        float theta = theta_d;
        for (int j = 0; j < 10; j++) {
            float theta2 = theta * theta, theta4 = theta2 * theta2, theta6 = theta4 * theta2, theta8 = theta4 * theta4;
            float k0_theta2 = c.k0 * theta2, k1_theta4 = c.k1 * theta4;
            float k2_theta6 = c.k2 * theta6, k3_theta8 = c.k3 * theta8;
            float theta_fix = (theta * (1 + k0_theta2 + k1_theta4 + k2_theta6 + k3_theta8) - theta_d) /
                              (1 + 3 * k0_theta2 + 5 * k1_theta4 + 7 * k2_theta6 + 9 * k3_theta8);
            theta = theta - theta_fix;
            if (fabsf(theta_fix) < precision) break;
        }
        scale = std::tan(theta) / theta_d;
    }
    ray[0] = pwx * scale; ray[1] = pwy * scale; ray[2] = 1.f;
}

// Right singular vector of the smallest singular value of the 4x4 matrix A (what Triangulate reads from
// cv::SVD::compute): eigenvector of A^T A by cyclic Jacobi rotations in double precision.
static void smallestRightSingularVector(const float A[16], double x[4]) {
    double M[4][4], V[4][4];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            double s = 0;
            for (int k = 0; k < 4; ++k) s += (double)A[k * 4 + i] * (double)A[k * 4 + j];
            M[i][j] = s;
            V[i][j] = i == j;
        }
    for (int sweep = 0; sweep < 30; ++sweep) {
        double off = 0;
        for (int p = 0; p < 4; ++p)
            for (int q = p + 1; q < 4; ++q) off += M[p][q] * M[p][q];
        if (off < 1e-300) break;
        for (int p = 0; p < 4; ++p)
            for (int q = p + 1; q < 4; ++q) {
                if (M[p][q] == 0) continue;
                const double theta = (M[q][q] - M[p][p]) / (2 * M[p][q]);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1));
                const double cs = 1 / std::sqrt(t * t + 1), sn = t * cs;
                for (int k = 0; k < 4; ++k) {
                    const double a = M[k][p], b = M[k][q];
                    M[k][p] = cs * a - sn * b; M[k][q] = sn * a + cs * b;
                }
                for (int k = 0; k < 4; ++k) {
                    const double a = M[p][k], b = M[q][k];
                    M[p][k] = cs * a - sn * b; M[q][k] = sn * a + cs * b;
                }
                for (int k = 0; k < 4; ++k) {
                    const double a = V[k][p], b = V[k][q];
                    V[k][p] = cs * a - sn * b; V[k][q] = sn * a + cs * b;
                }
            }
    }
    int m = 0;
    for (int i = 1; i < 4; ++i)
        if (M[i][i] < M[m][m]) m = i;
    for (int k = 0; k < 4; ++k) x[k] = V[k][m];
}

static inline void margin_update(float* margin, double value, double threshold) {
    if (!margin) return;
    const double rel = std::fabs(value - threshold) / std::max(std::fabs(threshold), 1e-12);
    *margin = std::min(*margin, (float)rel);
}

// KannalaBrandt8.cpp:338-407
float kb8TriangulateMatches(const Kb8& c1, const Kb8& c2, const KeyPoint& kp1, const KeyPoint& kp2, const float R12[9],
                            const float t12[3], float sigmaLevel, float unc, float p3D[3], float* margin) {
    if (margin) *margin = 1e30f;
    float r1[3], r2[3], r21[3];
    kb8Unproject(c1, kp1.x, kp1.y, r1);
    kb8Unproject(c2, kp2.x, kp2.y, r2);
    for (int i = 0; i < 3; ++i) r21[i] = R12[3 * i] * r2[0] + R12[3 * i + 1] * r2[1] + R12[3 * i + 2] * r2[2];
    double dot = 0, n1 = 0, n2 = 0;
    for (int i = 0; i < 3; ++i) { dot += (double)r1[i] * r21[i]; n1 += (double)r1[i] * r1[i]; n2 += (double)r21[i] * r21[i]; }
    const float cosParallaxRays = (float)(dot / (std::sqrt(n1) * std::sqrt(n2)));
    if (margin) *margin = std::min(*margin, (float)(std::fabs((double)cosParallaxRays - 0.9998) / 2e-4));   // relative to 1 - threshold
    if (cosParallaxRays > 0.9998) return -1;
    // Tcw1 = [I | 0], Tcw2 = [R21 | t21], R21 = R12^T, t21 = -R21 t12
    float R21[9], t21[3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) R21[3 * i + j] = R12[3 * j + i];
    for (int i = 0; i < 3; ++i) t21[i] = -(R21[3 * i] * t12[0] + R21[3 * i + 1] * t12[1] + R21[3 * i + 2] * t12[2]);
    const float p1x = r1[0], p1y = r1[1], p2x = r2[0], p2y = r2[1];
    float A[16] = {-1.f, 0.f, p1x, 0.f,
                   0.f, -1.f, p1y, 0.f,
                   p2x * R21[6] - R21[0], p2x * R21[7] - R21[1], p2x * R21[8] - R21[2], p2x * t21[2] - t21[0],
                   p2y * R21[6] - R21[3], p2y * R21[7] - R21[4], p2y * R21[8] - R21[5], p2y * t21[2] - t21[1]};
    double xh[4];
    smallestRightSingularVector(A, xh);
    float x3D[3] = {(float)(xh[0] / xh[3]), (float)(xh[1] / xh[3]), (float)(xh[2] / xh[3])};
    const float z1 = x3D[2];
    if (z1 <= 0) return -1;
    const float z2 = R21[6] * x3D[0] + R21[7] * x3D[1] + R21[8] * x3D[2] + t21[2];
    if (z2 <= 0) return -1;
    float uv1[2];
    kb8Project(c1, x3D, uv1);
    const float errX1 = uv1[0] - kp1.x, errY1 = uv1[1] - kp1.y;
    margin_update(margin, errX1 * errX1 + errY1 * errY1, 5.991 * sigmaLevel);
    if ((errX1 * errX1 + errY1 * errY1) > 5.991 * sigmaLevel) return -1;
    float x3D2[3];
    for (int i = 0; i < 3; ++i) x3D2[i] = R21[3 * i] * x3D[0] + R21[3 * i + 1] * x3D[1] + R21[3 * i + 2] * x3D[2] + t21[i];
    float uv2[2];
    kb8Project(c2, x3D2, uv2);
    const float errX2 = uv2[0] - kp2.x, errY2 = uv2[1] - kp2.y;
    margin_update(margin, errX2 * errX2 + errY2 * errY2, 5.991 * unc);
    if ((errX2 * errX2 + errY2 * errY2) > 5.991 * unc) return -1;
    p3D[0] = x3D[0]; p3D[1] = x3D[1]; p3D[2] = x3D[2];
    return z1;
}

// Frame.cc:1116-1156
int computeStereoFishEyeMatches(const Kb8& c1, const Kb8& c2, const KeyPoint* kpL, const uint8_t* descL, int nL, int monoLeft,
                                const KeyPoint* kpR, const uint8_t* descR, int nR, int monoRight, const float R12[9],
                                const float t12[3], const float* sigma2, int32_t* leftToRight, int32_t* rightToLeft,
                                float* depth, float* p3d, float* margins) {
    for (int i = 0; i < nL; ++i) { leftToRight[i] = -1; depth[i] = -1.f; p3d[3 * i] = p3d[3 * i + 1] = p3d[3 * i + 2] = 0.f; if (margins) margins[i] = 1e30f; }
    for (int i = 0; i < nR; ++i) rightToLeft[i] = -1;
    const int nq = nL - monoLeft, nt = nR - monoRight;
    if (nq <= 0) return 0;
    std::vector<int32_t> match(nq);
    knn2RatioMatch(descL + 32 * (size_t)monoLeft, nq, descR + 32 * (size_t)monoRight, nt, 0.7, match.data());
    int nMatches = 0;
    for (int q = 0; q < nq; ++q) {
        if (match[q] < 0) continue;
        const int iL = q + monoLeft, iR = match[q] + monoRight;
        float p[3], mg;
        const float d = kb8TriangulateMatches(c1, c2, kpL[iL], kpR[iR], R12, t12, sigma2[kpL[iL].octave], sigma2[kpR[iR].octave], p, &mg);
        if (margins) margins[iL] = std::min(mg, (float)(std::fabs(d - 0.0001f) / 0.0001f));
        if (d > 0.0001f) {
            leftToRight[iL] = iR;
            rightToLeft[iR] = iL;
            p3d[3 * iL] = p[0]; p3d[3 * iL + 1] = p[1]; p3d[3 * iL + 2] = p[2];
            depth[iL] = d;
            nMatches++;
        }
    }
    return nMatches;
}

}  // namespace orbref
