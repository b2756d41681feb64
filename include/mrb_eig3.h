/* mrb_eig3.h -- deterministic symmetric 3x3 eigen-solver shared by the CUDA kernels and the CPU oracle.
 *
 * The reference obtains surface normals from Eigen::SelfAdjointEigenSolver<Matrix3f> (dfields.cpp:293-309),
 * whose floating-point path (tridiagonalisation + implicit QL inside an unpinned Eigen version, README.md:5)
 * cannot be reproduced bit for bit and is not in /root/reference.  The feature sub-path is therefore
 * "parity unpinned" against the reference; this header DEFINES the arithmetic both of our
 * implementations use: a cyclic Jacobi iteration with a fixed sweep count, built from + - * / sqrtf only
 * (all correctly rounded on the host with -ffp-contract=off and on the device with -fmad=false), so that the
 * GPU result equals the oracle result bit for bit.
 */
#ifndef MRB_EIG3_H
#define MRB_EIG3_H

#include <math.h>

#ifdef __CUDACC__
#define MRB_HD __host__ __device__ __forceinline__
#else
#define MRB_HD static inline
#endif

#define MRB_EIG3_SWEEPS 10

/* a: symmetric matrix (row-major 3x3, destroyed); w: eigenvalues; v: eigenvectors in columns (v[r*3+c]). */
MRB_HD void mrb_eig3_jacobi(float a[9], float w[3], float v[9]) {
    int i, sweep, p, q, r;
    for (i = 0; i < 9; i++) v[i] = (i % 4 == 0) ? 1.0f : 0.0f;
    for (sweep = 0; sweep < MRB_EIG3_SWEEPS; sweep++) {
        for (p = 0; p < 2; p++) {
            for (q = p + 1; q < 3; q++) {
                float apq = a[p * 3 + q];
                if (apq == 0.0f) continue;
                {
                    float app = a[p * 3 + p], aqq = a[q * 3 + q];
                    float theta = (aqq - app) / (2.0f * apq);
                    float at = theta < 0.0f ? -theta : theta;
                    float t = 1.0f / (at + sqrtf(theta * theta + 1.0f));
                    float c, s;
                    if (theta < 0.0f) t = -t;
                    c = 1.0f / sqrtf(t * t + 1.0f);
                    s = t * c;
                    a[p * 3 + p] = app - t * apq;
                    a[q * 3 + q] = aqq + t * apq;
                    a[p * 3 + q] = 0.0f;
                    a[q * 3 + p] = 0.0f;
                    r = 3 - p - q;                       /* the third index */
                    {
                        float arp = a[r * 3 + p], arq = a[r * 3 + q];
                        float nrp = c * arp - s * arq, nrq = s * arp + c * arq;
                        a[r * 3 + p] = nrp; a[p * 3 + r] = nrp;
                        a[r * 3 + q] = nrq; a[q * 3 + r] = nrq;
                    }
                    for (r = 0; r < 3; r++) {
                        float vrp = v[r * 3 + p], vrq = v[r * 3 + q];
                        v[r * 3 + p] = c * vrp - s * vrq;
                        v[r * 3 + q] = s * vrp + c * vrq;
                    }
                }
            }
        }
    }
    w[0] = a[0]; w[1] = a[4]; w[2] = a[8];
}

/* getNormalVector (dfields.cpp:293-309): Eigen returns the eigenvalues in increasing order; the reference
 * then compares them truncated to int and keeps the first minimum, which for these covariances is always the
 * smallest eigenvalue.  Here: order the three eigenvalues increasingly (stable), run the same int-truncated
 * scan, normalise the chosen eigenvector. */
MRB_HD void mrb_normal_from_covariance(const float cov[9], float n[3]) {
    float a[9], w[3], v[9];
    int order[3] = {0, 1, 2};
    int i, j, index = 0;
    float mn = 10000.0f, len;
    for (i = 0; i < 9; i++) a[i] = cov[i];
    mrb_eig3_jacobi(a, w, v);
    for (i = 1; i < 3; i++)                              /* stable insertion sort by eigenvalue */
        for (j = i; j > 0 && w[order[j]] < w[order[j - 1]]; j--) { int t = order[j]; order[j] = order[j - 1]; order[j - 1] = t; }
    for (i = 0; i < 3; i++) {
        int eigenvalue = (int)w[order[i]];
        if ((float)eigenvalue < mn) { mn = (float)eigenvalue; index = i; }
    }
    n[0] = v[0 * 3 + order[index]]; n[1] = v[1 * 3 + order[index]]; n[2] = v[2 * 3 + order[index]];
    len = sqrtf(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
    n[0] = n[0] / len; n[1] = n[1] / len; n[2] = n[2] / len;
}

/* computePoint (mcubes.cpp:142-152): least-squares solution of  M p = vec  (M: n x 3 rows of unit normals) plus
 * the centroid shift.  The reference uses Eigen::JacobiSVD (+ setThreshold for edge features); our DEFINITION (shared
 * by oracle and GPU, parity unpinned against Eigen) goes through the 3x3 normal equations and the Jacobi solver above:
 *   A = M^T M = V diag(w) V^T,  sigma_i = sqrt(max(w_i, 0)),  p = sum_{kept i} v_i (v_i . M^T vec) / w_i,
 * keeping sigma_i > thr * sigma_max with thr = sigma_min / (sigma_max - 0.0001) for edges (this always drops the
 * smallest singular value: the solution is constrained to the edge line's normal plane) and
 * thr = 3 * FLT_EPSILON for corners (Eigen's default rank threshold for a 3-column matrix). */
MRB_HD void mrb_feature_point(const float* M, const float* vec, int n, const float shift[3], int is_edge, float out[3]) {
    float a[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, b[3] = {0, 0, 0}, w[3], v[9], sg[3];
    float smax, smin, thr, p[3] = {0, 0, 0};
    int i, r, c;
    for (i = 0; i < n; i++)
        for (r = 0; r < 3; r++) {
            for (c = 0; c < 3; c++) a[r * 3 + c] = a[r * 3 + c] + M[i * 3 + r] * M[i * 3 + c];
            b[r] = b[r] + M[i * 3 + r] * vec[i];
        }
    mrb_eig3_jacobi(a, w, v);
    for (i = 0; i < 3; i++) sg[i] = sqrtf(w[i] > 0.0f ? w[i] : 0.0f);
    smax = sg[0]; smin = sg[0];
    for (i = 1; i < 3; i++) { if (sg[i] > smax) smax = sg[i]; if (sg[i] < smin) smin = sg[i]; }
    thr = is_edge ? smin / (smax - 0.0001f) : 3.0f * 1.1920929e-7f;
    for (i = 0; i < 3; i++) {
        if (sg[i] > thr * smax) {
            float proj = v[0 * 3 + i] * b[0] + v[1 * 3 + i] * b[1] + v[2 * 3 + i] * b[2];
            float coef = proj / w[i];
            p[0] = p[0] + v[0 * 3 + i] * coef; p[1] = p[1] + v[1 * 3 + i] * coef; p[2] = p[2] + v[2 * 3 + i] * coef;
        }
    }
    out[0] = p[0] + shift[0]; out[1] = p[1] + shift[1]; out[2] = p[2] + shift[2];
}

#endif /* MRB_EIG3_H */
