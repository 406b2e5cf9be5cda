// svo_epnp.cuh -- EPnP pose from a minimal 5-point sample, host+device.
//
// cv::solvePnPRansac (called by the reference at src/vo.cpp:301-304) estimates every RANSAC
// hypothesis with SOLVEPNP_EPNP on 5 points whatever `flags` says (flags = SOLVEPNP_ITERATIVE is
// only used for the final refit on the inliers).  This restates OpenCV's calib3d epnp.cpp
// (Lepetit/Moreno-Noguer/Fua): PCA control points, barycentric coordinates, the 12x12 M^T M
// null space, the three beta approximations, 5 Gauss-Newton steps each, absolute orientation by
// SVD, best reprojection error wins.  All FP64.
#pragma once
#include "svo_core.cuh"

namespace svo {

// Cyclic Jacobi eigen-decomposition of a symmetric N x N matrix (row-major, destroyed).
// On return d[i] are the eigenvalues sorted in DESCENDING order and Vt row i is the matching
// unit eigenvector (i.e. the layout of cvSVD(..., CV_SVD_U_T) on a symmetric PSD matrix).
template <int N>
SVO_HD void jacobi_eigen_sym(double* A, double* Vt, double* d) {
    for (int i = 0; i < N; ++i)
        for (int j = 0; j < N; ++j) Vt[i * N + j] = (i == j) ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 60; ++sweep) {
        double off = 0.0, diag = 0.0;
        for (int i = 0; i < N; ++i) {
            diag += A[i * N + i] * A[i * N + i];
            for (int j = i + 1; j < N; ++j) off += A[i * N + j] * A[i * N + j];
        }
        if (off <= 1e-30 * diag || off == 0.0) break;
        for (int p = 0; p < N - 1; ++p)
            for (int q = p + 1; q < N; ++q) {
                const double apq = A[p * N + q];
                if (apq == 0.0) continue;
                const double app = A[p * N + p], aqq = A[q * N + q];
                const double theta = (aqq - app) / (2.0 * apq);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < N; ++k) {   // A <- A J
                    const double akp = A[k * N + p], akq = A[k * N + q];
                    A[k * N + p] = c * akp - s * akq;
                    A[k * N + q] = s * akp + c * akq;
                }
                for (int k = 0; k < N; ++k) {   // A <- J^T A
                    const double apk = A[p * N + k], aqk = A[q * N + k];
                    A[p * N + k] = c * apk - s * aqk;
                    A[q * N + k] = s * apk + c * aqk;
                }
                for (int k = 0; k < N; ++k) {   // rows of Vt are eigenvectors
                    const double vp = Vt[p * N + k], vq = Vt[q * N + k];
                    Vt[p * N + k] = c * vp - s * vq;
                    Vt[q * N + k] = s * vp + c * vq;
                }
            }
    }
    for (int i = 0; i < N; ++i) d[i] = A[i * N + i];
    for (int i = 0; i < N - 1; ++i) {   // selection sort, descending
        int m = i;
        for (int j = i + 1; j < N; ++j)
            if (d[j] > d[m]) m = j;
        if (m != i) {
            const double td = d[i];
            d[i] = d[m];
            d[m] = td;
            for (int k = 0; k < N; ++k) {
                const double tv = Vt[i * N + k];
                Vt[i * N + k] = Vt[m * N + k];
                Vt[m * N + k] = tv;
            }
        }
    }
}

// Least squares min ||A x - b|| for a ROWS x COLS system (row-major A, destroyed) by Householder QR.
// Returns false if A is (numerically) rank deficient.
template <int ROWS, int COLS>
SVO_HD bool householder_ls(double* A, double* b, double* x) {
#pragma unroll
    for (int k = 0; k < COLS; ++k) {
        double norm = 0.0;
        for (int i = k; i < ROWS; ++i) norm += A[i * COLS + k] * A[i * COLS + k];
        norm = sqrt(norm);
        if (norm == 0.0) return false;
        const double alpha = A[k * COLS + k] > 0 ? -norm : norm;
        // v = a_k - alpha e_k
        double v[ROWS];
        for (int i = 0; i < ROWS; ++i) v[i] = (i >= k) ? A[i * COLS + k] : 0.0;
        v[k] -= alpha;
        double vtv = 0.0;
        for (int i = k; i < ROWS; ++i) vtv += v[i] * v[i];
        if (vtv == 0.0) return false;
        for (int j = k; j < COLS; ++j) {
            double dot = 0.0;
            for (int i = k; i < ROWS; ++i) dot += v[i] * A[i * COLS + j];
            const double f = 2.0 * dot / vtv;
            for (int i = k; i < ROWS; ++i) A[i * COLS + j] -= f * v[i];
        }
        double dot = 0.0;
        for (int i = k; i < ROWS; ++i) dot += v[i] * b[i];
        const double f = 2.0 * dot / vtv;
        for (int i = k; i < ROWS; ++i) b[i] -= f * v[i];
    }
    for (int k = COLS - 1; k >= 0; --k) {
        double s = b[k];
        for (int j = k + 1; j < COLS; ++j) s -= A[k * COLS + j] * x[j];
        const double r = A[k * COLS + k];
        if (fabs(r) < 1e-300) return false;
        x[k] = s / r;
    }
    return true;
}

// SVD of a 3x3 matrix A = U diag(S) V^T by one-sided Jacobi.  U, V row-major.
SVO_HD void svd3x3(const double* Ain, double* U, double* S, double* V) {
    double A[9];
    for (int i = 0; i < 9; ++i) {
        A[i] = Ain[i];
        V[i] = (i % 4 == 0) ? 1.0 : 0.0;
    }
    for (int sweep = 0; sweep < 60; ++sweep) {
        bool changed = false;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                double a = 0, b = 0, c = 0;
                for (int r = 0; r < 3; ++r) {
                    a += A[r * 3 + p] * A[r * 3 + p];
                    b += A[r * 3 + q] * A[r * 3 + q];
                    c += A[r * 3 + p] * A[r * 3 + q];
                }
                if (fabs(c) <= DBL_EPSILON * sqrt(a * b)) continue;
                changed = true;
                const double zeta = (b - a) / (2.0 * c);
                const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double cs = 1.0 / sqrt(1.0 + t * t), sn = cs * t;
                for (int r = 0; r < 3; ++r) {
                    const double ap = A[r * 3 + p], aq = A[r * 3 + q];
                    A[r * 3 + p] = cs * ap - sn * aq;
                    A[r * 3 + q] = sn * ap + cs * aq;
                    const double vp = V[r * 3 + p], vq = V[r * 3 + q];
                    V[r * 3 + p] = cs * vp - sn * vq;
                    V[r * 3 + q] = sn * vp + cs * vq;
                }
            }
        if (!changed) break;
    }
    for (int k = 0; k < 3; ++k) S[k] = sqrt(A[k] * A[k] + A[3 + k] * A[3 + k] + A[6 + k] * A[6 + k]);
    // order columns by descending singular value
    for (int i = 0; i < 2; ++i) {
        int m = i;
        for (int j = i + 1; j < 3; ++j)
            if (S[j] > S[m]) m = j;
        if (m != i) {
            double ts = S[i]; S[i] = S[m]; S[m] = ts;
            for (int r = 0; r < 3; ++r) {
                double ta = A[r * 3 + i]; A[r * 3 + i] = A[r * 3 + m]; A[r * 3 + m] = ta;
                double tv = V[r * 3 + i]; V[r * 3 + i] = V[r * 3 + m]; V[r * 3 + m] = tv;
            }
        }
    }
    for (int k = 0; k < 3; ++k) {
        if (S[k] > 1e-12 * (S[0] > 0 ? S[0] : 1.0)) {
            for (int r = 0; r < 3; ++r) U[r * 3 + k] = A[r * 3 + k] / S[k];
        } else {   // complete the basis
            const int a = (k + 1) % 3, b = (k + 2) % 3;
            U[0 * 3 + k] = U[1 * 3 + a] * U[2 * 3 + b] - U[2 * 3 + a] * U[1 * 3 + b];
            U[1 * 3 + k] = U[2 * 3 + a] * U[0 * 3 + b] - U[0 * 3 + a] * U[2 * 3 + b];
            U[2 * 3 + k] = U[0 * 3 + a] * U[1 * 3 + b] - U[1 * 3 + a] * U[0 * 3 + b];
        }
    }
}

SVO_HD bool inv3x3(const double* m, double* o) {
    const double c0 = m[4] * m[8] - m[5] * m[7], c1 = m[5] * m[6] - m[3] * m[8], c2 = m[3] * m[7] - m[4] * m[6];
    const double det = m[0] * c0 + m[1] * c1 + m[2] * c2;
    if (det == 0.0 || !(fabs(det) > 0.0)) return false;
    const double id = 1.0 / det;
    o[0] = c0 * id; o[1] = (m[2] * m[7] - m[1] * m[8]) * id; o[2] = (m[1] * m[5] - m[2] * m[4]) * id;
    o[3] = c1 * id; o[4] = (m[0] * m[8] - m[2] * m[6]) * id; o[5] = (m[2] * m[3] - m[0] * m[5]) * id;
    o[6] = c2 * id; o[7] = (m[1] * m[6] - m[0] * m[7]) * id; o[8] = (m[0] * m[4] - m[1] * m[3]) * id;
    return true;
}

// ---------------------------------------------------------------------------
// Cooperative EPnP for one 5-point sample: executed by LANES lanes (32 on the device: one warp per
// RANSAC hypothesis; 1 in the host unit tests -- identical arithmetic, the lane loops are just serial).
// `Red` provides sync() and sumd() (sum of a double across the lanes).
struct EpnpWork {              // per-hypothesis scratch (shared memory on the device)
    double A[144];             // M^T M, diagonalised in place
    double Vt[144];            // rows = eigenvectors
    double M[120];             // 10 x 12
    double cs[12];             // (c, s) of the 6 concurrent rotations
    unsigned char pq[136];     // round-robin pair table [11][6][2]
    double dv[72];             // [4][6][3]
    double L[60], rho[6];
    double pws[15], us[10], alphas[20], cws[12];
    double Rs[3][9], ts[3][3], errs[3];
    int order[4];              // indices of the 4 smallest eigenvalues, ascending
    int ok;
};

struct SerialRedD {
    static SVO_HD void sync() {}
    static SVO_HD double sumd(double v) { return v; }
};

// round-robin pairing of 12 indices: round r in [0, 11), slot k in [0, 6)
SVO_HD void rr_pair(int r, int k, int& p, int& q) {
    int a, b;
    if (k == 0) { a = 11; b = r; }
    else { a = (r + k) % 11; b = (r - k + 11) % 11; }
    p = a < b ? a : b;
    q = a < b ? b : a;
}

// Jacobi rotation zeroing a_pq of a symmetric matrix: one sqrt, one division, one rsqrt.
SVO_HD void jacobi_cs(double app, double aqq, double apq, double& c, double& s) {
    c = 1.0;
    s = 0.0;
    if (apq != 0.0) {
        const double h = aqq - app;
        const double r = sqrt(h * h + 4.0 * apq * apq);
        const double t = (h >= 0 ? 2.0 : -2.0) * apq / (fabs(h) + r);
#if defined(__CUDA_ARCH__)
        c = rsqrt(t * t + 1.0);
#else
        c = 1.0 / sqrt(t * t + 1.0);
#endif
        s = t * c;
    }
}

// Parallel-order (round-robin) Jacobi on the 12x12 symmetric matrix w.A: 11 rounds of 6 disjoint
// rotations per sweep.  Per round: 6 lanes compute (c, s); then the 21 upper 2x2 blocks
// A_blk <- J_i^T A_blk J_j are updated in place (each block is owned by one lane, the mirror
// block is written too so the matrix stays exactly symmetric) together with the 6 x 12 row pairs of Vt.
template <int LANES, class Red>
SVO_HD void jacobi12_parallel(EpnpWork& w, int lane) {
    for (int i = lane; i < 144; i += LANES) w.Vt[i] = (i / 12 == i % 12) ? 1.0 : 0.0;
    for (int i = lane; i < 66; i += LANES) {
        int p, q;
        rr_pair(i / 6, i % 6, p, q);
        w.pq[2 * i] = (unsigned char)p;
        w.pq[2 * i + 1] = (unsigned char)q;
    }
    Red::sync();
    constexpr int NB = (21 + LANES - 1) / LANES;   // blocks per lane
    constexpr int NV = (72 + LANES - 1) / LANES;   // Vt items per lane
    // block l -> (ki <= kj) in row-major upper-triangular order, decoded once (packed ki | kj << 4)
    unsigned char blk[NB];
#pragma unroll
    for (int bi = 0; bi < NB; ++bi) {
        int ki = 0, rem = lane + bi * LANES;
        while (ki < 5 && rem >= 6 - ki) { rem -= 6 - ki; ++ki; }
        blk[bi] = (unsigned char)(ki | ((ki + rem) << 4));
    }
    for (int sweep = 0; sweep < 30; ++sweep) {
        double off = 0.0, diag = 0.0;
        for (int i = lane; i < 144; i += LANES) {
            const double v = w.A[i] * w.A[i];
            if (i / 12 == i % 12) diag += v; else off += v;
        }
        off = Red::sumd(off);
        diag = Red::sumd(diag);
#ifdef SVO_DEBUG_SWEEPS
        printf("sweep %d off %.3g diag %.3g\n", sweep, off, diag);
#endif
        if (off <= 1e-22 * diag || off == 0.0) break;
        for (int r = 0; r < 11; ++r) {
            const unsigned char* pq = w.pq + 12 * r;
            for (int k = lane; k < 6; k += LANES) {
                const int p = pq[2 * k], q = pq[2 * k + 1];
                jacobi_cs(w.A[p * 13], w.A[q * 13], w.A[p * 12 + q], w.cs[2 * k], w.cs[2 * k + 1]);
            }
            Red::sync();
#pragma unroll
            for (int bi = 0; bi < NB; ++bi) {
                const int l = lane + bi * LANES;
                if (l < 21) {
                    const int ki = blk[bi] & 15, kj = blk[bi] >> 4;
                    const int pi = pq[2 * ki], qi = pq[2 * ki + 1], pj = pq[2 * kj], qj = pq[2 * kj + 1];
                    const double ci = w.cs[2 * ki], si = w.cs[2 * ki + 1], cj = w.cs[2 * kj], sj = w.cs[2 * kj + 1];
                    const double a = w.A[pi * 12 + pj], b = w.A[pi * 12 + qj], c = w.A[qi * 12 + pj], d = w.A[qi * 12 + qj];
                    const double a1 = ci * a - si * c, b1 = ci * b - si * d, c1 = si * a + ci * c, d1 = si * b + ci * d;
                    double a2 = cj * a1 - sj * b1, b2 = sj * a1 + cj * b1, c2 = cj * c1 - sj * d1, d2 = sj * c1 + cj * d1;
                    if (ki == kj) { b2 = 0.0; c2 = 0.0; }
                    w.A[pi * 12 + pj] = a2; w.A[pi * 12 + qj] = b2; w.A[qi * 12 + pj] = c2; w.A[qi * 12 + qj] = d2;
                    w.A[pj * 12 + pi] = a2; w.A[qj * 12 + pi] = b2; w.A[pj * 12 + qi] = c2; w.A[qj * 12 + qi] = d2;
                }
            }
#pragma unroll
            for (int vi = 0; vi < NV; ++vi) {
                const int it = lane + vi * LANES;
                if (it < 72) {
                    const int k = it / 12, col = it - k * 12;
                    const int p = pq[2 * k], q = pq[2 * k + 1];
                    const double c = w.cs[2 * k], s = w.cs[2 * k + 1];
                    const double vp = w.Vt[p * 12 + col], vq = w.Vt[q * 12 + col];
                    w.Vt[p * 12 + col] = c * vp - s * vq;
                    w.Vt[q * 12 + col] = s * vp + c * vq;
                }
            }
            Red::sync();
        }
    }
}

// least squares with a run-time column count (<= 5), 6 rows; same instruction stream for the three
// beta approximations so that lanes 0..2 do not diverge
SVO_HD bool householder_ls6(double* A /*6 x 5 storage, row stride 5*/, double* b, double* x, int cols) {
    for (int k = 0; k < cols; ++k) {
        double norm = 0.0;
        for (int i = k; i < 6; ++i) norm += A[i * 5 + k] * A[i * 5 + k];
        norm = sqrt(norm);
        if (norm == 0.0) return false;
        const double alpha = A[k * 5 + k] > 0 ? -norm : norm;
        double v[6];
        for (int i = 0; i < 6; ++i) v[i] = (i >= k) ? A[i * 5 + k] : 0.0;
        v[k] -= alpha;
        double vtv = 0.0;
        for (int i = k; i < 6; ++i) vtv += v[i] * v[i];
        if (vtv == 0.0) return false;
        for (int j = k; j < cols; ++j) {
            double dot = 0.0;
            for (int i = k; i < 6; ++i) dot += v[i] * A[i * 5 + j];
            const double f = 2.0 * dot / vtv;
            for (int i = k; i < 6; ++i) A[i * 5 + j] -= f * v[i];
        }
        double dot = 0.0;
        for (int i = k; i < 6; ++i) dot += v[i] * b[i];
        const double f = 2.0 * dot / vtv;
        for (int i = k; i < 6; ++i) b[i] -= f * v[i];
    }
    for (int k = cols - 1; k >= 0; --k) {
        double s = b[k];
        for (int j = k + 1; j < cols; ++j) s -= A[k * 5 + j] * x[j];
        const double r = A[k * 5 + k];
        if (fabs(r) < 1e-300) return false;
        x[k] = s / r;
    }
    return true;
}

SVO_HD void epnp_gauss_newton(const double* L, const double* rho, double* betas) {
    for (int it = 0; it < 5; ++it) {
        double A[24], b[6], x[4];
        const double b0 = betas[0], b1 = betas[1], b2 = betas[2], b3 = betas[3];
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            const double* r = L + i * 10;
            double* a = A + i * 4;
            a[0] = 2 * r[0] * b0 + r[1] * b1 + r[3] * b2 + r[6] * b3;
            a[1] = r[1] * b0 + 2 * r[2] * b1 + r[4] * b2 + r[7] * b3;
            a[2] = r[3] * b0 + r[4] * b1 + 2 * r[5] * b2 + r[8] * b3;
            a[3] = r[6] * b0 + r[7] * b1 + r[8] * b2 + 2 * r[9] * b3;
            b[i] = rho[i] - (r[0] * b0 * b0 + r[1] * b0 * b1 + r[2] * b1 * b1 + r[3] * b0 * b2 + r[4] * b1 * b2 +
                             r[5] * b2 * b2 + r[6] * b0 * b3 + r[7] * b1 * b3 + r[8] * b2 * b3 + r[9] * b3 * b3);
        }
        if (!householder_ls<6, 4>(A, b, x)) return;   // OpenCV's qr_solve gives up on a singular A
#pragma unroll
        for (int i = 0; i < 4; ++i) betas[i] += x[i];
    }
}

// Orthogonal polar factor of a 3x3 matrix by scaled Newton iteration.  false if (near-)singular.
SVO_HD bool polar_orthogonal3(const double* M, double* Q) {
    double X[9], n2 = 0.0;
    for (int i = 0; i < 9; ++i) n2 += M[i] * M[i];
    if (!(n2 > 0.0)) return false;
    const double sc = sqrt(3.0 / n2);
    for (int i = 0; i < 9; ++i) X[i] = M[i] * sc;
    for (int it = 0; it < 40; ++it) {
        // cofactors: inv(X)^T = cof(X) / det(X)
        double C[9];
        C[0] = X[4] * X[8] - X[5] * X[7]; C[1] = X[5] * X[6] - X[3] * X[8]; C[2] = X[3] * X[7] - X[4] * X[6];
        C[3] = X[2] * X[7] - X[1] * X[8]; C[4] = X[0] * X[8] - X[2] * X[6]; C[5] = X[1] * X[6] - X[0] * X[7];
        C[6] = X[1] * X[5] - X[2] * X[4]; C[7] = X[2] * X[3] - X[0] * X[5]; C[8] = X[0] * X[4] - X[1] * X[3];
        const double det = X[0] * C[0] + X[1] * C[1] + X[2] * C[2];
        if (!(fabs(det) > 1e-9)) return false;   // X has unit-scale entries: rank deficient input
        const double id = 1.0 / det;
        double diff = 0.0;
        for (int i = 0; i < 9; ++i) {
            const double nx = 0.5 * (X[i] + C[i] * id);
            diff += fabs(nx - X[i]);
            X[i] = nx;
        }
        if (diff < 1e-15) break;
    }
    for (int i = 0; i < 9; ++i) Q[i] = X[i];
    return true;
}

// compute_ccs / compute_pcs / solve_for_sign / estimate_R_and_t / reprojection_error of epnp.cpp
SVO_HD double epnp_R_and_t(const EpnpWork& w, const Intrinsics& K, const double* betas, double* R9, double* t3) {
    double ccs[4][3], pcs[15];
    for (int i = 0; i < 4; ++i) ccs[i][0] = ccs[i][1] = ccs[i][2] = 0.0;
    for (int i = 0; i < 4; ++i) {
        const double* v = w.Vt + 12 * w.order[i];
        for (int j = 0; j < 4; ++j)
            for (int k = 0; k < 3; ++k) ccs[j][k] += betas[i] * v[3 * j + k];
    }
    for (int i = 0; i < 5; ++i) {
        const double* a = w.alphas + 4 * i;
        for (int j = 0; j < 3; ++j)
            pcs[3 * i + j] = a[0] * ccs[0][j] + a[1] * ccs[1][j] + a[2] * ccs[2][j] + a[3] * ccs[3][j];
    }
    if (pcs[2] < 0.0)
        for (int i = 0; i < 15; ++i) pcs[i] = -pcs[i];
    double pc0[3] = {0, 0, 0}, pw0[3] = {0, 0, 0};
    for (int i = 0; i < 5; ++i)
        for (int j = 0; j < 3; ++j) {
            pc0[j] += pcs[3 * i + j];
            pw0[j] += w.pws[3 * i + j];
        }
    for (int j = 0; j < 3; ++j) {
        pc0[j] /= 5;
        pw0[j] /= 5;
    }
    double abt[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < 5; ++i)
        for (int j = 0; j < 3; ++j)
            for (int k = 0; k < 3; ++k) abt[3 * j + k] += (pcs[3 * i + j] - pc0[j]) * (w.pws[3 * i + k] - pw0[k]);
    // R = U V^T of the SVD of ABt == the orthogonal polar factor of ABt (det = sign(det ABt)):
    // Newton iteration X <- (X + X^-T) / 2, quadratically convergent; SVD only for a (near-)singular ABt.
    if (!polar_orthogonal3(abt, R9)) {
        double U[9], S[3], V[9];
        svd3x3(abt, U, S, V);
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j)
                R9[i * 3 + j] = U[i * 3] * V[j * 3] + U[i * 3 + 1] * V[j * 3 + 1] + U[i * 3 + 2] * V[j * 3 + 2];
    }
    const double det = R9[0] * R9[4] * R9[8] + R9[1] * R9[5] * R9[6] + R9[2] * R9[3] * R9[7] - R9[2] * R9[4] * R9[6] -
                       R9[1] * R9[3] * R9[8] - R9[0] * R9[5] * R9[7];
    if (det < 0) {
        R9[6] = -R9[6];
        R9[7] = -R9[7];
        R9[8] = -R9[8];
    }
    for (int i = 0; i < 3; ++i) t3[i] = pc0[i] - (R9[i * 3] * pw0[0] + R9[i * 3 + 1] * pw0[1] + R9[i * 3 + 2] * pw0[2]);
    double sum2 = 0.0;
    for (int i = 0; i < 5; ++i) {
        const double* pw = w.pws + 3 * i;
        const double Xc = R9[0] * pw[0] + R9[1] * pw[1] + R9[2] * pw[2] + t3[0];
        const double Yc = R9[3] * pw[0] + R9[4] * pw[1] + R9[5] * pw[2] + t3[1];
        const double inv_Zc = 1.0 / (R9[6] * pw[0] + R9[7] * pw[1] + R9[8] * pw[2] + t3[2]);
        const double ue = K.cx + K.fx * Xc * inv_Zc, ve = K.cy + K.fy * Yc * inv_Zc;
        const double du = w.us[2 * i] - ue, dv = w.us[2 * i + 1] - ve;
        sum2 += sqrt(du * du + dv * dv);
    }
    return sum2 / 5;
}

// X: 5 x 3 object points (float32), uv: 5 x 2 pixel coordinates (float32).  Every lane returns the
// same pose; false when the sample is degenerate.
template <int LANES, class Red>
SVO_HD bool epnp5_solve(EpnpWork& w, const Intrinsics& K, const float* X, const float* uv, int lane, Pose& out) {
    if (lane == 0) {
        w.ok = 1;
        const double ifx = 1.0 / K.fx, ify = 1.0 / K.fy;
        for (int i = 0; i < 5; ++i) {
            for (int j = 0; j < 3; ++j) w.pws[3 * i + j] = (double)X[3 * i + j];
            // cv::undistortPoints (zero distortion) stores normalised coordinates as float32;
            // epnp::init_points maps them back to pixels in double
            const float xn = (float)(((double)uv[2 * i] - K.cx) * ifx);
            const float yn = (float)(((double)uv[2 * i + 1] - K.cy) * ify);
            w.us[2 * i] = (double)xn * K.fx + K.cx;
            w.us[2 * i + 1] = (double)yn * K.fy + K.cy;
        }
        // choose_control_points
        for (int j = 0; j < 3; ++j) {
            double s = 0.0;
            for (int i = 0; i < 5; ++i) s += w.pws[3 * i + j];
            w.cws[j] = s / 5;
        }
        double c[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, vt[9], dc[3];
        for (int i = 0; i < 5; ++i)
            for (int a = 0; a < 3; ++a)
                for (int b = 0; b < 3; ++b)
                    c[3 * a + b] += (w.pws[3 * i + a] - w.cws[a]) * (w.pws[3 * i + b] - w.cws[b]);
        jacobi_eigen_sym<3>(c, vt, dc);
        for (int i = 1; i < 4; ++i) {
            const double k = sqrt((dc[i - 1] > 0 ? dc[i - 1] : 0.0) / 5);
            for (int j = 0; j < 3; ++j) w.cws[3 * i + j] = w.cws[j] + k * vt[3 * (i - 1) + j];
        }
        // compute_barycentric_coordinates
        double cc[9], ci[9];
        for (int i = 0; i < 3; ++i)
            for (int j = 1; j < 4; ++j) cc[3 * i + j - 1] = w.cws[3 * j + i] - w.cws[i];
        if (!inv3x3(cc, ci)) w.ok = 0;
        else
            for (int i = 0; i < 5; ++i) {
                const double* pi = w.pws + 3 * i;
                double* a = w.alphas + 4 * i;
                for (int j = 0; j < 3; ++j)
                    a[1 + j] = ci[3 * j] * (pi[0] - w.cws[0]) + ci[3 * j + 1] * (pi[1] - w.cws[1]) +
                               ci[3 * j + 2] * (pi[2] - w.cws[2]);
                a[0] = 1.0 - a[1] - a[2] - a[3];
            }
        // rho
        for (int r = 0, a = 0, b = 1; r < 6; ++r) {
            double s = 0.0;
            for (int k = 0; k < 3; ++k) s += (w.cws[3 * a + k] - w.cws[3 * b + k]) * (w.cws[3 * a + k] - w.cws[3 * b + k]);
            w.rho[r] = s;
            ++b;
            if (b > 3) { ++a; b = a + 1; }
        }
    }
    Red::sync();
    if (!w.ok) return false;
    // fill_M
    for (int e = lane; e < 120; e += LANES) {
        const int row = e / 12, col = e - row * 12;
        const int i = row >> 1, k = col / 3, c3 = col - 3 * k;
        const double as = w.alphas[4 * i + k];
        double v;
        if ((row & 1) == 0) v = c3 == 0 ? as * K.fx : (c3 == 1 ? 0.0 : as * (K.cx - w.us[2 * i]));
        else v = c3 == 0 ? 0.0 : (c3 == 1 ? as * K.fy : as * (K.cy - w.us[2 * i + 1]));
        w.M[e] = v;
    }
    Red::sync();
    for (int e = lane; e < 144; e += LANES) {
        const int a = e / 12, b = e - a * 12;
        double s = 0.0;
        for (int r = 0; r < 10; ++r) s += w.M[r * 12 + a] * w.M[r * 12 + b];
        w.A[e] = s;
    }
    Red::sync();
    jacobi12_parallel<LANES, Red>(w, lane);
    if (lane == 0) {   // the 4 smallest eigenvalues, ascending (ut rows 11, 10, 9, 8 of cvSVD)
        bool used[12];
        for (int i = 0; i < 12; ++i) used[i] = false;
        for (int k = 0; k < 4; ++k) {
            int m = -1;
            for (int i = 0; i < 12; ++i)
                if (!used[i] && (m < 0 || w.A[i * 13] < w.A[m * 13])) m = i;
            used[m] = true;
            w.order[k] = m;
        }
    }
    Red::sync();
    for (int e = lane; e < 72; e += LANES) {   // dv[i][j][k]
        const int i = e / 18, j = (e - i * 18) / 3, k = e - i * 18 - j * 3;
        const int pa[6] = {0, 0, 0, 1, 1, 2}, pb[6] = {1, 2, 3, 2, 3, 3};
        const double* v = w.Vt + 12 * w.order[i];
        w.dv[e] = v[3 * pa[j] + k] - v[3 * pb[j] + k];
    }
    Red::sync();
    for (int e = lane; e < 60; e += LANES) {   // L_6x10
        const int r = e / 10, c = e - r * 10;
        const int ia[10] = {0, 0, 1, 0, 1, 2, 0, 1, 2, 3}, ib[10] = {0, 1, 1, 2, 2, 2, 3, 3, 3, 3};
        const double* a = w.dv + ia[c] * 18 + r * 3;
        const double* b = w.dv + ib[c] * 18 + r * 3;
        const double d = a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
        w.L[e] = (ia[c] == ib[c]) ? d : 2.0 * d;
    }
    Red::sync();
    for (int var = lane; var < 3; var += LANES) {
        // find_betas_approx_{1,2,3}: columns [B11 B12 B13 B14] / [B11 B12 B22] / [B11 B12 B22 B13 B23]
        const int ncol = var == 0 ? 4 : (var == 1 ? 3 : 5);
        const int colmap[3][5] = {{0, 1, 3, 6, 0}, {0, 1, 2, 0, 0}, {0, 1, 2, 3, 4}};
        double A[30], b[6], x[5] = {0, 0, 0, 0, 0}, be[4];
        for (int i = 0; i < 6; ++i) {
            for (int k = 0; k < 5; ++k) A[i * 5 + k] = k < ncol ? w.L[i * 10 + colmap[var][k]] : 0.0;
            b[i] = w.rho[i];
        }
        householder_ls6(A, b, x, ncol);
        if (var == 0) {
            if (x[0] < 0) { be[0] = sqrt(-x[0]); be[1] = -x[1] / be[0]; be[2] = -x[2] / be[0]; be[3] = -x[3] / be[0]; }
            else          { be[0] = sqrt(x[0]);  be[1] = x[1] / be[0];  be[2] = x[2] / be[0];  be[3] = x[3] / be[0]; }
        } else {
            if (x[0] < 0) { be[0] = sqrt(-x[0]); be[1] = (x[2] < 0) ? sqrt(-x[2]) : 0.0; }
            else          { be[0] = sqrt(x[0]);  be[1] = (x[2] > 0) ? sqrt(x[2]) : 0.0; }
            if (x[1] < 0) be[0] = -be[0];
            be[2] = var == 2 ? x[3] / be[0] : 0.0;
            be[3] = 0.0;
        }
        epnp_gauss_newton(w.L, w.rho, be);
        w.errs[var] = epnp_R_and_t(w, K, be, w.Rs[var], w.ts[var]);
    }
    Red::sync();
    int best = 0;
    if (w.errs[1] < w.errs[0]) best = 1;
    if (w.errs[2] < w.errs[best]) best = 2;
    bool finite = true;
    for (int i = 0; i < 9; ++i) {
        out.R[i] = w.Rs[best][i];
        finite = finite && isfinite(out.R[i]);
    }
    for (int i = 0; i < 3; ++i) {
        out.t[i] = w.ts[best][i];
        finite = finite && isfinite(out.t[i]);
    }
    return finite;
}

}  // namespace svo
