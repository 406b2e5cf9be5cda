// svo_epnp.cuh -- EPnP pose from a minimal 5-point sample, host+device.
//
// cv::solvePnPRansac (called by the reference at src/vo.cpp:301-304) estimates every RANSAC
// hypothesis with SOLVEPNP_EPNP on 5 points whatever `flags` says (flags = SOLVEPNP_ITERATIVE is
// only used for the final refit on the inliers).  Five points leave M^T M (12 x 12) with a
// two-dimensional exact null space, and EPnP's result depends on WHICH basis of it the SVD returns
// and on the sign / order conventions of the PCA that places the control points.  OpenCV runs all
// of its small SVDs (n < 25) through its own one-sided Jacobi (modules/core/src/lapack.cpp,
// JacobiSVDImpl_), a deterministic sequential algorithm -- so this file restates that algorithm and
// OpenCV's calib3d epnp.cpp OPERATION FOR OPERATION, in the same order and in plain IEEE double
// arithmetic (the library is built with --fmad=false; the host test build has no FMA contraction
// either).  The result is the same pose cv::solvePnP(SOLVEPNP_EPNP) returns, bit for bit in
// tests/test_host_cores.py, including the samples whose null-space basis is decided by rounding.
//
// Lane cooperation (device: LANES = 32, one warp per hypothesis; host tests: LANES = 1): every lane
// computes the sequential scalar chains redundantly (identical values in every lane), element-wise
// updates are split over the lanes.  The arithmetic does not depend on LANES.
#pragma once
#include "svo_core.cuh"

// phase timestamps for tools/epnp_bench.cu (compiled out everywhere else)
#ifndef SVO_EPNP_TICK
#define SVO_EPNP_TICK(k)
#endif

namespace svo {

// ---------------------------------------------------------------------------
// OpenCV core restatements

// cv::hypot<double> of lapack.cpp (NOT libm's hypot: different rounding)
SVO_HD double cv_hypot(double a, double b) {
    a = fabs(a);
    b = fabs(b);
    if (a > b) {
        b /= a;
        return a * sqrt(1 + b * b);
    }
    if (b > 0) {
        a /= b;
        return b * sqrt(1 + a * a);
    }
    return 0;
}

struct SerialRedD {
    static SVO_HD void sync() {}
    static SVO_HD void sync_all() {}
    static SVO_HD bool any(bool v) { return v; }
    static SVO_HD double sumd(double v) { return v; }
};

// JacobiSVDImpl_<double> as cv::SVD::compute runs it for an m x n matrix (m = M >= n = N).
//   At   N rows of length M: row i = column i of the input (OpenCV works on the transpose).  On return row i
//        is the i-th LEFT singular vector (column i of U), singular values descending.
//   Vt   N x N, row stride ldv, or nullptr (the V rotations never influence At / W): rows = right singular vectors
//   W    N singular values (during the sweeps: the cached squared row norms, exactly OpenCV's W[])
//   perm N ints of scratch
// Every lane executes the sequential scalar work redundantly on identical data (a warp is then ONE sequential
// machine without divergence; what bounds the time is the dependent FP64 chain, not the issue rate), keeps the two
// active rows in registers and stores identical values, so a lane only ever reads what it has written itself;
// the one hazard left -- a lane rewriting row j while another has not loaded it yet -- is closed by one warp
// barrier per pair.  The V rotations and the final row operations are split over the lanes by column.
// The skip test |p| <= eps sqrt(a b) is decided from p^2 against eps^2 a b and falls back to the literal
// expression only within 1e-6 relative of the threshold, so the decisions are OpenCV's, without a square root
// on the dependent path of every pair.
template <int M, int N, int LANES, class Red>
SVO_HD void cv_jacobi_svd(double* At, double* Vt, int ldv, double* W, int* perm, int lane) {
    constexpr int n = N;
    constexpr int UNROLL = N <= 3 ? N : 1;   // tiny systems: fully unrolled, everything in registers
    (void)UNROLL;
    const double eps = DBL_EPSILON * 10, minval = DBL_MIN, eps2 = eps * eps;
    if (Vt)
        for (int e = lane; e < n * n; e += LANES) Vt[(e / n) * ldv + (e % n)] = (e / n == e % n) ? 1.0 : 0.0;
#pragma unroll UNROLL
    for (int i = 0; i < n; ++i) {
        double sd = 0;
#pragma unroll
        for (int k = 0; k < M; ++k) {
            const double t = At[i * M + k];
            sd += t * t;
        }
        W[i] = sd;
    }
    Red::sync();
    const int max_iter = M > 30 ? M : 30;
    for (int iter = 0; iter < max_iter; ++iter) {
        bool changed = false;
#pragma unroll UNROLL
        for (int i = 0; i < n - 1; ++i) {
            double ai[M];
#pragma unroll
            for (int k = 0; k < M; ++k) ai[k] = At[i * M + k];
            double wi = W[i];
            bool dirty = false;
#pragma unroll UNROLL
            for (int j = i + 1; j < n; ++j) {
                double aj[M];
#pragma unroll
                for (int k = 0; k < M; ++k) aj[k] = At[j * M + k];
                const double wj = W[j];
                Red::sync();   // every lane holds row j before any lane rewrites it
                double p = 0;
#pragma unroll
                for (int k = 0; k < M; ++k) p += ai[k] * aj[k];
                const double lhs = p * p, rhs = eps2 * (wi * wj);
                bool skip;
                if (lhs > rhs * 1.000001 && rhs > 1e-280) skip = false;
                else if (lhs < rhs * 0.999999 && lhs > 1e-280) skip = true;
                else skip = fabs(p) <= eps * sqrt(wi * wj);
                if (skip) continue;
                p *= 2;
                const double beta = wi - wj, gamma = cv_hypot(p, beta);
                double c, s;
                if (beta < 0) {
                    const double delta = (gamma - beta) * 0.5;
                    s = sqrt(delta / gamma);
                    c = p / (gamma * s * 2);
                } else {
                    c = sqrt((gamma + beta) / (gamma * 2));
                    s = p / (gamma * c * 2);
                }
                double a = 0, b = 0;
#pragma unroll
                for (int k = 0; k < M; ++k) {
                    const double t0 = c * ai[k] + s * aj[k];
                    const double t1 = -s * ai[k] + c * aj[k];
                    ai[k] = t0;
                    At[j * M + k] = t1;
                    a += t0 * t0;
                    b += t1 * t1;
                }
                wi = a;
                W[j] = b;
                if (Vt) {
                    double* Vi = Vt + i * ldv;
                    double* Vj = Vt + j * ldv;
                    for (int k = lane; k < n; k += LANES) {
                        const double x = Vi[k], y = Vj[k];
                        Vi[k] = c * x + s * y;
                        Vj[k] = -s * x + c * y;
                    }
                }
                dirty = changed = true;
            }
            if (dirty) {
#pragma unroll
                for (int k = 0; k < M; ++k) At[i * M + k] = ai[k];
                W[i] = wi;
            }
        }
        if (!changed) break;
    }
    Red::sync();
    // singular values (the cached squared norms ARE the sequential sums OpenCV recomputes here), selection sort
    // (descending) by lane 0, row swaps split over the lanes
    if (lane == 0) {
#pragma unroll UNROLL
        for (int i = 0; i < n; ++i) W[i] = sqrt(W[i]);
#pragma unroll UNROLL
        for (int i = 0; i < n - 1; ++i) {
            int j = i;
#pragma unroll UNROLL
            for (int k = i + 1; k < n; ++k)
                if (W[j] < W[k]) j = k;
            perm[i] = j;
            if (i != j) {
                const double tw = W[i];
                W[i] = W[j];
                W[j] = tw;
            }
        }
    }
    Red::sync();
#pragma unroll UNROLL
    for (int i = 0; i < n - 1; ++i) {
        const int j = perm[i];
        if (i != j) {
            for (int k = lane; k < M; k += LANES) {
                const double t = At[i * M + k];
                At[i * M + k] = At[j * M + k];
                At[j * M + k] = t;
            }
            if (Vt)
                for (int k = lane; k < n; k += LANES) {
                    const double t = Vt[i * ldv + k];
                    Vt[i * ldv + k] = Vt[j * ldv + k];
                    Vt[j * ldv + k] = t;
                }
        }
    }
    Red::sync();
    // left singular vectors: rows scaled by 1 / w; a (numerically exact) zero singular value gets OpenCV's
    // pseudo-random completion (RNG 0x12345678, two Gram-Schmidt passes) -- lane 0, it practically never runs
    uint64_t rng = 0x12345678ull;
#pragma unroll UNROLL
    for (int i = 0; i < n; ++i) {
        double sd = W[i];
        if (sd <= minval) {
            Red::sync();   // rows j < i are complete
            if (lane == 0) {
                for (int ii = 0; ii < 100 && sd <= minval; ++ii) {
                    const double val0 = 1. / M;
                    for (int k = 0; k < M; ++k) At[i * M + k] = (cv_rng_next(rng) & 256) != 0 ? val0 : -val0;
                    for (int it = 0; it < 2; ++it)
                        for (int j = 0; j < i; ++j) {
                            sd = 0;
                            for (int k = 0; k < M; ++k) sd += At[i * M + k] * At[j * M + k];
                            double asum = 0;
                            for (int k = 0; k < M; ++k) {
                                const double t = At[i * M + k] - sd * At[j * M + k];
                                At[i * M + k] = t;
                                asum += fabs(t);
                            }
                            asum = asum > eps * 100 ? 1 / asum : 0;
                            for (int k = 0; k < M; ++k) At[i * M + k] *= asum;
                        }
                    sd = 0;
                    for (int k = 0; k < M; ++k) {
                        const double t = At[i * M + k];
                        sd += t * t;
                    }
                    sd = sqrt(sd);
                }
                const double s = sd > minval ? 1 / sd : 0.;
                for (int k = 0; k < M; ++k) At[i * M + k] *= s;
            }   // (the generator state only matters in lane 0)
            Red::sync();
            continue;
        }
        const double s = 1 / sd;
        for (int k = lane; k < M; k += LANES) At[i * M + k] *= s;
    }
    Red::sync();
}

// ---- the same algorithm for the larger systems (12 x 12, 6 x {3, 4, 5}), as a wavefront --------------------------
// OpenCV visits the pairs (0,1), (0,2), ..., (n-2,n-1) one after the other, but pair (i, j) only reads and writes rows
// i and j: it depends on (i, j-1) [row i] and (i-1, j) [row j] and on nothing else.  Executing every pair as soon as
// its two predecessors are done -- step i + j - 1 of its sweep, sweeps n steps apart -- gives each pair exactly the
// operands of the sequential order, hence bit-identical results, in ~n steps per sweep instead of n (n - 1) / 2.
// One lane runs one pair of the current step (<= 8 pairs are ever concurrent), serially and branch-free through the
// c / s computation so that the lanes stay converged.  A sweep is only known to be the last one (no rotation) when
// it ends; the pairs of the following sweep that have run by then all skipped (their rows had not changed since
// they were skipped the sweep before), so running them is harmless.
// Row stride LD (13 for the 12 x 12: the concurrent lanes then read distinct shared-memory banks).
template <int M, int LD>
SVO_HD bool cv_jacobi_pair(double* At, double* Vt, int ldv, double* W, int i, int j, int n) {
    const double eps = DBL_EPSILON * 10, eps2 = eps * eps;
    double ai[M], aj[M];
#pragma unroll
    for (int k = 0; k < M; ++k) {
        ai[k] = At[i * LD + k];
        aj[k] = At[j * LD + k];
    }
    const double wi = W[i], wj = W[j];
    double p = 0;
#pragma unroll
    for (int k = 0; k < M; ++k) p += ai[k] * aj[k];
    // |p| <= eps * sqrt(wi * wj), decided from the squares unless it is a close call
    const double lhs = p * p, rhs = eps2 * (wi * wj);
    bool skip;
    if (lhs > rhs * 1.000001 && rhs > 1e-280) skip = false;
    else if (lhs < rhs * 0.999999 && lhs > 1e-280) skip = true;
    else skip = fabs(p) <= eps * sqrt(wi * wj);
    if (skip) return false;
    p *= 2;
    const double beta = wi - wj;
    // gamma = cv_hypot(p, beta): |p| != 0 here, so the larger magnitude is positive
    const double pa = fabs(p), pb = fabs(beta);
    const double big = pa > pb ? pa : pb, small = pa > pb ? pb : pa;
    const double q = small / big;
    const double gamma = big * sqrt(1 + q * q);
    const bool neg = beta < 0;
    const double num = neg ? (gamma - beta) * 0.5 : (gamma + beta), den = neg ? gamma : gamma * 2;
    const double r = sqrt(num / den);
    const double o = p / (gamma * r * 2);
    const double c = neg ? o : r, s = neg ? r : o;
    double a = 0, b = 0;
#pragma unroll
    for (int k = 0; k < M; ++k) {
        const double t0 = c * ai[k] + s * aj[k];
        const double t1 = -s * ai[k] + c * aj[k];
        At[i * LD + k] = t0;
        At[j * LD + k] = t1;
        a += t0 * t0;
        b += t1 * t1;
    }
    W[i] = a;
    W[j] = b;
    if (Vt) {
        double* Vi = Vt + i * ldv;
        double* Vj = Vt + j * ldv;
        for (int k = 0; k < n; ++k) {
            const double x = Vi[k], y = Vj[k];
            Vi[k] = c * x + s * y;
            Vj[k] = -s * x + c * y;
        }
    }
    return true;
}

template <int M, int N, int LD, int LANES, class Red>
SVO_HD void cv_jacobi_svd_wave(double* At, double* Vt, int ldv, double* W, int* perm, int lane) {
    constexpr int n = N, LAST = 2 * N - 4;   // step of the last pair (n-2, n-1) within its sweep; sweeps start n steps apart
    const double eps = DBL_EPSILON * 10, minval = DBL_MIN;
    const int max_iter = M > 30 ? M : 30;
    if (Vt)
        for (int e = lane; e < n * n; e += LANES) Vt[(e / n) * ldv + (e % n)] = (e / n == e % n) ? 1.0 : 0.0;
    for (int i = lane; i < n; i += LANES) {
        double sd = 0;
#pragma unroll
        for (int k = 0; k < M; ++k) {
            const double t = At[i * LD + k];
            sd += t * t;
        }
        W[i] = sd;
    }
    Red::sync();
    int cur = 0;
    bool chg_cur = false, chg_next = false;
    for (int t = 0;; ++t) {
        const int r0 = t - n * cur, r1 = r0 - n;
        const int lo0 = r0 + 2 - n > 0 ? r0 + 2 - n : 0, c0 = (r0 >> 1) - lo0 + 1;
        const bool next_ok = r1 >= 0 && cur + 1 < max_iter;
        const int lo1 = r1 + 2 - n > 0 ? r1 + 2 - n : 0, c1 = next_ok ? (r1 >> 1) - lo1 + 1 : 0;
        bool rot_cur = false, rot_next = false;
        for (int q = lane; q < c0 + c1; q += LANES) {
            const bool in_cur = q < c0;
            const int i = in_cur ? lo0 + q : lo1 + (q - c0);
            const int j = (in_cur ? r0 : r1) + 1 - i;
            const bool rot = cv_jacobi_pair<M, LD>(At, Vt, ldv, W, i, j, n);
            rot_cur = rot_cur || (rot && in_cur);
            rot_next = rot_next || (rot && !in_cur);
        }
        Red::sync();
        chg_cur = Red::any(rot_cur) || chg_cur;
        chg_next = Red::any(rot_next) || chg_next;
        if (r0 == LAST) {
            if (!chg_cur || cur + 1 >= max_iter) break;
            ++cur;
            chg_cur = chg_next;
            chg_next = false;
        }
    }
    // singular values (the cached squared norms ARE the sequential sums OpenCV recomputes here), selection sort
    // (descending) by lane 0, row swaps split over the lanes
    if (lane == 0) {
        for (int i = 0; i < n; ++i) W[i] = sqrt(W[i]);
        for (int i = 0; i < n - 1; ++i) {
            int j = i;
            for (int k = i + 1; k < n; ++k)
                if (W[j] < W[k]) j = k;
            perm[i] = j;
            if (i != j) {
                const double tw = W[i];
                W[i] = W[j];
                W[j] = tw;
            }
        }
    }
    Red::sync();
    for (int i = 0; i < n - 1; ++i) {
        const int j = perm[i];
        if (i != j) {
            for (int k = lane; k < M; k += LANES) {
                const double t = At[i * LD + k];
                At[i * LD + k] = At[j * LD + k];
                At[j * LD + k] = t;
            }
            if (Vt)
                for (int k = lane; k < n; k += LANES) {
                    const double t = Vt[i * ldv + k];
                    Vt[i * ldv + k] = Vt[j * ldv + k];
                    Vt[j * ldv + k] = t;
                }
        }
    }
    Red::sync();
    // left singular vectors: rows scaled by 1 / w; a (numerically exact) zero singular value gets OpenCV's
    // pseudo-random completion (RNG 0x12345678, two Gram-Schmidt passes) -- lane 0, it practically never runs
    uint64_t rng = 0x12345678ull;
    for (int i = 0; i < n; ++i) {
        double sd = W[i];
        if (sd <= minval) {
            Red::sync();   // rows j < i are complete
            if (lane == 0) {
                for (int ii = 0; ii < 100 && sd <= minval; ++ii) {
                    const double val0 = 1. / M;
                    for (int k = 0; k < M; ++k) At[i * LD + k] = (cv_rng_next(rng) & 256) != 0 ? val0 : -val0;
                    for (int it = 0; it < 2; ++it)
                        for (int j = 0; j < i; ++j) {
                            sd = 0;
                            for (int k = 0; k < M; ++k) sd += At[i * LD + k] * At[j * LD + k];
                            double asum = 0;
                            for (int k = 0; k < M; ++k) {
                                const double t = At[i * LD + k] - sd * At[j * LD + k];
                                At[i * LD + k] = t;
                                asum += fabs(t);
                            }
                            asum = asum > eps * 100 ? 1 / asum : 0;
                            for (int k = 0; k < M; ++k) At[i * LD + k] *= asum;
                        }
                    sd = 0;
                    for (int k = 0; k < M; ++k) {
                        const double t = At[i * LD + k];
                        sd += t * t;
                    }
                    sd = sqrt(sd);
                }
                const double s = sd > minval ? 1 / sd : 0.;
                for (int k = 0; k < M; ++k) At[i * LD + k] *= s;
            }   // (the generator state only matters in lane 0)
            Red::sync();
            continue;
        }
        const double s = 1 / sd;
        for (int k = lane; k < M; k += LANES) At[i * LD + k] *= s;
    }
    Red::sync();
}

// SVBkSbImpl_<double> (cv::SVD::backSubst): x = V diag(1/w) U^T b, singular values <= 2 eps sum(w) dropped.
// U is given as Ut (row i = left singular vector i, length M), Vt N x N.  Serial.
//   vector right-hand side (cv::solve(..., DECOMP_SVD), nb == 1)
template <int M, int N>
SVO_HD void cv_svbksb_vec(const double* w, const double* Ut, const double* Vt, int ldv, const double* b, double* x) {
    constexpr int nm = M < N ? M : N;
    double threshold = 0;
#pragma unroll
    for (int i = 0; i < nm; ++i) threshold += w[i];
    threshold *= DBL_EPSILON * 2;
#pragma unroll
    for (int j = 0; j < N; ++j) x[j] = 0;
#pragma unroll
    for (int i = 0; i < nm; ++i) {
        double wi = w[i];
        if (fabs(wi) <= threshold) continue;
        wi = 1 / wi;
        double s = 0;
#pragma unroll
        for (int j = 0; j < M; ++j) s += Ut[i * M + j] * b[j];
        s *= wi;
#pragma unroll
        for (int j = 0; j < N; ++j) x[j] = x[j] + s * Vt[i * ldv + j];
    }
}
//   no right-hand side (cv::invert(..., DECOMP_SVD)): x (N x M, row-major) = pseudo-inverse
template <int M, int N>
SVO_HD void cv_svbksb_inv(const double* w, const double* Ut, const double* Vt, int ldv, double* x) {
    constexpr int nm = M < N ? M : N;
    double threshold = 0;
#pragma unroll
    for (int i = 0; i < nm; ++i) threshold += w[i];
    threshold *= DBL_EPSILON * 2;
#pragma unroll
    for (int j = 0; j < N * M; ++j) x[j] = 0;
#pragma unroll
    for (int i = 0; i < nm; ++i) {
        double wi = w[i];
        if (fabs(wi) <= threshold) continue;
        wi = 1 / wi;
        double buffer[M];
#pragma unroll
        for (int j = 0; j < M; ++j) buffer[j] = Ut[i * M + j] * wi;
#pragma unroll
        for (int r = 0; r < N; ++r) {
            const double s = Vt[i * ldv + r];
#pragma unroll
            for (int j = 0; j < M; ++j) x[r * M + j] = x[r * M + j] + s * buffer[j];
        }
    }
}

// epnp::qr_solve (Householder; including its column scan that never looks at the last row).  A: 6 x 4
// row-major, destroyed.  Leaves x untouched when a column is exactly zero, like the original.
SVO_HD void epnp_qr_solve(double* A, double* b, double* x) {
    constexpr int nr = 6, nc = 4;
    double A1[nc], A2[nc];
#pragma unroll
    for (int k = 0; k < nc; ++k) {
        double eta = fabs(A[k * nc + k]);
#pragma unroll
        for (int i = k + 1; i < nr; ++i) {
            const double elt = fabs(A[(i - 1) * nc + k]);
            if (eta < elt) eta = elt;
        }
        if (eta == 0) return;
        double sum2 = 0.0;
        const double inv_eta = 1. / eta;
#pragma unroll
        for (int i = k; i < nr; ++i) {
            A[i * nc + k] *= inv_eta;
            sum2 += A[i * nc + k] * A[i * nc + k];
        }
        double sigma = sqrt(sum2);
        if (A[k * nc + k] < 0) sigma = -sigma;
        A[k * nc + k] += sigma;
        A1[k] = sigma * A[k * nc + k];
        A2[k] = -eta * sigma;
#pragma unroll
        for (int j = k + 1; j < nc; ++j) {
            double sum = 0;
#pragma unroll
            for (int i = k; i < nr; ++i) sum += A[i * nc + k] * A[i * nc + j];
            const double tau = sum / A1[k];
#pragma unroll
            for (int i = k; i < nr; ++i) A[i * nc + j] -= tau * A[i * nc + k];
        }
    }
#pragma unroll
    for (int j = 0; j < nc; ++j) {
        double tau = 0;
#pragma unroll
        for (int i = j; i < nr; ++i) tau += A[i * nc + j] * b[i];
        tau /= A1[j];
#pragma unroll
        for (int i = j; i < nr; ++i) b[i] -= tau * A[i * nc + j];
    }
    x[nc - 1] = b[nc - 1] / A2[nc - 1];
#pragma unroll
    for (int i = nc - 2; i >= 0; --i) {
        double sum = 0;
#pragma unroll
        for (int j = i + 1; j < nc; ++j) sum += A[i * nc + j] * x[j];
        x[i] = (b[i] - sum) / A2[i];
    }
}

// ---------------------------------------------------------------------------
#define SVO_EPNP_LDA 13
struct EpnpWork {              // per-hypothesis scratch (shared memory on the device)
    double A[12 * SVO_EPNP_LDA];   // M^T M (row stride 13) -> rows = left singular vectors (ut of epnp::compute_pose)
    double M[120];             // 10 x 12
    double W[12];
    double dv[72];             // [4][6][3]
    double L[60], rho[6];
    double pws[15], us[10], alphas[20], cws[12];
    double vAt[3][30], vVt[3][25], vW[3][5];   // the three beta approximations: SVD of the 6 x {4, 3, 5} systems
    double Rs[3][9], ts[3][3], errs[3];
    int perm[12], vperm[3][5];
    int ok;
};

SVO_HD double epnp_dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

// epnp::gauss_newton: 5 iterations of compute_A_and_b_gauss_newton + qr_solve
SVO_HD void epnp_gauss_newton(const double* L, const double* rho, double* betas) {
    double x[4] = {0, 0, 0, 0};
    for (int it = 0; it < 5; ++it) {
        double A[24], b[6];
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            const double* r = L + i * 10;
            double* a = A + i * 4;
            a[0] = 2 * r[0] * betas[0] + r[1] * betas[1] + r[3] * betas[2] + r[6] * betas[3];
            a[1] = r[1] * betas[0] + 2 * r[2] * betas[1] + r[4] * betas[2] + r[7] * betas[3];
            a[2] = r[3] * betas[0] + r[4] * betas[1] + 2 * r[5] * betas[2] + r[8] * betas[3];
            a[3] = r[6] * betas[0] + r[7] * betas[1] + r[8] * betas[2] + 2 * r[9] * betas[3];
            b[i] = rho[i] - (r[0] * betas[0] * betas[0] + r[1] * betas[0] * betas[1] + r[2] * betas[1] * betas[1] +
                             r[3] * betas[0] * betas[2] + r[4] * betas[1] * betas[2] + r[5] * betas[2] * betas[2] +
                             r[6] * betas[0] * betas[3] + r[7] * betas[1] * betas[3] + r[8] * betas[2] * betas[3] +
                             r[9] * betas[3] * betas[3]);
        }
        epnp_qr_solve(A, b, x);
#pragma unroll
        for (int i = 0; i < 4; ++i) betas[i] += x[i];
    }
}

// compute_ccs / compute_pcs / solve_for_sign / estimate_R_and_t / reprojection_error of epnp.cpp
SVO_HD double epnp_R_and_t(const EpnpWork& w, const Intrinsics& K, const double* betas, double* R9, double* t3) {
    double ccs[4][3], pcs[15];
#pragma unroll
    for (int i = 0; i < 4; ++i) ccs[i][0] = ccs[i][1] = ccs[i][2] = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const double* v = w.A + SVO_EPNP_LDA * (11 - i);
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int k = 0; k < 3; ++k) ccs[j][k] += betas[i] * v[3 * j + k];
    }
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        const double* a = w.alphas + 4 * i;
#pragma unroll
        for (int j = 0; j < 3; ++j)
            pcs[3 * i + j] = a[0] * ccs[0][j] + a[1] * ccs[1][j] + a[2] * ccs[2][j] + a[3] * ccs[3][j];
    }
    if (pcs[2] < 0.0) {
#pragma unroll
        for (int i = 0; i < 15; ++i) pcs[i] = -pcs[i];
    }
    double pc0[3] = {0, 0, 0}, pw0[3] = {0, 0, 0};
#pragma unroll
    for (int i = 0; i < 5; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            pc0[j] += pcs[3 * i + j];
            pw0[j] += w.pws[3 * i + j];
        }
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        pc0[j] /= 5;
        pw0[j] /= 5;
    }
    // cvSVD(ABt, D, U, V): R = U V^T, third row negated when det < 0.  At = ABt^T, accumulated in place.
    double At[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, Vt[9], D[3];
    int perm[3];
#pragma unroll
    for (int i = 0; i < 5; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
#pragma unroll
            for (int k = 0; k < 3; ++k) At[3 * k + j] += (pcs[3 * i + j] - pc0[j]) * (w.pws[3 * i + k] - pw0[k]);
    cv_jacobi_svd<3, 3, 1, SerialRedD>(At, Vt, 3, D, perm, 0);
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)   // U[i][k] = At[k][i], V[j][k] = Vt[k][j]
            R9[i * 3 + j] = At[0 + i] * Vt[0 + j] + At[3 + i] * Vt[3 + j] + At[6 + i] * Vt[6 + j];
    const double det = R9[0] * R9[4] * R9[8] + R9[1] * R9[5] * R9[6] + R9[2] * R9[3] * R9[7] - R9[2] * R9[4] * R9[6] -
                       R9[1] * R9[3] * R9[8] - R9[0] * R9[5] * R9[7];
    if (det < 0) {
        R9[6] = -R9[6];
        R9[7] = -R9[7];
        R9[8] = -R9[8];
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) t3[i] = pc0[i] - epnp_dot3(R9 + 3 * i, pw0);
    double sum2 = 0.0;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        const double* pw = w.pws + 3 * i;
        const double Xc = epnp_dot3(R9, pw) + t3[0];
        const double Yc = epnp_dot3(R9 + 3, pw) + t3[1];
        const double inv_Zc = 1.0 / (epnp_dot3(R9 + 6, pw) + t3[2]);
        const double ue = K.cx + K.fx * Xc * inv_Zc, ve = K.cy + K.fy * Yc * inv_Zc;
        const double u = w.us[2 * i], v = w.us[2 * i + 1];
        sum2 += sqrt((u - ue) * (u - ue) + (v - ve) * (v - ve));
    }
    return sum2 / 5;
}

// find_betas_approx_{1,2,3} + gauss_newton + compute_R_and_t for ONE of the three approximations, executed by the
// LANES lanes of one warp: columns [B11 B12 B13 B14] / [B11 B12 B22] / [B11 B12 B22 B13 B23] of L_6x10,
// cv::solve(..., DECOMP_SVD).  The scalar tail (back substitution, Gauss-Newton, R and t) runs redundantly in every lane.
template <int VAR, int LANES, class Red>
SVO_HD void epnp_variant(EpnpWork& w, const Intrinsics& K, int lane) {
    constexpr int ncol = VAR == 0 ? 4 : (VAR == 1 ? 3 : 5);
    const int colmap[3][5] = {{0, 1, 3, 6, 0}, {0, 1, 2, 0, 0}, {0, 1, 2, 3, 4}};
    double* At = w.vAt[VAR];
    double* Vt = w.vVt[VAR];
    for (int e = lane; e < ncol * 6; e += LANES) At[e] = w.L[(e % 6) * 10 + colmap[VAR][e / 6]];
    Red::sync();
    cv_jacobi_svd_wave<6, ncol, 6, LANES, Red>(At, Vt, 5, w.vW[VAR], w.vperm[VAR], lane);
    double x[5] = {0, 0, 0, 0, 0}, be[4];
    cv_svbksb_vec<6, ncol>(w.vW[VAR], At, Vt, 5, w.rho, x);
    SVO_EPNP_TICK(5)
    if (VAR == 0) {
        if (x[0] < 0) { be[0] = sqrt(-x[0]); be[1] = -x[1] / be[0]; be[2] = -x[2] / be[0]; be[3] = -x[3] / be[0]; }
        else          { be[0] = sqrt(x[0]);  be[1] = x[1] / be[0];  be[2] = x[2] / be[0];  be[3] = x[3] / be[0]; }
    } else {
        if (x[0] < 0) { be[0] = sqrt(-x[0]); be[1] = (x[2] < 0) ? sqrt(-x[2]) : 0.0; }
        else          { be[0] = sqrt(x[0]);  be[1] = (x[2] > 0) ? sqrt(x[2]) : 0.0; }
        if (x[1] < 0) be[0] = -be[0];
        be[2] = VAR == 2 ? x[3] / be[0] : 0.0;
        be[3] = 0.0;
    }
    epnp_gauss_newton(w.L, w.rho, be);
    SVO_EPNP_TICK(6)
    double R9[9], t3[3];
    const double err = epnp_R_and_t(w, K, be, R9, t3);
    if (lane == 0) {
        w.errs[VAR] = err;
        for (int i = 0; i < 9; ++i) w.Rs[VAR][i] = R9[i];
        for (int i = 0; i < 3; ++i) w.ts[VAR][i] = t3[i];
    }
}

// X: 5 x 3 object points (float32), uv: 5 x 2 pixel coordinates (float32).  Executed by GROUPS warps of LANES lanes
// (device: 3 x 32, one block per hypothesis -- the three beta approximations run side by side; host tests: 1 x 1).
// `Red` provides sync() (the lanes of one group) and sync_all() (all groups).  Every lane returns the same pose;
// false when the result is not finite.
template <int LANES, int GROUPS, class Red>
SVO_HD bool epnp5_solve(EpnpWork& w, const Intrinsics& K, const float* X, const float* uv, int lane, int group,
                        Pose& out) {
    constexpr int n = 5;
    const int tid = group * LANES + lane;
    constexpr int NT = LANES * GROUPS;
    SVO_EPNP_TICK(0)
    if (tid == 0) {
        w.ok = 1;
        const double ifx = 1.0 / K.fx, ify = 1.0 / K.fy;
        for (int i = 0; i < n; ++i) {
            for (int j = 0; j < 3; ++j) w.pws[3 * i + j] = (double)X[3 * i + j];
            // cv::undistortPoints (zero distortion) stores normalised coordinates as float32;
            // epnp::init_points maps them back to pixels in double
            const float xn = (float)(((double)uv[2 * i] - K.cx) * ifx);
            const float yn = (float)(((double)uv[2 * i + 1] - K.cy) * ify);
            w.us[2 * i] = (double)xn * K.fx + K.cx;
            w.us[2 * i + 1] = (double)yn * K.fy + K.cy;
        }
        // choose_control_points: centroid + PCA axes (SVD of PW0^T PW0, rows of U^T)
        double cws[12];
        for (int j = 0; j < 3; ++j) cws[j] = 0;
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < 3; ++j) cws[j] += w.pws[3 * i + j];
        for (int j = 0; j < 3; ++j) cws[j] /= n;
        double c[9], dc[3];
        int perm[3];
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
            for (int b = a; b < 3; ++b) {   // cv::mulTransposed: upper triangle, sequential over the rows, mirrored
                double s = 0;
                for (int k = 0; k < n; ++k) s += (w.pws[3 * k + a] - cws[a]) * (w.pws[3 * k + b] - cws[b]);
                c[3 * a + b] = s;
                c[3 * b + a] = s;
            }
        cv_jacobi_svd<3, 3, 1, SerialRedD>(c, nullptr, 0, dc, perm, 0);   // symmetric: the transpose is the matrix itself
#pragma unroll
        for (int i = 1; i < 4; ++i) {
            const double k = sqrt(dc[i - 1] / n);
#pragma unroll
            for (int j = 0; j < 3; ++j) cws[3 * i + j] = cws[j] + k * c[3 * (i - 1) + j];
        }
        // compute_barycentric_coordinates: cvInvert(CC, CC_inv, CV_SVD)
        double cct[9], vt[9], wd[3], ci[9];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 1; j < 4; ++j) cct[(j - 1) * 3 + i] = cws[3 * j + i] - cws[i];   // transposed: row = column of CC
        cv_jacobi_svd<3, 3, 1, SerialRedD>(cct, vt, 3, wd, perm, 0);
        cv_svbksb_inv<3, 3>(wd, cct, vt, 3, ci);
        for (int i = 0; i < n; ++i) {
            const double* pi = w.pws + 3 * i;
            double* a = w.alphas + 4 * i;
#pragma unroll
            for (int j = 0; j < 3; ++j)
                a[1 + j] = ci[3 * j] * (pi[0] - cws[0]) + ci[3 * j + 1] * (pi[1] - cws[1]) + ci[3 * j + 2] * (pi[2] - cws[2]);
            a[0] = 1.0 - a[1] - a[2] - a[3];
        }
        // compute_rho
        const int pa[6] = {0, 0, 0, 1, 1, 2}, pb[6] = {1, 2, 3, 2, 3, 3};
#pragma unroll
        for (int r = 0; r < 6; ++r) {
            const double* p1 = cws + 3 * pa[r];
            const double* p2 = cws + 3 * pb[r];
            w.rho[r] = (p1[0] - p2[0]) * (p1[0] - p2[0]) + (p1[1] - p2[1]) * (p1[1] - p2[1]) +
                       (p1[2] - p2[2]) * (p1[2] - p2[2]);
        }
        for (int i = 0; i < 12; ++i) w.cws[i] = cws[i];
    }
    Red::sync_all();
    SVO_EPNP_TICK(1)
    // fill_M
    for (int e = tid; e < 120; e += NT) {
        const int row = e / 12, col = e - row * 12;
        const int i = row >> 1, k = col / 3, c3 = col - 3 * k;
        const double as = w.alphas[4 * i + k];
        double v;
        if ((row & 1) == 0) v = c3 == 0 ? as * K.fx : (c3 == 1 ? 0.0 : as * (K.cx - w.us[2 * i]));
        else v = c3 == 0 ? 0.0 : (c3 == 1 ? as * K.fy : as * (K.cy - w.us[2 * i + 1]));
        w.M[e] = v;
    }
    Red::sync_all();
    for (int e = tid; e < 144; e += NT) {   // M^T M, sequential over the 10 rows
        const int a = e / 12, b = e - a * 12;
        double s = 0.0;
#pragma unroll
        for (int r = 0; r < 10; ++r) s += w.M[r * 12 + a] * w.M[r * 12 + b];
        w.A[a * SVO_EPNP_LDA + b] = s;
    }
    Red::sync_all();
    SVO_EPNP_TICK(2)
    if (group == 0) cv_jacobi_svd_wave<12, 12, SVO_EPNP_LDA, LANES, Red>(w.A, nullptr, 0, w.W, w.perm, lane);
    Red::sync_all();
    SVO_EPNP_TICK(3)
    // compute_L_6x10: v[i] = ut row 11 - i
    for (int e = tid; e < 72; e += NT) {   // dv[i][j][k]
        const int i = e / 18, j = (e - i * 18) / 3, k = e - i * 18 - j * 3;
        const int pa[6] = {0, 0, 0, 1, 1, 2}, pb[6] = {1, 2, 3, 2, 3, 3};
        const double* v = w.A + SVO_EPNP_LDA * (11 - i);
        w.dv[e] = v[3 * pa[j] + k] - v[3 * pb[j] + k];
    }
    Red::sync_all();
    for (int e = tid; e < 60; e += NT) {   // L_6x10
        const int r = e / 10, c = e - r * 10;
        const int ia[10] = {0, 0, 1, 0, 1, 2, 0, 1, 2, 3}, ib[10] = {0, 1, 1, 2, 2, 2, 3, 3, 3, 3};
        const double d = epnp_dot3(w.dv + ia[c] * 18 + r * 3, w.dv + ib[c] * 18 + r * 3);
        w.L[e] = (ia[c] == ib[c]) ? d : 2.0 * d;
    }
    Red::sync_all();
    SVO_EPNP_TICK(4)
    for (int var = group; var < 3; var += GROUPS) {
        if (var == 0) epnp_variant<0, LANES, Red>(w, K, lane);
        else if (var == 1) epnp_variant<1, LANES, Red>(w, K, lane);
        else epnp_variant<2, LANES, Red>(w, K, lane);
    }
    Red::sync_all();
    SVO_EPNP_TICK(7)
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
