/* CPU oracle for the soft-DTW dynamic programme.  TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the reference's numba-CPU recurrences
 *   forward : soft_dtw_cuda.py:185-206 (compute_softdtw)
 *   backward: soft_dtw_cuda.py:209-240 (compute_softdtw_backward)
 * in float64, which is what the reference's CPU path computes in (numpy
 * float64 arrays).  Built by oracle/Makefile into
 * oracle/libsdtw_oracle.so and called through ctypes from oracle/sdtw_oracle.py.
 * Only tests/, smoke() and bench.py's cpu_baseline leg may load it.
 *
 * Array shapes: D (B,N,M); R and E (B,N+2,M+2), all C-contiguous doubles.
 */
#include <math.h>
#include <stddef.h>

#define IDX(b, i, j) (((size_t)(b) * (size_t)(n + 2) + (size_t)(i)) * (size_t)(m + 2) + (size_t)(j))

/* soft_dtw_cuda.py:190-206.  R is fully written here (inf border, R[0,0]=0). */
void sdtw_oracle_forward(const double *D, int nb, int n, int m, double gamma,
                         double bandwidth, double *R)
{
    for (int b = 0; b < nb; ++b) {
        for (int i = 0; i < n + 2; ++i)
            for (int j = 0; j < m + 2; ++j)
                R[IDX(b, i, j)] = INFINITY;
        R[IDX(b, 0, 0)] = 0.0;
        for (int j = 1; j <= m; ++j) {
            for (int i = 1; i <= n; ++i) {
                if (bandwidth > 0.0 && fabs((double)(i - j)) > bandwidth)
                    continue; /* Sakoe-Chiba pruning, :197-198 */
                double r0 = -R[IDX(b, i - 1, j - 1)] / gamma;
                double r1 = -R[IDX(b, i - 1, j)] / gamma;
                double r2 = -R[IDX(b, i, j - 1)] / gamma;
                double rmax = fmax(fmax(r0, r1), r2);
                double rsum = exp(r0 - rmax) + exp(r1 - rmax) + exp(r2 - rmax);
                double softmin = -gamma * (log(rsum) + rmax);
                R[IDX(b, i, j)] =
                    D[((size_t)b * n + (size_t)(i - 1)) * m + (size_t)(j - 1)] + softmin;
            }
        }
    }
}

/* soft_dtw_cuda.py:215-240.  R is modified in place exactly as the reference
 * does (border -> -inf, corner copy, inf -> -inf); E is fully written. */
void sdtw_oracle_backward(const double *D, double *R, int nb, int n, int m,
                          double gamma, double bandwidth, double *E)
{
    for (int b = 0; b < nb; ++b) {
        for (int i = 0; i < n + 2; ++i)
            for (int j = 0; j < m + 2; ++j)
                E[IDX(b, i, j)] = 0.0;
        E[IDX(b, n + 1, m + 1)] = 1.0;
        for (int i = 0; i < n + 2; ++i)
            R[IDX(b, i, m + 1)] = -INFINITY;
        for (int j = 0; j < m + 2; ++j)
            R[IDX(b, n + 1, j)] = -INFINITY;
        R[IDX(b, n + 1, m + 1)] = R[IDX(b, n, m)];
        for (int j = m; j >= 1; --j) {
            for (int i = n; i >= 1; --i) {
                if (isinf(R[IDX(b, i, j)]))
                    R[IDX(b, i, j)] = -INFINITY;
                if (bandwidth > 0.0 && fabs((double)(i - j)) > bandwidth)
                    continue;
                /* padded D: zero outside [1..n]x[1..m] (:219-221) */
                double d_dn = (i + 1 <= n) ? D[((size_t)b * n + (size_t)i) * m + (size_t)(j - 1)] : 0.0;
                double d_rt = (j + 1 <= m) ? D[((size_t)b * n + (size_t)(i - 1)) * m + (size_t)j] : 0.0;
                double d_dg = (i + 1 <= n && j + 1 <= m)
                                  ? D[((size_t)b * n + (size_t)i) * m + (size_t)j]
                                  : 0.0;
                double r = R[IDX(b, i, j)];
                double a = exp((R[IDX(b, i + 1, j)] - r - d_dn) / gamma);
                double bb = exp((R[IDX(b, i, j + 1)] - r - d_rt) / gamma);
                double c = exp((R[IDX(b, i + 1, j + 1)] - r - d_dg) / gamma);
                E[IDX(b, i, j)] = E[IDX(b, i + 1, j)] * a + E[IDX(b, i, j + 1)] * bb +
                                  E[IDX(b, i + 1, j + 1)] * c;
            }
        }
    }
}
