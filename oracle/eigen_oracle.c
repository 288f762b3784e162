/*
 * eigen_oracle.c -- CPU restatement of the reference's reversible eigen-decomposition.  TEST INFRASTRUCTURE ONLY
 * (same rules as phylokernel_oracle.c: only tests/, smoke() and bench.py's CPU legs may use it).
 *
 * Follows ModelMarkov::decomposeRateMatrixRev (model/modelmarkov.cpp:1602-1632) ->
 * EigenDecomposition::eigensystem_sym (utils/eigendecomposition.cpp:165-317):
 *   1. normalise the state frequencies to sum 1                                        (:188-196)
 *   2. computeRateMatrix: a[i][j] = pi_j r_ij, diagonal = -row sum, scaled so that the
 *      expected rate sum_i pi_i sum_{j!=i} a[i][j] equals total_num_subst               (:478-519)
 *   3. eliminateZero: drop states with pi <= ZERO_FREQ (1e-10, eigendecomposition.h:24) (:521-555)
 *   4. symmetrizeRateMatrix: s[i][j] = a[j][i] sqrt(pi_j)/sqrt(pi_i) for j < i          (:557-578)
 *   5. tred2 (Householder tridiagonalisation) + tqli (implicit QL), the published EISPACK /
 *      Numerical Recipes algorithms, restated here from their textbook description       (:218-220)
 *   6. evec[i][j] = v[i][j]/sqrt(pi_i), inv_evec[i][j] = v[j][i] sqrt(pi_j); dropped states get
 *      identity rows/columns and eigenvalue 0                                           (:227-251)
 * Pinned against the reference's own eigensystem_sym through oracle/_ref (tests/test_eigen.py).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define ORC_ZERO_FREQ 1e-10

static double pythag(double a, double b) {
    double absa = fabs(a), absb = fabs(b);
    if (absa > absb) return absa * sqrt(1.0 + (absb / absa) * (absb / absa));
    return absb == 0.0 ? 0.0 : absb * sqrt(1.0 + (absa / absb) * (absa / absb));
}

/* Householder reduction of the symmetric matrix a[n][n] (row-major) to tridiagonal form; on return a holds the
 * orthogonal transformation, d the diagonal, e the sub-diagonal (e[0] = 0). */
static void householder_tridiagonal(double *a, int n, double *d, double *e) {
#define A(i, j) a[(size_t)(i) * n + (j)]
    for (int i = n - 1; i > 0; i--) {
        int l = i - 1;
        double h = 0.0, scale = 0.0;
        if (l > 0) {
            for (int k = 0; k <= l; k++) scale += fabs(A(i, k));
            if (scale == 0.0) {
                e[i] = A(i, l);
            } else {
                for (int k = 0; k <= l; k++) {
                    A(i, k) /= scale;
                    h += A(i, k) * A(i, k);
                }
                double f = A(i, l);
                double g = f >= 0.0 ? -sqrt(h) : sqrt(h);
                e[i] = scale * g;
                h -= f * g;
                A(i, l) = f - g;
                f = 0.0;
                for (int j = 0; j <= l; j++) {
                    A(j, i) = A(i, j) / h;
                    g = 0.0;
                    for (int k = 0; k <= j; k++) g += A(j, k) * A(i, k);
                    for (int k = j + 1; k <= l; k++) g += A(k, j) * A(i, k);
                    e[j] = g / h;
                    f += e[j] * A(i, j);
                }
                double hh = f / (h + h);
                for (int j = 0; j <= l; j++) {
                    f = A(i, j);
                    e[j] = g = e[j] - hh * f;
                    for (int k = 0; k <= j; k++) A(j, k) -= f * e[k] + g * A(i, k);
                }
            }
        } else {
            e[i] = A(i, l);
        }
        d[i] = h;
    }
    d[0] = 0.0;
    e[0] = 0.0;
    for (int i = 0; i < n; i++) {
        int l = i - 1;
        if (d[i] != 0.0) {
            for (int j = 0; j <= l; j++) {
                double g = 0.0;
                for (int k = 0; k <= l; k++) g += A(i, k) * A(k, j);
                for (int k = 0; k <= l; k++) A(k, j) -= g * A(k, i);
            }
        }
        d[i] = A(i, i);
        A(i, i) = 1.0;
        for (int j = 0; j <= l; j++) A(j, i) = A(i, j) = 0.0;
    }
#undef A
}

/* implicit-shift QL on the tridiagonal (d, e); z (the Householder transformation) becomes the eigenvectors in
 * columns.  Returns 0, or -1 when an eigenvalue does not converge in 100 iterations. */
static int ql_implicit(double *d, double *e, int n, double *z) {
#define Z(i, j) z[(size_t)(i) * n + (j)]
    for (int i = 1; i < n; i++) e[i - 1] = e[i];
    e[n - 1] = 0.0;
    for (int l = 0; l < n; l++) {
        int iter = 0, m;
        do {
            for (m = l; m < n - 1; m++) {
                double dd = fabs(d[m]) + fabs(d[m + 1]);
                if (fabs(e[m]) + dd == dd) break;
            }
            if (m != l) {
                if (iter++ == 100) return -1;
                double g = (d[l + 1] - d[l]) / (2.0 * e[l]);
                double r = pythag(g, 1.0);
                g = d[m] - d[l] + e[l] / (g + (g >= 0.0 ? fabs(r) : -fabs(r)));
                double s = 1.0, c = 1.0, p = 0.0;
                int i;
                for (i = m - 1; i >= l; i--) {
                    double f = s * e[i], b = c * e[i];
                    e[i + 1] = (r = pythag(f, g));
                    if (r == 0.0) {
                        d[i + 1] -= p;
                        e[m] = 0.0;
                        break;
                    }
                    s = f / r;
                    c = g / r;
                    g = d[i + 1] - p;
                    r = (d[i] - g) * s + 2.0 * c * b;
                    d[i + 1] = g + (p = s * r);
                    g = c * r - b;
                    for (int k = 0; k < n; k++) {
                        f = Z(k, i + 1);
                        Z(k, i + 1) = s * Z(k, i) + c * f;
                        Z(k, i) = c * Z(k, i) - s * f;
                    }
                }
                if (r == 0.0 && i >= l) continue;
                d[l] -= p;
                e[l] = g;
                e[m] = 0.0;
            }
        } while (m != l);
    }
#undef Z
    return 0;
}

/* exch: upper triangle, row-major, n(n-1)/2 entries (ModelMarkov::rates with half_matrix); freqs[n].
 * eval[n], evec[n][n] (row = state), inv_evec[n][n] (row = eigen index). */
int orc_eigen_rev(int n, const double *exch, const double *freqs, double total_num_subst, double *eval, double *evec,
                  double *inv_evec) {
    double *pi = malloc(sizeof(double) * n), *a = calloc((size_t)n * n, sizeof(double));
    double *m = malloc(sizeof(double) * n);
    double sum = 0.0;
    for (int i = 0; i < n; i++) sum += freqs[i];
    for (int i = 0; i < n; i++) pi[i] = freqs[i] * (1.0 / sum);
    /* decomposeRateMatrixRev :1609-1616: a rate touching a (near-)zero frequency state is dropped */
    for (int i = 0, k = 0; i < n; i++)
        for (int j = i + 1; j < n; j++, k++) {
            double r = (freqs[i] <= ORC_ZERO_FREQ || freqs[j] <= ORC_ZERO_FREQ) ? 0.0 : exch[k];
            a[(size_t)i * n + j] = r;
            a[(size_t)j * n + i] = r;
        }
    /* computeRateMatrix */
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) a[(size_t)i * n + j] = pi[j] * a[(size_t)i * n + j];
    sum = 0.0;
    for (int i = 0; i < n; i++) {
        double t = 0.0;
        for (int j = 0; j < n; j++)
            if (j != i) t += a[(size_t)i * n + j];
        m[i] = t;
        sum += t * pi[i];
    }
    double delta = total_num_subst / sum;
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) a[(size_t)i * n + j] = (i != j) ? delta * a[(size_t)i * n + j] : delta * (-m[i]);
    /* eliminateZero */
    int *map = malloc(sizeof(int) * n), nn = 0;
    for (int i = 0; i < n; i++)
        if (pi[i] > ORC_ZERO_FREQ) map[nn++] = i;
    double *s = malloc(sizeof(double) * (size_t)nn * nn), *rt = malloc(sizeof(double) * nn);
    double *d = malloc(sizeof(double) * nn), *e = malloc(sizeof(double) * nn);
    for (int i = 0; i < nn; i++)
        for (int j = 0; j < nn; j++) s[(size_t)i * nn + j] = a[(size_t)map[i] * n + map[j]];
    /* symmetrizeRateMatrix */
    for (int i = 0; i < nn; i++) rt[i] = sqrt(pi[map[i]]);
    for (int i = 0; i < nn; i++) {
        double tmp = 1.0 / rt[i];
        for (int j = 0; j < i; j++) {
            s[(size_t)j * nn + i] *= rt[j] * tmp;
            s[(size_t)i * nn + j] = s[(size_t)j * nn + i];
        }
    }
    householder_tridiagonal(s, nn, d, e);
    int rc = ql_implicit(d, e, nn, s);
    /* back to all states */
    for (int i = 0; i < n; i++) {
        eval[i] = 0.0;
        for (int j = 0; j < n; j++) evec[(size_t)i * n + j] = inv_evec[(size_t)i * n + j] = (i == j) ? 1.0 : 0.0;
    }
    for (int i = 0; i < nn; i++) {
        eval[map[i]] = d[i];
        for (int j = 0; j < nn; j++) {
            evec[(size_t)map[i] * n + map[j]] = s[(size_t)i * nn + j] / rt[i];
            inv_evec[(size_t)map[i] * n + map[j]] = s[(size_t)j * nn + i] * rt[j];
        }
    }
    free(pi); free(a); free(m); free(map); free(s); free(rt); free(d); free(e);
    return rc;
}
