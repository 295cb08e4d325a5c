/* TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * Kraft's SLSQP, restated from the published algorithm (see slsqp_kraft.h for the sources and the scope).
 * Matrices are column-major with explicit leading dimensions and the routines keep Lawson & Hanson's / Kraft's
 * argument conventions (1-based positions in the comments, 0-based storage), so that each loop can be checked
 * against the publication.  Dot products and norms accumulate in long double.
 */
#include "slsqp_kraft.h"

#include <cmath>
#include <cstring>

namespace kraft {
namespace {

const double EPMACH = 2.22e-16;

double ddot(int n, const double* x, int incx, const double* y, int incy) {
    long double s = 0;
    for (int i = 0; i < n; ++i) s += (long double)x[(long)i * incx] * y[(long)i * incy];
    return (double)s;
}
double dnrm2(int n, const double* x, int incx) {
    double xmax = 0;
    for (int i = 0; i < n; ++i) {
        const double a = std::fabs(x[(long)i * incx]);
        if (xmax < a) xmax = a;
    }
    if (xmax == 0) return 0;
    const double scale = 1.0 / xmax;
    long double s = 0;
    for (int i = 0; i < n; ++i) {
        const double xs = scale * x[(long)i * incx];
        s += (long double)xs * xs;
    }
    return xmax * std::sqrt((double)s);
}
void daxpy(int n, double a, const double* x, int incx, double* y, int incy) {
    if (a == 0) return;
    for (int i = 0; i < n; ++i) y[(long)i * incy] += a * x[(long)i * incx];
}
void dcopy(int n, const double* x, int incx, double* y, int incy) {
    for (int i = 0; i < n; ++i) y[(long)i * incy] = x[(long)i * incx];
}
void dscal(int n, double a, double* x, int incx) {
    for (int i = 0; i < n; ++i) x[(long)i * incx] *= a;
}
/* Givens rotation (BLAS level 1 rotg / rot) */
void rotg(double* da, double* db, double* c, double* s) {
    double roe = *db;
    if (std::fabs(*da) > std::fabs(*db)) roe = *da;
    const double scale = std::fabs(*da) + std::fabs(*db);
    double r, z;
    if (scale == 0) {
        *c = 1; *s = 0; r = 0;
    } else {
        const double p = *da / scale, q = *db / scale;
        r = scale * std::sqrt(p * p + q * q);
        if (roe < 0) r = -r;
        *c = *da / r;
        *s = *db / r;
    }
    z = *s;
    if (std::fabs(*c) > 0 && std::fabs(*c) <= *s) z = 1.0 / *c;
    *da = r;
    *db = z;
}
void rot(int n, double* x, int incx, double* y, int incy, double c, double s) {
    for (int i = 0; i < n; ++i) {
        double& xi = x[(long)i * incx];
        double& yi = y[(long)i * incy];
        const double t = c * xi + s * yi;
        yi = c * yi - s * xi;
        xi = t;
    }
}

}  // namespace

/* Lawson & Hanson, "Solving Least Squares Problems" (1974), algorithm H12: construction (mode 1) and application
 * (mode 2) of the Householder transformation Q = I + u u^T / b.  u is the pivot vector, element j at u[(j-1) iue];
 * c holds ncv vectors, element i of vector k at c[(i-1) ice + (k-1) icv]. */
void h12(int mode, int lpivot, int l1, int m, double* u, int iue, double* up, double* c, int ice, int icv, int ncv) {
    if (0 >= lpivot || lpivot >= l1 || l1 > m) return;
    auto U = [&](int j) -> double& { return u[(long)(j - 1) * iue]; };
    double cl = std::fabs(U(lpivot));
    if (mode != 2) {
        for (int j = l1; j <= m; ++j) {
            const double sm = std::fabs(U(j));
            if (sm > cl) cl = sm;
        }
        if (cl <= 0) return;
        const double clinv = 1.0 / cl;
        double d = U(lpivot) * clinv;
        double sm = d * d;
        for (int j = l1; j <= m; ++j) {
            d = U(j) * clinv;
            sm += d * d;
        }
        cl *= std::sqrt(sm);
        if (U(lpivot) > 0) cl = -cl;
        *up = U(lpivot) - cl;
        U(lpivot) = cl;
    } else if (cl <= 0) {
        return;
    }
    if (ncv <= 0) return;
    double b = *up * U(lpivot);
    if (b >= 0) return;
    b = 1.0 / b;
    long i2 = 1 - icv + (long)ice * (lpivot - 1);
    const long incr = (long)ice * (l1 - lpivot);
    for (int j = 1; j <= ncv; ++j) {
        i2 += icv;
        long i3 = i2 + incr;
        long i4 = i3;
        double sm = c[i2 - 1] * *up;
        for (int i = l1; i <= m; ++i) {
            sm += c[i3 - 1] * U(i);
            i3 += ice;
        }
        if (sm == 0) continue;
        sm *= b;
        c[i2 - 1] += sm * *up;
        for (int i = l1; i <= m; ++i) {
            c[i4 - 1] += sm * U(i);
            i4 += ice;
        }
    }
}

/* Lawson & Hanson's NNLS: min || A x - b ||  s.t.  x >= 0.  A is m x n (leading dimension mda) and is overwritten
 * by Q A, b by Q b; w receives the dual vector, z and index are work space.  mode: 1 ok, 2 bad dimensions,
 * 3 more than 3 n iterations. */
void nnls(double* a, int mda, int m, int n, double* b, double* x, double* rnorm, double* w, double* z, int* index, int* mode) {
    const double factor = 0.01;
    auto A = [&](int i, int j) -> double& { return a[(long)(i - 1) + (long)(j - 1) * mda]; };
    *mode = 2;
    if (m <= 0 || n <= 0) return;
    *mode = 1;
    int iter = 0;
    const int itmax = 3 * n;
    for (int i = 1; i <= n; ++i) index[i - 1] = i;
    int iz1 = 1;
    const int iz2 = n;
    int nsetp = 0, npp1 = 1;
    for (int i = 0; i < n; ++i) x[i] = 0;
    double up = 0;
    for (;;) {
        /* step two: dual variables of the zero set */
        if (iz1 > iz2 || nsetp >= m) break;
        for (int iz = iz1; iz <= iz2; ++iz) {
            const int j = index[iz - 1];
            w[j - 1] = ddot(m - nsetp, &A(npp1, j), 1, &b[npp1 - 1], 1);
        }
        int j = 0, izmax = 0;
        bool found = false;
        for (;;) {
            /* step three: the most positive dual variable */
            double wmax = 0;
            for (int iz = iz1; iz <= iz2; ++iz) {
                const int jj = index[iz - 1];
                if (w[jj - 1] <= wmax) continue;
                wmax = w[jj - 1];
                izmax = iz;
            }
            if (wmax <= 0) break;
            const int iz = izmax;
            j = index[iz - 1];
            /* step four: test column j for linear dependency on the set P */
            const double asave = A(npp1, j);
            h12(1, npp1, npp1 + 1, m, &A(1, j), 1, &up, z, 1, 1, 0);
            const double unorm = dnrm2(nsetp, &A(1, j), 1);
            const double t = factor * std::fabs(A(npp1, j));
            const double d = unorm + t;
            if (d - unorm > 0) {
                dcopy(m, b, 1, z, 1);
                h12(2, npp1, npp1 + 1, m, &A(1, j), 1, &up, z, 1, 1, 1);
                if (z[npp1 - 1] / A(npp1, j) > 0) { found = true; break; }
            }
            A(npp1, j) = asave;
            w[j - 1] = 0;
        }
        if (!found) break;
        /* step five: move column j from the zero set to P */
        dcopy(m, z, 1, b, 1);
        index[izmax - 1] = index[iz1 - 1];
        index[iz1 - 1] = j;
        ++iz1;
        nsetp = npp1;
        ++npp1;
        for (int jz = iz1; jz <= iz2; ++jz) {
            const int jj = index[jz - 1];
            h12(2, nsetp, npp1, m, &A(1, j), 1, &up, &A(1, jj), 1, mda, 1);
        }
        for (int i = npp1; i <= m; ++i) A(i, j) = 0;
        w[j - 1] = 0;
        /* step six: solve the triangular system of the set P; steps seven to eleven: keep x feasible */
        for (;;) {
            int jj = 0;
            for (int ip = nsetp; ip >= 1; --ip) {
                if (ip != nsetp) daxpy(ip, -z[ip], &A(1, jj), 1, z, 1);
                jj = index[ip - 1];
                z[ip - 1] /= A(ip, jj);
            }
            ++iter;
            if (iter > itmax) { *mode = 3; goto solution; }
            double alpha = 1.0;
            jj = 0;
            for (int ip = 1; ip <= nsetp; ++ip) {
                if (z[ip - 1] > 0) continue;
                const int l = index[ip - 1];
                const double t = -x[l - 1] / (z[ip - 1] - x[l - 1]);
                if (alpha < t) continue;
                alpha = t;
                jj = ip;
            }
            for (int ip = 1; ip <= nsetp; ++ip) {
                const int l = index[ip - 1];
                x[l - 1] = (1.0 - alpha) * x[l - 1] + alpha * z[ip - 1];
            }
            if (jj == 0) break;  /* back to step two */
            /* step eleven: move coefficient index(jj) from P to the zero set */
            int i = index[jj - 1];
            for (;;) {
                x[i - 1] = 0;
                ++jj;
                for (int jc = jj; jc <= nsetp; ++jc) {
                    const int ii = index[jc - 1];
                    index[jc - 2] = ii;
                    double c, s;
                    rotg(&A(jc - 1, ii), &A(jc, ii), &c, &s);
                    const double t = A(jc - 1, ii);
                    rot(n, &A(jc - 1, 1), mda, &A(jc, 1), mda, c, s);
                    A(jc - 1, ii) = t;
                    A(jc, ii) = 0;
                    rot(1, &b[jc - 2], 1, &b[jc - 1], 1, c, s);
                }
                npp1 = nsetp;
                --nsetp;
                --iz1;
                index[iz1 - 1] = i;
                if (nsetp <= 0) { *mode = 3; goto solution; }
                bool again = false;
                for (jj = 1; jj <= nsetp; ++jj) {
                    i = index[jj - 1];
                    if (x[i - 1] <= 0) { again = true; break; }
                }
                if (!again) break;
            }
            dcopy(m, b, 1, z, 1);
        }
    }
solution:
    {
        const int k = npp1 < m ? npp1 : m;
        *rnorm = dnrm2(m - nsetp, &b[k - 1], 1);
        if (npp1 > m) for (int i = 0; i < n; ++i) w[i] = 0;
    }
}

/* Least distance programming (Lawson & Hanson ch. 23):  min || x ||  s.t.  G x >= h,  G m x n (leading dimension
 * mg).  On return w[0..m) holds the Lagrange multipliers.  mode: 1 ok, 2 bad dimension, 3 NNLS iterations,
 * 4 constraints incompatible.  w needs (n + 1) (m + 2) + 2 m doubles, index m ints. */
void ldp(const double* g, int mg, int m, int n, const double* h, double* x, double* xnorm, double* w, int* index, int* mode) {
    *mode = 2;
    if (n <= 0) return;
    *mode = 1;
    for (int i = 0; i < n; ++i) x[i] = 0;
    *xnorm = 0;
    if (m == 0) return;
    /* dual problem: E = [G^T; h^T] ((n+1) x m), f = (0, ..., 0, 1) */
    long iw = 0;
    for (int j = 1; j <= m; ++j) {
        for (int i = 1; i <= n; ++i) w[iw++] = g[(long)(j - 1) + (long)(i - 1) * mg];
        w[iw++] = h[j - 1];
    }
    const long jf = iw;  /* f */
    for (int i = 1; i <= n; ++i) w[iw++] = 0;
    w[iw++] = 1.0;
    const int n1 = n + 1;
    const long iz = iw;
    const long iy = iz + n1;
    const long iwdual = iy + m;
    double rnorm = 0;
    nnls(w, n1, n1, m, &w[jf], &w[iy], &rnorm, &w[iwdual], &w[iz], index, mode);
    if (*mode != 1) return;
    *mode = 4;
    if (rnorm <= 0) return;
    double fac = 1.0 - ddot(m, h, 1, &w[iy], 1);
    {
        const double d = 1.0 + fac;
        if (!(d - 1.0 > 0)) return;
    }
    *mode = 1;
    fac = 1.0 / fac;
    for (int j = 1; j <= n; ++j) x[j - 1] = fac * ddot(m, &g[(long)(j - 1) * mg], 1, &w[iy], 1);
    *xnorm = dnrm2(n, x, 1);
    /* Lagrange multipliers of the primal problem */
    for (int i = 0; i < m; ++i) w[i] = 0;
    daxpy(m, fac, &w[iy], 1, w, 1);
}

namespace {

/* min || E x - f ||  s.t.  G x >= h  (Lawson & Hanson ch. 23, "LSI"): QR of E, change of variables to a least
 * distance problem.  E me x n (ld le), G mg x n (ld lg).  mode 5: E singular. */
void lsi(double* e, double* f, double* g, double* h, int le, int me, int lg, int mg, int n, double* x, double* xnorm, double* w, int* jw, int* mode) {
    auto E = [&](int i, int j) -> double& { return e[(long)(i - 1) + (long)(j - 1) * le]; };
    auto G = [&](int i, int j) -> double& { return g[(long)(i - 1) + (long)(j - 1) * lg]; };
    double t = 0;
    for (int i = 1; i <= n; ++i) {
        const int j = i + 1 < n ? i + 1 : n;
        h12(1, i, i + 1, me, &E(1, i), 1, &t, &E(1, j), 1, le, n - i);
        h12(2, i, i + 1, me, &E(1, i), 1, &t, f, 1, 1, 1);
    }
    *mode = 5;
    for (int i = 1; i <= mg; ++i) {
        for (int j = 1; j <= n; ++j) {
            if (!(std::fabs(E(j, j)) >= EPMACH)) return;
            G(i, j) = (G(i, j) - ddot(j - 1, &G(i, 1), lg, &E(1, j), 1)) / E(j, j);
        }
        h[i - 1] -= ddot(n, &G(i, 1), lg, f, 1);
    }
    ldp(g, lg, mg, n, h, x, xnorm, w, jw, mode);
    if (*mode != 1) return;
    daxpy(n, 1.0, f, 1, x, 1);
    for (int i = n; i >= 1; --i) {
        const int j = i + 1 < n ? i + 1 : n;
        x[i - 1] = (x[i - 1] - ddot(n - i, &E(i, j), le, &x[j - 1], 1)) / E(i, i);
    }
    const int j = n + 1 < me ? n + 1 : me;
    t = dnrm2(me - n, &f[j - 1], 1);
    *xnorm = std::sqrt(*xnorm * *xnorm + t * t);
}

/* min || E x - f ||  s.t.  C x = d,  G x >= h  ("LSEI"): the equality constraints are eliminated by Householder
 * transformations of C from the right.  C mc x n (ld lc), E me x n (ld le), G mg x n (ld lg).
 * w[0..mc) receives the multipliers of the equalities, w[mc..mc+mg) those of the inequalities.
 * mode: 1 ok, 2 mc > n, 3/4 from NNLS/LDP, 5 E singular, 6 C singular, 7 unconstrained problem (not provided here). */
void lsei(double* c, double* d, double* e, double* f, double* g, double* h, int lc, int mc, int le, int me, int lg, int mg, int n, double* x,
          double* xnrm, double* w, int* jw, int* mode) {
    auto C = [&](int i, int j) -> double& { return c[(long)(i - 1) + (long)(j - 1) * lc]; };
    auto E = [&](int i, int j) -> double& { return e[(long)(i - 1) + (long)(j - 1) * le]; };
    auto G = [&](int i, int j) -> double& { return g[(long)(i - 1) + (long)(j - 1) * lg]; };
    *mode = 2;
    if (mc > n) return;
    const int l = n - mc;
    const int mc1 = mc + 1;
    /* work space behind the multipliers and LDP's area, 1-based positions as in the publication */
    const long iw = (long)(l + 1) * (mg + 2) + 2 * (long)mg + mc;
    const long ie = iw + mc + 1;
    const long jf = ie + (long)me * l;
    const long ig = jf + me;
    auto W = [&](long k) -> double& { return w[k - 1]; };
    /* triangularise C and apply the factors to E and G */
    for (int i = 1; i <= mc; ++i) {
        const int j = i + 1 < lc ? i + 1 : lc;
        h12(1, i, i + 1, n, &C(i, 1), lc, &W(iw + i), &C(j, 1), lc, 1, mc - i);
        h12(2, i, i + 1, n, &C(i, 1), lc, &W(iw + i), e, le, 1, me);
        h12(2, i, i + 1, n, &C(i, 1), lc, &W(iw + i), g, lg, 1, mg);
    }
    /* solve C x = d and modify f */
    *mode = 6;
    for (int i = 1; i <= mc; ++i) {
        if (std::fabs(C(i, i)) < EPMACH) return;
        x[i - 1] = (d[i - 1] - ddot(i - 1, &C(i, 1), lc, x, 1)) / C(i, i);
    }
    *mode = 1;
    for (int k = 0; k < mg; ++k) W(mc1 + k) = 0;
    if (mc != n) {
        for (int i = 1; i <= me; ++i) W(jf - 1 + i) = f[i - 1] - ddot(mc, &E(i, 1), le, x, 1);
        /* store the transformed E and G */
        for (int i = 1; i <= me; ++i) dcopy(l, &E(i, mc1), le, &W(ie - 1 + i), me);
        for (int i = 1; i <= mg; ++i) dcopy(l, &G(i, mc1), lg, &W(ig - 1 + i), mg);
        if (mg <= 0) {
            /* least squares without inequality constraints (Lawson & Hanson's HFTI): cannot occur on this path --
             * the bounds always contribute 2 n inequality rows */
            *mode = 7;
            return;
        }
        for (int i = 1; i <= mg; ++i) h[i - 1] -= ddot(mc, &G(i, 1), lg, x, 1);
        lsi(&W(ie), &W(jf), &W(ig), h, me, me, mg, mg, l, &x[mc1 - 1], xnrm, &W(mc1), jw, mode);
        if (mc == 0) return;
        const double t = dnrm2(mc, x, 1);
        *xnrm = std::sqrt(*xnrm * *xnrm + t * t);
        if (*mode != 1) return;
    }
    /* solution of the original problem and Lagrange multipliers */
    for (int i = 1; i <= me; ++i) f[i - 1] = ddot(n, &E(i, 1), le, x, 1) - f[i - 1];
    for (int i = 1; i <= mc; ++i) d[i - 1] = ddot(me, &E(1, i), 1, f, 1) - ddot(mg, &G(1, i), 1, &W(mc1), 1);
    for (int i = mc; i >= 1; --i) h12(2, i, i + 1, n, &C(i, 1), lc, &W(iw + i), x, 1, 1, 1);
    for (int i = mc; i >= 1; --i) {
        const int j = i + 1 < lc ? i + 1 : lc;
        W(i) = (d[i - 1] - ddot(mc - i, &C(j, i), 1, &W(j), 1)) / C(i, i);
    }
}

struct LsqWork {
    std::vector<double> w;
    std::vector<int> jw;
};

/* The QP sub-problem  min 1/2 x^T B x + g^T x  s.t.  a_j^T x - b_j = 0 (j <= meq), >= 0 (j > meq), xl <= x <= xu
 * with B = L D L^T as the least squares problem  min || D^{1/2} L^T x + D^{-1/2} L^{-1} g ||  ("LSQ").
 * l: packed column-wise unit lower triangle with D on its diagonal, nl entries; called with n variables or, for the
 * augmented problem, n + 1 (then l[nl-1] is the weight of the extra variable).  a is la x (n+1).  y receives the
 * multipliers of the m general constraints. */
void lsq(int m, int meq, int n, int nl, int la, const double* l, const double* g, const double* a, const double* b, const double* xl,
         const double* xu, double* x, double* y, LsqWork& ws, int* mode) {
    const int n1 = n + 1;
    const int mineq = m - meq;
    const int m1 = mineq + n + n;
    int n2 = n1 * n / 2 + 1;
    n2 = (n2 == nl) ? 0 : 1;
    const int n3 = n - n2;
    const int lc = meq > 1 ? meq : 1;
    /* areas: E n x n, f n, C meq x n, d meq, G m1 x n, h m1, then LSEI's work space */
    const long ie = 1;
    const long jf = ie + (long)n * n;
    const long ic = jf + n;
    const long id = ic + (long)meq * n;
    const long ig = id + meq;
    const long ih = ig + (long)m1 * n;
    const long iw = ih + m1;
    const long lsei_w = 2 * (long)meq + (long)n * n + n + (long)m1 * n + (long)(n - meq + 1) * (m1 + 2) + 2 * (long)m1 + 16;
    const long need = iw + lsei_w + n;
    if ((long)ws.w.size() < need) ws.w.resize(need);
    if ((long)ws.jw.size() < m1 + 1) ws.jw.resize(m1 + 1);
    std::fill(ws.w.begin(), ws.w.begin() + need, 0.0);
    double* w = ws.w.data();
    auto W = [&](long k) -> double& { return w[k - 1]; };
    /* recover E and f from L and g */
    long i2 = 1, i3 = 1, i4 = 1;
    for (int i = 1; i <= n3; ++i) {
        const int i1 = n1 - i;
        const double diag = std::sqrt(l[i2 - 1]);
        for (int k = 0; k < i1; ++k) W(i3 + k) = 0;
        dcopy(i1 - n2, &l[i2 - 1], 1, &W(i3), n);
        dscal(i1 - n2, diag, &W(i3), n);
        W(i3) = diag;
        W(jf - 1 + i) = (g[i - 1] - ddot(i - 1, &W(i4), 1, &W(jf), 1)) / diag;
        i2 += i1 - n2;
        i3 += n1;
        i4 += n;
    }
    if (n2 == 1) {
        W(i3) = l[nl - 1];
        for (int k = 0; k < n3; ++k) W(i4 + k) = 0;
        W(jf - 1 + n) = 0;
    }
    dscal(n, -1.0, &W(jf), 1);
    if (meq > 0) {
        for (int i = 1; i <= meq; ++i) dcopy(n, &a[i - 1], la, &W(ic - 1 + i), meq);
        dcopy(meq, b, 1, &W(id), 1);
        dscal(meq, -1.0, &W(id), 1);
    }
    if (mineq > 0) {
        for (int i = 1; i <= mineq; ++i) dcopy(n, &a[meq + i - 1], la, &W(ig - 1 + i), m1);
        dcopy(mineq, &b[meq], 1, &W(ih), 1);
        dscal(mineq, -1.0, &W(ih), 1);
    }
    /* the bounds as 2 n inequality rows:  x >= xl,  -x >= -xu */
    {
        const long ip = ig + mineq;
        for (int i = 1; i <= n; ++i) W(ip - 1 + i + (long)(i - 1) * m1) = 1.0;
        const long im1 = ip + n;
        for (int i = 1; i <= n; ++i) W(im1 - 1 + i + (long)(i - 1) * m1) = -1.0;
        const long il = ih + mineq;
        const long iu = il + n;
        dcopy(n, xl, 1, &W(il), 1);
        dcopy(n, xu, 1, &W(iu), 1);
        dscal(n, -1.0, &W(iu), 1);
    }
    double xnorm = 0;
    lsei(&W(ic), &W(id), &W(ie), &W(jf), &W(ig), &W(ih), lc, meq, n, n, m1, m1, n, x, &xnorm, &W(iw), ws.jw.data(), mode);
    if (*mode == 1) {
        dcopy(m, &W(iw), 1, y, 1);
        /* keep the step inside the box against roundoff (documented NLopt behaviour: bounds are never violated) */
        for (int i = 0; i < n; ++i) {
            if (x[i] < xl[i]) x[i] = xl[i];
            else if (x[i] > xu[i]) x[i] = xu[i];
        }
    }
}

}  // namespace

/* Rank-one modification of an LDL^T factorisation,  A <- A + sigma z z^T  (Fletcher & Powell 1974, the composite-t
 * method Kraft uses).  a: packed column-wise, D on the diagonal positions; z is destroyed; w: work space for
 * sigma < 0. */
void ldl_update(int n, double* a, double* z, double sigma, double* w) {
    if (sigma == 0) return;
    long ij = 1;
    double t = 1.0 / sigma;
    if (sigma < 0) {
        /* prepare the negative update */
        for (int i = 1; i <= n; ++i) w[i - 1] = z[i - 1];
        for (int i = 1; i <= n; ++i) {
            const double v = w[i - 1];
            t += v * v / a[ij - 1];
            for (int j = i + 1; j <= n; ++j) {
                ++ij;
                w[j - 1] -= v * a[ij - 1];
            }
            ++ij;
        }
        if (t >= 0) t = EPMACH / sigma;
        for (int i = 1; i <= n; ++i) {
            const int j = n + 1 - i;
            ij -= i;
            const double u = w[j - 1];
            w[j - 1] = t;
            t -= u * u / a[ij - 1];
        }
    }
    for (int i = 1; i <= n; ++i) {
        const double v = z[i - 1];
        const double delta = v / a[ij - 1];
        double tp;
        if (sigma < 0) tp = w[i - 1];
        else tp = t + delta * v;
        const double alpha = tp / t;
        a[ij - 1] = alpha * a[ij - 1];
        if (i == n) return;
        const double beta = delta / tp;
        if (alpha > 4.0) {
            const double gamma = t / tp;
            for (int j = i + 1; j <= n; ++j) {
                ++ij;
                const double u = a[ij - 1];
                a[ij - 1] = gamma * u + beta * z[j - 1];
                z[j - 1] -= v * u;
            }
        } else {
            for (int j = i + 1; j <= n; ++j) {
                ++ij;
                z[j - 1] -= v * a[ij - 1];
                a[ij - 1] += beta * z[j - 1];
            }
        }
        ++ij;
        t = tp;
    }
}

/* ---- the SQP iteration with NLopt's driver logic around it ---- */
Result slsqp_minimize(const Problem& p, double* xbest, double* minf, Stats* st_out) {
    Stats st;
    const int n = p.n, meq = p.meq, m = p.meq + p.mineq;
    const int n1 = n + 1, la = m > 1 ? m : 1;
    const int n2 = n1 * n / 2, n3 = n2 + 1;
    const double alfmin = 0.1;
    if (n <= 0 || (int)p.lb.size() != n || (int)p.ub.size() != n) return R_INVALID_ARGS;
    for (int i = 0; i < n; ++i)
        if (p.lb[i] > p.ub[i] || xbest[i] < p.lb[i] || xbest[i] > p.ub[i]) return R_INVALID_ARGS;  /* NLopt rejects such a start */

    std::vector<double> x(xbest, xbest + n), x0(n), xprev(xbest, xbest + n), g(n1, 0.0), a((size_t)la * n1, 0.0), c(la, 0.0), r(m + 2 * n1 + 2, 0.0),
        l(n3 + 1, 0.0), mu(la, 0.0), s(n1, 0.0), u(n1, 0.0), v(n1, 0.0), wl(n1, 0.0), ctmp(n, 0.0);
    LsqWork ws;
    double f = 0, fprev = HUGE_VAL;
    *minf = HUGE_VAL;
    bool feasible = false, feasible_cur = false;
    double infeas = HUGE_VAL, infeas_cur = HUGE_VAL;
    Result ret = R_SUCCESS;

    /* objective (+ gradients) at x, with NLopt's feasibility bookkeeping (equalities |h| <= tol, inequalities
     * g >= -tol in Kraft's sign) */
    auto evaluate = [&](bool want_grad) {
        f = p.f(x.data(), want_grad ? g.data() : nullptr);
        ++st.evals;
        feasible_cur = true;
        infeas_cur = 0;
        if (!std::isfinite(f)) return;
        for (int j = 0; j < m; ++j) {
            c[j] = p.c(j, x.data(), want_grad ? ctmp.data() : nullptr);
            if (want_grad) for (int i = 0; i < n; ++i) a[(size_t)j + (size_t)i * la] = ctmp[i];
            const double viol = j < meq ? std::fabs(c[j]) : -c[j];
            if (viol > infeas_cur) infeas_cur = viol;
            const double tol = (size_t)j < p.ctol.size() ? p.ctol[j] : 0.0;
            feasible_cur = feasible_cur && viol <= tol;
        }
    };
    auto keep_best = [&]() {
        if (std::isfinite(f) && ((f < *minf && (feasible_cur || !feasible)) || (!feasible && infeas_cur < infeas))) {
            *minf = f;
            feasible = feasible_cur;
            infeas = infeas_cur;
            std::memcpy(xbest, x.data(), sizeof(double) * n);
        }
    };
    auto out_of_evals = [&]() { return p.maxeval > 0 && st.evals >= p.maxeval; };
    auto penalty = [&](const std::vector<double>& weight) {  /* sum_j mu_j max(-c_j, c_j for equalities) */
        double h = 0;
        for (int j = 0; j < m; ++j) {
            const double h3 = j < meq ? c[j] : 0.0;
            const double t = -c[j] > h3 ? -c[j] : h3;
            h += weight[j] * t;
        }
        return h;
    };

    evaluate(true);
    keep_best();
    if (out_of_evals()) ret = R_MAXEVAL_REACHED;

    int ireset = 0;
    bool reset = true;
    while (ret == R_SUCCESS) {
        if (reset) {
            /* reset the BFGS matrix to the identity, at most five times */
            ++ireset;
            if (ireset > 5) {
                /* positive directional derivative in the line search.  Kraft's relaxed convergence test is void with
                 * acc = 0; NLopt then reports "roundoff limited", or -- when the current point is feasible -- applies its
                 * own x tolerance against the start of this iteration */
                ret = R_ROUNDOFF_LIMITED;
                if (feasible_cur) {
                    bool same = true, small = !p.xtol_abs.empty();
                    for (int i = 0; i < n; ++i) {
                        if (x[i] != x0[i]) same = false;
                        if (small && std::fabs(x[i] - x0[i]) >= p.xtol_abs[i]) small = false;
                    }
                    if (same || small) ret = R_XTOL_REACHED;
                }
                break;
            }
            ++st.resets;
            std::fill(l.begin(), l.end(), 0.0);
            for (int i = 1, j = 1; i <= n; ++i) {
                l[j - 1] = 1.0;
                j += n1 - i;
            }
            reset = false;
        }
        /* search direction as the solution of the QP sub-problem */
        for (int i = 0; i < n; ++i) { u[i] = p.lb[i] - x[i]; v[i] = p.ub[i] - x[i]; }
        double h4 = 1.0;
        int mode = 1;
        lsq(m, meq, n, n3, la, l.data(), g.data(), a.data(), c.data(), u.data(), v.data(), s.data(), r.data(), ws, &mode);
        if (mode == 6 && n == meq) mode = 4;
        if (mode == 4) {
            /* augmented problem for an inconsistent linearisation: one more variable delta in [0, 1] that relaxes the
             * constraints, weighted by rho = 100, 1000, ... */
            ++st.augmented;
            for (int j = 0; j < m; ++j) a[(size_t)j + (size_t)n * la] = j < meq ? -c[j] : (-c[j] > 0 ? -c[j] : 0.0);
            std::fill(s.begin(), s.end(), 0.0);
            g[n] = 0;
            l[n3 - 1] = 100.0;
            s[n] = 1.0;
            u[n] = 0;
            v[n] = 1.0;
            int incons = 0;
            for (;;) {
                lsq(m, meq, n1, n3, la, l.data(), g.data(), a.data(), c.data(), u.data(), v.data(), s.data(), r.data(), ws, &mode);
                h4 = 1.0 - s[n];
                if (mode == 4) {
                    l[n3 - 1] *= 10.0;
                    if (++incons > 5) break;
                    continue;
                }
                break;
            }
        }
        st.lsq_mode = mode;
        if (mode != 1) {
            /* NLopt: modes 3 (NNLS iterations) and 4 (incompatible constraints) -> FAILURE "bug: more than iter SQP
             * iterations"; 5, 6, 7 (singular sub-problem) -> ROUNDOFF_LIMITED */
            ret = (mode == 3 || mode == 4) ? R_FAILURE : R_ROUNDOFF_LIMITED;
            break;
        }
        /* Lagrangian gradient at the old point, multipliers for the L1 merit function */
        for (int i = 0; i < n; ++i) v[i] = g[i] - ddot(m, &a[(size_t)i * la], 1, r.data(), 1);
        const double f0 = f;
        x0 = x;
        const double gs = ddot(n, g.data(), 1, s.data(), 1);
        for (int j = 0; j < m; ++j) {
            const double h3 = std::fabs(r[j]);
            const double half = (mu[j] + h3) / 2.0;
            mu[j] = h3 > half ? h3 : half;
        }
        const double h1 = penalty(mu);
        const double t0 = f0 + h1;
        double h3 = gs - h1 * h4;
        if (h3 >= 0) { reset = true; continue; }
        /* inexact line search */
        int line = 0;
        double alpha = 1.0;
        bool have_grad = false;
        for (;;) {
            ++line;
            h3 = alpha * h3;
            dscal(n, alpha, s.data(), 1);
            for (int i = 0; i < n; ++i) {
                double xi = x0[i] + s[i];
                if (xi < p.lb[i]) xi = p.lb[i];
                else if (xi > p.ub[i]) xi = p.ub[i];
                x[i] = xi;
            }
            have_grad = line == 1;  /* the first trial is evaluated with gradients, later ones without */
            evaluate(have_grad);
            keep_best();
            if (out_of_evals()) { ret = R_MAXEVAL_REACHED; break; }
            const double t = f + penalty(mu);
            const double dh = t - t0;
            if (dh <= h3 / 10.0 || line > 10) break;
            const double cand = h3 / (2.0 * (h3 - dh));
            alpha = cand > alfmin ? cand : alfmin;
        }
        if (ret != R_SUCCESS) break;
        /* line search complete: gradients at the accepted point, NLopt's convergence tests between iterations */
        ++st.iters;
        if (!have_grad) { evaluate(true); }
        keep_best();
        if (!std::isinf(fprev)) {
            bool same = true, small = !p.xtol_abs.empty();
            for (int i = 0; i < n; ++i) {
                if (x[i] != xprev[i]) same = false;
                if (small && std::fabs(x[i] - xprev[i]) >= p.xtol_abs[i]) small = false;
            }
            if (same || small) ret = R_XTOL_REACHED;
        }
        fprev = f;
        xprev = x;
        if (ret == R_SUCCESS && out_of_evals()) ret = R_MAXEVAL_REACHED;
        if (ret != R_SUCCESS) break;
        /* BFGS update of the LDL^T factors with Powell's damping */
        for (int i = 0; i < n; ++i) u[i] = g[i] - ddot(m, &a[(size_t)i * la], 1, r.data(), 1) - v[i];
        {   /* v = L D L^T s */
            long k = 0;
            for (int i = 1; i <= n; ++i) {
                double h = 0;
                ++k;
                for (int j = i + 1; j <= n; ++j) {
                    ++k;
                    h += l[k - 1] * s[j - 1];
                }
                v[i - 1] = s[i - 1] + h;
            }
            k = 1;
            for (int i = 1; i <= n; ++i) {
                v[i - 1] = l[k - 1] * v[i - 1];
                k += n1 - i;
            }
            for (int i = n; i >= 1; --i) {
                double h = 0;
                k = i;
                for (int j = 1; j <= i - 1; ++j) {
                    h += l[k - 1] * v[j - 1];
                    k += n - j;
                }
                v[i - 1] += h;
            }
        }
        double b1 = ddot(n, s.data(), 1, u.data(), 1);
        const double b2 = ddot(n, s.data(), 1, v.data(), 1);
        const double b3 = 0.2 * b2;
        if (b1 < b3) {
            const double th = (b2 - b3) / (b2 - b1);
            b1 = b3;
            dscal(n, th, u.data(), 1);
            daxpy(n, 1.0 - th, v.data(), 1, u.data(), 1);
        }
        ldl_update(n, l.data(), u.data(), 1.0 / b1, wl.data());
        ldl_update(n, l.data(), v.data(), -1.0 / b2, wl.data());
    }
    if (st_out) *st_out = st;
    return ret;
}

}  // namespace kraft
