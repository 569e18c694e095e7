/* nsoracle.c -- CPU oracle of the nsb200 hot path in plain C.  TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library (through oracle/native.py).  The product never links or calls it.
 *
 * It restates what the reference asks dolfin 2019.1.0 to do for one assemble() call
 * (SURVEY.md App. A.4): loop over the cells in mesh order, gather geometry and coefficient
 * dofs, evaluate the element tensor by numerical quadrature the way an FFC-generated
 * tabulate_tensor does, and add it into a CSR matrix by searching each row for the column
 * (PETSc MatSetValues, ADD_VALUES).  Forms (index notation of SURVEY.md App. B.2):
 *   momentum residual / Jacobian   source/ns_ipcs_solver.py:106-147 (F :126-130, J :136),
 *                                  source/ns_bdf_solver.py:68-88, source/ns_solver_base.py:913-936
 *                                  with the term builders source/ns_solver_base.py:370-399,
 *                                  :478-499 (Picard), :662-673 (viscous, reduced form)
 *   pressure Poisson a / L         source/ns_ipcs_solver.py:160-164
 *   velocity correction a / L      source/ns_ipcs_solver.py:183-187
 *
 * Deliberately generic: basis tabulations and the quadrature rule are INPUTS (computed by
 * oracle/fem.py), nothing is shared with the closed-form tables of the CUDA kernels.
 * Parity unpinned against FEniCS itself (see oracle/__init__.py).
 *
 * Threading: OpenMP over cells with atomic adds (the reference itself is one thread; the
 * threads only make the timed CPU baseline as fast as this box allows).  With one thread the
 * summation order is dolfin's cell order.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define MAXD 3
#define MAXNL2 10
#define MAXNW (MAXD * MAXNL2 + MAXD + 1)

typedef struct {
    int32_t d;              /* 2 | 3 */
    int32_t nc;             /* cells */
    int32_t nl2;            /* scalar P2 dofs per cell: 6 | 10 */
    int32_t nq;             /* quadrature points */
    const double* coords;   /* [nv][d] */
    const int32_t* cells;   /* [nc][d+1] ascending (UFC ordered mesh) */
    const double* qw;       /* [nq] reference weights (sum = 1/d!) */
    const double* phi2;     /* [nq][nl2] */
    const double* dphi2;    /* [nq][nl2][d] reference gradients */
    const double* phi1;     /* [nq][d+1] */
    const double* dphi1;    /* [d+1][d] constant reference gradients */
} nso_mesh;

typedef struct {
    double inv_k, a0, a1, a2, c_conv, c_p, c_visc;
} nso_coef;

int nso_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void nso_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* K[j][i] = d xi_j / d x_i, returns |det J| */
static double geometry(const nso_mesh* m, int c, double K[MAXD][MAXD]) {
    const int d = m->d;
    const int32_t* v = m->cells + (size_t)c * (d + 1);
    double J[MAXD][MAXD];
    for (int i = 0; i < d; ++i)
        for (int j = 0; j < d; ++j) J[i][j] = m->coords[(size_t)v[j + 1] * d + i] - m->coords[(size_t)v[0] * d + i];
    double det;
    if (d == 2) {
        det = J[0][0] * J[1][1] - J[0][1] * J[1][0];
        K[0][0] = J[1][1] / det;  K[0][1] = -J[0][1] / det;
        K[1][0] = -J[1][0] / det; K[1][1] = J[0][0] / det;
    } else {
        const double c00 = J[1][1] * J[2][2] - J[1][2] * J[2][1];
        const double c01 = J[1][2] * J[2][0] - J[1][0] * J[2][2];
        const double c02 = J[1][0] * J[2][1] - J[1][1] * J[2][0];
        det = J[0][0] * c00 + J[0][1] * c01 + J[0][2] * c02;
        K[0][0] = c00 / det;
        K[0][1] = (J[0][2] * J[2][1] - J[0][1] * J[2][2]) / det;
        K[0][2] = (J[0][1] * J[1][2] - J[0][2] * J[1][1]) / det;
        K[1][0] = c01 / det;
        K[1][1] = (J[0][0] * J[2][2] - J[0][2] * J[2][0]) / det;
        K[1][2] = (J[0][2] * J[1][0] - J[0][0] * J[1][2]) / det;
        K[2][0] = c02 / det;
        K[2][1] = (J[0][1] * J[2][0] - J[0][0] * J[2][1]) / det;
        K[2][2] = (J[0][0] * J[1][1] - J[0][1] * J[1][0]) / det;
    }
    return fabs(det);
}

/* PETSc-style insertion: find column in the sorted row, add */
static inline int csr_add(const int64_t* rowptr, const int32_t* col, double* vals, int64_t row, int32_t c, double v) {
    int64_t lo = rowptr[row], hi = rowptr[row + 1] - 1;
    while (lo <= hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (col[mid] == c) {
#pragma omp atomic
            vals[mid] += v;
            return 0;
        }
        if (col[mid] < c) lo = mid + 1; else hi = mid - 1;
    }
    return -1;
}

/* ------------------------------------------------------------------------------------------
 * Momentum (+ continuity) residual and Jacobian, IPCS tentative velocity and monolithic forms.
 *   vdofs [nc][d*nl2]  global index of the velocity local dof b*nl2+I (cell-local order
 *                      [u_0 | u_1 | (u_2)]) in u, u1, u2, b and the matrix
 *   pdofs [nc][d+1]    global index of the pressure local dofs in p (and in b / the matrix
 *                      when NSO_MIXED: then u and p are the same mixed vector)
 *   flags              NSO_VEC | NSO_MAT | NSO_PICARD | NSO_BDF | NSO_MIXED
 * b and vals must be zeroed by the caller.  Returns 0, or -1 if a column is not in the pattern.
 * ------------------------------------------------------------------------------------------ */
#define NSO_VEC 1
#define NSO_MAT 2
#define NSO_PICARD 4
#define NSO_BDF 8
#define NSO_MIXED 32

int nso_momentum(const nso_mesh* m, const int32_t* vdofs, const int32_t* pdofs, int flags, const nso_coef* cf,
                 const double* u, const double* u1, const double* u2, const double* p, double* b,
                 const int64_t* rowptr, const int32_t* col, double* vals) {
    const int d = m->d, nl = m->nl2, nv = d * nl, np1 = d + 1, nq = m->nq;
    const int do_vec = flags & NSO_VEC, do_mat = flags & NSO_MAT, newton = !(flags & NSO_PICARD);
    const int bdf = flags & NSO_BDF, mixed = flags & NSO_MIXED;
    const int nw = mixed ? nv + np1 : nv;
    const double mass_coef = bdf ? cf->a0 * cf->inv_k : 0.0;
    int err = 0;
#pragma omp parallel for schedule(static) reduction(| : err)
    for (int c = 0; c < m->nc; ++c) {
        double K[MAXD][MAXD];
        const double detJ = geometry(m, c, K);
        const int32_t* vd = vdofs + (size_t)c * nv;
        const int32_t* pd = pdofs + (size_t)c * np1;
        double U[MAXNL2][MAXD], Uacc[MAXNL2][MAXD], P[MAXD + 1];
        for (int I = 0; I < nl; ++I)
            for (int a = 0; a < d; ++a) {
                const int32_t g = vd[a * nl + I];
                U[I][a] = u[g];
                Uacc[I][a] = bdf ? cf->a0 * u[g] + cf->a1 * u1[g] + cf->a2 * u2[g] : 0.0;
            }
        for (int M = 0; M < np1; ++M) P[M] = p ? p[pd[M]] : 0.0;
        double Fe[MAXNW];
        double Ae[MAXNW * MAXNW];
        if (do_mat) memset(Ae, 0, sizeof(double) * (size_t)nw * nw);
        for (int i = 0; i < nw; ++i) Fe[i] = 0.0;
        for (int q = 0; q < nq; ++q) {
            const double w = m->qw[q] * detJ;
            const double* ph = m->phi2 + (size_t)q * nl;
            const double* dph = m->dphi2 + (size_t)q * nl * d;
            const double* ps = m->phi1 + (size_t)q * np1;
            double g2[MAXNL2][MAXD];
            for (int I = 0; I < nl; ++I)
                for (int i = 0; i < d; ++i) {
                    double s = 0.0;
                    for (int j = 0; j < d; ++j) s += dph[I * d + j] * K[j][i];
                    g2[I][i] = s;
                }
            double uq[MAXD], aq[MAXD], gu[MAXD][MAXD], pq = 0.0;
            for (int a = 0; a < d; ++a) {
                uq[a] = 0.0; aq[a] = 0.0;
                for (int j = 0; j < d; ++j) gu[a][j] = 0.0;
            }
            for (int I = 0; I < nl; ++I)
                for (int a = 0; a < d; ++a) {
                    uq[a] += ph[I] * U[I][a];
                    aq[a] += ph[I] * Uacc[I][a];
                    for (int j = 0; j < d; ++j) gu[a][j] += g2[I][j] * U[I][a];
                }
            for (int M = 0; M < np1; ++M) pq += ps[M] * P[M];
            if (do_vec) {
                double divu = 0.0;
                for (int a = 0; a < d; ++a) divu += gu[a][a];
                for (int bb = 0; bb < d; ++bb) {
                    double conv = 0.0;
                    for (int j = 0; j < d; ++j) conv += gu[bb][j] * uq[j];
                    for (int I = 0; I < nl; ++I) {
                        double visc = 0.0;
                        for (int j = 0; j < d; ++j) visc += gu[bb][j] * g2[I][j];
                        Fe[bb * nl + I] += w * ((cf->inv_k * aq[bb] + cf->c_conv * conv) * ph[I]
                                                - cf->c_p * pq * g2[I][bb] + cf->c_visc * visc);
                    }
                }
                if (mixed)
                    for (int M = 0; M < np1; ++M) Fe[nv + M] += w * (-cf->c_p) * divu * ps[M];
            }
            if (do_mat) {
                for (int I = 0; I < nl; ++I)
                    for (int J = 0; J < nl; ++J) {
                        double adv = 0.0, vis = 0.0;
                        for (int j = 0; j < d; ++j) {
                            adv += g2[J][j] * uq[j];
                            vis += g2[I][j] * g2[J][j];
                        }
                        const double S = w * (mass_coef * ph[I] * ph[J] + cf->c_conv * adv * ph[I] + cf->c_visc * vis);
                        for (int bb = 0; bb < d; ++bb) {
                            Ae[(size_t)(bb * nl + I) * nw + bb * nl + J] += S;
                            if (newton)
                                for (int a = 0; a < d; ++a)
                                    Ae[(size_t)(bb * nl + I) * nw + a * nl + J] +=
                                        w * cf->c_conv * ph[J] * gu[bb][a] * ph[I];
                        }
                    }
                if (mixed)
                    for (int I = 0; I < nl; ++I)
                        for (int bb = 0; bb < d; ++bb)
                            for (int M = 0; M < np1; ++M) {
                                const double v = w * (-cf->c_p) * g2[I][bb] * ps[M];
                                Ae[(size_t)(bb * nl + I) * nw + nv + M] += v;
                                Ae[(size_t)(nv + M) * nw + bb * nl + I] += v;
                            }
            }
        }
        if (do_vec)
            for (int i = 0; i < nw; ++i) {
                const int32_t g = i < nv ? vd[i] : pd[i - nv];
#pragma omp atomic
                b[g] += Fe[i];
            }
        if (do_mat) {
            for (int i = 0; i < nw; ++i) {
                const int32_t gi = i < nv ? vd[i] : pd[i - nv];
                for (int j = 0; j < nw; ++j) {
                    const int32_t gj = j < nv ? vd[j] : pd[j - nv];
                    err |= csr_add(rowptr, col, vals, gi, gj, Ae[(size_t)i * nw + j]);
                }
            }
        }
    }
    return err ? -1 : 0;
}

/* ------------------------------------------------------------------------------------------
 * Scalar forms on P1 / P2 (IPCS steps 2 and 3, projections).
 *   which = 0: P1 stiffness  int grad psi_M . grad psi_N          source/ns_ipcs_solver.py:160
 *           1: P1 mass       int psi_M psi_N                      dlfn.project, ns_solver_base.py:1168
 *           2: vector-P2 mass delta_ab int phi_I phi_J on V       source/ns_ipcs_solver.py:183
 * dofs [nc][nloc] are the row = column dofs (pdofs for 0/1, vdofs for 2).
 * ------------------------------------------------------------------------------------------ */
int nso_scalar_matrix(const nso_mesh* m, int which, const int32_t* dofs, const int64_t* rowptr,
                      const int32_t* col, double* vals) {
    const int d = m->d, nl = m->nl2, np1 = d + 1, nq = m->nq;
    int err = 0;
#pragma omp parallel for schedule(static) reduction(| : err)
    for (int c = 0; c < m->nc; ++c) {
        double K[MAXD][MAXD];
        const double detJ = geometry(m, c, K);
        if (which == 2) {
            double Me[MAXNL2][MAXNL2];
            memset(Me, 0, sizeof Me);
            for (int q = 0; q < nq; ++q) {
                const double w = m->qw[q] * detJ;
                const double* ph = m->phi2 + (size_t)q * nl;
                for (int I = 0; I < nl; ++I)
                    for (int J = 0; J < nl; ++J) Me[I][J] += w * ph[I] * ph[J];
            }
            const int32_t* vd = dofs + (size_t)c * d * nl;
            /* dolfin inserts the whole d*nl x d*nl element tensor, zeros included */
            for (int bb = 0; bb < d; ++bb)
                for (int I = 0; I < nl; ++I)
                    for (int a = 0; a < d; ++a)
                        for (int J = 0; J < nl; ++J)
                            err |= csr_add(rowptr, col, vals, vd[bb * nl + I], vd[a * nl + J],
                                           a == bb ? Me[I][J] : 0.0);
        } else {
            double g1[MAXD + 1][MAXD], Ae[MAXD + 1][MAXD + 1];
            for (int M = 0; M < np1; ++M)
                for (int i = 0; i < d; ++i) {
                    double s = 0.0;
                    for (int j = 0; j < d; ++j) s += m->dphi1[M * d + j] * K[j][i];
                    g1[M][i] = s;
                }
            memset(Ae, 0, sizeof Ae);
            for (int q = 0; q < nq; ++q) {
                const double w = m->qw[q] * detJ;
                const double* ps = m->phi1 + (size_t)q * np1;
                for (int M = 0; M < np1; ++M)
                    for (int N = 0; N < np1; ++N) {
                        if (which == 0) {
                            double s = 0.0;
                            for (int i = 0; i < d; ++i) s += g1[M][i] * g1[N][i];
                            Ae[M][N] += w * s;
                        } else {
                            Ae[M][N] += w * ps[M] * ps[N];
                        }
                    }
            }
            const int32_t* pd = dofs + (size_t)c * np1;
            for (int M = 0; M < np1; ++M)
                for (int N = 0; N < np1; ++N) err |= csr_add(rowptr, col, vals, pd[M], pd[N], Ae[M][N]);
        }
    }
    return err ? -1 : 0;
}

/* Poisson rhs  int grad(p^n).grad(q) - (a0/k) div(u*) q      source/ns_ipcs_solver.py:161-164 */
int nso_poisson_rhs(const nso_mesh* m, const int32_t* vdofs, const int32_t* pdofs, const double* pn,
                    const double* ustar, double a0_over_k, double* b) {
    const int d = m->d, nl = m->nl2, np1 = d + 1, nq = m->nq;
#pragma omp parallel for schedule(static)
    for (int c = 0; c < m->nc; ++c) {
        double K[MAXD][MAXD];
        const double detJ = geometry(m, c, K);
        const int32_t* vd = vdofs + (size_t)c * d * nl;
        const int32_t* pd = pdofs + (size_t)c * np1;
        double g1[MAXD + 1][MAXD], gp[MAXD], be[MAXD + 1];
        for (int M = 0; M < np1; ++M) {
            be[M] = 0.0;
            for (int i = 0; i < d; ++i) {
                double s = 0.0;
                for (int j = 0; j < d; ++j) s += m->dphi1[M * d + j] * K[j][i];
                g1[M][i] = s;
            }
        }
        for (int i = 0; i < d; ++i) {
            gp[i] = 0.0;
            for (int M = 0; M < np1; ++M) gp[i] += g1[M][i] * pn[pd[M]];
        }
        for (int q = 0; q < nq; ++q) {
            const double w = m->qw[q] * detJ;
            const double* dph = m->dphi2 + (size_t)q * nl * d;
            const double* ps = m->phi1 + (size_t)q * np1;
            double divu = 0.0;
            for (int I = 0; I < nl; ++I)
                for (int a = 0; a < d; ++a) {
                    double g = 0.0;
                    for (int j = 0; j < d; ++j) g += dph[I * d + j] * K[j][a];
                    divu += g * ustar[vd[a * nl + I]];
                }
            for (int M = 0; M < np1; ++M) {
                double s = 0.0;
                for (int i = 0; i < d; ++i) s += gp[i] * g1[M][i];
                be[M] += w * (s - a0_over_k * divu * ps[M]);
            }
        }
        for (int M = 0; M < np1; ++M) {
#pragma omp atomic
            b[pd[M]] += be[M];
        }
    }
    return 0;
}

/* correction rhs  int u*.w - (k/a0) grad(p - p_old).w      source/ns_ipcs_solver.py:184-187 */
int nso_correction_rhs(const nso_mesh* m, const int32_t* vdofs, const int32_t* pdofs, const double* ustar,
                       const double* pnew, const double* pold, double k_over_a0, double* b) {
    const int d = m->d, nl = m->nl2, np1 = d + 1, nq = m->nq;
#pragma omp parallel for schedule(static)
    for (int c = 0; c < m->nc; ++c) {
        double K[MAXD][MAXD];
        const double detJ = geometry(m, c, K);
        const int32_t* vd = vdofs + (size_t)c * d * nl;
        const int32_t* pd = pdofs + (size_t)c * np1;
        double gp[MAXD], be[MAXD * MAXNL2];
        for (int i = 0; i < d; ++i) {
            gp[i] = 0.0;
            for (int M = 0; M < np1; ++M) {
                double s = 0.0;
                for (int j = 0; j < d; ++j) s += m->dphi1[M * d + j] * K[j][i];
                gp[i] += s * (pnew[pd[M]] - pold[pd[M]]);
            }
        }
        for (int i = 0; i < d * nl; ++i) be[i] = 0.0;
        for (int q = 0; q < nq; ++q) {
            const double w = m->qw[q] * detJ;
            const double* ph = m->phi2 + (size_t)q * nl;
            for (int bb = 0; bb < d; ++bb) {
                double uq = 0.0;
                for (int I = 0; I < nl; ++I) uq += ph[I] * ustar[vd[bb * nl + I]];
                const double f = uq - k_over_a0 * gp[bb];
                for (int I = 0; I < nl; ++I) be[bb * nl + I] += w * f * ph[I];
            }
        }
        for (int i = 0; i < d * nl; ++i) {
#pragma omp atomic
            b[vd[i]] += be[i];
        }
    }
    return 0;
}

/* y = A x for the CSR oracle matrices (used by the algorithm-matched CPU baseline) */
void nso_spmv(int64_t n, const int64_t* rowptr, const int32_t* col, const double* vals, const double* x, double* y) {
#pragma omp parallel for schedule(static, 1024)
    for (int64_t i = 0; i < n; ++i) {
        double s = 0.0;
        for (int64_t k = rowptr[i]; k < rowptr[i + 1]; ++k) s += vals[k] * x[col[k]];
        y[i] = s;
    }
}

/* ---------------------------------------------------------------------------------------------
 * Krylov solvers in C/OpenMP: the algorithm-matched CPU arm of bench.py (SURVEY.md 8d variant (b)) --
 * the same Jacobi-preconditioned CG / right-Jacobi restarted GMRES with CGS-2 as oracle/krylov.py
 * (which stays the readable statement, and the two are compared in tests/test_oracle.py), without the
 * numpy temporaries, so that the CPU baseline is bound by the host's memory bandwidth and not by the
 * interpreter.  Stopping rule of include/nsb200.h: ||r|| <= max(rtol ||b||, atol).
 * ------------------------------------------------------------------------------------------- */
static double par_dot(int64_t n, const double* x, const double* y) {
    double s = 0.0;
#pragma omp parallel for reduction(+ : s) schedule(static)
    for (int64_t i = 0; i < n; ++i) s += x[i] * y[i];
    return s;
}

static double par_sum(int64_t n, const double* x) {
    double s = 0.0;
#pragma omp parallel for reduction(+ : s) schedule(static)
    for (int64_t i = 0; i < n; ++i) s += x[i];
    return s;
}

/* Jacobi-PCG.  nullspace != 0: constant null space, rhs and solution projected to zero mean.
 * Returns the iteration count (negative: not converged); *resid = final ||r||. */
int nso_pcg(int64_t n, const int64_t* rowptr, const int32_t* col, const double* vals, const double* b_in, double* x,
            double rtol, double atol, int maxit, int nullspace, double* resid) {
    double* w = (double*)malloc(sizeof(double) * 6 * (size_t)n);
    if (!w) return -1000000;
    double *b = w, *r = w + n, *z = w + 2 * n, *p = w + 3 * n, *q = w + 4 * n, *dinv = w + 5 * n;
    const double shift = nullspace ? par_sum(n, b_in) / (double)n : 0.0;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        double d = 1.0;
        for (int64_t k = rowptr[i]; k < rowptr[i + 1]; ++k)
            if (col[k] == i) d = vals[k];
        dinv[i] = 1.0 / d;
        b[i] = b_in[i] - shift;
    }
    nso_spmv(n, rowptr, col, vals, x, q);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        r[i] = b[i] - q[i];
        z[i] = dinv[i] * r[i];
        p[i] = z[i];
    }
    double rz = par_dot(n, r, z), rr = par_dot(n, r, r);
    const double bn = sqrt(par_dot(n, b, b));
    const double tol = fmax(rtol * bn, atol);
    int it = 0;
    while (sqrt(rr) > tol && it < maxit) {
        nso_spmv(n, rowptr, col, vals, p, q);
        const double alpha = rz / par_dot(n, p, q);
        double rz_new = 0.0, rr_new = 0.0;
#pragma omp parallel for reduction(+ : rz_new, rr_new) schedule(static)
        for (int64_t i = 0; i < n; ++i) {
            x[i] += alpha * p[i];
            const double ri = r[i] - alpha * q[i];
            r[i] = ri;
            const double zi = dinv[i] * ri;
            z[i] = zi;
            rz_new += ri * zi;
            rr_new += ri * ri;
        }
        const double beta = rz_new / rz;
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n; ++i) p[i] = z[i] + beta * p[i];
        rz = rz_new;
        rr = rr_new;
        ++it;
    }
    if (nullspace) {
        const double m = par_sum(n, x) / (double)n;
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n; ++i) x[i] -= m;
    }
    if (resid) *resid = sqrt(rr);
    free(w);
    return sqrt(rr) <= tol ? it : -it - 1;
}

/* right-Jacobi-preconditioned GMRES(restart), CGS-2 Arnoldi.  Returns iterations (negative: not converged). */
int nso_gmres(int64_t n, const int64_t* rowptr, const int32_t* col, const double* vals, const double* b, double* x,
              double rtol, double atol, int maxit, int restart, double* resid) {
    const int m = restart;
    double* V = (double*)malloc(sizeof(double) * (size_t)(m + 4) * (size_t)n);
    double* H = (double*)calloc((size_t)(m + 1) * m, sizeof(double));
    double* sm = (double*)malloc(sizeof(double) * (size_t)(5 * m + 8));
    if (!V || !H || !sm) { free(V); free(H); free(sm); return -1000000; }
    double *w = V + (size_t)(m + 1) * n, *t = w + n, *dinv = t + n;
    double *cs = sm, *sn = sm + m, *g = sm + 2 * m, *y = sm + 3 * m + 1, *h = sm + 4 * m + 1;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        double d = 1.0;
        for (int64_t k = rowptr[i]; k < rowptr[i + 1]; ++k)
            if (col[k] == i) d = vals[k];
        dinv[i] = 1.0 / d;
    }
    const double tol = fmax(rtol * sqrt(par_dot(n, b, b)), atol);
    int it = 0;
    double rn = 0.0;
    for (;;) {
        nso_spmv(n, rowptr, col, vals, x, t);
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n; ++i) w[i] = b[i] - t[i];
        rn = sqrt(par_dot(n, w, w));
        if (rn <= tol || it >= maxit) break;
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n; ++i) V[i] = w[i] / rn;
        memset(g, 0, sizeof(double) * (m + 1));
        g[0] = rn;
        int j = 0;
        while (j < m && it < maxit) {
            const double* vj = V + (size_t)j * n;
#pragma omp parallel for schedule(static)
            for (int64_t i = 0; i < n; ++i) t[i] = dinv[i] * vj[i];
            nso_spmv(n, rowptr, col, vals, t, w);
            double* Hj = H + (size_t)j * (m + 1);
            for (int k = 0; k <= j; ++k) Hj[k] = 0.0;
            for (int pass = 0; pass < 2; ++pass) {
                for (int k = 0; k <= j; ++k) h[k] = par_dot(n, V + (size_t)k * n, w);
#pragma omp parallel for schedule(static)
                for (int64_t i = 0; i < n; ++i) {
                    double wi = w[i];
                    for (int k = 0; k <= j; ++k) wi -= h[k] * V[(size_t)k * n + i];
                    w[i] = wi;
                }
                for (int k = 0; k <= j; ++k) Hj[k] += h[k];
            }
            const double wn = sqrt(par_dot(n, w, w));
            Hj[j + 1] = wn;
            double* vn = V + (size_t)(j + 1) * n;
#pragma omp parallel for schedule(static)
            for (int64_t i = 0; i < n; ++i) vn[i] = w[i] / wn;
            for (int k = 0; k < j; ++k) {
                const double a = cs[k] * Hj[k] + sn[k] * Hj[k + 1];
                Hj[k + 1] = -sn[k] * Hj[k] + cs[k] * Hj[k + 1];
                Hj[k] = a;
            }
            const double dd = hypot(Hj[j], Hj[j + 1]);
            cs[j] = Hj[j] / dd; sn[j] = Hj[j + 1] / dd;
            Hj[j] = dd; Hj[j + 1] = 0.0;
            g[j + 1] = -sn[j] * g[j];
            g[j] = cs[j] * g[j];
            ++it; ++j;
            if (fabs(g[j]) <= tol) break;
        }
        for (int i = j - 1; i >= 0; --i) {
            double s = g[i];
            for (int k = i + 1; k < j; ++k) s -= H[(size_t)k * (m + 1) + i] * y[k];
            y[i] = s / H[(size_t)i * (m + 1) + i];
        }
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
            double s = 0.0;
            for (int k = 0; k < j; ++k) s += y[k] * V[(size_t)k * n + i];
            x[i] += dinv[i] * s;
        }
    }
    if (resid) *resid = rn;
    free(V); free(H); free(sm);
    return rn <= tol ? it : -it - 1;
}
