/*
 * dg1d_oracle.c -- CPU restatement of the explicit RK-DG hot path of cpraveen/fem dg1d/python.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load it, and only as the checker or as the
 * timed CPU baseline.  The product (fem_b200/) never links, imports or executes anything here.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks every function below against golden
 * vectors produced by executing the unmodified reference (tests/golden/gen_golden.py), per stage
 * (residual, RK update, limiter) and at final time.  The flux family of the C sibling (csib_*,
 * dg1d/c/euler/src/flux.c) is pinned against that file compiled unmodified (oracle/_ref, built by
 * oracle/Makefile) and against vectors and solver runs recorded from it (tests/golden/csib*.npz).
 * The Euler Roe flux (orc_euler_roe_flux: no dg1d/python or dg1d/c counterpart) restates
 * PDE::roe_flux of dg2d/models/euler/pde.h:209-285 for dim=1, normal=+1, and is pinned against that header
 * compiled unmodified with stand-ins for its deal.II types (oracle/ref_dg2d_roe.cc, oracle/dealii_standin).
 * One exception is "parity unpinned" and says so where it is defined: the scalar inflow/outflow boundary
 * (orc_scalar_*_bc: the rule of dg1d/scalar_legendre/dg.h:367-413, whose code needs all of deal.II).
 *
 * Layout is the reference's: one C-contiguous (nc, nd) array per conserved variable
 * (dgsolver.py:102-104, euler.py:151-159).  Euler arrays are passed as u[3][nc][nd].
 *
 * Floating-point contract ("canonical arithmetic").  numpy evaluates array expressions one
 * IEEE-754 operation at a time (no contraction), Python scalars likewise.  The only fused
 * operations in the reference come from BLAS: `a.dot(b)` on length<=5 vectors is
 *      s = a0*b0; s = fma(a_i, b_i, s)            (ddot)
 * and `A @ x` on (k+1)x(k+1) matrices is, per row, for k <= 2,
 *      t = fma(a0, x0, a1*x1); t = fma(a_i, x_i, t), i >= 2   (dgemv, OpenBLAS 0.3.30 Haswell)
 * (measured bit-for-bit in the build container; for k >= 3 OpenBLAS uses a SIMD order that is
 * not reproduced, the same row rule is used and agrees to a few ulp).  Everything else is one
 * rounding per operation, in the reference's source order.  This file is compiled with
 * -ffp-contract=off so the compiler cannot add or remove fusions; fma() appears only where
 * stated above.  `x**2` on numpy scalars goes through libm pow(), which is not always the
 * correctly-rounded square (0.08% of inputs differ by 1 ulp); x*x is used here.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define MAXND 8

enum { PDE_LINEAR = 0, PDE_BURGER = 1, PDE_EULER = 2, PDE_VARADV = 3 };
enum { FLUX_CENTRAL = 0, FLUX_UPWIND = 1, FLUX_ROE = 2, FLUX_GODUNOV = 3, FLUX_RUSANOV = 4,
       FLUX_LF = 5, FLUX_ECUSP = 6, FLUX_HLLC = 7, FLUX_AUSMDV = 8, FLUX_LFC = 9 };

typedef struct {
    int k;                       /* degree; nd = Nq = k+1 */
    double Vf[MAXND * MAXND];    /* Vf[q*nd + j] = phi_j(xg_q)            dgsolver.py:74  euler.py:58 */
    double Vg[MAXND * MAXND];    /* Vg[i*nd + q] = phi_i'(xg_q) * wg_q    dgsolver.py:78  euler.py:59 */
    double bm[MAXND], bp[MAXND]; /* phi_j(-1), phi_j(+1)                  dgsolver.py:87-90 */
    double ark[3], brk[3];       /* dgsolver.py:7-8 (brk = 1.0 - ark, F6) */
    double sqrt3, isqrt3;        /* euler.py:11-12 */
    double gamma;                /* gas.py:3 */
    double xg[MAXND];            /* Gauss nodes, dgsolver.py:71 (varadv needs x) */
    double xmin, dx;             /* mesh, dgsolver.py:37-39 (varadv needs x) */
    double speed_const;          /* varadv.py:15-17: max_speed is a constant the host evaluates */
} orc_ops;

/* ---- canonical small BLAS ------------------------------------------------------------- */
static inline double dotv(int n, const double *a, const double *b)
{
    double s = a[0] * b[0];
    for (int i = 1; i < n; ++i) s = fma(a[i], b[i], s);
    return s;
}

static inline double mvrow(int n, const double *a, const double *x)
{
    if (n == 1) return a[0] * x[0];
    double t = fma(a[0], x[0], a[1] * x[1]);
    for (int i = 2; i < n; ++i) t = fma(a[i], x[i], t);
    return t;
}

static inline double sgn(double a) { return (a > 0.0) ? 1.0 : ((a < 0.0) ? -1.0 : 0.0); }

/* limiter.py:3-12 */
static inline double minmod(double a, double b, double c, double Mdx2)
{
    if (fabs(a) < Mdx2) return a;
    double sa = sgn(a), sb = sgn(b), sc = sgn(c);
    if (sa == sb && sb == sc) {
        double m = fabs(a);
        if (fabs(b) < m) m = fabs(b);
        if (fabs(c) < m) m = fabs(c);
        return sa * m;
    }
    return 0.0;
}

/* ---- scalar PDE plugins: linadv.py, burger.py ------------------------------------------ */
/* varadv.py:3-4: 1.0 + (1.0 - x**2)**5.  numpy/libm evaluate the fifth power with pow(); here t*t*t*t*t by
 * squaring (<= 2 ulp from pow): varadv parity is tolerance-level, not bitwise. */
static inline double va_speed(double x)
{
    double t = 1.0 - x * x;
    double t2 = t * t;
    return 1.0 + (t2 * t2) * t;
}

static inline double sc_flux_x(int pde, double x, double u)
{
    if (pde == PDE_VARADV) return va_speed(x) * u; /* varadv.py:7-8 */
    return (pde == PDE_LINEAR) ? u : 0.5 * (u * u); /* linadv.py:2-3, burger.py:4-5 */
}

static inline double sc_flux(int pde, double u) { return sc_flux_x(pde, 0.0, u); }

static inline double sc_numflux(int pde, int flux, double x, double ul, double ur)
{
    if (pde == PDE_VARADV) { /* varadv.py:11-13: max(a,0.0)*ul + min(a,0.0)*ur */
        double a = va_speed(x);
        double ap = (0.0 > a) ? 0.0 : a, am = (0.0 < a) ? 0.0 : a;
        return ap * ul + am * ur;
    }
    if (pde == PDE_LINEAR) {
        if (flux == FLUX_CENTRAL) return 0.5 * (ul + ur); /* linadv.py:6-7 */
        return ul;                                        /* linadv.py:10-11 */
    }
    if (flux == FLUX_CENTRAL) return 0.5 * (sc_flux(pde, ul) + sc_flux(pde, ur)); /* burger.py:7-8 */
    if (flux == FLUX_ROE) {                                                      /* burger.py:16-18 */
        double a = fabs((ul + ur) / 2);
        return 0.5 * (sc_flux(pde, ul) + sc_flux(pde, ur)) - 0.5 * a * (ur - ul);
    }
    /* godunov, burger.py:11-14; python max/min keep the first argument on ties */
    double u1 = (ul > 0.0) ? ul : 0.0;
    double u2 = (ur < 0.0) ? ur : 0.0;
    double f1 = sc_flux(pde, u1), f2 = sc_flux(pde, u2);
    return (f2 > f1) ? f2 : f1;
}

/* dgsolver.py:170-203.  u1,res: (nc,nd).  Periodic. */
void orc_scalar_residual(const orc_ops *op, int pde, int flux, long nc, const double *u1, double *res)
{
    const int nd = op->k + 1;
    double nVg[MAXND * MAXND];
    for (int i = 0; i < nd * nd; ++i) nVg[i] = -op->Vg[i]; /* `-Vg @ f` is (-Vg) @ f */
    double *ff = (double *)malloc(sizeof(double) * (size_t)nc); /* ff[i] = flux at left face of cell i */
#pragma omp parallel for schedule(static)
    for (long i = 0; i < nc; ++i) {
        double u[MAXND], f[MAXND];
        const double xc = op->xmin + i * op->dx + 0.5 * op->dx; /* dgsolver.py:180-181 */
        for (int q = 0; q < nd; ++q) u[q] = mvrow(nd, op->Vf + q * nd, u1 + i * nd);
        for (int q = 0; q < nd; ++q) f[q] = sc_flux_x(pde, xc + 0.5 * op->dx * op->xg[q], u[q]);
        for (int m = 0; m < nd; ++m) res[i * nd + m] = mvrow(nd, nVg + m * nd, f);
        long il = (i == 0) ? nc - 1 : i - 1;
        double ul = dotv(nd, u1 + il * nd, op->bp);
        double ur = dotv(nd, u1 + i * nd, op->bm);
        ff[i] = sc_numflux(pde, flux, op->xmin + i * op->dx, ul, ur); /* dgsolver.py:190,199 (i = 0: xmin) */
    }
#pragma omp parallel for schedule(static)
    for (long i = 0; i < nc; ++i) {
        double fl = ff[i], fr = ff[(i == nc - 1) ? 0 : i + 1];
        for (int m = 0; m < nd; ++m) {
            double r = res[i * nd + m];
            if (i == nc - 1) { /* periodic face is scattered first: dgsolver.py:191 */
                r = r + fr * op->bp[m];
                r = r - fl * op->bm[m];
            } else {
                r = r - fl * op->bm[m];
                r = r + fr * op->bp[m];
            }
            res[i * nd + m] = r;
        }
    }
    free(ff);
}

/* dgsolver.py:272 / euler.py:235-237, n doubles */
void orc_rk_update(const orc_ops *op, int rk, long n, const double *u0, double *u1, const double *res, double lam)
{
    const double a = op->ark[rk], b = op->brk[rk];
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i) u1[i] = a * u0[i] + b * (u1[i] - lam * res[i]);
}

/* dgsolver.py:206-228 (the caller applies the `limit == 'yes'` gate) */
void orc_scalar_limiter(const orc_ops *op, long nc, double *u1, double Mdx2)
{
    const int nd = op->k + 1;
    const double sqrt3 = op->sqrt3, isqrt3 = 1.0 / sqrt3;
#pragma omp parallel for schedule(static)
    for (long i = 0; i < nc; ++i) {
        double ul = u1[((i == 0) ? nc - 1 : i - 1) * nd];
        double ur = u1[((i == nc - 1) ? 0 : i + 1) * nd];
        double du = isqrt3 * minmod(sqrt3 * u1[i * nd + 1], u1[i * nd] - ul, ur - u1[i * nd], Mdx2);
        if (fabs(du - u1[i * nd + 1]) > 1.0e-6) {
            u1[i * nd + 1] = du;
            for (int m = 2; m < nd; ++m) u1[i * nd + m] = 0.0;
        }
    }
}

/* linadv.py:13, burger.py:21-22 */
double orc_scalar_max_speed(int pde, int k, long nc, const double *u1)
{
    if (pde == PDE_LINEAR) return 1.0;
    if (pde == PDE_VARADV) return -1.0; /* constant: the caller uses orc_ops.speed_const */
    const int nd = k + 1;
    double m = 0.0;
    int nan = 0; /* ndarray.max() propagates NaN */
#pragma omp parallel for reduction(max : m) reduction(| : nan) schedule(static)
    for (long i = 0; i < nc; ++i) {
        double a = fabs(u1[i * nd]);
        if (a != a) nan = 1;
        if (a > m) m = a;
    }
    return nan ? NAN : m;
}

/* dgsolver.py:255-278 (plotting omitted).  u0,u1,res: (nc,nd) work arrays owned by the caller.
 * Returns the number of steps; *t_out the final time; dt_log (may be NULL) receives min(it,cap) dts. */
long orc_scalar_solve(const orc_ops *op, int pde, int flux, int limit, long nc, double cfl, double dx,
                      double Tf, double Mdx2, double *u0, double *u1, double *res, long max_steps,
                      double *t_out, double *dt_log, long dt_cap)
{
    const int nd = op->k + 1;
    long it = 0;
    double t = 0.0;
    while (t < Tf && (max_steps < 0 || it < max_steps)) {
        double ms = (pde == PDE_VARADV) ? op->speed_const : orc_scalar_max_speed(pde, op->k, nc, u1);
        double dt = cfl * dx / ms;
        double lam = dt / dx;
        if (t + dt > Tf) {
            dt = Tf - t;
            lam = dt / dx;
        }
        memcpy(u0, u1, sizeof(double) * (size_t)nc * nd);
        for (int rk = 0; rk < 3; ++rk) {
            orc_scalar_residual(op, pde, flux, nc, u1, res);
            orc_rk_update(op, rk, nc * nd, u0, u1, res, lam);
            if (limit) orc_scalar_limiter(op, nc, u1, Mdx2);
        }
        if (dt_log && it < dt_cap) dt_log[it] = dt;
        t += dt;
        it += 1;
    }
    if (t_out) *t_out = t;
    return it;
}

/* ---- Scalar PDEs on a mesh with closed ends: the C++ twin's inflow/outflow boundaries ----------------
 * dg1d/scalar_legendre/dg.h:367-413: an inflow face takes the physical flux of boundary_condition(stage_time), an
 * outflow face the physical flux of the cell's own trace.  The twin needs deal.II and cannot be built here: this part
 * of the oracle is PARITY UNPINNED; it restates the twin's boundary rule inside the Python reference's
 * arithmetic (same volume term, traces, interior order (vol - fL*bm) + fR*bp for every cell), and leaves the two
 * boundary cells unlimited as euler.py:242 does for its copy boundaries (the twin's limiter has no non-periodic form). */
enum { BC_INFLOW = 3, BC_OUTFLOW = 1 };

void orc_scalar_residual_bc(const orc_ops *op, int pde, int flux, long nc, const double *u1, double *res, int bcl, int bcr,
                            double gl, double gr)
{
    const int nd = op->k + 1;
    double nVg[MAXND * MAXND];
    for (int i = 0; i < nd * nd; ++i) nVg[i] = -op->Vg[i];
    double *ff = (double *)malloc(sizeof(double) * (size_t)(nc + 1)); /* ff[i] = flux at face i, i = 0..nc */
#pragma omp parallel for schedule(static)
    for (long i = 0; i < nc; ++i) {
        double u[MAXND], f[MAXND];
        const double xc = op->xmin + i * op->dx + 0.5 * op->dx;
        for (int q = 0; q < nd; ++q) u[q] = mvrow(nd, op->Vf + q * nd, u1 + i * nd);
        for (int q = 0; q < nd; ++q) f[q] = sc_flux_x(pde, xc + 0.5 * op->dx * op->xg[q], u[q]);
        for (int m = 0; m < nd; ++m) res[i * nd + m] = mvrow(nd, nVg + m * nd, f);
        const double xf = op->xmin + i * op->dx;
        const double ur = dotv(nd, u1 + i * nd, op->bm);
        if (i > 0) {
            ff[i] = sc_numflux(pde, flux, xf, dotv(nd, u1 + (i - 1) * nd, op->bp), ur);
        } else if (bcl == BC_INFLOW) {
            ff[0] = sc_flux_x(pde, xf, gl); /* dg.h:372-376 */
        } else {
            ff[0] = sc_numflux(pde, flux, xf, ur, ur); /* outflow: numflux(u,u) is the physical flux (dg.h:378-382) */
        }
    }
    {
        const double xf = op->xmin + nc * op->dx;
        const double ul = dotv(nd, u1 + (nc - 1) * nd, op->bp);
        ff[nc] = (bcr == BC_INFLOW) ? sc_flux_x(pde, xf, gr) : sc_numflux(pde, flux, xf, ul, ul); /* dg.h:397-407 */
    }
#pragma omp parallel for schedule(static)
    for (long i = 0; i < nc; ++i)
        for (int m = 0; m < nd; ++m) {
            double r = res[i * nd + m];
            r = r - ff[i] * op->bm[m];
            r = r + ff[i + 1] * op->bp[m];
            res[i * nd + m] = r;
        }
    free(ff);
}

void orc_scalar_limiter_bc(const orc_ops *op, long nc, double *u1, double Mdx2)
{
    const int nd = op->k + 1;
    const double sqrt3 = op->sqrt3, isqrt3 = 1.0 / sqrt3;
#pragma omp parallel for schedule(static)
    for (long i = 1; i < nc - 1; ++i) { /* averages are never changed, so the loop order does not matter */
        double ul = u1[(i - 1) * nd], ur = u1[(i + 1) * nd];
        double du = isqrt3 * minmod(sqrt3 * u1[i * nd + 1], u1[i * nd] - ul, ur - u1[i * nd], Mdx2);
        if (fabs(du - u1[i * nd + 1]) > 1.0e-6) {
            u1[i * nd + 1] = du;
            for (int m = 2; m < nd; ++m) u1[i * nd + m] = 0.0;
        }
    }
}

/* The time loop of dgsolver.py:255-278 with the twin's stage times (dg.h:539, :602): t, t+dt, then
 * ark[2]*t + brk[2]*((t+dt)+dt) ... evaluated with the reference's ark/brk.  bdry[(3*step + rk)*2 + {0,1}] are the
 * prescribed left/right states of each stage, filled by the caller for `nsteps_table` steps (the caller knows dt in
 * advance only for constant wave speed; otherwise it calls this function one step at a time with max_steps). */
long orc_scalar_solve_bc(const orc_ops *op, int pde, int flux, int limit, long nc, double cfl, double dx, double Tf,
                         double Mdx2, double *u0, double *u1, double *res, long max_steps, double *t_io, int bcl, int bcr,
                         const double *bdry)
{
    const int nd = op->k + 1;
    long it = 0;
    double t = *t_io;
    while (t < Tf && (max_steps < 0 || it < max_steps)) {
        double ms = (pde == PDE_VARADV) ? op->speed_const : orc_scalar_max_speed(pde, op->k, nc, u1);
        double dt = cfl * dx / ms;
        double lam = dt / dx;
        if (t + dt > Tf) {
            dt = Tf - t;
            lam = dt / dx;
        }
        memcpy(u0, u1, sizeof(double) * (size_t)nc * nd);
        for (int rk = 0; rk < 3; ++rk) {
            const double *b = bdry + (3 * it + rk) * 2;
            orc_scalar_residual_bc(op, pde, flux, nc, u1, res, bcl, bcr, b[0], b[1]);
            orc_rk_update(op, rk, nc * nd, u0, u1, res, lam);
            if (limit) orc_scalar_limiter_bc(op, nc, u1, Mdx2);
        }
        t += dt;
        it += 1;
    }
    *t_io = t;
    return it;
}

/* ---- Euler: gas.py ------------------------------------------------------------------------ */
static inline double eu_pressure(double g, double rho, double mom, double ene)
{
    return (g - 1.0) * (ene - 0.5 * (mom * mom) / rho); /* gas.py:7-8 */
}

static inline void eu_flux(double g, const double u[3], double f[3])
{
    double pre = eu_pressure(g, u[0], u[1], u[2]); /* gas.py:11-16 */
    f[0] = u[1];
    f[1] = pre + (u[1] * u[1]) / u[0];
    f[2] = (u[2] + pre) * u[1] / u[0];
}

static inline double eu_max_eig(double g, const double u[3])
{
    double vel = u[1] / u[0]; /* gas.py:20-24 */
    double pre = eu_pressure(g, u[0], u[1], u[2]);
    return fabs(vel) + sqrt(g * pre / u[0]);
}

/* gas.py:79-89 */
static inline void eu_rusanov(double g, const double ul[3], const double ur[3], double f[3])
{
    double fl[3], fr[3];
    eu_flux(g, ul, fl);
    eu_flux(g, ur, fr);
    double laml = eu_max_eig(g, ul), lamr = eu_max_eig(g, ur);
    double lam = (lamr > laml) ? lamr : laml; /* python max(laml, lamr) */
    for (int v = 0; v < 3; ++v) f[v] = 0.5 * (fl[v] + fr[v]) - 0.5 * lam * (ur[v] - ul[v]);
}

/* Roe flux with entropy fix delta = 0.1 c: dg2d/models/euler/pde.h:209-285 restricted to dim=1,
 * normal = (+1).  dg1d/python ships no Roe flux; pinned bit for bit against PDE::roe_flux<2> itself
 * (normal (1,0), zero y-momentum) compiled from the reference header: oracle/ref_dg2d_roe.cc, tests/golden/csibvec_*.npz
 * key 'roe', tests/golden/dg2d_roe_*.npz. */
void orc_euler_roe_flux(double g, const double ul[3], const double ur[3], double flux[3])
{
    /* con2prim, pde.h:62-78 */
    const double rho_l = ul[0], rho_r = ur[0];
    const double v_l = ul[1] / rho_l, v_r = ur[1] / rho_r;
    const double p_l = (g - 1.0) * (ul[2] - 0.5 * rho_l * (v_l * v_l));
    const double p_r = (g - 1.0) * (ur[2] - 0.5 * rho_r * (v_r * v_r));
    const double rho_l_sqrt = sqrt(rho_l), rho_r_sqrt = sqrt(rho_r);
    const double fact_l = rho_l_sqrt / (rho_l_sqrt + rho_r_sqrt);
    const double fact_r = 1.0 - fact_l;
    const double velocity = v_l * fact_l + v_r * fact_r;
    const double v2 = velocity * velocity;
    const double vel_normal = velocity;
    const double dv = v_r - v_l;
    const double v_dot_dv = velocity * dv;
    const double v2_l = v_l * v_l, v2_r = v_r * v_r;
    const double h_l = g * p_l / (rho_l * (g - 1.0)) + 0.5 * v2_l;
    const double h_r = g * p_r / (rho_r * (g - 1.0)) + 0.5 * v2_r;
    const double density = rho_l_sqrt * rho_r_sqrt;
    const double h = h_l * fact_l + h_r * fact_r;
    const double c = sqrt((g - 1.0) * (h - 0.5 * v2));
    const double drho = rho_r - rho_l;
    const double dp = p_r - p_l;
    const double dvn = v_r - v_l;
    const double a1 = (dp - density * c * dvn) / (2.0 * c * c);
    const double a2 = drho - dp / (c * c);
    const double a3 = (dp + density * c * dvn) / (2.0 * c * c);
    double l1 = fabs(vel_normal - c);
    double l2 = fabs(vel_normal);
    double l3 = fabs(vel_normal + c);
    const double delta = 0.1 * c;
    if (l1 < delta) l1 = 0.5 * (l1 * l1 / delta + delta);
    if (l3 < delta) l3 = 0.5 * (l3 * l3 / delta + delta);
    const double D0 = l1 * a1 + l2 * a2 + l3 * a3;
    const double D2 = l1 * a1 * (h - c * vel_normal) + l2 * a2 * 0.5 * v2
                    + l2 * density * (v_dot_dv - vel_normal * dvn) + l3 * a3 * (h + c * vel_normal);
    flux[0] = 0.5 * (rho_l * v_l + rho_r * v_r - D0);
    flux[2] = 0.5 * (rho_l * h_l * v_l + rho_r * h_r * v_r - D2);
    const double p_avg = 0.5 * (p_l + p_r);
    const double D1 = (velocity - c) * l1 * a1 + velocity * l2 * a2 + (dv - dvn) * l2 * density
                    + (velocity + c) * l3 * a3;
    flux[1] = p_avg + 0.5 * (ul[1] * v_l + ur[1] * v_r) - 0.5 * D1;
}

/* ---- The flux family of the C sibling, dg1d/c/euler/src/flux.c ------------------------
 * Pinned against that file itself, compiled where it lies into oracle/_ref/libcsib_euler_flux.so
 * (oracle/Makefile, target ref), and against tests/golden/csib_*.npz generated from it.  The sibling fixes
 * gamma = 1.4 as a macro; here it is the run-time g (identical bits for g = 1.4: every constant expression
 * of flux.c is one rounded operation).  Association follows C: a*b/c/d = ((a*b)/c)/d. */
typedef struct { double d, v, p; } csib_prim;

static inline csib_prim csib_primitive(double g, const double u[3])
{
    csib_prim w; /* flux.c:52-54 and every other flux there: d, u = m/d, p = (g-1)(E - 0.5 d u u) */
    w.d = u[0];
    w.v = u[1] / w.d;
    w.p = (g - 1.0) * (u[2] - 0.5 * w.d * w.v * w.v);
    return w;
}

static inline double csib_enthalpy(double g, csib_prim w) /* flux.c:85 */
{
    return g * w.p / w.d / (g - 1.0) + 0.5 * w.v * w.v;
}

/* flux.c:29-42 (LFFlux) with flux.c:47-68 (MaxEigVal) */
static void csib_lf(double g, const double ul[3], const double ur[3], double f[3])
{
    double fl[3], fr[3];
    eu_flux(g, ul, fl); /* flux.c:10-19 == gas.py:11-16 bit for bit (0.5*m*m == 0.5*(m*m)) */
    eu_flux(g, ur, fr);
    const csib_prim l = csib_primitive(g, ul), r = csib_primitive(g, ur);
    const double ll = fabs(l.v) + sqrt(g * l.p / l.d), lr = fabs(r.v) + sqrt(g * r.p / r.d);
    const double lam = (ll > lr) ? ll : lr;
    for (int i = 0; i < 3; ++i) f[i] = 0.5 * (fl[i] + fr[i] - lam * (ur[i] - ul[i]));
}

/* flux.c:139-166 (RoeAverage): a conserved state with Roe-averaged u, h and d = sqrt(dl dr) */
static void csib_roe_average(double g, const double ul[3], const double ur[3], double U[3])
{
    const csib_prim l = csib_primitive(g, ul), r = csib_primitive(g, ur);
    const double hl = csib_enthalpy(g, l), hr = csib_enthalpy(g, r);
    const double d1 = sqrt(l.d), d2 = sqrt(r.d), d3 = 1.0 / (d1 + d2);
    const double u = (d1 * l.v + d2 * r.v) * d3;
    const double h = (d1 * hl + d2 * hr) * d3;
    const double d = sqrt(l.d * r.d);
    const double p = (g - 1.0) * (h - 0.5 * u * u) * d / g;
    U[0] = d;
    U[1] = d * u;
    U[2] = p / (g - 1.0) + 0.5 * d * u * u;
}

/* project.c:66-106 (EigMat): same matrices as gas.py:37-76 but with chained divisions */
static void csib_eigmat(double g, const double U[3], double R[3][3], double Ri[3][3])
{
    const double g1 = g - 1.0, g2 = g1 / 2.0;
    const csib_prim w = csib_primitive(g, U);
    const double d = w.d, v = w.v;
    const double c = sqrt(g * w.p / d);
    const double h = c * c / g1 + 0.5 * v * v;
    const double f = d / c / 2.0;
    Ri[0][0] = 1.0 - g2 * v * v / c / c;
    Ri[1][0] = (g2 * v * v - v * c) / d / c;
    Ri[2][0] = -(g2 * v * v + v * c) / d / c;
    Ri[0][1] = g1 * v / c / c;
    Ri[1][1] = (c - g1 * v) / d / c;
    Ri[2][1] = (c + g1 * v) / d / c;
    Ri[0][2] = -g1 / c / c;
    Ri[1][2] = g1 / d / c;
    Ri[2][2] = -g1 / d / c;
    R[0][0] = 1.0;
    R[1][0] = v;
    R[2][0] = v * v / 2.0;
    R[0][1] = f;
    R[1][1] = (v + c) * f;
    R[2][1] = (h + v * c) * f;
    R[0][2] = -f;
    R[1][2] = -(v - c) * f;
    R[2][2] = -(h - v * c) * f;
}

/* project.c:109-122 (Multi): in place, accumulating from 0.0 */
static void csib_multi(double M[3][3], double x[3])
{
    const double t[3] = {x[0], x[1], x[2]};
    for (int i = 0; i < 3; ++i) {
        double a = 0.0;
        for (int j = 0; j < 3; ++j) a += M[i][j] * t[j];
        x[i] = a;
    }
}

/* flux.c:105-134 (LFCFlux): Lax-Friedrichs in the characteristic variables of the Roe-averaged state */
static void csib_lfc(double g, const double ul_in[3], const double ur_in[3], double f[3])
{
    double ul[3] = {ul_in[0], ul_in[1], ul_in[2]}, ur[3] = {ur_in[0], ur_in[1], ur_in[2]};
    double fl[3], fr[3], U[3], R[3][3], Ri[3][3], lam[3];
    eu_flux(g, ul, fl);
    eu_flux(g, ur, fr);
    csib_roe_average(g, ul, ur, U);
    csib_eigmat(g, U, R, Ri);
    csib_multi(Ri, fl);
    csib_multi(Ri, fr);
    csib_multi(Ri, ul);
    csib_multi(Ri, ur);
    const csib_prim w = csib_primitive(g, U);
    const double a = sqrt(g * w.p / w.d);
    lam[0] = fabs(w.v);
    lam[1] = fabs(w.v + a);
    lam[2] = fabs(w.v - a);
    for (int i = 0; i < 3; ++i) f[i] = 0.5 * (fl[i] + fr[i] - lam[i] * (ur[i] - ul[i]));
    csib_multi(R, f);
}

/* flux.c:171-236 (ECUSPFlux, Jameson's E-CUSP with Roe-averaged Mach number) */
static void csib_ecusp(double g, const double ul[3], const double ur[3], double f[3])
{
    const double eps = 0.01;
    double fl[3], fr[3];
    eu_flux(g, ul, fl);
    eu_flux(g, ur, fr);
    const csib_prim l = csib_primitive(g, ul), r = csib_primitive(g, ur);
    const double hl = csib_enthalpy(g, l), hr = csib_enthalpy(g, r);
    const double d1 = sqrt(l.d), d2 = sqrt(r.d), d3 = 1.0 / (d1 + d2);
    const double u = (d1 * l.v + d2 * r.v) * d3;
    const double h = (d1 * hl + d2 * hr) * d3;
    const double a = sqrt((g - 1.0) * (h - 0.5 * u * u));
    const double m = u / a;
    double alpha = fabs(m), beta, m1;
    if (alpha >= 1.0) {
        beta = (m > 0.0) ? 1.0 : -1.0;
    } else if (m >= 0.0) {
        m1 = 2.0 * m - 1.0;
        beta = (m1 > 0.0) ? m1 : 0.0;
    } else {
        m1 = 2.0 * m + 1.0;
        beta = (m1 < 0.0) ? m1 : 0.0;
    }
    if (alpha < eps) alpha = 0.5 * (eps + m * m / eps);
    const double du = r.v - l.v;
    const double fp[3] = {0.0, r.p - l.p, r.p * r.v - l.p * l.v};
    for (int i = 0; i < 3; ++i) {
        const double ua = 0.5 * (ul[i] + ur[i]);
        const double dv = alpha * a * (ur[i] - ul[i]) + beta * du * ua + beta * fp[i];
        f[i] = 0.5 * (fl[i] + fr[i] - dv);
    }
}

/* flux.c:241-302 (HLLCFlux).  Kept exactly as the sibling computes it: the star-region update is
 * S*f*(Us - U) with Us = (1, u*, ...), not S*(f*Us - U), so it is not a consistent flux unless the
 * density is 1; the branch order (and so the NaN behaviour) is the sibling's too. */
static void csib_hllc(double g, const double ul[3], const double ur[3], double f[3])
{
    const csib_prim l = csib_primitive(g, ul), r = csib_primitive(g, ur);
    const double al = sqrt(g * l.p / l.d), ar = sqrt(g * r.p / r.d);
    const double da = 0.5 * (l.d + r.d), aa = 0.5 * (al + ar);
    const double ps = 0.5 * (l.p + r.p - (r.v - l.v) * da * aa);
    const double us = 0.5 * (l.v + r.v - (r.p - l.p) / da / aa);
    double ql, qr;
    if (ps <= l.p) ql = 1.0;
    else ql = sqrt(1.0 + 0.5 * (g + 1.0) * (ps / l.p - 1.0) / g);
    if (ps <= r.p) qr = 1.0;
    else qr = sqrt(1.0 + 0.5 * (g + 1.0) * (ps / r.p - 1.0) / g);
    const double sl = l.v - al * ql, sr = r.v + ar * qr;
    if (0.0 <= sl) {
        eu_flux(g, ul, f);
    } else if (sl <= 0.0 && 0.0 <= us) {
        double fl[3];
        eu_flux(g, ul, fl);
        const double fac = l.d * (sl - l.v) / (sl - us);
        const double star[3] = {1.0, us, ul[2] / l.d + (us - l.v) * (us + l.p / l.d / (sl - l.v))};
        for (int i = 0; i < 3; ++i) f[i] = fl[i] + sl * fac * (star[i] - ul[i]);
    } else if (us <= 0.0 && 0.0 <= sr) {
        double fr[3];
        eu_flux(g, ur, fr);
        const double fac = r.d * (sr - r.v) / (sr - us);
        const double star[3] = {1.0, us, ur[2] / r.d + (us - r.v) * (us + r.p / r.d / (sr - r.v))};
        for (int i = 0; i < 3; ++i) f[i] = fr[i] + sr * fac * (star[i] - ur[i]);
    } else {
        eu_flux(g, ur, f);
    }
}

/* flux.c:307-406 (AUSMDVFlux, Wada & Liou, with the shock-fix blending K = 10) */
static void csib_ausmdv(double g, const double ql[3], const double qr[3], double f[3])
{
    const double g1 = g - 1.0;
    const double dl = ql[0], vl = ql[1] / ql[0], pl = g1 * (ql[2] - 0.5 * ql[1] * ql[1] / ql[0]);
    const double dr = qr[0], vr = qr[1] / qr[0], pr = g1 * (qr[2] - 0.5 * qr[1] * qr[1] / qr[0]);
    const double cl = sqrt(g * pl / dl), hl = cl * cl / g1 + 0.5 * vl * vl;
    const double cr = sqrt(g * pr / dr), hr = cr * cr / g1 + 0.5 * vr * vr;
    const double cm = (cl > cr) ? cl : cr;
    const double ml = fabs(vl) / cm, mr = fabs(vr) / cm;
    double t1 = pl / dl, t2 = pr / dr;
    const double wl = 2 * t1 / (t1 + t2), wr = 2 * t2 / (t1 + t2);
    double ulp, urm, plp, prm;
    t1 = vl + cm;
    t2 = vl + fabs(vl);
    if (ml <= 1.0) ulp = wl * (0.25 * t1 * t1 / cm - 0.5 * t2) + 0.5 * t2;
    else ulp = 0.5 * t2;
    t1 = vr - cm;
    t2 = vr - fabs(vr);
    if (mr <= 1.0) urm = wr * (-0.25 * t1 * t1 / cm - 0.5 * t2) + 0.5 * t2;
    else urm = 0.5 * t2;
    t1 = vl / cm + 1.0;
    t2 = 2.0 - vl / cm;
    if (ml <= 1.0) plp = 0.25 * pl * t1 * t1 * t2;
    else plp = 0.5 * pl * (vl + fabs(vl)) / (vl);
    t1 = vr / cm - 1.0;
    t2 = 2.0 + vr / cm;
    if (mr <= 1.0) prm = 0.25 * pr * t1 * t1 * t2;
    else prm = 0.5 * pr * (vr - fabs(vr)) / (vr);
    const double phalf = plp + prm;
    const double pmin = (pl < pr) ? pl : pr;
    double s = fabs(pr - pl) / pmin;
    double amin = 10.0 * s;
    amin = (amin < 1.0) ? amin : 1.0;
    s = 0.5 * amin;
    f[0] = ulp * dl + urm * dr;
    f[2] = 0.5 * (f[0] * (hl + hr) - fabs(f[0]) * (hr - hl));
    const double fv = dl * vl * ulp + dr * vr * urm;
    const double fd = 0.5 * (f[0] * (vl + vr) - fabs(f[0]) * (vr - vl));
    const double fn = (0.5 + s) * fv + (0.5 - s) * fd;
    f[1] = fn + phalf;
}

/* one entry for the tests: flux = FLUX_LF .. FLUX_LFC */
void orc_euler_csib_flux(double g, int flux, const double ul[3], const double ur[3], double f[3])
{
    switch (flux) {
    case FLUX_LF: csib_lf(g, ul, ur, f); break;
    case FLUX_ECUSP: csib_ecusp(g, ul, ur, f); break;
    case FLUX_HLLC: csib_hllc(g, ul, ur, f); break;
    case FLUX_AUSMDV: csib_ausmdv(g, ul, ur, f); break;
    case FLUX_LFC: csib_lfc(g, ul, ur, f); break;
    default: f[0] = f[1] = f[2] = NAN;
    }
}
static inline void eu_numflux(double g, int flux, const double ul[3], const double ur[3], double f[3])
{
    if (flux == FLUX_ROE) orc_euler_roe_flux(g, ul, ur, f);
    else if (flux >= FLUX_LF && flux <= FLUX_LFC) orc_euler_csib_flux(g, flux, ul, ur, f);
    else eu_rusanov(g, ul, ur, f);
}

void orc_euler_rusanov_flux(double g, const double ul[3], const double ur[3], double f[3])
{
    eu_rusanov(g, ul, ur, f);
}

/* euler.py:192-233.  u1,res: [3][nc][nd].  Copy ("neumann") boundary faces. */
void orc_euler_residual(const orc_ops *op, int flux, long nc, const double *u1, double *res)
{
    const int nd = op->k + 1;
    const double g = op->gamma;
    const long S = nc * nd;
    double nVg[MAXND * MAXND];
    for (int i = 0; i < nd * nd; ++i) nVg[i] = -op->Vg[i];
    double *ff = (double *)malloc(sizeof(double) * 3 * (size_t)(nc + 1)); /* face j = left face of cell j */
#pragma omp parallel for schedule(static)
    for (long i = 0; i < nc; ++i) {
        double uq[3][MAXND], fq[3][MAXND];
        for (int v = 0; v < 3; ++v)
            for (int q = 0; q < nd; ++q) uq[v][q] = mvrow(nd, op->Vf + q * nd, u1 + v * S + i * nd);
        for (int q = 0; q < nd; ++q) {
            double u[3] = {uq[0][q], uq[1][q], uq[2][q]}, f[3];
            eu_flux(g, u, f);
            fq[0][q] = f[0]; fq[1][q] = f[1]; fq[2][q] = f[2];
        }
        for (int v = 0; v < 3; ++v)
            for (int m = 0; m < nd; ++m) res[v * S + i * nd + m] = mvrow(nd, nVg + m * nd, fq[v]);
        double ul[3], ur[3];
        for (int v = 0; v < 3; ++v) {
            ur[v] = dotv(nd, u1 + v * S + i * nd, op->bm);
            ul[v] = (i == 0) ? ur[v] : dotv(nd, u1 + v * S + (i - 1) * nd, op->bp); /* euler.py:203-205 */
        }
        eu_numflux(g, flux, ul, ur, ff + 3 * i);
        if (i == nc - 1) { /* euler.py:227-230 */
            double w[3];
            for (int v = 0; v < 3; ++v) w[v] = dotv(nd, u1 + v * S + i * nd, op->bp);
            eu_numflux(g, flux, w, w, ff + 3 * nc);
        }
    }
#pragma omp parallel for schedule(static)
    for (long i = 0; i < nc; ++i)
        for (int v = 0; v < 3; ++v) {
            double fl = ff[3 * i + v], fr = ff[3 * (i + 1) + v];
            for (int m = 0; m < nd; ++m) {
                double r = res[v * S + i * nd + m];
                r = r - fl * op->bm[m];
                r = r + fr * op->bp[m];
                res[v * S + i * nd + m] = r;
            }
        }
    free(ff);
}

/* gas.py:37-76: L (rows = characteristic fields), R (columns = eigenvectors), from cell average */
static inline void eu_eigmat(double gam, const double u[3], double R[3][3], double L[3][3])
{
    const double g1 = gam - 1.0, g2 = 0.5 * g1;
    const double d = u[0];
    const double v = u[1] / d;
    const double p = eu_pressure(gam, u[0], u[1], u[2]);
    const double c = sqrt(gam * p / d);
    const double h = (c * c) / g1 + 0.5 * (v * v);
    const double f = 0.5 * d / c;
    L[0][0] = 1.0 - g2 * (v * v) / (c * c);
    L[1][0] = (g2 * (v * v) - v * c) / (d * c);
    L[2][0] = -(g2 * (v * v) + v * c) / (d * c);
    L[0][1] = g1 * v / (c * c);
    L[1][1] = (c - g1 * v) / (d * c);
    L[2][1] = (c + g1 * v) / (d * c);
    L[0][2] = -g1 / (c * c);
    L[1][2] = g1 / (d * c);
    L[2][2] = -g1 / (d * c);
    R[0][0] = 1.0;
    R[1][0] = v;
    R[2][0] = 0.5 * (v * v);
    R[0][1] = f;
    R[1][1] = (v + c) * f;
    R[2][1] = (h + v * c) * f;
    R[0][2] = -f;
    R[1][2] = -(v - c) * f;
    R[2][2] = -(h - v * c) * f;
}

void orc_euler_eigmat(double gam, const double u[3], double *R9, double *L9)
{
    eu_eigmat(gam, u, (double(*)[3])R9, (double(*)[3])L9);
}

/* euler.py:239-273 (caller applies nothing: the k>0 gate is here).  Cells 1..nc-2 only.
 * Mode 0 is never written, so the loop is order-independent and is parallelised. */
void orc_euler_limiter(const orc_ops *op, int char_lim, long nc, double *u1, double Mdx2)
{
    const int nd = op->k + 1;
    if (op->k <= 0) return;
    const long S = nc * nd;
    const double sqrt3 = op->sqrt3, isqrt3 = op->isqrt3;
#pragma omp parallel for schedule(static)
    for (long i = 1; i < nc - 1; ++i) {
        double dul[3], dur[3], uc[3], du[3], dun[3];
        for (int v = 0; v < 3; ++v) {
            const double *a = u1 + v * S;
            dul[v] = a[i * nd] - a[(i - 1) * nd];
            dur[v] = a[(i + 1) * nd] - a[i * nd];
            uc[v] = a[i * nd];
            du[v] = a[i * nd + 1];
        }
        double R[3][3], L[3][3];
        if (char_lim == 1) {
            eu_eigmat(op->gamma, uc, R, L);
            double t0[3], t1[3], t2[3];
            for (int r = 0; r < 3; ++r) {
                t0[r] = mvrow(3, L[r], du);
                t1[r] = mvrow(3, L[r], dul);
                t2[r] = mvrow(3, L[r], dur);
            }
            for (int r = 0; r < 3; ++r) { du[r] = t0[r]; dul[r] = t1[r]; dur[r] = t2[r]; }
        }
        int lim = 0;
        for (int j = 0; j < 3; ++j) {
            dun[j] = isqrt3 * minmod(sqrt3 * du[j], dul[j], dur[j], Mdx2);
            if (fabs(dun[j] - du[j]) > 1.0e-6) lim = 1;
        }
        if (lim) {
            if (char_lim == 1) {
                double t[3];
                for (int r = 0; r < 3; ++r) t[r] = mvrow(3, R[r], dun);
                for (int r = 0; r < 3; ++r) dun[r] = t[r];
            }
            for (int v = 0; v < 3; ++v) {
                u1[v * S + i * nd + 1] = dun[v];
                for (int m = 2; m < nd; ++m) u1[v * S + i * nd + m] = 0.0;
            }
        }
    }
}

/* gas.py:27-31 */
double orc_euler_max_speed(const orc_ops *op, long nc, const double *u1)
{
    const int nd = op->k + 1;
    const long S = nc * nd;
    const double g = op->gamma;
    double m = -INFINITY;
    int nan = 0; /* ndarray.max() propagates NaN: the reference's loop then ends with t = nan */
#pragma omp parallel for reduction(max : m) reduction(| : nan) schedule(static)
    for (long i = 0; i < nc; ++i) {
        double u[3] = {u1[i * nd], u1[S + i * nd], u1[2 * S + i * nd]};
        double lam = eu_max_eig(g, u);
        if (lam != lam) nan = 1;
        if (lam > m) m = lam;
    }
    return nan ? NAN : m;
}

/* euler.py:180-274 (plotting omitted).  u0,u1,res: [3][nc][nd]. */
long orc_euler_solve(const orc_ops *op, int flux, int char_lim, long nc, double cfl, double dx, double Tf,
                     double Mdx2, double *u0, double *u1, double *res, long max_steps, double *t_out,
                     double *dt_log, long dt_cap)
{
    const int nd = op->k + 1;
    const long n = 3 * nc * nd;
    long it = 0;
    double t = 0.0;
    while (t < Tf && (max_steps < 0 || it < max_steps)) {
        double dt = cfl * dx / orc_euler_max_speed(op, nc, u1);
        double lam = dt / dx;
        if (t + dt > Tf) {
            dt = Tf - t;
            lam = dt / dx;
        }
        memcpy(u0, u1, sizeof(double) * (size_t)n);
        for (int rk = 0; rk < 3; ++rk) {
            orc_euler_residual(op, flux, nc, u1, res);
            orc_rk_update(op, rk, n, u0, u1, res, lam);
            orc_euler_limiter(op, char_lim, nc, u1, Mdx2);
        }
        if (dt_log && it < dt_cap) dt_log[it] = dt;
        t += dt;
        it += 1;
    }
    if (t_out) *t_out = t;
    return it;
}

/* IC projection, dgsolver.py:117-122 / euler.py:162-169: u[i,j] = 0.5 * val.dot(Vf[:,j]*wg).
 * vals: (nc, Nq) point values; out: (nc, nd). */
void orc_project(const orc_ops *op, const double *wg, long nc, const double *vals, double *out)
{
    const int nd = op->k + 1;
    double W[MAXND][MAXND];
    for (int j = 0; j < nd; ++j)
        for (int q = 0; q < nd; ++q) W[j][q] = op->Vf[q * nd + j] * wg[q];
#pragma omp parallel for schedule(static)
    for (long i = 0; i < nc; ++i)
        for (int j = 0; j < nd; ++j) out[i * nd + j] = 0.5 * dotv(nd, vals + i * nd, W[j]);
}

/* dgsolver.py:232-252: error norm with an ng-point rule; V: (ng, nd), exact: (nc, ng).  Left-to-right sums. */
double orc_error_norm(int nd, int ng, const double *V, const double *w, long nc, double dx, const double *u1,
                      const double *exact)
{
    double err = 0.0;
    for (long i = 0; i < nc; ++i) {
        double e = 0.0;
        for (int q = 0; q < ng; ++q) {
            double un = dotv(nd, V + q * nd, u1 + i * nd);
            double d = un - exact[i * ng + q];
            e = e + (d * d) * w[q];
        }
        err += 0.5 * dx * e;
    }
    return sqrt(err);
}

/* Uniform-point evaluation: what the reference gives to its plotting calls (the display itself is left out).
 * Scalar, dgsolver.py:131-139 / :157-163: per cell f = Vu @ u1[i,:], umin = min(umin, f.min()), umax likewise, starting
 * from 1e20 / -1e20.  Euler, euler.py:83-100 / :121-138: r = Vu.dot(rho[i,:]); m = Vu.dot(mom[i,:]); v = m/r;
 * E = Vu.dot(ene[i,:]); pre = pressure(r,m,E) (gas.py:7-8 on arrays: one rounding per operation, m**2 = m*m), with the
 * running min/max of r, v and pre.  Vu: (nu, nd); u1: (nvar, nc, nd); out: (nvar, nc, nu) = f, or r, v, pre;
 * minmax: (nvar, 2).  numpy's min/max propagate NaN, so does this. */
static inline void run_minmax(double x, double *mn, double *mx)
{
    if (isnan(x) || isnan(*mn)) *mn = NAN; else if (x < *mn) *mn = x;
    if (isnan(x) || isnan(*mx)) *mx = NAN; else if (x > *mx) *mx = x;
}

void orc_eval_points(const orc_ops *op, int nvar, int nu, const double *Vu, long nc, const double *u1, double *out, double *minmax)
{
    const int nd = op->k + 1;
    for (int q = 0; q < nvar; ++q) { minmax[2 * q] = 1.0e20; minmax[2 * q + 1] = -1.0e20; }
    for (long i = 0; i < nc; ++i)
        for (int p = 0; p < nu; ++p) {
            if (nvar == 1) {
                const double f = mvrow(nd, Vu + p * nd, u1 + i * nd);
                out[i * nu + p] = f;
                run_minmax(f, &minmax[0], &minmax[1]);
            } else {
                const double r = mvrow(nd, Vu + p * nd, u1 + (0 * nc + i) * nd);
                const double m = mvrow(nd, Vu + p * nd, u1 + (1 * nc + i) * nd);
                const double E = mvrow(nd, Vu + p * nd, u1 + (2 * nc + i) * nd);
                const double v = m / r;
                const double pre = eu_pressure(op->gamma, r, m, E);
                out[(0 * nc + i) * nu + p] = r;
                out[(1 * nc + i) * nu + p] = v;
                out[(2 * nc + i) * nu + p] = pre;
                run_minmax(r, &minmax[0], &minmax[1]);
                run_minmax(v, &minmax[2], &minmax[3]);
                run_minmax(pre, &minmax[4], &minmax[5]);
            }
        }
}

int orc_sizeof_ops(void) { return (int)sizeof(orc_ops); }
