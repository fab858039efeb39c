// Numerical fluxes of the C sibling of the Python reference (dg1d/c/euler/src/flux.c), offered as
// further values of the Euler -num_flux option: lf, ecusp, hllc, ausmdv, lfc (dg1d/c/euler/src/dg.h:7-11).
// Arithmetic: one rounding per operation in the sibling's own order (the library is built with -fmad=false and nvcc's
// '/' and sqrt() are correctly rounded), so results are bit-identical to flux.c compiled by gcc -- checked against
// tests/golden/csibvec_flux_pairs.npz and the csib_* solver goldens.  The sibling fixes gamma = 1.4 (dg.h:20);
// here it is the handle's gamma.  These fluxes are not on the headline path: they use plain divisions.
// Included by dg1d_kernels.cuh inside namespace dg1d, after eu_flux().
#pragma once

struct CsibPrim {
    double d, v, p;
};

// d, u = m/d, p = (g-1)(E - 0.5 d u u): flux.c:52-54, 83-85, 143-145, ...
__device__ __forceinline__ CsibPrim csib_primitive(double g, const double (&u)[3])
{
    CsibPrim w;
    w.d = u[0];
    w.v = u[1] / w.d;
    w.p = (g - 1.0) * (u[2] - 0.5 * w.d * w.v * w.v);
    return w;
}

__device__ __forceinline__ double csib_enthalpy(double g, const CsibPrim &w)  // flux.c:85
{
    return g * w.p / w.d / (g - 1.0) + 0.5 * w.v * w.v;
}

// flux.c:10-19 (EulerFlux) is gas.py:11-16 bit for bit; eu_flux() is the literal form of the latter
__device__ __forceinline__ void csib_lf(double g, const double (&ul)[3], const double (&ur)[3], double (&f)[3])
{
    // flux.c:29-42 with MaxEigVal, flux.c:47-68
    double fl[3], fr[3];
    eu_flux(g, ul, fl);
    eu_flux(g, ur, fr);
    const CsibPrim l = csib_primitive(g, ul), r = csib_primitive(g, ur);
    const double ll = fabs(l.v) + sqrt(g * l.p / l.d), lr = fabs(r.v) + sqrt(g * r.p / r.d);
    const double lam = (ll > lr) ? ll : lr;
#pragma unroll
    for (int i = 0; i < 3; ++i) f[i] = 0.5 * (fl[i] + fr[i] - lam * (ur[i] - ul[i]));
}

// flux.c:139-166: conserved state with Roe-averaged velocity and enthalpy, density sqrt(dl dr)
__device__ __forceinline__ void csib_roe_average(double g, const double (&ul)[3], const double (&ur)[3], double (&U)[3])
{
    const CsibPrim l = csib_primitive(g, ul), r = csib_primitive(g, ur);
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

// project.c:66-106 (EigMat): the matrices of gas.py:37-76 with the sibling's chained divisions
__device__ __forceinline__ void csib_eigmat(double g, const double (&U)[3], double (&R)[3][3], double (&Ri)[3][3])
{
    const double g1 = g - 1.0, g2 = g1 / 2.0;
    const CsibPrim w = csib_primitive(g, U);
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

// project.c:109-122 (Multi): x <- M x, each row accumulated from 0.0
__device__ __forceinline__ void csib_multi(const double (&M)[3][3], double (&x)[3])
{
    const double t0 = x[0], t1 = x[1], t2 = x[2];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        double a = 0.0;
        a += M[i][0] * t0;
        a += M[i][1] * t1;
        a += M[i][2] * t2;
        x[i] = a;
    }
}

// flux.c:105-134 (LFCFlux): Lax-Friedrichs on the characteristic variables of the Roe-averaged state
__device__ __forceinline__ void csib_lfc(double g, const double (&ul_in)[3], const double (&ur_in)[3], double (&f)[3])
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
    const CsibPrim w = csib_primitive(g, U);
    const double a = sqrt(g * w.p / w.d);
    lam[0] = fabs(w.v);
    lam[1] = fabs(w.v + a);
    lam[2] = fabs(w.v - a);
#pragma unroll
    for (int i = 0; i < 3; ++i) f[i] = 0.5 * (fl[i] + fr[i] - lam[i] * (ur[i] - ul[i]));
    csib_multi(R, f);
}

// flux.c:171-236 (ECUSPFlux)
__device__ __forceinline__ void csib_ecusp(double g, const double (&ul)[3], const double (&ur)[3], double (&f)[3])
{
    const double eps = 0.01;
    double fl[3], fr[3];
    eu_flux(g, ul, fl);
    eu_flux(g, ur, fr);
    const CsibPrim l = csib_primitive(g, ul), r = csib_primitive(g, ur);
    const double hl = csib_enthalpy(g, l), hr = csib_enthalpy(g, r);
    const double d1 = sqrt(l.d), d2 = sqrt(r.d), d3 = 1.0 / (d1 + d2);
    const double u = (d1 * l.v + d2 * r.v) * d3;
    const double h = (d1 * hl + d2 * hr) * d3;
    const double a = sqrt((g - 1.0) * (h - 0.5 * u * u));
    const double m = u / a;
    double alpha = fabs(m), beta;
    if (alpha >= 1.0) {
        beta = (m > 0.0) ? 1.0 : -1.0;
    } else if (m >= 0.0) {
        const double m1 = 2.0 * m - 1.0;
        beta = (m1 > 0.0) ? m1 : 0.0;
    } else {
        const double m1 = 2.0 * m + 1.0;
        beta = (m1 < 0.0) ? m1 : 0.0;
    }
    if (alpha < eps) alpha = 0.5 * (eps + m * m / eps);
    const double du = r.v - l.v;
    const double fp[3] = {0.0, r.p - l.p, r.p * r.v - l.p * l.v};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const double ua = 0.5 * (ul[i] + ur[i]);
        const double dv = alpha * a * (ur[i] - ul[i]) + beta * du * ua + beta * fp[i];
        f[i] = 0.5 * (fl[i] + fr[i] - dv);
    }
}

// flux.c:241-302 (HLLCFlux), exactly as the sibling computes it: S*f*(Us - U) in the star regions, which is a
// consistent flux only where the density is 1, and the sibling's branch order (NaN falls to the last).
__device__ __forceinline__ void csib_hllc(double g, const double (&ul)[3], const double (&ur)[3], double (&f)[3])
{
    const CsibPrim l = csib_primitive(g, ul), r = csib_primitive(g, ur);
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
#pragma unroll
        for (int i = 0; i < 3; ++i) f[i] = fl[i] + sl * fac * (star[i] - ul[i]);
    } else if (us <= 0.0 && 0.0 <= sr) {
        double fr[3];
        eu_flux(g, ur, fr);
        const double fac = r.d * (sr - r.v) / (sr - us);
        const double star[3] = {1.0, us, ur[2] / r.d + (us - r.v) * (us + r.p / r.d / (sr - r.v))};
#pragma unroll
        for (int i = 0; i < 3; ++i) f[i] = fr[i] + sr * fac * (star[i] - ur[i]);
    } else {
        eu_flux(g, ur, f);
    }
}

// flux.c:307-406 (AUSMDVFlux; shock-fix blending with K = 10)
__device__ __forceinline__ void csib_ausmdv(double g, const double (&ql)[3], const double (&qr)[3], double (&f)[3])
{
    const double g1 = g - 1.0;
    const double dl = ql[0], vl = ql[1] / ql[0], pl = g1 * (ql[2] - 0.5 * ql[1] * ql[1] / ql[0]);
    const double dr = qr[0], vr = qr[1] / qr[0], pr = g1 * (qr[2] - 0.5 * qr[1] * qr[1] / qr[0]);
    const double cl = sqrt(g * pl / dl), hl = cl * cl / g1 + 0.5 * vl * vl;
    const double cr = sqrt(g * pr / dr), hr = cr * cr / g1 + 0.5 * vr * vr;
    const double cm = (cl > cr) ? cl : cr;
    const double ml = fabs(vl) / cm, mr = fabs(vr) / cm;
    double t1 = pl / dl, t2 = pr / dr;
    const double wl = 2.0 * t1 / (t1 + t2), wr = 2.0 * t2 / (t1 + t2);
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
