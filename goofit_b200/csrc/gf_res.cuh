// gf_res.cuh — three-body resonance shapes B_i(m12,m13,m23) evaluated from a node record.
// Arithmetic follows /root/reference/src/PDFs/physics/resonances/{Resonance,RBW,GS,NonRes,Gauss}.cu
// operation for operation (same association order) so cached waves agree to the last few ulp.
#pragma once
#include "gf_kmatrix.cuh"
#include "gf_math.cuh"

struct gf_decay3 { // Amp3Body node consts: Amp3Body.cu:144-149
    double M, d1, d2, d3, rad, mrad;
    double inv_M; // 1 / M (gf_load_decay3)
};

enum { GF_PAIR_12 = 0, GF_PAIR_13 = 1, GF_PAIR_23 = 2 }; // DalitzPlotHelpers.h:81

// gf_cm_mom / gf_cm_mom_mother with 1/s (and sqrt(s)) supplied by the caller, who shares them over everything a pair needs
__device__ __forceinline__ double gf_cm_mom_s(double rs, double inv_s, double a, double b) {
    double k1 = 1 - (a + b) * (a + b) * inv_s;
    k1 = k1 >= 0 ? sqrt(k1) : 1;
    double k2 = 1 - (a - b) * (a - b) * inv_s;
    k2 = k2 >= 0 ? sqrt(k2) : 1;
    return 0.5 * rs * k1 * k2;
}
// For a pair mass inside the Dalitz plot both k1 and k2 are negative (s < (dm - d3m)^2 < (dm + d3m)^2), i.e. both factors are 1:
// the square root is taken of the selected argument (sqrt(1) = 1), not selected after a square root of a negative number, which
// sent every event through the out-of-line NaN path of sqrt (ncu: 6 of the 9 calls per event-warp of k_amp3_waves)
__device__ __forceinline__ double gf_cm_mom_mother_s(double s, double inv_s, double dm, double inv_dm, double d3m) {
    double k1 = 1 - (dm + d3m) * (dm + d3m) * inv_s;
    double k2 = 1 - (dm - d3m) * (dm - d3m) * inv_s;
    if(k1 < 0 && k2 < 0) return 0.5 * s * inv_dm;
    k1 = sqrt(k1 >= 0 ? k1 : 1.0);
    k2 = sqrt(k2 >= 0 ? k2 : 1.0);
    return 0.5 * s * k1 * k2 * inv_dm;
}
// Resonance.cu:50-63
__device__ inline double gf_cm_mom(double s, double a, double b) {
    double k1 = 1 - (a + b) * (a + b) / s;
    k1 = k1 >= 0 ? sqrt(k1) : 1;
    double k2 = 1 - (a - b) * (a - b) / s;
    k2 = k2 >= 0 ? sqrt(k2) : 1;
    return 0.5 * sqrt(s) * k1 * k2;
}
// Resonance.cu:65-78
__device__ inline double gf_cm_mom_mother(double s, double dm, double d3m) {
    double k1 = 1 - (dm + d3m) * (dm + d3m) / s;
    k1 = k1 >= 0 ? sqrt(k1) : 1;
    double k2 = 1 - (dm - d3m) * (dm - d3m) / s;
    k2 = k2 >= 0 ? sqrt(k2) : 1;
    return 0.5 * s * k1 * k2 / dm;
}
// Resonance.cu:80-89
__device__ inline double gf_damp_sq(double q, int spin, double r) {
    double sq = r * r * q * q;
    double a = 2 * sq;
    double d = sq - 3;
    double b = 13 * sq / (d * d) + 9 * sq * sq; // pow(square-3, 2)
    return spin == 2 ? b : a;
}
// Resonance.cu:91-100
__device__ inline double gf_damp_sq_norm(double q, int spin, double r) {
    double sq = r * r * q * q;
    double a = 1 + sq;
    double b = a + 8 + 2 * sq + sq * sq;
    return spin == 2 ? b : a;
}
// Resonance.cu:102-153 (Zemach factor, EvtDalitzReso convention)
__device__ inline double gf_spin3(int spin, const gf_decay3 &D, double m12, double m13, double m23, int cyc) {
    if(spin == 0) return 1;
    double mA = cyc == GF_PAIR_12 ? D.d1 : (cyc == GF_PAIR_13 ? D.d3 : D.d2);
    double mB = cyc == GF_PAIR_12 ? D.d2 : (cyc == GF_PAIR_13 ? D.d1 : D.d3);
    double mC = cyc == GF_PAIR_12 ? D.d3 : (cyc == GF_PAIR_13 ? D.d2 : D.d1);
    double sAC = cyc == GF_PAIR_12 ? m13 : (cyc == GF_PAIR_13 ? m23 : m12);
    double sBC = cyc == GF_PAIR_12 ? m23 : (cyc == GF_PAIR_13 ? m12 : m13);
    double sAB = cyc == GF_PAIR_12 ? m12 : (cyc == GF_PAIR_13 ? m13 : m23);
    double mf = 1.0 / sAB;
    double s = -1;
    s *= ((sBC - sAC) + (mf * (D.M * D.M - mC * mC) * (mA * mA - mB * mB)));
    if(spin == 2) {
        s *= s;
        double t1 = D.M * D.M - mC * mC;
        double x = ((sAB - (2 * D.M * D.M) - (2 * mC * mC)) + mf * (t1 * t1));
        double t2 = mA * mA - mB * mB;
        x *= ((sAB - (2 * mA * mA) - (2 * mB * mB)) + mf * (t2 * t2));
        x /= 3;
        s -= x;
    }
    return s;
}

// the same with mf = 1 / sAB supplied (shared per pair)
__device__ __forceinline__ double gf_spin3_mf(int spin, const gf_decay3 &D, double m12, double m13, double m23, int cyc, double mf) {
    if(spin == 0) return 1;
    double mA = cyc == GF_PAIR_12 ? D.d1 : (cyc == GF_PAIR_13 ? D.d3 : D.d2);
    double mB = cyc == GF_PAIR_12 ? D.d2 : (cyc == GF_PAIR_13 ? D.d1 : D.d3);
    double mC = cyc == GF_PAIR_12 ? D.d3 : (cyc == GF_PAIR_13 ? D.d2 : D.d1);
    double sAC = cyc == GF_PAIR_12 ? m13 : (cyc == GF_PAIR_13 ? m23 : m12);
    double sBC = cyc == GF_PAIR_12 ? m23 : (cyc == GF_PAIR_13 ? m12 : m13);
    double sAB = cyc == GF_PAIR_12 ? m12 : (cyc == GF_PAIR_13 ? m13 : m23);
    double s = -1;
    s *= ((sBC - sAC) + (mf * (D.M * D.M - mC * mC) * (mA * mA - mB * mB)));
    if(spin == 2) {
        s *= s;
        double t1 = D.M * D.M - mC * mC;
        double x = ((sAB - (2 * D.M * D.M) - (2 * mC * mC)) + mf * (t1 * t1));
        double t2 = mA * mA - mB * mB;
        x *= ((sAB - (2 * mA * mA) - (2 * mB * mB)) + mf * (t2 * t2));
        x *= (1.0 / 3.0);
        s -= x;
    }
    return s;
}

// RBW.cu:9-86 ; nterm = 1 (plain) or 2 (symmetrised: second term with the next cyclic index)
__device__ inline double2 gf_res_rbw(int nterm, double mass, double width, int spin, int cyc, bool norm,
                                     const gf_decay3 &D, double m12, double m13, double m23) {
    double2 result = make_double2(0.0, 0.0);
    const double mass2 = mass * mass;
    for(int it = 0; it < nterm; ++it) {
        double s = cyc == GF_PAIR_12 ? m12 : (cyc == GF_PAIR_13 ? m13 : m23);
        double da = cyc == GF_PAIR_23 ? D.d2 : D.d1;
        double db = cyc == GF_PAIR_12 ? D.d2 : D.d3;
        double bach = cyc == GF_PAIR_12 ? D.d3 : (cyc == GF_PAIR_13 ? D.d2 : D.d1);
        double fr = 1, frM = 1;
        double q = gf_cm_mom(s, da, db);
        double q0 = gf_cm_mom(mass2, da, db);
        double qM, qM0 = 0;
        if(norm) {
            qM = gf_cm_mom_mother(s, D.M, bach);
            qM0 = gf_cm_mom_mother(mass2, D.M, bach);
        } else {
            qM = gf_cm_mom(D.M * D.M, sqrt(s), bach);
        }
        if(spin != 0) {
            if(norm) {
                fr = gf_damp_sq_norm(q0, spin, D.rad) / gf_damp_sq_norm(q, spin, D.rad);
                frM = gf_damp_sq_norm(qM0, spin, D.mrad) / gf_damp_sq_norm(qM, spin, D.mrad);
            } else {
                fr = gf_damp_sq(q, spin, D.rad);
                frM = gf_damp_sq(qM, spin, D.mrad);
            }
        }
        double A = mass2 - s;
        double B = mass2 * width * pow(q / q0, 2.0 * spin + 1) * fr / sqrt(s);
        double C = 1.0 / (A * A + B * B);
        double re = A * C, im = B * C;
        double f = sqrt(fr);
        re *= f; im *= f;
        f = sqrt(frM);
        re *= f; im *= f;
        f = gf_spin3(spin, D, m12, m13, m23, cyc);
        re *= f; im *= f;
        result.x += re;
        result.y += im;
        if(nterm > 1) cyc = (cyc + 1) % 3;
    }
    return result;
}

// Gounaris-Sakurai helpers, Resonance.cu:5-48
__device__ inline double gf_gs_h(double s, double a, double b) {
    const double pi = 3.14159265359;
    double sm = a + b, rs = sqrt(s), k = gf_cm_mom(s, a, b);
    return ((2 / pi) * (k / rs) * log((rs + 2 * k) / (sm)));
}
__device__ inline double gf_gs_dhds(double s, double a, double b) {
    const double pi = 3.14159265359;
    double k = gf_cm_mom(s, a, b);
    return gf_gs_h(s, a, b) * (1.0 / (8.0 * (k * k)) - 1.0 / (2.0 * s)) + 1.0 / (2.0 * pi * s);
}
__device__ inline double gf_gs_f(double s, double m2, double gam, double a, double b) {
    double k = gf_cm_mom(s, a, b), k0 = gf_cm_mom(m2, a, b);
    double f = gam * m2 / (k0 * k0 * k0);
    f *= ((k * k) * (gf_gs_h(s, a, b) - gf_gs_h(m2, a, b)) + (m2 - s) * (k0 * k0) * gf_gs_dhds(m2, a, b));
    return f;
}
__device__ inline double gf_gs_d(double s, double a, double b) {
    const double pi = 3.14159265359;
    double sm = a + b, sm24 = sm * sm / 4.0, m = sqrt(s), k = gf_cm_mom(s, a, b);
    return 3.0 / pi * sm24 / (k * k) * log((m + 2 * k) / sm) + m / (2 * pi * k) - sm24 * m / (pi * (k * k * k));
}
// GS.cu:9-75
__device__ inline double2 gf_res_gs(double mass, double width, int spin, int cyc, bool norm, const gf_decay3 &D,
                                    double m12, double m13, double m23) {
    double s = cyc == GF_PAIR_12 ? m12 : (cyc == GF_PAIR_13 ? m13 : m23);
    double bach = cyc == GF_PAIR_12 ? D.d3 : (cyc == GF_PAIR_13 ? D.d2 : D.d1);
    double da = cyc == GF_PAIR_23 ? D.d2 : D.d1;
    double db = cyc == GF_PAIR_12 ? D.d2 : D.d3;
    double fr = 1, frD = 1;
    double m2 = mass * mass;
    double q = gf_cm_mom(s, da, db), q0 = gf_cm_mom(m2, da, db);
    double qM, qM0 = 0;
    if(norm) {
        qM = gf_cm_mom_mother(s, D.M, bach);
        qM0 = gf_cm_mom_mother(m2, D.M, bach);
    } else {
        qM = gf_cm_mom(D.M * D.M, sqrt(s), bach);
    }
    if(spin != 0) {
        if(norm) {
            fr = gf_damp_sq_norm(q0, spin, D.rad);
            fr /= gf_damp_sq_norm(q, spin, D.rad);
            frD = gf_damp_sq_norm(qM0, spin, D.mrad);
            frD /= gf_damp_sq_norm(qM, spin, D.mrad);
        } else {
            fr = gf_damp_sq(q, spin, D.rad);
            frD = gf_damp_sq(qM, spin, D.mrad);
        }
    }
    double Dn = (1.0 + gf_gs_d(m2, D.d2, D.d3) * width / sqrt(m2));
    double E = m2 - s + gf_gs_f(s, m2, width, D.d2, D.d3);
    double F = sqrt(m2) * width * pow(q / q0, 2.0 * spin + 1) * fr;
    Dn /= (E * E + F * F);
    double re = Dn * E, im = Dn * F;
    double f = sqrt(fr);
    re *= f; im *= f;
    f = sqrt(frD);
    re *= f; im *= f;
    f = gf_spin3(spin, D, m12, m13, m23, cyc);
    re *= f; im *= f;
    return make_double2(re, im);
}
// Gauss.cu:9-28 (uses the mass, not the squared mass)
__device__ inline double2 gf_res_gauss(double mass, double width, int cyc, double m12, double m13, double m23) {
    double x = sqrt(cyc == GF_PAIR_12 ? m12 : (cyc == GF_PAIR_13 ? m13 : m23));
    x -= mass;
    x /= width;
    x *= x;
    double r = exp(-0.5 * x);
    r /= width;
    return make_double2(r, 0.0);
}

// complex helpers for the shapes below ((re, im) in a double2); division as in gf_amp4.cuh
__device__ inline double2 gf_cmul(double2 a, double2 b) { return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
__device__ inline double2 gf_cscale(double2 a, double s) { return make_double2(a.x * s, a.y * s); }
__device__ inline double2 gf_cdiv(double2 a, double2 b) {
    double denom = b.x * b.x + b.y * b.y;
    return make_double2((a.x * b.x + a.y * b.y) / denom, (a.y * b.x - a.x * b.y) / denom);
}
// LASS.cu:9-117 (K pi S-wave: resonance + effective-range background)
__device__ inline double2 gf_res_lass(const double *p, int spin, int cyc, bool norm, const gf_decay3 &D, double m12,
                                      double m13, double m23) {
    double resmass = p[0], reswidth = p[1], _a = p[2], _r = p[3], _R = p[4], _phiR = p[5], _B = p[6], _phiB = p[7];
    double s = cyc == GF_PAIR_12 ? m12 : (cyc == GF_PAIR_13 ? m13 : m23);
    double bach = cyc == GF_PAIR_12 ? D.d3 : (cyc == GF_PAIR_13 ? D.d2 : D.d1);
    double da = cyc == GF_PAIR_23 ? D.d2 : D.d1;
    double db = cyc == GF_PAIR_12 ? D.d2 : D.d3;
    double fr = 1, frD = 1;
    resmass *= resmass;
    double q = gf_cm_mom(s, da, db), q0 = gf_cm_mom(resmass, da, db);
    double qM, qM0 = 0;
    if(norm) {
        qM = gf_cm_mom_mother(s, D.M, bach);
        qM0 = gf_cm_mom_mother(resmass, D.M, bach);
    } else {
        qM = gf_cm_mom(D.M * D.M, sqrt(s), bach);
    }
    if(spin != 0) {
        if(norm) {
            fr = gf_damp_sq_norm(q0, spin, D.rad) / gf_damp_sq_norm(q, spin, D.rad);
            frD = gf_damp_sq_norm(qM0, spin, D.mrad) / gf_damp_sq_norm(qM, spin, D.mrad);
        } else {
            fr = gf_damp_sq(q, spin, D.rad);
            frD = gf_damp_sq(qM, spin, D.mrad);
        }
    }
    double g = reswidth * pow(q / q0, 2.0 * spin + 1) * fr / sqrt(s);
    double cot_deltaB = (1.0 / (_a * q)) + 0.5 * _r * q;
    double qcot_deltaB = (1.0 / _a) + 0.5 * _r * q * q;
    double2 expi2deltaB = gf_cdiv(make_double2(qcot_deltaB, q), make_double2(qcot_deltaB, -q));
    double2 resT = gf_cscale(make_double2(cos(_phiR + 2 * _phiB), sin(_phiR + 2 * _phiB)), _R);
    double2 prop = gf_cdiv(make_double2(1, 0), make_double2(resmass - s, sqrt(resmass) * g));
    resT = gf_cmul(resT, gf_cmul(gf_cscale(prop, resmass * reswidth / q0), expi2deltaB));
    double2 bg = gf_cscale(gf_cscale(gf_cscale(make_double2(cos(_phiB), sin(_phiB)), _B), cos(_phiB) + cot_deltaB * sin(_phiB)), sqrt(s));
    bg = gf_cdiv(bg, make_double2(qcot_deltaB, -q));
    resT.x += bg.x;
    resT.y += bg.y;
    resT = gf_cscale(resT, sqrt(fr));
    resT = gf_cscale(resT, sqrt(frD));
    resT = gf_cscale(resT, gf_spin3(spin, D, m12, m13, m23, cyc));
    return resT;
}
// FLATTE.cu:18-74 (coupled pi pi / K K channels; symmDP adds the m12 <-> m13 swapped term)
__device__ inline double2 gf_res_flatte(const double *p, int cyc, int doSwap, double m12, double m13, double m23) {
    const double resmass = p[0], g1 = p[1], g2 = p[2] * g1;
    const double pipmass = 0.13957018, pi0mass = 0.1349766, kpmass = 0.493677, k0mass = 0.497614;
    const double twopimasssq = 4 * pipmass * pipmass, twopi0masssq = 4 * pi0mass * pi0mass;
    const double twokmasssq = 4 * kpmass * kpmass, twok0masssq = 4 * k0mass * k0mass;
    double2 ret = make_double2(0.0, 0.0);
    for(int i = 0; i < 1 + doSwap; i++) {
        double rpr = 0, rpi = 0, rkr = 0, rki = 0;
        double s = cyc == GF_PAIR_12 ? m12 : (cyc == GF_PAIR_13 ? m13 : m23);
        if(s >= twopimasssq) rpr += (2. / 3) * sqrt(1 - twopimasssq / s); else rpi += (2. / 3) * sqrt(-1 + twopimasssq / s);
        if(s >= twopi0masssq) rpr += (1. / 3) * sqrt(1 - twopi0masssq / s); else rpi += (1. / 3) * sqrt(-1 + twopi0masssq / s);
        if(s >= twokmasssq) rkr += 0.5 * sqrt(1 - twokmasssq / s); else rki += 0.5 * sqrt(-1 + twokmasssq / s);
        if(s >= twok0masssq) rkr += 0.5 * sqrt(1 - twok0masssq / s); else rki += 0.5 * sqrt(-1 + twok0masssq / s);
        double A = (resmass * resmass - s) + resmass * (rpi * g1 + rki * g2);
        double B = resmass * (rpr * g1 + rkr * g2);
        double C = 1.0 / (A * A + B * B);
        ret.x += A * C;
        ret.y += B * C;
        if(doSwap) {
            double t = m12;
            m12 = m13;
            m13 = t;
        }
    }
    return ret;
}

// RhoOmegaMix.cu:9-135 : rho (Gounaris-Sakurai) times (1 + B * omega), B = (real + i imag) * delta * (m_rho + m_omega).
// Upstream writes the omega term out as RBW.cu's plain Breit-Wigner and the rho term as GS.cu's, operation for
// operation, so both are taken from the functions above. params: omega mass,width, rho mass,width, real, imag, delta.
__device__ inline double2 gf_res_rhoomega(const double *p, int spin, int cyc, bool norm, const gf_decay3 &D, double m12,
                                          double m13, double m23) {
    double2 omega = gf_res_rbw(1, p[0], p[1], spin, cyc, norm, D, m12, m13, m23);
    double2 rho = gf_res_gs(p[2], p[3], spin, cyc, norm, D, m12, m13, m23);
    double Delta = p[6] * (p[2] + p[0]);
    double2 mix = gf_cmul(make_double2(p[4] * Delta, p[5] * Delta), omega);
    mix.x += 1.0;
    return gf_cmul(rho, mix);
}

// Spline.cu:10-87 : complex cubic spline in the pair's squared mass (plus, with doSwap, the same spline in the second
// pair). consts: cyc, doSwap, nKnots, knot[nKnots], then the 2*nKnots second derivatives (re, im per knot) the host
// computed when the resonance was built -- upstream keeps those in one __device__ array filled by the constructor only
// (Spline.cu:8,112), so they do NOT follow the knot values during a fit; params: re, im per knot.
__device__ inline double2 gf_res_spline(const double *p, const double *c, double m12, double m13, double m23) {
    const int cyc = (int)c[0], runs = 1 + (int)c[1], n = (int)c[2];
    const double *knot = c + 3, *d2 = c + 3 + n;
    double mAB = m12, mAC = m13;
    if(cyc == GF_PAIR_13) { mAB = m13; mAC = m12; }
    if(cyc == GF_PAIR_23) { mAB = m23; mAC = m12; }
    double re = 0, im = 0;
    for(int it = 0; it < runs; ++it) {
        int hi = 0;
        while(hi < n && !(mAB < knot[hi])) hi++;
        if(hi > 0 && hi < n) {
            int lo = hi - 1;
            double dm = knot[hi] - knot[lo];
            double aa = (knot[hi] - mAB) / dm, bb = 1 - aa;
            double aa3 = aa * aa * aa, bb3 = bb * bb * bb;
            re = re + aa * p[2 * lo] + bb * p[2 * hi] + ((aa3 - aa) * d2[2 * lo] + (bb3 - bb) * d2[2 * hi]) * (dm * dm) / 6.0;
            im = im + aa * p[2 * lo + 1] + bb * p[2 * hi + 1]
                 + ((aa3 - aa) * d2[2 * lo + 1] + (bb3 - bb) * d2[2 * hi + 1]) * (dm * dm) / 6.0;
        }
        mAB = mAC;
    }
    return make_double2(re, im);
}

// Dispatch on the resonance record (Amp3Body_TD.cu:36-39 getResonanceAmplitude)
__device__ inline double2 gf_resonance(const gf_node &nd, const double *params, const double *consts,
                                       const gf_decay3 &D, double m12, double m13, double m23) {
    const double *p = params + nd.par_off;
    const double *c = consts + nd.con_off;
    switch(nd.op) {
    case GF_OP_RES_RBW: return gf_res_rbw(1, p[0], p[1], (int)c[0], (int)c[1], c[2] != 0.0, D, m12, m13, m23);
    case GF_OP_RES_RBW_SYM: return gf_res_rbw(2, p[0], p[1], (int)c[0], (int)c[1], c[2] != 0.0, D, m12, m13, m23);
    case GF_OP_RES_GS: return gf_res_gs(p[0], p[1], (int)c[0], (int)c[1], c[2] != 0.0, D, m12, m13, m23);
    case GF_OP_RES_NONRES: return make_double2(1.0, 0.0); // NonRes.cu:8-11
    case GF_OP_RES_GAUSS: return gf_res_gauss(p[0], p[1], (int)c[0], m12, m13, m23);
    case GF_OP_RES_LASS: return gf_res_lass(p, (int)c[0], (int)c[1], c[2] != 0.0, D, m12, m13, m23);
    case GF_OP_RES_FLATTE: return gf_res_flatte(p, (int)c[0], (int)c[1], m12, m13, m23);
    case GF_OP_RES_SPLINE: return gf_res_spline(p, c, m12, m13, m23);
    case GF_OP_RES_RHOOMEGA: return gf_res_rhoomega(p, (int)c[0], (int)c[1], c[2] != 0.0, D, m12, m13, m23);
    case GF_OP_RES_KMATRIX: { // resonances/kMatrix.cu:19-121: s is the pair's squared mass
        const int cyc = (int)c[0];
        const gf_kc r = gf_kmatrix(cyc == GF_PAIR_12 ? m12 : (cyc == GF_PAIR_13 ? m13 : m23), p);
        return make_double2(r.re, r.im);
    }
    default: return make_double2(0.0, 0.0);
    }
}

// out-of-line copy for the rare resonance kinds, so that their log/pow-heavy code does not set the register budget of
// the cache kernels' main loop
__device__ __noinline__ double2 gf_resonance_cold(const gf_node *nd, const double *params, const double *consts,
                                                  const gf_decay3 *D, double m12, double m13, double m23) {
    return gf_resonance(*nd, params, consts, *D, m12, m13, m23);
}

// ---- shared-work evaluation of all resonances of one Amp3Body ------------------------------------------------------
// The reference evaluates every resonance from scratch (one thrust::transform per resonance). Of that work
//   * q0, qM0 and the numerators of the normalised Blatt-Weisskopf ratios depend on the resonance parameters only
//     -> gf_res_inv, computed once per block into shared memory;
//   * s, sqrt(s), q(s), qM(s) and the Zemach factors depend on the event and the resonant PAIR only, and the 16
//     resonances of examples/dalitz share 3 pairs -> gf_pair_inv, computed once per event and pair.
// Every shared quantity is produced by the same operations as above, so it is bit-identical to the unshared path; the
// one deliberate change is pow(q/q0, 2*spin+1) -> a multiply chain (<= 2 ulp from libm's pow; spin 0 is exact).
struct gf_res_inv {
    int op, spin, cyc, norm, nterm, pad;
    double mass, width, mass2;
    double q0[2], qM0[2], frn[2], frMn[2]; // per symmetrisation term (cyc, cyc+1)
    double inv_q0, mw;                      // 1/q0 and mass^2 * width of the first term
    double sfr;                             // sqrt(frn[0] * frMn[0]) (normalised form)
};
struct gf_pair_inv {
    double s, rs, inv_rs, q, qMn, qMp, z1, z2; // z1/z2: Zemach factor for spin 1 / 2
    // Blatt-Weisskopf pieces per spin 1, 2: normalised form -> 1/F(q), 1/F(qM); plain form -> F(q), F(qM)
    double bw1, bw2, bwM1, bwM2;
    double sbw1, sbw2; // sqrt(bw1 * bwM1), sqrt(bw2 * bwM2) (normalised form)
};
struct gf_pair_need { // what the flagged resonances need of each pair (uniform over the grid)
    int used, norm, plain, maxspin;
};

__device__ inline void gf_pair_masses(int cyc, const gf_decay3 &D, double &da, double &db, double &bach) {
    da = cyc == GF_PAIR_23 ? D.d2 : D.d1;
    db = cyc == GF_PAIR_12 ? D.d2 : D.d3;
    bach = cyc == GF_PAIR_12 ? D.d3 : (cyc == GF_PAIR_13 ? D.d2 : D.d1);
}
__device__ inline void gf_res_prepare(gf_res_inv &R, const gf_node &nd, const double *params, const double *consts,
                                      const gf_decay3 &D) {
    const double *p = params + nd.par_off;
    const double *c = consts + nd.con_off;
    R.op = nd.op;
    R.spin = 0; R.cyc = 0; R.norm = 0; R.nterm = 1; R.pad = 0;
    R.mass = R.width = R.mass2 = 0;
    R.sfr = 1;
    if(nd.op != GF_OP_RES_RBW && nd.op != GF_OP_RES_RBW_SYM) return; // only plain RBWs take the shared-work path
    R.mass = p[0]; R.width = p[1]; R.mass2 = p[0] * p[0];
    R.spin = (int)c[0]; R.cyc = (int)c[1]; R.norm = c[2] != 0.0;
    R.nterm = nd.op == GF_OP_RES_RBW_SYM ? 2 : 1;
    int cyc = R.cyc;
    for(int t = 0; t < R.nterm; ++t) {
        double da, db, bach;
        gf_pair_masses(cyc, D, da, db, bach);
        R.q0[t] = gf_cm_mom(R.mass2, da, db);
        R.qM0[t] = R.norm ? gf_cm_mom_mother(R.mass2, D.M, bach) : 0.0;
        R.frn[t] = R.spin ? gf_damp_sq_norm(R.q0[t], R.spin, D.rad) : 1.0;
        R.frMn[t] = R.spin ? gf_damp_sq_norm(R.qM0[t], R.spin, D.mrad) : 1.0;
        cyc = (cyc + 1) % 3;
    }
    R.inv_q0 = 1.0 / R.q0[0];
    R.mw = R.mass2 * R.width;
    R.sfr = sqrt(R.frn[0] * R.frMn[0]);
}
__device__ inline void gf_pair_prepare(gf_pair_inv &Q, const gf_pair_need &need, int cyc, const gf_decay3 &D, double m12,
                                       double m13, double m23) {
    double da, db, bach;
    gf_pair_masses(cyc, D, da, db, bach);
    Q.s = cyc == GF_PAIR_12 ? m12 : (cyc == GF_PAIR_13 ? m13 : m23);
    Q.rs = sqrt(Q.s);
    // 1/s once per pair: it serves the two-body momenta (four divisions upstream), the Zemach factors (two) and 1/sqrt(s)
    const double inv_s = gf_rcp(Q.s);
    Q.inv_rs = Q.rs * inv_s;
    Q.q = gf_cm_mom_s(Q.rs, inv_s, da, db);
    Q.qMn = need.norm ? gf_cm_mom_mother_s(Q.s, inv_s, D.M, D.inv_M, bach) : 0.0;
    Q.qMp = need.plain ? gf_cm_mom(D.M * D.M, Q.rs, bach) : 0.0;
    Q.z1 = Q.z2 = 1.0;
    Q.bw1 = Q.bw2 = Q.bwM1 = Q.bwM2 = 1.0;
    Q.sbw1 = Q.sbw2 = 1.0;
    if(need.maxspin >= 1) Q.z1 = gf_spin3_mf(1, D, m12, m13, m23, cyc, inv_s);
    if(need.maxspin >= 2) Q.z2 = gf_spin3_mf(2, D, m12, m13, m23, cyc, inv_s);
    if(need.norm && need.maxspin >= 1) {
        Q.bw1 = gf_rcp(gf_damp_sq_norm(Q.q, 1, D.rad));
        Q.bwM1 = gf_rcp(gf_damp_sq_norm(Q.qMn, 1, D.mrad));
        Q.sbw1 = sqrt(Q.bw1 * Q.bwM1);
    }
    if(need.norm && need.maxspin >= 2) {
        Q.bw2 = gf_rcp(gf_damp_sq_norm(Q.q, 2, D.rad));
        Q.bwM2 = gf_rcp(gf_damp_sq_norm(Q.qMn, 2, D.mrad));
        Q.sbw2 = sqrt(Q.bw2 * Q.bwM2);
    }
}
__device__ inline double gf_pow_odd(double r, int spin) { // r^(2 spin + 1)
    const double r2 = r * r, o1 = r * r2, o2 = o1 * r2; // the same products in the same order as the loop, selected instead of branched on
    if(spin > 2) {
        double out = o2;
        for(int k = 2; k < spin; ++k) out *= r2;
        return out;
    }
    return spin == 0 ? r : (spin == 1 ? o1 : o2);
}
// one plain-RBW term from the shared pieces (RBW.cu:9-86 with the event/pair part hoisted into P). Against the
// reference's operation sequence this trades four divisions and one square root per resonance for multiplications by
// reciprocals that are shared per pair (1/sqrt(s), 1/F(q), 1/F(qM)) or per resonance (1/q0), and takes
// sqrt(fr * frM) once instead of sqrt(fr) * sqrt(frM): each substitution moves the result by at most an ulp or two.
// the normalised form (and spin 0) with the spin known at compile time: the same operations in the same order as the general
// function below, without its selects and loops (the spin is uniform over the warp, so the dispatch is one branch)
template <int SPIN>
__device__ __forceinline__ double2 gf_rbw_term_n(const gf_res_inv &R, const gf_pair_inv &P) {
    const double A = R.mass2 - P.s;
    const double x = P.q * R.inv_q0;
    double pw = x;
    if(SPIN >= 1) {
        const double x2 = x * x;
        pw = x * x2;
        if(SPIN == 2) pw = pw * x2;
    }
    double B = R.mw * pw;
    if(SPIN != 0) B = B * (R.frn[0] * (SPIN == 2 ? P.bw2 : P.bw1));
    B = B * P.inv_rs;
    const double C = gf_rcp(A * A + B * B);
    const double f = SPIN == 0 ? C : (R.sfr * (SPIN == 2 ? P.sbw2 : P.sbw1)) * (SPIN == 2 ? P.z2 : P.z1) * C;
    return make_double2(A * f, B * f);
}
__device__ inline double2 gf_rbw_term(const gf_res_inv &R, const gf_pair_inv &P, const gf_decay3 &D) {
    const int spin = R.spin;
    if(spin == 0) return gf_rbw_term_n<0>(R, P);
    if(R.norm && spin == 1) return gf_rbw_term_n<1>(R, P);
    if(R.norm && spin == 2) return gf_rbw_term_n<2>(R, P);
    double fr = 1, frM = 1;
    if(spin != 0) {
        if(R.norm) {
            fr = R.frn[0] * (spin == 2 ? P.bw2 : P.bw1); // as Resonance.cu:91-100: anything but spin 2 takes the spin-1 form
            frM = R.frMn[0] * (spin == 2 ? P.bwM2 : P.bwM1);
        } else {
            // not to be evaluated speculatively: with no plain-form resonance on the pair qMp is 0, and the division inside
            // gf_damp_sq then takes its out-of-line path (0 / x) on every event for a value that is thrown away
            asm volatile("");
            fr = gf_damp_sq(P.q, spin, D.rad);
            frM = gf_damp_sq(P.qMp, spin, D.mrad);
        }
    }
    double A = R.mass2 - P.s;
    double B = R.mw * gf_pow_odd(P.q * R.inv_q0, spin) * fr * P.inv_rs;
    double C = gf_rcp(A * A + B * B);
    // sqrt(fr frM) z C: spin 0 has fr = frM = z = 1 (exactly C); in the normalised form the square root splits into a factor per
    // resonance and one per pair and spin, both taken once (<= 2 ulp from the square root of the product)
    double f;
    if(spin == 0)
        f = C;
    else if(R.norm)
        f = (R.sfr * (spin == 2 ? P.sbw2 : P.sbw1)) * (spin == 2 ? P.z2 : P.z1) * C;
    else
        f = sqrt(fr * frM) * (spin == 2 ? P.z2 : P.z1) * C;
    return make_double2(A * f, B * f);
}
#ifndef GF_PAIR_UNROLL
#define GF_PAIR_UNROLL 0 // measured: unrolling the three pairs keeps all of their state live (468 B of spills, 205 local loads per warp-event); 2.50 -> 2.07 ms without
#endif
#ifndef GF_RBW_ILP
#define GF_RBW_ILP 1 // measured on 10^7 events x 16 resonances: 1: 2.35 ms, 2: 2.49 ms, 4: 2.37 ms (tools/ab_compare.sh): the chains already overlap across warps
#endif
#define GF_MAX_RES 64
struct gf_amp3_shared {
    gf_res_inv res[GF_MAX_RES];
    gf_pair_need need[3];
    int n_pair[3], n_other, n_flat;
    unsigned char by_pair[3][GF_MAX_RES]; // flagged plain RBWs grouped by resonant pair
    unsigned char other[GF_MAX_RES];      // flagged GS / Gauss / symmetrised RBW / ...: evaluated unshared, out of line
    unsigned char flat[GF_MAX_RES];       // flagged NonRes (NonRes.cu:8-11: the constant 1): no call, so that the ordinary Dalitz
                                          // model makes no out-of-line call per event (the spills around it were ~160 local
                                          // loads and stores per event in the round-2 profile)
};
// block-cooperative: fill S for the flagged resonances; call from all threads, then __syncthreads()
__device__ inline void gf_amp3_prepare(gf_amp3_shared &S, const gf_sprog &P, const int32_t *res_nodes, int nres,
                                       unsigned long long mask, const gf_decay3 &D) {
    for(int i = threadIdx.x; i < nres; i += blockDim.x)
        if((mask >> i) & 1ull) gf_res_prepare(S.res[i], P.nodes[res_nodes[i]], P.params, P.consts, D);
    __syncthreads();
    if(threadIdx.x < 3) {
        gf_pair_need n = {0, 0, 0, 0};
        const int c = threadIdx.x;
        int cnt = 0;
        for(int i = 0; i < nres; ++i) {
            if(!((mask >> i) & 1ull)) continue;
            const gf_res_inv &R = S.res[i];
            if(R.op != GF_OP_RES_RBW || R.cyc != c) continue;
            S.by_pair[c][cnt++] = (unsigned char)i;
            n.used = 1;
            if(R.norm) n.norm = 1; else n.plain = 1;
            if(R.spin > n.maxspin) n.maxspin = R.spin;
        }
        S.need[c] = n;
        S.n_pair[c] = cnt;
    } else if(threadIdx.x == 3) {
        int cnt = 0, nf = 0;
        for(int i = 0; i < nres; ++i) {
            if(!((mask >> i) & 1ull) || S.res[i].op == GF_OP_RES_RBW) continue;
            if(S.res[i].op == GF_OP_RES_NONRES) S.flat[nf++] = (unsigned char)i;
            else S.other[cnt++] = (unsigned char)i;
        }
        S.n_other = cnt;
        S.n_flat = nf;
    }
}
// every flagged resonance at one Dalitz point; out(i, value) receives each amplitude exactly once
template <class Out>
__device__ inline void gf_amp3_point(const gf_amp3_shared &S, const gf_sprog &P, const int32_t *res_nodes,
                                     const gf_decay3 &D, bool inside, double m12, double m13, double m23, const Out &out) {
#if GF_PAIR_UNROLL
#pragma unroll
#else
#pragma unroll 1
#endif
    for(int c = 0; c < 3; ++c) {
        const int n = S.n_pair[c];
        if(n == 0) continue;
        gf_pair_inv Q;
        if(inside) gf_pair_prepare(Q, S.need[c], c, D, m12, m13, m23);
        // GF_RBW_ILP terms are formed together: each is a dependent chain (A^2 + B^2 -> reciprocal -> square root), and the
        // resonances of a pair are independent of each other, so interleaving them hides the FP64 latency one chain leaves exposed
        int k = 0;
        for(; k + GF_RBW_ILP <= n; k += GF_RBW_ILP) {
            double2 v[GF_RBW_ILP];
#pragma unroll
            for(int u = 0; u < GF_RBW_ILP; ++u) v[u] = inside ? gf_rbw_term(S.res[S.by_pair[c][k + u]], Q, D) : make_double2(0.0, 0.0);
#pragma unroll
            for(int u = 0; u < GF_RBW_ILP; ++u) out(S.by_pair[c][k + u], v[u]);
        }
        for(; k < n; ++k) {
            const int i = S.by_pair[c][k];
            out(i, inside ? gf_rbw_term(S.res[i], Q, D) : make_double2(0.0, 0.0));
        }
    }
    for(int k = 0; k < S.n_flat; ++k) out(S.flat[k], inside ? make_double2(1.0, 0.0) : make_double2(0.0, 0.0));
    for(int k = 0; k < S.n_other; ++k) {
        const int i = S.other[k];
        out(i, inside ? gf_resonance_cold(&P.nodes[res_nodes[i]], P.params, P.consts, &D, m12, m13, m23) : make_double2(0.0, 0.0));
    }
}

