// gf_res.cuh — three-body resonance shapes B_i(m12,m13,m23) evaluated from a node record.
// Arithmetic follows /root/reference/src/PDFs/physics/resonances/{Resonance,RBW,GS,NonRes,Gauss}.cu
// operation for operation (same association order) so cached waves agree to the last few ulp.
#pragma once

struct gf_decay3 { // Amp3Body node consts: Amp3Body.cu:144-149
    double M, d1, d2, d3, rad, mrad;
};

enum { GF_PAIR_12 = 0, GF_PAIR_13 = 1, GF_PAIR_23 = 2 }; // DalitzPlotHelpers.h:81

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
    default: return make_double2(0.0, 0.0);
    }
}
