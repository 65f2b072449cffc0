// gf_amp4.cuh — four-body kinematics, lineshapes and spin factors evaluated from node records.
// Arithmetic follows the reference operation for operation (same association order):
//   get4Vecs / getmass        src/PDFs/physics/DalitzPlotHelpers.cu:61-276
//   BW / SBW / lass_MINT      src/PDFs/physics/lineshapes/BW.cu:14-148, Common.cu:8-51
//   bugg_MINT / bugg_MINT3    src/PDFs/physics/lineshapes/Bugg.cu:11-128
//   spin factors              src/PDFs/physics/SpinFactors.cu:83-568, detail/SpinHelper.h
#pragma once
#include "gf_kmatrix.cuh"
#include "gf_math.cuh"

enum gf_ls_kind { GF_LS_RBW = 0, GF_LS_SBW, GF_LS_LASS, GF_LS_BUGG, GF_LS_BUGG3, GF_LS_ONE, GF_LS_NONRES, GF_LS_FLATTE, GF_LS_GLASS, GF_LS_FOCUS, GF_LS_KMATRIX, GF_LS_GSPLINE };
// same order as the reference's SF_4Body enum (SpinFactors.h:18-35)
enum gf_sf_kind {
    GF_SF_DtoPP1_PtoSP2_StoP3P4 = 0, GF_SF_DtoPP1_PtoVP2_VtoP3P4, GF_SF_DtoV1V2_S, GF_SF_DtoV1V2_P, GF_SF_DtoV1V2_D,
    GF_SF_DtoAP1_AtoVP2_VtoP3P4, GF_SF_DtoAP1_AtoVP2Dwave_VtoP3P4, GF_SF_DtoVS_VtoP1P2_StoP3P4,
    GF_SF_DtoV1P1_V1toV2P2_V2toP3P4, GF_SF_DtoAP1_AtoSP2_StoP3P4, GF_SF_DtoTP1_TtoVP2_VtoP3P4,
    GF_SF_FF_12_34_L1, GF_SF_FF_12_34_L2, GF_SF_FF_123_4_L1, GF_SF_FF_123_4_L2, GF_SF_ONE
};

struct gf_cplx {
    double re, im;
};
__device__ inline gf_cplx gf_c(double r, double i) { return gf_cplx{r, i}; }
__device__ inline gf_cplx operator+(gf_cplx a, gf_cplx b) { return gf_c(a.re + b.re, a.im + b.im); }
__device__ inline gf_cplx operator-(gf_cplx a, gf_cplx b) { return gf_c(a.re - b.re, a.im - b.im); }
__device__ inline gf_cplx operator*(gf_cplx a, gf_cplx b) { return gf_c(a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re); }
__device__ inline gf_cplx operator*(double s, gf_cplx a) { return gf_c(s * a.re, s * a.im); }
__device__ inline gf_cplx operator*(gf_cplx a, double s) { return gf_c(a.re * s, a.im * s); }
// complex / complex as libcu++'s cuda::std::complex computes it for finite operands: it pre-scales the divisor by a
// power of two (exact), so the textbook form below yields the same bits.
__device__ inline gf_cplx operator/(gf_cplx a, gf_cplx b) {
    double denom = b.re * b.re + b.im * b.im;
    return gf_c((a.re * b.re + a.im * b.im) / denom, (a.im * b.re - a.re * b.im) / denom);
}
__device__ inline gf_cplx operator/(double a, gf_cplx b) { return gf_c(a, 0.0) / b; }

// ---- kinematics ----------------------------------------------------------------------------------
// vecs[16]: (px,py,pz,E) of the four daughters
__device__ inline void gf_get4vecs(double *V, double m12, double m34, double cos12, double cos34, double phi, double M,
                                   double m1, double m2, double m3, double m4) {
    double E1 = (m12 * m12 + m1 * m1 - m2 * m2) / (2 * m12);
    double E2 = (m12 * m12 - m1 * m1 + m2 * m2) / (2 * m12);
    double E3 = (m34 * m34 + m3 * m3 - m4 * m4) / (2 * m34);
    double E4 = (m34 * m34 - m3 * m3 + m4 * m4) / (2 * m34);
    double p1 = sqrt(E1 * E1 - m1 * m1);
    double p3 = sqrt(E3 * E3 - m3 * m3);
    double sin12 = sqrt(1 - cos12 * cos12);
    double sin34 = sqrt(1 - cos34 * cos34);
    double ED1 = (M * M + m12 * m12 - m34 * m34) / (2 * m12);
    double PD1 = sqrt(ED1 * ED1 - M * M);
    double beta1 = PD1 / ED1;
    double gamma1 = 1.0 / sqrt(1 - beta1 * beta1);
    double ED2 = (M * M - m12 * m12 + m34 * m34) / (2 * m34);
    double PD2 = sqrt(ED2 * ED2 - M * M);
    double beta2 = -PD2 / ED2;
    double gamma2 = 1.0 / sqrt(1 - beta2 * beta2);
    V[0] = cos12 * p1;  V[4] = -cos12 * p1; V[8] = -cos34 * p3; V[12] = cos34 * p3;
    V[1] = sin12 * p1;  V[5] = -sin12 * p1; V[9] = -sin34 * p3; V[13] = sin34 * p3;
    V[2] = 0; V[6] = 0; V[10] = 0; V[14] = 0;
    V[3] = E1; V[7] = E2; V[11] = E3; V[15] = E4;
    double tE = V[3], tX = V[0];
    V[3] = gamma1 * (tE + beta1 * tX);
    V[0] = gamma1 * (tX + beta1 * tE);
    tE = V[7]; tX = V[4];
    V[7] = gamma1 * (tE + beta1 * tX);
    V[4] = gamma1 * (tX + beta1 * tE);
    tE = V[11]; tX = V[8];
    V[11] = gamma2 * (tE + beta2 * tX);
    V[8]  = gamma2 * (tX + beta2 * tE);
    tE = V[15]; tX = V[12];
    V[15] = gamma2 * (tE + beta2 * tX);
    V[12] = gamma2 * (tX + beta2 * tE);
    double cosphi = cos(phi), sinphi = sin(phi);
    V[2] = sinphi * V[1];
    V[1] = cosphi * V[1];
    V[6] = sinphi * V[5];
    V[5] = cosphi * V[5];
}
__device__ inline double gf_mass2p(const double *a, const double *b) {
    return sqrt(-((a[0] + b[0]) * (a[0] + b[0])) - ((a[1] + b[1]) * (a[1] + b[1])) - ((a[2] + b[2]) * (a[2] + b[2]))
                + ((a[3] + b[3]) * (a[3] + b[3])));
}
__device__ inline double gf_mass3p(const double *a, const double *b, const double *c) {
    return sqrt(-((a[0] + b[0] + c[0]) * (a[0] + b[0] + c[0])) - ((a[1] + b[1] + c[1]) * (a[1] + b[1] + c[1]))
                - ((a[2] + b[2] + c[2]) * (a[2] + b[2] + c[2])) + ((a[3] + b[3] + c[3]) * (a[3] + b[3] + c[3])));
}
// DP4Pair: 0 M_12, 1 M_34, 2 M_13, 3 M_14, 4 M_23, 5 M_24, 6 M_12_3 ... 17 M_34_2 (DalitzPlotHelpers.h:56-76)
__device__ inline double gf_getmass(int pair, double &d1, double &d2, const double *v, double m1, double m2, double m3, double m4) {
    const double *P1 = v, *P2 = v + 4, *P3 = v + 8, *P4 = v + 12;
    switch(pair) {
    case 2: d1 = m1; d2 = m3; return gf_mass2p(P1, P3);
    case 3: d1 = m1; d2 = m4; return gf_mass2p(P1, P4);
    case 4: d1 = m2; d2 = m3; return gf_mass2p(P2, P3);
    case 5: d1 = m2; d2 = m4; return gf_mass2p(P2, P4);
    case 6: d1 = gf_mass2p(P1, P2); d2 = m3; return gf_mass3p(P1, P2, P3);
    case 7: d1 = gf_mass2p(P1, P3); d2 = m2; return gf_mass3p(P1, P2, P3);
    case 8: d1 = gf_mass2p(P2, P3); d2 = m1; return gf_mass3p(P1, P2, P3);
    case 9: d1 = gf_mass2p(P1, P2); d2 = m4; return gf_mass3p(P1, P2, P4);
    case 10: d1 = gf_mass2p(P1, P4); d2 = m2; return gf_mass3p(P1, P2, P4);
    case 11: d1 = gf_mass2p(P2, P4); d2 = m1; return gf_mass3p(P1, P2, P4);
    case 12: d1 = gf_mass2p(P1, P3); d2 = m4; return gf_mass3p(P1, P3, P4);
    case 13: d1 = gf_mass2p(P1, P4); d2 = m3; return gf_mass3p(P1, P3, P4);
    case 14: d1 = gf_mass2p(P3, P4); d2 = m1; return gf_mass3p(P1, P3, P4);
    case 15: d1 = gf_mass2p(P2, P3); d2 = m4; return gf_mass3p(P2, P3, P4);
    case 16: d1 = gf_mass2p(P2, P4); d2 = m3; return gf_mass3p(P2, P3, P4);
    case 17: d1 = gf_mass2p(P3, P4); d2 = m2; return gf_mass3p(P2, P3, P4);
    default: d1 = 0; d2 = 0; return 0;
    }
}

// constant daughters of a pair as gf_getmass reports them; for three-body pairs (>= 6) d1 is the event-dependent two-body
// sub-mass and is left untouched here
__device__ inline void gf_pair_daughters(int pair, double &d1, double &d2, double m1, double m2, double m3, double m4) {
    switch(pair) {
    case 2: d1 = m1; d2 = m3; break;
    case 3: d1 = m1; d2 = m4; break;
    case 4: d1 = m2; d2 = m3; break;
    case 5: d1 = m2; d2 = m4; break;
    case 6: case 13: case 16: d1 = 0; d2 = m3; break;
    case 7: case 10: case 17: d1 = 0; d2 = m2; break;
    case 8: case 11: case 14: d1 = 0; d2 = m1; break;
    case 9: case 12: case 15: d1 = 0; d2 = m4; break;
    default: d1 = 0; d2 = 0; break;
    }
}

// ---- lineshapes -------------------------------------------------------------------------------------
#define GF_MPIPLUS 0.139570
#define GF_MKPLUS 0.493677
#define GF_META 0.547862

__device__ inline double gf_bl_prime(double z2, double z02, int L) { // Common.cu:8-21
    if(L == 0) return 1.0;
    if(L == 1) return (1 + z02) / (1 + z2);
    if(L == 2) return (z02 * z02 + 3 * z02 + 9) / (z2 * z2 + 3 * z2 + 9);
    return 0;
}
__device__ inline double gf_bl(double z2, int L) { // Common.cu:23-36
    if(L == 0) return 1.0;
    if(L == 1) return 2 * z2 / (1 + z2);
    if(L == 2) return (13 * z2 * z2) / (z2 * z2 + 3 * z2 + 9);
    return 0;
}
__device__ inline double gf_bl2(double z2, int L) { // Common.cu:38-51
    if(L == 0) return 1.0;
    if(L == 1) return 1.0 / (1 + z2);
    if(L == 2) return 1.0 / (z2 * z2 + 3 * z2 + 9);
    return 0;
}
// Parameter-only part of BW / SBW / lass_MINT (BW.cu:19-27,62-63; :77-82,106-107; :127-131): what does not depend on the event
// while the pair's daughters are constants. The lineshape kernels evaluate it once per block for such pairs (three-body
// pairs have an event-dependent d1 and evaluate it per event).
struct gf_bw_inv {
    double mpsq, mmsq;
    double prSq;     // |num2 / (4 mass^2)|
    double prSq2;    // max(num2 / (4 mass^2), 0)
    double inv_prSq; // 1 / prSq
    double k;        // normalisation of the relativistic (RBW) or simple (SBW) Breit-Wigner
};
__device__ inline gf_bw_inv gf_bw_invariants(bool sbw, double m1, double m2, double mass, double width) {
    gf_bw_inv I;
    I.mpsq = (m1 + m2) * (m1 + m2);
    I.mmsq = (m1 - m2) * (m1 - m2);
    double num2 = (mass * mass - I.mpsq) * (mass * mass - I.mmsq);
    double pr = num2 / (4 * mass * mass);
    I.prSq2 = pr < 0 ? 0 : pr;
    I.prSq = fabs(pr);
    I.inv_prSq = 1.0 / I.prSq;
    if(sbw) {
        double gamma = sqrt(mass * mass * (mass * mass + width * width));
        I.k = mass * width * gamma / sqrt(mass * mass + gamma);
    } else {
        double gamma = mass * sqrt((mass * mass + width * width));
        I.k = (2.0 * sqrt(2.0) / M_PI) * mass * width * gamma / sqrt(mass * mass + gamma);
    }
    return I;
}
// Blatt-Weisskopf pieces of one event for orbital L, all from ONE reciprocal of the shared denominator D(z2) = 1 + z2 (L = 1)
// or z2^2 + 3 z2 + 9 (L = 2):  BL' = N(z02) / D(z2) (Common.cu:8-21), BL = 2 z2 / D or 13 z2^2 / D (:23-36), BL2 = 1 / D (:38-51)
struct gf_bl_set { double inv_d, z2; int L; };
__device__ __forceinline__ gf_bl_set gf_bl_prepare(double z2, int L) {
    gf_bl_set B;
    B.z2 = z2;
    B.L = L;
    B.inv_d = L == 0 ? 1.0 : gf_rcp(L == 1 ? 1 + z2 : z2 * z2 + 3 * z2 + 9);
    return B;
}
__device__ __forceinline__ double gf_bl_prime_s(const gf_bl_set &B, double z02) {
    if(B.L == 0) return 1.0;
    if(B.L == 1) return (1 + z02) * B.inv_d;
    if(B.L == 2) return (z02 * z02 + 3 * z02 + 9) * B.inv_d;
    return 0;
}
__device__ __forceinline__ double gf_bl_s(const gf_bl_set &B) {
    if(B.L == 0) return 1.0;
    if(B.L == 1) return 2 * B.z2 * B.inv_d;
    if(B.L == 2) return (13 * B.z2 * B.z2) * B.inv_d;
    return 0;
}
__device__ __forceinline__ double gf_bl2_s(const gf_bl_set &B) { return B.L <= 2 ? (B.L == 0 ? 1.0 : B.inv_d) : 0; }
// BW.cu:14-69. Against the reference's operation sequence the per-event part shares reciprocals: 1/Mpair serves num / (4 M^2)
// and mass / Mpair, 1 / prSq is parameter-only, and the Blatt-Weisskopf ratios share the reciprocal of their common
// denominator: three reciprocals (branch-free, gf_math.cuh) and two square roots per event instead of six divisions and two. Each substitution moves the result by
// an ulp or two (the amplitudes are compared with the reference at 1e-10, the NLL at 1e-12: tests/test_parity_gpu.py).
__device__ inline gf_cplx gf_ls_bw_pt(double Mpair, double mass, double width, int orbital, int FF, double meson_radius, const gf_bw_inv &I) {
    const int to2Lplus1 = 2 * orbital + 1;
    double M2 = Mpair * Mpair;
    double num = (M2 - I.mpsq) * (M2 - I.mmsq);
    double rM = gf_rcp(Mpair);
    double pABSq = num * (0.25 * (rM * rM));
    double pratio = sqrt(pABSq * I.inv_prSq);
    double pr2J1 = 1;
    for(int i = 0; i < to2Lplus1; i++) pr2J1 *= pratio;
    double mratio = mass * rM;
    double r2 = meson_radius * meson_radius;
    const gf_bl_set B = gf_bl_prepare(pABSq * r2, orbital);
    double thisFR = gf_bl_prime_s(B, I.prSq * r2);
    double frFactor = 1;
    if(0 != orbital && 0 != FF) {
        frFactor = (FF == 1 ? gf_bl_s(B) : gf_bl_prime_s(B, I.prSq2 * r2));
        frFactor = (FF == 3 ? gf_bl2_s(B) : frFactor);
    }
    double GofM = width * pr2J1 * mratio * thisFR;
    gf_cplx bw = gf_c(mass * mass - M2, mass * GofM);
    double den = (mass * mass - M2) * (mass * mass - M2) + mass * GofM * mass * GofM;
    return ((sqrt(I.k * frFactor)) * gf_rcp(den)) * bw;
}
// gf_ls_bw_pt with the orbital momentum known at compile time (k_amp4_update dispatches on it once per lineshape, uniformly over
// the grid): the same operations in the same order, without the loop over 2L+1 factors and the selects on L
template <int L>
__device__ __forceinline__ gf_cplx gf_ls_bw_pt_t(double Mpair, double mass, double width, int FF, double r2, const gf_bw_inv &I) {
    const double M2 = Mpair * Mpair;
    const double num = (M2 - I.mpsq) * (M2 - I.mmsq);
    const double rM = gf_rcp(Mpair);
    const double pABSq = num * (0.25 * (rM * rM));
    const double pratio = sqrt(pABSq * I.inv_prSq);
    double pr2J1 = pratio;
    if(L >= 1) pr2J1 = (pr2J1 * pratio) * pratio;
    if(L >= 2) pr2J1 = (pr2J1 * pratio) * pratio;
    const double mratio = mass * rM;
    const double z2 = pABSq * r2;
    const double inv_d = L == 0 ? 1.0 : gf_rcp(L == 1 ? 1 + z2 : z2 * z2 + 3 * z2 + 9);
    auto prime = [&](double z02) { return L == 0 ? 1.0 : (L == 1 ? (1 + z02) * inv_d : (z02 * z02 + 3 * z02 + 9) * inv_d); };
    const double thisFR = prime(I.prSq * r2);
    double frFactor = 1;
    if(L != 0 && FF != 0) {
        frFactor = FF == 1 ? (L == 1 ? 2 * z2 * inv_d : (13 * z2 * z2) * inv_d) : prime(I.prSq2 * r2);
        frFactor = FF == 3 ? inv_d : frFactor;
    }
    const double GofM = width * pr2J1 * mratio * thisFR;
    const double dm = mass * mass - M2;
    const gf_cplx bw = gf_c(dm, mass * GofM);
    const double den = dm * dm + mass * GofM * mass * GofM;
    return ((sqrt(I.k * frFactor)) * gf_rcp(den)) * bw;
}
__device__ inline gf_cplx gf_ls_bw(double Mpair, double m1, double m2, double mass, double width, int orbital, int FF,
                                   double meson_radius) {
    return gf_ls_bw_pt(Mpair, mass, width, orbital, FF, meson_radius, gf_bw_invariants(false, m1, m2, mass, width));
}
// BW.cu:72-115
__device__ inline gf_cplx gf_ls_sbw_pt(double Mpair, double mass, double width, int orbital, int FF, double meson_radius, const gf_bw_inv &I) {
    double M2 = Mpair * Mpair;
    double frFactor = 1;
    if(0 != orbital && 0 != FF) {
        double num = (M2 - I.mpsq) * (M2 - I.mmsq);
        double pABSq = num * (0.25 * gf_rcp(M2));
        double r2 = meson_radius * meson_radius;
        const gf_bl_set B = gf_bl_prepare(pABSq * r2, orbital);
        frFactor = (FF == 1 ? gf_bl_s(B) : gf_bl_prime_s(B, I.prSq2 * r2));
        frFactor = (FF == 3 ? gf_bl2_s(B) : frFactor);
    }
    double GofM = width;
    gf_cplx bw = gf_c(mass * mass - M2, mass * GofM);
    double den = (mass * mass - M2) * (mass * mass - M2) + mass * GofM * mass * GofM;
    return ((sqrt(I.k * frFactor)) * gf_rcp(den)) * bw;
}
__device__ inline gf_cplx gf_ls_sbw(double Mpair, double m1, double m2, double mass, double width, int orbital, int FF,
                                    double meson_radius) {
    return gf_ls_sbw_pt(Mpair, mass, width, orbital, FF, meson_radius, gf_bw_invariants(true, m1, m2, mass, width));
}
// BW.cu:118-148
__device__ inline gf_cplx gf_ls_lass_pt(double Mpair, double resmass, double reswidth, int orbital, int FF, double meson_radius,
                                        const gf_bw_inv &I) {
    double rMass2 = Mpair * Mpair;
    double a = 2.07, r = 3.32;
    double num = (rMass2 - I.mpsq) * (rMass2 - I.mmsq);
    double pABSq = num / (4 * rMass2);
    double prSq = I.prSq;
    double y = 2.0 * a * sqrt(pABSq);
    double x = 2.0 + a * r * pABSq;
    double cot = x / y;
    gf_cplx phaseshift = gf_c((cot * cot - 1) / (1 + cot * cot), 2 * cot / (1 + cot * cot));
    gf_cplx den = gf_c(sqrt(pABSq) * cot, (-1.) * sqrt(pABSq));
    double SF = Mpair * sqrt(prSq) / (resmass * resmass * reswidth);
    gf_cplx BG = SF / den;
    return BG + phaseshift * gf_ls_bw_pt(Mpair, resmass, reswidth, orbital, FF, meson_radius, I);
}
__device__ inline gf_cplx gf_ls_lass(double Mpair, double m1, double m2, double resmass, double reswidth, int orbital,
                                     int FF, double meson_radius) {
    return gf_ls_lass_pt(Mpair, resmass, reswidth, orbital, FF, meson_radius, gf_bw_invariants(false, m1, m2, resmass, reswidth));
}
// Bugg.cu:11-36
__device__ inline gf_cplx gf_bugg_rho2(double s, double m) {
    double rho_squared = 1. - 4. * m * m / s;
    gf_cplx rv = (rho_squared >= 0) ? gf_c(1, 0) : gf_c(0, 1);
    rho_squared = (rho_squared >= 0) ? sqrt(rho_squared) : sqrt(-rho_squared);
    return rho_squared * rv;
}
__device__ inline double gf_bugg_j1(double s, double m) {
    double rho_pipi = gf_bugg_rho2(s, m).re;
    double rv = 2.;
    rv += (rho_pipi > 0.) ? rho_pipi * log((1. - rho_pipi) / (1. + rho_pipi)) : 0;
    return rv / M_PI;
}
__device__ inline double gf_bugg_gamma4pi(double s, double mpi, double g_4pi, double M, double lambda_4pi, double s0_4pi) {
    return (s < (16. * mpi * mpi)) ? 0
                                   : g_4pi * (1. / (1 + exp(lambda_4pi * (s0_4pi - s))))
                                         / (1. / (1 + exp(lambda_4pi * (s0_4pi - M * M))));
}
// Bugg.cu:39-84 (mass/width constructor arguments are ignored by the reference)
__device__ inline gf_cplx gf_ls_bugg(double Mpair) {
    double s = Mpair * Mpair;
    double M = 0.953, b1 = 1.302, b2 = 0.340, A = 2.426, g_4pi = 0.011, g_2K = 0.6, g_2eta = 0.2, alpha = 1.3, sA = 0.41;
    double s0_4pi = 7.082 / 2.845, lambda_4pi = 2.845;
    double g1sq = (b1 + b2 * s) * exp(-(s - M * M) / A);
    double z = gf_bugg_j1(s, GF_MPIPLUS) - gf_bugg_j1(M * M, GF_MPIPLUS);
    gf_cplx gamma_2pi = gf_c(g1sq * (s - sA * GF_MPIPLUS * GF_MPIPLUS) / (M * M - sA * GF_MPIPLUS * GF_MPIPLUS)
                                 * gf_bugg_rho2(s, GF_MPIPLUS).re, 0);
    gf_cplx gamma_2K = (g_2K * g1sq * s / (M * M)
                        * exp((-1) * alpha * sqrt((s - 4. * GF_MKPLUS * GF_MKPLUS) * (s - 4. * GF_MKPLUS * GF_MKPLUS))))
                       * gf_bugg_rho2(s, GF_MKPLUS);
    gf_cplx gamma_2eta = (g_2eta * g1sq * s / (M * M)
                          * exp((-1) * alpha * sqrt((s - 4. * GF_META * GF_META) * (s - 4. * GF_META * GF_META))))
                         * gf_bugg_rho2(s, GF_META);
    gf_cplx gamma_4pi = gf_c(gf_bugg_gamma4pi(s, GF_MPIPLUS, g_4pi, M, lambda_4pi, s0_4pi), 0);
    gf_cplx Gamma_tot = gamma_2pi + gamma_2K + gamma_2eta + gamma_4pi;
    gf_cplx den = gf_c(M * M - s - M * g1sq * (s - sA * GF_MPIPLUS * GF_MPIPLUS) / (M * M - sA * GF_MPIPLUS * GF_MPIPLUS) * z, 0)
                  - (gf_c(0, 1) * M) * Gamma_tot;
    gf_cplx rv = 1.0 / den;
    return rv * sqrt(1000.0);
}
// Bugg.cu:87-128
__device__ inline gf_cplx gf_ls_bugg3(double Mpair) {
    double s = Mpair * Mpair;
    double M = 0.953, b1 = 1.302, b2 = 0.340, A = 2.426, g_4pi = 0.011, g_2K = 0.6, g_2eta = 0.2, alpha = 1.3;
    double s0_4pi = 7.082 / 2.845, lambda_4pi = 2.845;
    double sA = 0.41 * GF_MPIPLUS * GF_MPIPLUS;
    double g1sq = M * (b1 + b2 * s) * exp(-(s - M * M) / A);
    double z = gf_bugg_j1(s, GF_MPIPLUS) - gf_bugg_j1(M * M, GF_MPIPLUS);
    double adlerZero = (s - sA) / (M * M - sA);
    double mk4 = 4. * GF_MKPLUS * GF_MKPLUS, me4 = 4. * GF_META * GF_META;
    double tmp1 = s > mk4 ? s - mk4 : mk4 - s;
    double tmp2 = s > me4 ? s - me4 : me4 - s;
    gf_cplx gamma_2pi = gf_c(g1sq * adlerZero * gf_bugg_rho2(s, GF_MPIPLUS).re, 0);
    gf_cplx gamma_2K = (g_2K * g1sq * s / (M * M) * exp((-1) * alpha * tmp1)) * gf_bugg_rho2(s, GF_MKPLUS);
    gf_cplx gamma_2eta = (g_2eta * g1sq * s / (M * M) * exp((-1) * alpha * tmp2)) * gf_bugg_rho2(s, GF_META);
    gf_cplx gamma_4pi = gf_c(gf_bugg_gamma4pi(s, GF_MPIPLUS, g_4pi, M, lambda_4pi, s0_4pi), 0);
    gf_cplx Gamma_tot = gamma_2pi + gamma_2K + gamma_2eta + gamma_4pi;
    gf_cplx den = gf_c(M * M - s - adlerZero * g1sq * z, 0) - gf_c(0, 1) * Gamma_tot;
    return 1.0 / den;
}
// NonRes.cu:8-25 : the Blatt-Weisskopf barrier alone
__device__ inline gf_cplx gf_ls_nonres(double Mpair, double m1, double m2, int orbital, double meson_radius) {
    double M2 = Mpair * Mpair;
    double mpsq = (m1 + m2) * (m1 + m2), mmsq = (m1 - m2) * (m1 - m2);
    double pABSq = (M2 - mpsq) * (M2 - mmsq) / (4 * M2);
    return gf_c(sqrt(gf_bl2(pABSq * meson_radius * meson_radius, orbital)), 0.0);
}
// Flatte.cu:8-12 : sqrt(1 - (2 m0/m)^2), continued below threshold
__device__ inline gf_cplx gf_flatte_asqrt(double m0, double m) {
    double a2 = 1 - (2 * m0 / m) * (2 * m0 / m);
    return a2 > 0 ? gf_c(sqrt(a2), 0) : gf_c(0, sqrt(-a2));
}
// Flatte.cu:14-55 : f0(980) with the couplings hard-wired upstream (gPi 0.165, gK/gPi 4.21)
__device__ inline gf_cplx gf_ls_flatte(double Mpair, double m1, double m2, double resmass, int orbital, double meson_radius) {
    double rMass2 = Mpair * Mpair;
    const double gPi = .165, gK = gPi * 4.21, mPi0 = .1349766, mK0 = .497648;
    double mpsq = (m1 + m2) * (m1 + m2), mmsq = (m1 - m2) * (m1 - m2);
    double pABSq = (rMass2 - mpsq) * (rMass2 - mmsq) / (4 * rMass2);
    gf_cplx Gpipi = (1. / 3.) * gf_flatte_asqrt(mPi0, Mpair) + (2. / 3.) * gf_flatte_asqrt(GF_MPIPLUS, Mpair);
    gf_cplx GKK = (1. / 2.) * gf_flatte_asqrt(mK0, Mpair) + (1. / 2.) * gf_flatte_asqrt(GF_MKPLUS, Mpair);
    gf_cplx width = gPi * Gpipi + gK * GKK;
    double fr = gf_bl2(pABSq * meson_radius * meson_radius, orbital);
    return sqrt(fr) / (gf_c(resmass * resmass - rMass2, 0) - gf_c(0, resmass) * width);
}
// GLASS.cu:12-60 : generalised LASS; params mass,width,a,r,phiF,phiR,F (R fixed to 1 upstream)
__device__ inline gf_cplx gf_ls_glass(double Mpair, double m1, double m2, const double *p, int orbital, double meson_radius) {
    double resmass = p[0], reswidth = p[1], a = p[2], r = p[3], phiF = p[4], phiR = p[5], F = p[6];
    double rMass2 = Mpair * Mpair;
    double mpsq = (m1 + m2) * (m1 + m2), mmsq = (m1 - m2) * (m1 - m2);
    double num = (rMass2 - mpsq) * (rMass2 - mmsq);
    double num2 = (resmass * resmass - mpsq) * (resmass * resmass - mmsq);
    double pABSq = num / (4 * rMass2);
    double prSq = fabs(num2 / (4 * resmass * resmass));
    double pratio = sqrt(pABSq / prSq);
    double pr2J1 = 1;
    for(int i = 0; i < 2 * orbital + 1; i++) pr2J1 *= pratio;
    double mratio = resmass / Mpair;
    double r2 = meson_radius * meson_radius;
    double thisFR = gf_bl_prime(pABSq * r2, prSq * r2, orbital);
    double GofM = reswidth * pr2J1 * mratio * thisFR;
    double y = 2.0 * a * sqrt(pABSq);
    double x = 2.0 + a * r * pABSq;
    double scattphase = phiF + atan(y / x);
    double resphase = phiR + atan(resmass * GofM / (resmass * resmass - rMass2));
    double rho = 1.0 / sqrt(pABSq / rMass2);
    double ss, cs, s2, c2;
    sincos(scattphase, &ss, &cs);
    sincos(resphase + 2 * scattphase, &s2, &c2);
    return ((F * ss) * gf_c(cs, ss) + (1.0 * sin(resphase)) * gf_c(c2, s2)) * rho;
}
// FOCUS.cu:12-17 : two-body phase space, continued below threshold (principal square root of a real number)
__device__ inline gf_cplx gf_focus_phsp(double s, double m0, double m1) {
    double mp = m0 + m1, mm = m0 - m1;
    double a2 = (1.0 - mp * mp / s) * (1.0 - mm * mm / s);
    return a2 >= 0 ? gf_c(sqrt(a2), 0) : gf_c(0, sqrt(-a2));
}
// FOCUS.cu:19-78 : K pi S-wave K-matrix; mod 0 = Kpi, 1 = KEta, anything else = I 3/2
#define GF_METAP 0.96778
__device__ inline gf_cplx gf_ls_focus(double Mpair, int mod) {
    const gf_cplx I = gf_c(0, 1);
    double s = Mpair * Mpair;
    const double sNorm = GF_MKPLUS * GF_MKPLUS + GF_MPIPLUS * GF_MPIPLUS, s12 = 0.23, s32 = 0.27;
    double I12 = (s - s12) / sNorm, I32 = (s - s32) / sNorm;
    gf_cplx rho1 = gf_focus_phsp(s, GF_MKPLUS, GF_MPIPLUS), rho2 = gf_focus_phsp(s, GF_MKPLUS, GF_METAP);
    const double pmass = 1.7919, g0 = 0.31072, g1 = -0.02323;
    double X = s / sNorm - 1;
    double k00 = g0 * g0 / (pmass - s) + 0.79299 - 0.15099 * X + 0.00811 * (X * X);
    double k11 = g1 * g1 / (pmass - s) + 0.17054 - 0.0219 * X + 0.00085655 * (X * X);
    double k01 = g0 * g1 / (pmass - s) + 0.15040 - 0.038266 * X + 0.0022596 * (X * X);
    double K11 = I12 * k00, K12 = I12 * k01, K22 = I12 * k11;
    double K32 = I32 * (-0.22147 + 0.026637 * X - 0.00092057 * (X * X));
    double detK = K11 * K22 - K12 * K12;
    gf_cplx del = gf_c(1., 0) - rho1 * rho2 * detK - I * (rho1 * K11 + rho2 * K22);
    if(mod == 0) return (gf_c(K11, 0) - I * rho2 * detK) / del;
    if(mod == 1) return K12 / del;
    return 1. / (gf_c(1., 0) - I * rho2 * K32);
}
// which kinds have a parameter-only part worth keeping per block (gf_bw_inv), and whether it is the SBW form
__device__ __host__ inline bool gf_ls_has_inv(int kind) { return kind == GF_LS_RBW || kind == GF_LS_SBW || kind == GF_LS_LASS; }
// as gf_lineshape below for the kinds of gf_ls_has_inv, with the parameter-only part supplied by the caller
__device__ inline gf_cplx gf_lineshape_inv(int kind, const double *p, const double *c, double Mpair, const gf_bw_inv &I) {
    switch(kind) {
    case GF_LS_RBW: return gf_ls_bw_pt(Mpair, p[0], p[1], (int)c[1], (int)c[2], c[3], I);
    case GF_LS_SBW: return gf_ls_sbw_pt(Mpair, p[0], p[1], (int)c[1], (int)c[2], c[3], I);
    default: return gf_ls_lass_pt(Mpair, p[0], p[1], (int)c[1], (int)c[2], c[3], I);
    }
}
// GSpline.cu:31-95 (AmpGen's GSpline.EFF): a Breit-Wigner whose running width follows a cubic spline in s.
// consts = Mpair, L, FF, radius, s_min, s_max, nBins, n, then the n constants the constructor registers after them (the
// curvature terms). As upstream, getSpline reads the knot values from those nBins constants and the curvatures from the nBins AFTER them;
// upstream registers only nBins constants there, so its second read runs past its own block into whatever follows in the flat
// constant table. Here a read past the record's n constants yields 0.
__device__ inline double gf_gspline_at(double x, bool continued, const double *c) {
    const double s_min = c[4], s_max = c[5];
    const int nBins = (int)c[6], n = (int)c[7], start = 8;
    auto C = [&](int k) { return (k >= start && k < start + n) ? c[k] : 0.0; };
    if(x <= s_min) return continued ? C(start) : 0;
    if(x >= s_max) return continued ? C(start + nBins - 1) : 0;
    const double spacing = (s_max - s_min) / (nBins - 1.);
    const double dx = fmod((x - s_min), spacing);
    const int bin = (int)((x - s_min) / spacing);
    const double m_x_0 = C(start + bin), m_x_1 = C(start + bin + 1), m_xf_0 = C(start + bin + nBins), m_xf_1 = C(start + bin + nBins + 1);
    return m_x_0 + dx * ((m_x_1 - m_x_0) / spacing - (m_xf_1 + 2 * m_xf_0) * spacing / 6) + dx * dx * m_xf_0 + dx * dx * dx * (m_xf_1 - m_xf_0) / (6 * spacing);
}
__device__ inline gf_cplx gf_ls_gspline(double Mpair, double m1, double m2, const double *p, const double *c) {
    const double mass = p[0], width = p[1], radius = c[3];
    const double s = Mpair * Mpair, s1 = m1 * m1, s2 = m2 * m2;
    const double q2 = fabs(s / 4. - (s1 + s2) / 2. + (s1 - s2) * (s1 - s2) / (4 * s));
    const double BF = exp(-q2 * (radius * radius) / 2);
    const double width_shape = width * gf_gspline_at(s, true, c);
    const double width_norm = width * gf_gspline_at(mass * mass, false, c);
    const double gamma = mass * sqrt(mass * mass + width * width);
    const double k = 2 * sqrt(2.) * mass * width * gamma / (M_PI * sqrt(mass * mass + gamma));
    const double norm = sqrt(k) * BF;
    const double running_width = width_norm == 0 ? 0 : width * width_shape / width_norm;
    return norm / gf_c(mass * mass - s, -mass * running_width);
}
// dispatch on the lineshape record: params mass,width ; consts Mpair,L,FF,radius ; ints[0] kind
__device__ inline gf_cplx gf_lineshape(int kind, const double *p, const double *c, double Mpair, double d1, double d2) {
    switch(kind) {
    case GF_LS_RBW: return gf_ls_bw(Mpair, d1, d2, p[0], p[1], (int)c[1], (int)c[2], c[3]);
    case GF_LS_SBW: return gf_ls_sbw(Mpair, d1, d2, p[0], p[1], (int)c[1], (int)c[2], c[3]);
    case GF_LS_LASS: return gf_ls_lass(Mpair, d1, d2, p[0], p[1], (int)c[1], (int)c[2], c[3]);
    case GF_LS_BUGG: return gf_ls_bugg(Mpair);
    case GF_LS_BUGG3: return gf_ls_bugg3(Mpair);
    case GF_LS_NONRES: return gf_ls_nonres(Mpair, d1, d2, (int)c[1], c[3]);
    case GF_LS_FLATTE: return gf_ls_flatte(Mpair, d1, d2, p[0], (int)c[1], c[3]);
    case GF_LS_GLASS: return gf_ls_glass(Mpair, d1, d2, p, (int)c[1], c[3]);
    case GF_LS_FOCUS: return gf_ls_focus(Mpair, (int)c[4]);
    case GF_LS_KMATRIX: { // lineshapes/kMatrix.cu:18-119
        const gf_kc r = gf_kmatrix(Mpair * Mpair, p);
        return gf_c(r.re, r.im);
    }
    case GF_LS_GSPLINE: return gf_ls_gspline(Mpair, d1, d2, p, c);
    default: return gf_c(1.0, 0.0);
    }
}

// ---- Lorentz helpers (detail/SpinHelper.h) ---------------------------------------------------------------
struct gf_lv {
    double x, y, z, e;
};
__device__ inline gf_lv gf_v(double x, double y, double z, double e) { return gf_lv{x, y, z, e}; }
__device__ inline gf_lv operator+(gf_lv a, gf_lv b) { return gf_v(a.x + b.x, a.y + b.y, a.z + b.z, a.e + b.e); }
__device__ inline gf_lv operator-(gf_lv a, gf_lv b) { return gf_v(a.x - b.x, a.y - b.y, a.z - b.z, a.e - b.e); }
__device__ inline gf_lv operator*(gf_lv a, double s) { return gf_v(a.x * s, a.y * s, a.z * s, a.e * s); }
__device__ inline gf_lv operator*(double s, gf_lv a) { return gf_v(a.x * s, a.y * s, a.z * s, a.e * s); }
__device__ inline double gf_dot(gf_lv a, gf_lv b) { return a.e * b.e - a.x * b.x - a.y * b.y - a.z * b.z; }
__device__ inline double gf_mag2(gf_lv a) { return gf_dot(a, a); }
__device__ inline double gf_m(gf_lv a) { return sqrt(gf_mag2(a)); }
// Daughter p (uniform over the grid) of the four four-vectors. Written as selects over compile-time offsets: a runtime-indexed
// read would force the sixteen components into local memory for the whole spin-factor kernel (4x10^7 local loads per 10^7
// events in the round-1 profile); four predicated moves per component keep them in registers.
__device__ __forceinline__ gf_lv gf_load_lv(const double *V, int p) {
    return gf_v(p == 0 ? V[0] : (p == 1 ? V[4] : (p == 2 ? V[8] : V[12])), p == 0 ? V[1] : (p == 1 ? V[5] : (p == 2 ? V[9] : V[13])),
                p == 0 ? V[2] : (p == 1 ? V[6] : (p == 2 ? V[10] : V[14])), p == 0 ? V[3] : (p == 1 ? V[7] : (p == 2 ? V[11] : V[15])));
}
// ZTspin1: q - (q.p) p / mR^2   (SpinHelper.h:77-84)
__device__ inline gf_lv gf_zt1(gf_lv q, gf_lv p, double mR) { return q - gf_dot(q, p) * p * (1. / (mR * mR)); }
// SpinSumV::Dot (SpinHelper.h:86-97)
__device__ inline gf_lv gf_ssv_dot(gf_lv p, double mR, gf_lv rhs) { return -1.0 * rhs + p * (gf_dot(p, rhs) / (mR * mR)); }
__device__ inline double gf_ssv_sandwich(gf_lv p, double mR, gf_lv l, gf_lv r) { return gf_dot(l, gf_ssv_dot(p, mR, r)); }
// SpinSumT::Sandwich (SpinHelper.h:99-113)
__device__ inline double gf_sst_sandwich(gf_lv p, double mR, gf_lv lm, gf_lv ln, gf_lv ra, gf_lv rb) {
    double manb = gf_ssv_sandwich(p, mR, lm, ra) * gf_ssv_sandwich(p, mR, ln, rb);
    double mbna = gf_ssv_sandwich(p, mR, lm, rb) * gf_ssv_sandwich(p, mR, ln, ra);
    double mnab = gf_ssv_sandwich(p, mR, lm, ln) * gf_ssv_sandwich(p, mR, ra, rb);
    return (1. / 2.) * (manb + mbna) - (1. / 3.) * mnab;
}
struct gf_lm { // rows X, Y, Z, E (SpinHelper.h:115-208)
    gf_lv r[4];
};
__device__ inline gf_lm gf_outer(gf_lv p) { // SymmLorentzMatrix(const gpuLVec p), SpinHelper.h:250-270
    gf_lm m;
    m.r[0] = gf_v(p.x * p.x, p.x * p.y, p.x * p.z, p.x * p.e);
    m.r[1] = gf_v(p.y * p.x, p.y * p.y, p.y * p.z, p.y * p.e);
    m.r[2] = gf_v(p.z * p.x, p.z * p.y, p.z * p.z, p.z * p.e);
    m.r[3] = gf_v(p.e * p.x, p.e * p.y, p.e * p.z, p.e * p.e);
    return m;
}
// ZTspin2 (SpinHelper.h:383-404): tt - (1/3) t.Mag2() (g - uu/mR^2)
__device__ inline gf_lm gf_zt2(gf_lv q, gf_lv p, double mR) {
    gf_lv t = gf_zt1(q, p, mR);
    gf_lm tt = gf_outer(t), uu = gf_outer(p), g, out;
    double inv = 1. / (mR * mR);
    for(int i = 0; i < 4; i++) uu.r[i] = uu.r[i] * inv; // uu /= mR^2 multiplies by the reciprocal
    g.r[0] = gf_v(-1, 0, 0, 0);
    g.r[1] = gf_v(0, -1, 0, 0);
    g.r[2] = gf_v(0, 0, -1, 0);
    g.r[3] = gf_v(0, 0, 0, 1);
    double f = (1. / 3.) * gf_mag2(t);
    for(int i = 0; i < 4; i++) out.r[i] = tt.r[i] - (g.r[i] - uu.r[i]) * f;
    return out;
}
// SymmLorentzMatrix::Contract (SpinHelper.h:292-296)
__device__ inline gf_lv gf_contract(const gf_lm &m, gf_lv v) { return v.e * m.r[3] - v.x * m.r[0] - v.y * m.r[1] - v.z * m.r[2]; }
// LeviCivita with three vectors (SpinHelper.h:424-442)
__device__ inline gf_lv gf_levi3(gf_lv a, gf_lv b, gf_lv c) {
    gf_lv v;
    v.e = -1. * a.x * (b.y * c.z - b.z * c.y) - a.y * (b.z * c.x - b.x * c.z) - a.z * (b.x * c.y - b.y * c.x);
    v.z = a.x * (b.e * c.y - b.y * c.e) + a.y * (b.x * c.e - b.e * c.x) + a.e * (b.y * c.x - b.x * c.y);
    v.y = a.x * (b.z * c.e - b.e * c.z) + a.z * (b.e * c.x - b.x * c.e) + a.e * (b.x * c.z - b.z * c.x);
    v.x = a.y * (b.e * c.z - b.z * c.e) + a.z * (b.y * c.e - b.e * c.y) + a.e * (b.z * c.y - b.y * c.z);
    return v;
}

// ---- spin factors (SpinFactors.cu:83-568, ZEMACH branch) ---------------------------------------------------
// consts: mD0 (or mother radius for the FF_* kinds), P0..P3 = which daughter plays which role
__device__ inline double gf_spinfactor(int kind, const double *c, const double *V) {
    if(kind == GF_SF_ONE || kind == GF_SF_DtoPP1_PtoSP2_StoP3P4) return 1.0;
    gf_lv P1 = gf_load_lv(V, (int)c[1]), P2 = gf_load_lv(V, (int)c[2]), P3 = gf_load_lv(V, (int)c[3]), P4 = gf_load_lv(V, (int)c[4]);
    switch(kind) {
    case GF_SF_FF_12_34_L1:
    case GF_SF_FF_12_34_L2:
    case GF_SF_FF_123_4_L1:
    case GF_SF_FF_123_4_L2: {
        bool two = (kind == GF_SF_FF_12_34_L1 || kind == GF_SF_FF_12_34_L2);
        double m1 = two ? gf_mag2(P1 + P2) : gf_mag2(P1 + P2 + P3);
        double m2 = two ? gf_mag2(P3 + P4) : gf_mag2(P4);
        double s = gf_mag2(P1 + P2 + P3 + P4);
        double q2 = s / 4. - (m1 + m2) / 2. + (m1 - m2) * (m1 - m2) / (4 * s);
        double z2 = q2 * c[0] * c[0];
        bool l1 = (kind == GF_SF_FF_12_34_L1 || kind == GF_SF_FF_123_4_L1);
        double ff = l1 ? 1.0 / (1 + z2) : 1.0 / (z2 * z2 + 3 * z2 + 9);
        return sqrt(ff);
    }
    case GF_SF_DtoPP1_PtoVP2_VtoP3P4: {
        gf_lv pV = P3 + P4, qV = P3 - P4, pP = pV + P2, qP = pV - P2;
        gf_lv LP = gf_zt1(qP, pP, gf_m(pP)), LV = gf_zt1(qV, pV, gf_m(pV));
        return gf_dot(LP, LV);
    }
    case GF_SF_DtoV1V2_S: {
        gf_lv pV1 = P1 + P2, qV1 = P1 - P2, pV2 = P3 + P4, qV2 = P3 - P4;
        double MV1 = sqrt(gf_dot(pV1, pV1)), MV2 = sqrt(gf_dot(pV2, pV2));
        return (gf_dot(qV1, qV2) - gf_dot(qV1, pV1) * gf_dot(pV1, qV2) / (MV1 * MV1)
                - gf_dot(qV1, pV2) * gf_dot(pV2, qV2) / (MV2 * MV2)
                + gf_dot(qV1, pV1) * gf_dot(pV1, pV2) * gf_dot(pV2, qV2) / (MV1 * MV1 * MV2 * MV2));
    }
    case GF_SF_DtoV1V2_P: {
        gf_lv pV1 = P1 + P2, qV1 = P2 - P1, pV2 = P3 + P4, qV2 = P4 - P3, pD = pV1 + pV2, qD = pV1 - pV2;
        double MV1 = sqrt(gf_dot(pV1, pV1)), MV2 = sqrt(gf_dot(pV2, pV2));
        gf_lv LD = gf_zt1(qD, pD, gf_m(pD)), LV1 = gf_zt1(qV1, pV1, MV1), LV2 = gf_zt1(qV2, pV2, MV2);
        return -gf_dot(gf_levi3(pD, LD, LV1), LV2);
    }
    case GF_SF_DtoV1V2_D: {
        gf_lv pV1 = P1 + P2, qV1 = P1 - P2, pV2 = P3 + P4, qV2 = P3 - P4, pD = pV1 + pV2, qD = pV1 - pV2;
        double mD = gf_m(pD);
        gf_lv tV1 = gf_zt1(qV1, pV1, gf_m(pV1)), tV2 = gf_zt1(qV2, pV2, gf_m(pV2));
        gf_lm tD = gf_zt2(qD, pD, mD);
        return gf_dot(tV1, gf_contract(tD, tV2));
    }
    case GF_SF_DtoV1P1_V1toV2P2_V2toP3P4: {
        gf_lv pV1 = P2 + P3 + P4, pV2 = P3 + P4, qV1 = (P3 + P4) - P2, qV2 = P3 - P4;
        double MV1 = gf_m(pV1), MV2 = gf_m(pV2);
        gf_lv pD = pV1 + P1, qD = pV1 - P1;
        gf_lv LD = gf_zt1(qD, pD, gf_m(pD)), LV1 = gf_zt1(qV1, pV1, MV1), LV2 = gf_zt1(qV2, pV2, MV2);
        gf_lv tmp = gf_ssv_dot(pV1, MV1, LD);
        return gf_dot(gf_levi3(pV1, LV1, LV2), tmp);
    }
    case GF_SF_DtoVS_VtoP1P2_StoP3P4: {
        gf_lv pS = P3 + P4, pV = P1 + P2, qV = P1 - P2, pD = pV + pS, qD = pV - pS;
        gf_lv LD = gf_zt1(qD, pD, gf_m(pD)), LV = gf_zt1(qV, pV, gf_m(pV));
        return gf_dot(LD, LV);
    }
    case GF_SF_DtoAP1_AtoSP2_StoP3P4: {
        gf_lv pS = P3 + P4, pA = P2 + pS, qA = pS - P2, pD = pA + P1, qD = pA - P1;
        gf_lv LD = gf_zt1(qD, pD, gf_m(pD)), LA = gf_zt1(qA, pA, gf_m(pA));
        return gf_dot(LD, LA);
    }
    case GF_SF_DtoAP1_AtoVP2_VtoP3P4: {
        gf_lv pV = P3 + P4, qV = P3 - P4, pA = P2 + pV, pD = P1 + pA, qD = pA - P1;
        gf_lv LB = gf_zt1(qD, pD, gf_m(pD)), LV = gf_zt1(qV, pV, gf_m(pV));
        gf_lv tmp = gf_ssv_dot(pA, gf_m(pA), LV);
        return -(gf_dot(LB, tmp));
    }
    case GF_SF_DtoAP1_AtoVP2Dwave_VtoP3P4: {
        gf_lv pV = P3 + P4, qV = P3 - P4, pA = P2 + pV, qA = P2 - pV, pD = P1 + pA, qD = pA - P1;
        gf_lv LD = gf_zt1(qD, pD, gf_m(pD)), LV = gf_zt1(qV, pV, gf_m(pV));
        gf_lm LA = gf_zt2(qA, pA, gf_m(pA));
        return gf_dot(LD, gf_contract(LA, LV));
    }
    case GF_SF_DtoTP1_TtoVP2_VtoP3P4: {
        gf_lv pV = P3 + P4, qV = P3 - P4, pT = P2 + pV, qT = pV - P2, pD = P1 + pT, qD = pT - P1;
        double mT = gf_m(pT);
        gf_lv tV = gf_zt1(qV, pV, gf_m(pV)), tD = gf_zt1(qD, pD, gf_m(pD)), t1T = gf_zt1(qT, pT, mT);
        gf_lv tmp = gf_levi3(t1T, pT, gf_ssv_dot(pT, mT, tV));
        return gf_sst_sandwich(pT, mT, tD, tD, tmp, t1T)
               + 1. / 3. * gf_dot(tD, tD) / (gf_dot(pT, pT)) * gf_sst_sandwich(pT, mT, pD, pD, tmp, t1T);
    }
    default: return 1.0;
    }
}
