"""Model builders shared by the tests, bench.py and smoke(): the same trees oracle/ref_driver.cpp builds
with the reference's classes (examples/dalitz/dalitz.cpp:124-322 etc.), here with goofit_b200's."""
import math

import numpy as np

import goofit_b200 as gf

MD0 = 1.86484
MD02 = MD0 * MD0
MD02INV = 1.0 / MD02
PIP = 0.13957018
PI0 = 0.1349766


class Keep:
    """Holds Python references so C++ objects referenced by raw pointer stay alive."""

    def __init__(self):
        self.objs = []

    def __call__(self, o):
        self.objs.append(o)
        return o


def dataset_from_columns(observables, columns):
    d = gf.UnbinnedDataSet(*observables)
    d.from_matrix(np.ascontiguousarray(np.vstack([np.asarray(c, dtype=np.float64) for c in columns])))
    return d


class Dalitz:
    """D0 -> pi+ pi- pi0, 16 resonances (15 RBW + NonRes), constant efficiency."""

    def __init__(self, float_masses=False, grid=240, eff_kind=0):
        k = self.keep = Keep()
        self.m12 = gf.Observable("m12", 0, 3)
        self.m13 = gf.Observable("m13", 0, 3)
        self.evt = gf.EventNumber("eventNumber")
        self.m12.numbins = grid
        self.m13.numbins = grid
        d = self.info = gf.DecayInfo3()
        d.motherMass, d.daug1Mass, d.daug2Mass, d.daug3Mass = MD0, PI0, PIP, PIP
        d.meson_radius, d.mother_meson_radius = 1.5, 5.0
        V = gf.Variable
        rhoM, rhoW = V("rho_mass", 0.7758, 0.01, 0.7, 0.8), V("rho_width", 0.1503, 0.01, 0.1, 0.2)
        m1450, w1450 = V("rhop_1450_mass", 1.465, 0.01, 1.0, 2.0), V("rhop_1450_width", 0.400, 0.01, 0.01, 5.0)
        m1700, w1700 = V("rhop_1700_mass", 1.720, 0.01, 1.6, 1.9), V("rhop_1700_width", 0.250, 0.01, 0.1, 1.0)
        P12, P13, P23 = 0, 1, 2
        rs = [
            ("rhop", 1, 0, rhoM, rhoW, 1, P12),
            ("rho0", 0.565, 0.164, rhoM, rhoW, 1, P23),
            ("rhom", 0.714, -0.025, rhoM, rhoW, 1, P13),
            ("rhop_1450", -0.174, -0.117, m1450, w1450, 1, P12),
            ("rho0_1450", 0.325, 0.057, m1450, w1450, 1, P23),
            ("rhom_1450", 0.788, 0.226, m1450, w1450, 1, P13),
            ("rhop_1700", 2.151, -0.658, m1700, w1700, 1, P12),
            ("rho0_1700", 2.400, -0.734, m1700, w1700, 1, P23),
            ("rhom_1700", 1.286, -1.532, m1700, w1700, 1, P13),
            ("f0_980", 0.008 * (-MD02), -0.013 * (-MD02), V("f0_980_mass", 0.980, 0.01, 0.8, 1.2), V("f0_980_width", 0.044, 0.001, 0.001, 0.08), 0, P23),
            ("f0_1370", -0.058 * (-MD02), 0.026 * (-MD02), V("f0_1370_mass", 1.434, 0.01, 1.2, 1.6), V("f0_1370_width", 0.173, 0.01, 0.01, 0.4), 0, P23),
            ("f0_1500", 0.057 * (-MD02), 0.012 * (-MD02), V("f0_1500_mass", 1.507, 0.01, 1.3, 1.7), V("f0_1500_width", 0.109, 0.01, 0.01, 0.3), 0, P23),
            ("f0_1710", 0.070 * (-MD02), 0.087 * (-MD02), V("f0_1710_mass", 1.714, 0.01, 1.5, 2.9), V("f0_1710_width", 0.140, 0.01, 0.01, 0.5), 0, P23),
            ("f2_1270", -1.027 * (-MD02INV), -0.162 * (-MD02INV), V("f2_1270_mass", 1.2754, 0.01, 1.0, 1.5), V("f2_1270_width", 0.1851, 0.01, 0.01, 0.4), 2, P23),
            ("f0_600", 0.068 * (-MD02), 0.010 * (-MD02), V("f0_600_mass", 0.500, 0.01, 0.3, 0.7), V("f0_600_width", 0.400, 0.01, 0.2, 0.6), 0, P23),
        ]
        self.coefs, self.masses, self.widths, self.resonances = [], [], [], []
        nre, nim = V("nonr_amp_real", -0.5595, 0.001, 0, 0), V("nonr_amp_imag", 0.108761, 0.1, 0, 0)
        self.coefs += [nre, nim]
        self.resonances.append(k(gf.Resonances.NonRes("nonr", nre, nim)))
        for i, (n, re, im, m, w, spin, cyc) in enumerate(rs):
            vre = V(n + "_amp_real", re) if i == 0 else V(n + "_amp_real", re, 0.001, 0, 0)
            vim = V(n + "_amp_imag", im) if i == 0 else V(n + "_amp_imag", im, 0.1, 0, 0)
            self.coefs += [vre, vim]
            self.masses.append(m)
            self.widths.append(w)
            self.resonances.append(k(gf.Resonances.RBW(n, vre, vim, m, w, spin, cyc)))
        if not float_masses:
            for r in self.resonances:
                r.setParameterConstantness(True)
        for r in self.resonances:
            d.add_resonance(r)
        one, zero = V("constantOne", 1), V("constantZero", 0)
        if eff_kind == 1:  # veto bands on two of the Dalitz variables (oracle make_dalitz eff_kind 1)
            vetos = [gf.VetoInfo(V("vetoLo1", 0.9), V("vetoHi1", 1.1), gf.PAIR_12), gf.VetoInfo(V("vetoLo2", 0.2), V("vetoHi2", 0.3), gf.PAIR_23)]
            self.eff = k(gf.DalitzVetoPdf("veto", self.m12, self.m13, V("vetoM", MD0), V("veto1", PI0), V("veto2", PIP), V("veto3", PIP), vetos))
        elif eff_kind == 2:  # the D -> KS0 h h efficiency polynomial
            c = [V("sq%d" % i, v) for i, v in enumerate((0.1, -0.2, 0.3, 0.15, 0.05, 1.0, 0.01, -0.02))]
            self.eff = k(gf.SquareDalitzEffPdf("sqeff", self.m12, self.m13, *c))
        else:
            self.eff = k(gf.PolynomialPdf("constantEff", [self.m12, self.m13], [one], [zero, zero], 0))
        self.signal = gf.Amp3Body("signalPDF", self.m12, self.m13, self.evt, d, self.eff)

    def dataset(self, m12, m13):
        n = len(m12)
        return dataset_from_columns([self.m12, self.m13, self.evt], [m12, m13, np.arange(n, dtype=np.float64)])


class Zoo3:
    """The other three-body shapes on one Dalitz plot: LASS (spin 0 and 1), FLATTE (plain and swapped), GS, Gauss, symmetrised RBW, spin-2 RBW,
    RhoOmegaMix (normalised and plain barrier factors), Spline (swapped);
    every shape parameter free (mirrors oracle/ref_driver.cpp case_zoo3)."""

    def __init__(self, grid=90):
        k = self.keep = Keep()
        self.m12 = gf.Observable("m12", 0, 3)
        self.m13 = gf.Observable("m13", 0, 3)
        self.evt = gf.EventNumber("eventNumber")
        self.m12.numbins = grid
        self.m13.numbins = grid
        d = self.info = gf.DecayInfo3()
        d.motherMass, d.daug1Mass, d.daug2Mass, d.daug3Mass = MD0, PI0, PIP, PIP
        d.meson_radius, d.mother_meson_radius = 1.5, 5.0
        V = gf.Variable
        self.pars = []

        def F(n, v, lo, hi):
            x = V(n, v, 0.001, lo, hi)
            self.pars.append(x)
            return x

        P12, P13, P23 = 0, 1, 2
        R = gf.Resonances
        knots = [0.1, 0.45, 0.8, 1.2, 1.6, 2.0, 2.5, 2.95]
        sp_re = [F("sp_re%d" % i, 0.4 * math.cos(1.3 * i) + 0.1 * i, -5, 5) for i in range(len(knots))]
        sp_im = [F("sp_im%d" % i, 0.3 * math.sin(0.9 * i) - 0.05 * i, -5, 5) for i in range(len(knots))]
        self.resonances = [
            R.Spline("spline", V("sp_r", 0.6), V("sp_i", -0.2), knots, sp_re, sp_im, P13, True),
            R.LASS("lass", V("lass_r", 0.8), V("lass_i", -0.3), F("lass_m", 1.425, 1.3, 1.5), F("lass_w", 0.27, 0.1, 0.4), F("lass_a", 2.07, 1, 3),
                   F("lass_r_", 3.32, 2, 4), F("lass_R", 1.0, 0.5, 1.5), F("lass_phiR", -1.9, -3, 3), F("lass_B", 0.96, 0.5, 1.5),
                   F("lass_phiB", 0.0017, -1, 1), 0, P12, True),
            R.LASS("lass1", V("lass1_r", 0.2), V("lass1_i", 0.1), F("lass1_m", 0.896, 0.8, 1.0), F("lass1_w", 0.05, 0.01, 0.1), F("lass1_a", 1.8, 1, 3),
                   F("lass1_r_", 1.76, 1, 4), F("lass1_R", 1.0, 0.5, 1.5), F("lass1_phiR", 0.3, -3, 3), F("lass1_B", 0.6, 0.1, 1.5),
                   F("lass1_phiB", -0.2, -1, 1), 1, P13, False),
            R.FLATTE("f0", V("f0_r", 0.5), V("f0_i", 0.4), F("f0_m", 0.965, 0.9, 1.0), F("f0_g1", 0.165, 0.1, 0.3), F("f0_rg", 4.21, 3, 5), P23, False),
            R.FLATTE("f0s", V("f0s_r", -0.3), V("f0s_i", 0.2), F("f0s_m", 0.98, 0.9, 1.05), F("f0s_g1", 0.2, 0.1, 0.3), F("f0s_rg", 3.5, 3, 5), P12,
                     True),
            R.GS("rho_gs", V("gs_r", 1.0), V("gs_i", 0.0), F("gs_m", 0.7758, 0.7, 0.85), F("gs_w", 0.1503, 0.1, 0.2), 1, P23),
            R.Gauss("gauss", V("ga_r", 0.1), V("ga_i", 0.05), F("ga_m", 1.2, 1.0, 1.4), F("ga_s", 0.1, 0.05, 0.3), P13),
            R.RBW("rho_sym", V("rs_r", 0.3), V("rs_i", -0.2), F("rs_m", 0.7758, 0.7, 0.85), F("rs_w", 0.15, 0.1, 0.2), 1, P12, True, True),
            R.RBW("f2_plain", V("f2_r", 0.2), V("f2_i", 0.3), F("f2_m", 1.2754, 1.2, 1.35), F("f2_w", 0.1851, 0.1, 0.3), 2, P23, False),
            R.RhoOmegaMix("rho_omega", V("ro_r", 0.7), V("ro_i", 0.1), F("om_m", 0.78265, 0.75, 0.8), F("om_w", 0.00849, 0.005, 0.02),
                          F("ro_m", 0.7690, 0.7, 0.85), F("ro_w", 0.1509, 0.1, 0.2), F("mix_re", 0.00185, -0.1, 0.1),
                          F("mix_im", -0.0007, -0.1, 0.1), F("mix_delta", 1.0, 0.5, 1.5), 1, P23, True, False),
            R.RhoOmegaMix("rho_omega_u", V("rou_r", -0.2), V("rou_i", 0.3), F("omu_m", 0.78265, 0.75, 0.8), F("omu_w", 0.00849, 0.005, 0.02),
                          F("rou_m", 0.7690, 0.7, 0.85), F("rou_w", 0.1509, 0.1, 0.2), F("mixu_re", 0.003, -1, 1), F("mixu_im", -0.002, -1, 1),
                          F("mixu_delta", 0.8, 0.5, 1.5), 1, P12, False, False),
        ]
        for r in self.resonances:
            d.add_resonance(k(r))
        one, zero = V("one", 1), V("zero", 0)
        self.eff = k(gf.PolynomialPdf("constantEff", [self.m12, self.m13], [one], [zero, zero], 0))
        self.signal = gf.Amp3Body("zoo3", self.m12, self.m13, self.evt, d, self.eff)

    def dataset(self, m12, m13):
        n = len(m12)
        return dataset_from_columns([self.m12, self.m13, self.evt], [m12, m13, np.arange(n, dtype=np.float64)])


def in_dalitz(m12, m13):
    """vectorised Dalitz boundary for the D0 -> pi0 pi+ pi+ kinematics above (DalitzPlotHelpers.cu:14-59)"""
    bigM, dm1, dm2, dm3 = MD0, PI0, PIP, PIP
    m23 = bigM * bigM + dm1 * dm1 + dm2 * dm2 + dm3 * dm3 - m12 - m13
    ok = (m23 >= 0) & (m23 <= (bigM - dm1) ** 2) & (m12 >= (dm1 + dm2) ** 2) & (m12 <= (bigM - dm3) ** 2)
    with np.errstate(invalid="ignore", divide="ignore"):
        s = np.sqrt(m12)
        e1 = 0.5 * (m12 - dm2 * dm2 + dm1 * dm1) / s
        e3 = 0.5 * (bigM * bigM - m12 - dm3 * dm3) / s
        r1 = np.sqrt(e1 * e1 - dm1 * dm1)
        r3 = np.sqrt(e3 * e3 - dm3 * dm3)
        es = (e1 + e3) ** 2
        ok &= (m13 >= es - (r1 + r3) ** 2) & (m13 <= es - (r1 - r3) ** 2)
    return ok


def dalitz_uniform_events(n, seed):
    """n points uniform inside the Dalitz boundary (numpy Generator; synthetic bench/test input)"""
    rng = np.random.default_rng(seed)
    a = np.empty(0)
    b = np.empty(0)
    while len(a) < n:
        x = rng.uniform(0, 3, size=int((n - len(a)) * 3.2) + 1024)
        y = rng.uniform(0, 3, size=len(x))
        ok = in_dalitz(x, y)
        a = np.concatenate([a, x[ok]])
        b = np.concatenate([b, y[ok]])
    return np.ascontiguousarray(a[:n]), np.ascontiguousarray(b[:n])


class DP4:
    """D0 -> K- pi+ pi+ pi-: 8 Bose-symmetrised amplitudes (examples/DP4/DP4.cu:51-232), AddPdf(1.0; Amp4Body, flat bkg)."""

    def __init__(self, mc_events=0, float_ls=False):
        k = self.keep = Keep()
        O, V = gf.Observable, gf.Variable
        self.m12, self.m34 = O("m12", 0, 3), O("m34", 0, 3)
        self.cos12, self.cos34, self.phi = O("cos12", -1, 1), O("cos34", -1, 1), O("phi", -3.5, 3.5)
        self.evt = gf.EventNumber("eventNumber", 0, 2147483647)
        mD0, piP, Km = 1.8645, 0.13957018, 0.493677
        info = self.info = gf.DecayInfo4()
        info.meson_radius = 1.5
        info.particle_masses = [mD0, piP, piP, Km, piP]
        self.lsvars = []

        def LV(n, v, lo, hi):
            x = V(n, v, 0.001, lo, hi) if float_ls else V(n, v)
            self.lsvars.append(x)
            return x

        RhoMass, RhoWidth = LV("rho_mass", 0.77526, 0.6, 0.9), LV("rho_width", 0.1478, 0.05, 0.3)
        KstarM, KstarW = LV("KstarM", 0.89581, 0.8, 1.0), LV("KstarW", 0.0474, 0.01, 0.1)
        f600M, f600W = LV("f600M", 0.519, 0.3, 0.7), LV("f600W", 0.454, 0.2, 0.7)
        a1M, a1W = LV("a1M", 1.23, 1.0, 1.5), LV("a1W", 0.42, 0.2, 0.7)
        K1M, K1W = LV("K1M", 1.272, 1.1, 1.4), LV("K1W", 0.09, 0.03, 0.2)
        K1430M, K1430W = LV("K1430M", 1.414, 1.2, 1.6), LV("K1430W", 0.29, 0.1, 0.5)
        S = gf.SF_4Body

        def SF2(kind, a, b):
            return [k(gf.SpinFactor("SF", kind, mD0, *a)), k(gf.SpinFactor("SF", kind, mD0, *b))]

        SFKRS = SF2(S.DtoV1V2_V1toP1P2_V2toP3P4_S, (0, 1, 2, 3), (3, 1, 2, 0))
        SFKRP = SF2(S.DtoV1V2_V1toP1P2_V2toP3P4_P, (0, 1, 2, 3), (3, 1, 2, 0))
        SFKRD = SF2(S.DtoV1V2_V1toP1P2_V2toP3P4_D, (0, 1, 2, 3), (3, 1, 2, 0))
        SFKF = SF2(S.DtoVS_VtoP1P2_StoP3P4, (2, 3, 0, 1), (2, 0, 3, 1))
        SFKK = SF2(S.DtoAP1_AtoSP2_StoP3P4, (0, 1, 3, 2), (3, 1, 0, 2))
        SFK1R = SF2(S.DtoAP1_AtoVP2_VtoP3P4, (3, 2, 0, 1), (0, 2, 3, 1))
        SFA1R = SF2(S.DtoAP1_AtoVP2_VtoP3P4, (2, 3, 0, 1), (2, 0, 3, 1))
        SFA1RD = SF2(S.DtoAP1_AtoVP2Dwave_VtoP3P4, (2, 3, 0, 1), (2, 0, 3, 1))
        L = gf.Lineshapes
        rho = lambda p: k(L.RBW("rho(770)", RhoMass, RhoWidth, 1, int(p)))
        kst = lambda p: k(L.RBW("K*(892)bar", KstarM, KstarW, 1, int(p)))
        P = gf.DP4Pair
        LSKRS = [rho(P.M_12), kst(P.M_34), rho(P.M_24), kst(P.M_13)]
        LSKRP = [rho(P.M_12), kst(P.M_34), rho(P.M_24), kst(P.M_13)]
        LSKRD = [rho(P.M_12), kst(P.M_34), rho(P.M_24), kst(P.M_13)]
        LSKF = [kst(P.M_34), k(L.Bugg("f600", f600M, f600W, 0, int(P.M_12))), kst(P.M_13), k(L.Bugg("f600", f600M, f600W, 0, int(P.M_24)))]
        LSKK = [k(L.SBW("K(1)(1270)bar", K1M, K1W, 1, int(P.M_34_2))), k(L.LASS("K(0)*(1430)bar", K1430M, K1430W, 0, int(P.M_34))),
                k(L.SBW("K(1)(1270)bar2", K1M, K1W, 1, int(P.M_13_2))), k(L.LASS("K(0)*(1430)bar2", K1430M, K1430W, 0, int(P.M_13)))]
        LSK1R = [k(L.SBW("K(1)(1270)bar", K1M, K1W, 0, int(P.M_12_3))), rho(P.M_12), k(L.SBW("K(1)(1270)bar", K1M, K1W, 0, int(P.M_24_3))), rho(P.M_24)]
        LSA1R = [k(L.SBW("a(1)(1260)+", a1M, a1W, 0, int(P.M_12_4))), rho(P.M_12), k(L.SBW("a(1)(1260)+", a1M, a1W, 0, int(P.M_24_1))), rho(P.M_24)]
        LSA1RD = [k(L.SBW("a(1)(1260)+", a1M, a1W, 2, int(P.M_12_4))), rho(P.M_12), k(L.SBW("a(1)(1260)+", a1M, a1W, 2, int(P.M_24_1))), rho(P.M_24)]
        self.coefs = []

        def Cf(n, v, fl):
            x = V(n, v, 0.001, 0, 0) if fl else V(n, v)
            self.coefs.append(x)
            return x

        A = gf.Amplitude
        AS = k(A("K*(892)rho(770)_S", Cf("amp_real1", -0.1, True), Cf("amp_imag1", 0.1, True), LSKRS, SFKRS, 2))
        AP = k(A("K*(892)rho(770)_P", Cf("amp_real2", -0.02, True), Cf("amp_imag2", -0.07, True), LSKRP, SFKRP, 2))
        AD = k(A("K*(892)rho(770)_D", Cf("amp_real3", -0.4, True), Cf("amp_imag3", 0.4, True), LSKRD, SFKRD, 2))
        KF = k(A("KF", Cf("amp_real4", 0.0120787, False), Cf("amp_imag4", -0.0332525, False), LSKF, SFKF, 2))
        KK = k(A("LSKK", Cf("amp_real5", 0.0109033, False), Cf("amp_imag5", -0.00186219, False), LSKK, SFKK, 2))
        K1R = k(A("LSK1R", Cf("amp_real6", -0.10728, False), Cf("amp_imag6", -0.130213, False), LSK1R, SFK1R, 2))
        A1R = k(A("LSA1R", Cf("amp_real7", 1.0, False), Cf("amp_imag7", 0.0, False), LSA1R, SFA1R, 2))
        A1RD = k(A("LSA1RD", Cf("amp_real8", -0.94921, False), Cf("amp_imag8", -1.73407, False), LSA1RD, SFA1RD, 2))
        for a in (KF, AS, AP, AD, KK, K1R, A1R, A1RD):
            info.add_amplitude(a)
        self.observables = [self.m12, self.m34, self.cos12, self.cos34, self.phi, self.evt]
        self.eff = k(gf.PolynomialPdf("constantEff", self.observables, [V("constantOne", 1)], [V("constantZero", 0)] * 2, 0))
        self.dp = gf.Amp4Body("test", self.observables, info, self.eff, mc_events)
        self.backgr = k(gf.PolynomialPdf("backgr", self.m12, [V("constant", 0.1)]))
        self.signal = gf.AddPdf("signal", V("constant2", 1.0), self.dp, self.backgr)

    def dataset(self, cols5):
        n = len(cols5[0])
        return dataset_from_columns(self.observables, list(cols5) + [np.arange(n, dtype=np.float64)])


class Zoo4:
    """The four-body lineshapes DP4 does not use (One, NonRes, Flatte, GLASS, FOCUS Kpi/KEta/I32) and the nine remaining spin
    factors, one permutation each (mirrors oracle/ref_driver.cpp case_zoo4)."""

    def __init__(self):
        k = self.keep = Keep()
        O, V = gf.Observable, gf.Variable
        self.m12, self.m34 = O("m12", 0, 3), O("m34", 0, 3)
        self.cos12, self.cos34, self.phi = O("cos12", -1, 1), O("cos34", -1, 1), O("phi", -3.5, 3.5)
        self.evt = gf.EventNumber("eventNumber", 0, 2147483647)
        mD0, piP, Km = 1.8645, 0.13957018, 0.493677
        info = self.info = gf.DecayInfo4()
        info.meson_radius = 1.5
        info.particle_masses = [mD0, piP, piP, Km, piP]
        self.pars = {}

        def F(n, v, lo, hi):
            self.pars[n] = V(n, v, 0.001, lo, hi)
            return self.pars[n]

        dm, dw = F("dummy_m", 1.0, 0.5, 1.5), F("dummy_w", 0.1, 0.05, 0.5)
        k0m, k0w = F("K0s1430_m", 1.425, 1.3, 1.5), F("K0s1430_w", 0.27, 0.1, 0.4)
        glass = [F("glass_a", 2.07, 1, 3), F("glass_r", 3.32, 2, 4), F("glass_phiF", 0.2, -1, 1), F("glass_phiR", -0.3, -1, 1),
                 F("glass_F", 0.9, 0.5, 1.5)]
        f0m = F("f0_m", 0.965, 0.9, 1.0)
        rhoM, rhoW = F("rho_m", 0.77526, 0.6, 0.9), F("rho_w", 0.1478, 0.05, 0.3)
        ksM, ksW = F("Kstar_m", 0.89581, 0.8, 1.0), F("Kstar_w", 0.0474, 0.01, 0.1)
        S, L, P = gf.SF_4Body, gf.Lineshapes, gf.DP4Pair
        SF = lambda kind, *idx: k(gf.SpinFactor("SF", kind, mD0, *idx))
        Mod = L.FOCUS.Mod
        L1 = [k(L.FOCUS("focus_kpi", Mod.Kpi, dm, dw, 0, int(P.M_34))), k(L.NonRes("nonres_p", dm, dw, 1, int(P.M_12)))]
        S1 = [SF(S.DtoPP1_PtoSP2_StoP3P4, 0, 1, 2, 3), SF(S.FF_12_34_L1, 0, 1, 2, 3)]
        L2 = [k(L.FOCUS("focus_keta", Mod.KEta, dm, dw, 0, int(P.M_13))), k(L.FOCUS("focus_i32", Mod.I32, dm, dw, 0, int(P.M_24)))]
        S2 = [SF(S.DtoPP1_PtoVP2_VtoP3P4, 3, 2, 0, 1), SF(S.FF_12_34_L2, 0, 1, 2, 3)]
        L3 = [k(L.GLASS("glass", k0m, k0w, 0, int(P.M_34), gf.FF.BL2, 1.5, glass)), k(L.One("one", dm, dw, 0, int(P.M_12)))]
        S3 = [SF(S.DtoV1P1_V1toV2P2_V2toP3P4, 0, 1, 2, 3), SF(S.FF_123_4_L1, 0, 1, 2, 3)]
        L4 = [k(L.Flatte("f0_980", f0m, dw, 0, int(P.M_12))), k(L.NonRes("nonres_d", dm, dw, 2, int(P.M_34_2)))]
        S4 = [SF(S.DtoTP1_TtoVP2_VtoP3P4, 2, 3, 0, 1), SF(S.FF_123_4_L2, 0, 1, 2, 3)]
        L5 = [k(L.RBW("rho", rhoM, rhoW, 1, int(P.M_12))), k(L.RBW("Kstar", ksM, ksW, 1, int(P.M_34)))]
        S5 = [SF(S.ONE, 0, 1, 2, 3)]
        A = gf.Amplitude
        for n, re, im, ls, sf in (("A1", 1.0, 0.0, L1, S1), ("A2", 0.4, -0.7, L2, S2), ("A3", -0.3, 0.5, L3, S3), ("A4", 0.8, 0.2, L4, S4),
                                  ("A5", -0.05, 0.02, L5, S5)):
            info.add_amplitude(k(A(n, V(n.lower() + "r", re), V(n.lower() + "i", im), ls, sf, 1)))
        self.observables = [self.m12, self.m34, self.cos12, self.cos34, self.phi, self.evt]
        self.eff = k(gf.PolynomialPdf("constantEff", self.observables, [V("constantOne", 1)], [V("constantZero", 0)] * 2, 0))
        self.dp = gf.Amp4Body("zoo4", self.observables, info, self.eff, 0)

    def dataset(self, cols5):
        n = len(cols5[0])
        return dataset_from_columns(self.observables, list(cols5) + [np.arange(n, dtype=np.float64)])


class TDDP4:
    """Time-dependent D0 -> K- pi+ pi+ pi- (examples/TDDP4/TDDP4.cu:38-156): three Bose-symmetrised K* rho amplitudes for
    each of A and B, TruthResolution or ThreeGaussResolution, AddPdf(1.0; Amp4Body_TD, flat bkg) — the tree
    oracle/ref_driver.cpp make_tddp4() builds with the reference's classes."""

    def __init__(self, three_gauss=False, float_ls=False, float_mix=True, shared=False, mc_to_gen=0, norm_seed=7, norm_events=None):
        k = self.keep = Keep()
        O, V = gf.Observable, gf.Variable
        self.m12, self.m34 = O("m12", 0, 3), O("m34", 0, 3)
        self.cos12, self.cos34, self.phi = O("cos12", -1, 1), O("cos34", -1, 1), O("phi", -3.5, 3.5)
        self.evt = gf.EventNumber("eventNumber", 0, 2147483647)
        self.dtime, self.sigmat = O("dtime", 0, 10), O("sigmat", -3, 3)
        mD0, piP, Km = 1.8645, 0.13957018, 0.493677
        if float_mix:
            tau, xm, ym = V("tau", 0.4101, 0.001, 0.100, 0.900), V("xmixing", 0.005, 0.001, -0.15, 0.15), V("ymixing", 0.01, 0.001, -0.15, 0.15)
        else:
            tau, xm, ym = V("tau", 0.4101), V("xmixing", 0.005), V("ymixing", 0.01)
        self.mix = [tau, xm, ym]
        info = self.info = gf.DecayInfo4t(tau, xm, ym, V("SqWStoRSrate", 1.0 / np.sqrt(300.0)))
        info.meson_radius = 1.5
        info.particle_masses = [mD0, piP, piP, Km, piP]
        self.lsvars = []

        def LV(n, v, lo, hi):
            x = V(n, v, 0.001, lo, hi) if float_ls else V(n, v)
            self.lsvars.append(x)
            return x

        RhoMass, RhoWidth = LV("rho_mass", 0.77526, 0.7, 0.85), LV("rho_width", 0.1478, 0.1, 0.2)
        KstarM, KstarW = LV("KstarM", 0.89581, 0.85, 0.95), LV("KstarW", 0.0474, 0.03, 0.07)
        S, L, P, FF = gf.SF_4Body, gf.Lineshapes, gf.DP4Pair, gf.FF

        def SF2(kind):
            return [k(gf.SpinFactor("SF", kind, mD0, 0, 1, 2, 3)), k(gf.SpinFactor("SF", kind, mD0, 3, 1, 2, 0))]

        def LS4():
            return [k(L.RBW("rho(770)", RhoMass, RhoWidth, 1, int(P.M_12), FF.BL2)), k(L.RBW("K*(892)bar", KstarM, KstarW, 1, int(P.M_34), FF.BL2)),
                    k(L.RBW("rho(770)", RhoMass, RhoWidth, 1, int(P.M_24), FF.BL2)), k(L.RBW("K*(892)bar", KstarM, KstarW, 1, int(P.M_13), FF.BL2))]

        A = gf.Amplitude
        Sa = k(A("K*(892)rho(770)_S", V("amp_real1", 1.0), V("amp_imag1", 0.0), LS4(), SF2(S.DtoV1V2_V1toP1P2_V2toP3P4_S), 2))
        Pa = k(A("K*(892)rho(770)_P", V("amp_real2", 0.526), V("amp_imag2", -0.626), LS4(), SF2(S.DtoV1V2_V1toP1P2_V2toP3P4_P), 2))
        Da = k(A("K*(892)rho(770)_D", V("amp_real3", 26.537), V("amp_imag3", 12.284), LS4(), SF2(S.DtoV1V2_V1toP1P2_V2toP3P4_D), 2))
        SB = k(A("B_K*(892)rho(770)_S", V("Bamp_real1", 1.0), V("Bamp_imag1", 0.0), LS4(), SF2(S.DtoV1V2_V1toP1P2_V2toP3P4_S), 2))
        PB = k(A("B_K*(892)rho(770)_P", V("Bamp_real2", -0.145), V("Bamp_imag2", 0.86), LS4(), SF2(S.DtoV1V2_V1toP1P2_V2toP3P4_P), 2))
        DB = k(A("B_K*(892)rho(770)_D", V("Bamp_real3", 24.343), V("Bamp_imag3", 5.329), LS4(), SF2(S.DtoV1V2_V1toP1P2_V2toP3P4_D), 2))
        for a in (Sa, Pa, Da):
            info.add_amplitude_B(a)
        for a in ((SB, Pa, DB) if shared else (SB, PB, DB)):
            info.add_amplitude(a)
        if three_gauss:
            vals = [("coreFrac", 0.8), ("tailFrac", 0.15), ("coreBias", 0.05), ("coreScale", 1.0), ("tailBias", 0.1), ("tailScale", 2.0),
                    ("outlBias", -0.05), ("outlScale", 5.0), ("selBias", 0.02), ("cbo", 0.01), ("cws", 1.1), ("cbo2", -0.02), ("cws2", 0.9)]
            self.res = k(gf.ThreeGaussResolution(*[V(n, v) for n, v in vals]))
        else:
            self.res = k(gf.TruthResolution())
        self.observables = [self.m12, self.m34, self.cos12, self.cos34, self.phi, self.evt, self.dtime, self.sigmat]
        self.eff = k(gf.PolynomialPdf("constantEff", self.observables, [V("constantOne", 1)], [V("constantZero", 0)] * 2, 0))
        if norm_events is not None:  # caller-supplied batches (e.g. weighted normalisation events)
            self.dp = gf.Amp4Body_TD("test", self.observables, info, self.res, self.eff, None, norm_events)
        else:
            self.dp = gf.Amp4Body_TD("test", self.observables, info, self.res, self.eff, None, norm_seed, mc_to_gen)
        self.backgr = k(gf.PolynomialPdf("backgr", self.m12, [V("constant", 0.1)]))
        self.signal = gf.AddPdf("signal", V("constant2", 1.0), self.dp, self.backgr)

    def dataset(self, cols7):
        """cols7 = m12, m34, cos12, cos34, phi, dtime, sigmat"""
        n = len(cols7[0])
        return dataset_from_columns(self.observables, list(cols7[:5]) + [np.arange(n, dtype=np.float64)] + list(cols7[5:7]))


def zach_model(dm_values, step=0.00003, numbins=270):
    """D* -> D0 pi delta-M shape of examples/zachFit/zachFit.cpp:143-230, one resolution component:
    AddPdf(bkg_frac; Argus, KinLimitBW (x) ScaledGaussian) with setIntegrationConstants(0.1395, 0.1665, step)."""
    V = gf.Variable
    dm = gf.Observable("dm", 0.1395, 0.1665)
    dm.numbins = numbins
    d = dataset_from_columns([dm], [dm_values])
    o = dict(dm=dm, d=d)
    o["mean"], o["width"] = V("kl_mean", 0.1454, 0.0001, 0.1450, 0.1458), V("kl_width", 0.0004, 0.0001, 0.0001, 0.001)
    o["kl"] = gf.KinLimitBWPdf("klbw", dm, o["mean"], o["width"])
    o["kl"].setMasses(1.8645, 0.13957)
    o["bkg"] = gf.ArgusPdf("argus_dm", dm, V("arg_m0", 0.13957, 0.0001, 0.1, 0.2), V("arg_slope", -1.0, 0.1, -100, 10), False, V("arg_power", 0.5, 0.01, 0.1, 2))
    o["rs"] = V("res_sigma", 0.00025, 0.00001, 0.0001, 0.001)
    o["res"] = gf.ScaledGaussianPdf("resolution", dm, V("dummyzero", 0), o["rs"], V("delta", 0.00002, 0.00001, -0.001, 0.001), V("epsilon", 0.05, 0.01, -0.5, 0.5))
    o["conv"] = gf.ConvolutionPdf("conv", dm, o["kl"], o["res"])
    o["conv"].setIntegrationConstants(0.1395, 0.1665, step)
    o["frac"] = V("bkg_frac", 0.3, 0.01, 0, 1)
    o["total"] = gf.AddPdf("total", o["frac"], o["bkg"], o["conv"])
    return o


class ZachFull:
    """The complete zachFit tree (examples/zachFit/zachFit.cpp:143-230), MC-stage variables fixed at their start values:
    total = AddPdf(bkg_frac; Argus bkg, AddPdf(gfrac1, gfrac2, afrac; conv1, conv2, Argus, conv3)) with
    conv_i = KinLimitBW(mean_i, width_bw) (x) ScaledGaussian(0, sigma_i, delta, epsilon), 2.7e-6 integration step and a 2700-bin
    normalisation grid. johnson=True swaps the convolved model for JohnsonSUPdf (the variant BASELINE.json words).
    Mirrors oracle/ref_driver.cpp make_zach()."""

    def __init__(self, johnson=False):
        k = self.keep = Keep()
        V = gf.Variable
        dm = self.dm = gf.Observable("dm", 0.1395, 0.1665)
        dm.numbins = 2700
        mean = [V("kpi_mc_mean1", 0.145402), V("kpi_mc_mean2", 0.145465), V("kpi_mc_mean3", 0.145404)]
        sigma = [V("kpi_mc_sigma1", 0.00010), V("kpi_mc_sigma2", 0.00075), V("kpi_mc_sigma3", 0.00020)]
        pimass, aslope, apower = V("kpi_mc_pimass", 0.13957), V("kpi_mc_aslope", -20.0), V("kpi_mc_apower", 1.3)
        gfrac1, gfrac2, afrac = V("kpi_mc_gfrac1", 0.65), V("kpi_mc_gfrac2", 0.02), V("kpi_mc_afrac", 0.005)
        argus = k(gf.ArgusPdf("argus", dm, pimass, aslope, False, apower))
        dummyzero = V("kpi_rd_dummyzero", 0)
        delta, epsilon = V("kpi_rd_delta", 0.000002, -0.00005, 0.00005), V("kpi_rd_epsilon", 0.05, -0.1, 0.2)
        width_bw = V("kpi_rd_width_bw", 0.0001, 0.00001, 0.0005)
        convs = []
        jg, jd = V("kpi_rd_jsu_gamma", -0.3), V("kpi_rd_jsu_delta", 1.1)
        for i in range(3):
            res = k(gf.ScaledGaussianPdf("resolution%d" % (i + 1), dm, dummyzero, sigma[i], delta, epsilon))
            if johnson:
                model = k(gf.JohnsonSUPdf("jsu%d" % (i + 1), dm, mean[i], width_bw, jg, jd))
            else:
                model = k(gf.KinLimitBWPdf("rbw%d" % (i + 1), dm, mean[i], width_bw))
            c = k(gf.ConvolutionPdf("signal%d" % (i + 1), dm, model, res))
            c.setIntegrationConstants(0.1395, 0.1665, 0.0000027)
            convs.append(c)
        self.convs = convs
        signal = k(gf.AddPdf("signal", [gfrac1, gfrac2, afrac], [convs[0], convs[1], argus, convs[2]]))
        slope = V("kpi_rd_slope", -1.0, 0.1, -35.0, 25.0)
        bkg = k(gf.ArgusPdf("bkg", dm, pimass, slope, False))
        bkg_frac = V("kpi_rd_bkg_frac", 0.03, 0.0, 0.3)
        self.total = gf.AddPdf("total", [bkg_frac], [bkg, signal])
        self.free = [width_bw, delta, epsilon, slope, bkg_frac]

    def dataset(self, dm_values):
        return dataset_from_columns([self.dm], [dm_values])


def kmatrix_pars():
    """Anisovich-Sarantsev pi pi S-wave K-matrix (the values of oracle/ref_driver.cpp KMatrixPars)"""
    V = gf.Variable
    br, bi = [8.5, 12.2, 29.2, 10.8, 3.1], [-4.2, 9.6, -2.1, -5.3, 7.7]
    fr, fi = [-4.7, -23.2, -1.9, -14.7, 5.0], [6.4, -10.9, 15.3, 2.0, -3.3]
    fs = [0.23399, 0.15044, -0.20545, 0.32825, 0.35412]
    g = [[0.22889, -0.55377, 0.00000, -0.39899, -0.34639], [0.94128, 0.55095, 0.00000, 0.39065, 0.31503],
         [0.36856, 0.23888, 0.55639, 0.18340, 0.18681], [0.33650, 0.40907, 0.85679, 0.19906, -0.00984],
         [0.18171, -0.17558, -0.79658, -0.00355, 0.22358]]
    pm = [0.65100, 1.20360, 1.55817, 1.21000, 1.82206]
    K = dict(sA0=V("sA0", -0.15), sA=V("sA", 1.0), s0_prod=V("s0_prod", -0.07, 0.001, -5, 5), s0_scatt=V("s0_scatt", -3.92637))
    K["beta_r"] = [V("beta_r%d" % i, br[i], 0.01, -100, 100) for i in range(5)]
    K["beta_i"] = [V("beta_i%d" % i, bi[i], 0.01, -100, 100) for i in range(5)]
    K["fprod_r"] = [V("fprod_r%d" % i, fr[i]) for i in range(5)]
    K["fprod_i"] = [V("fprod_i%d" % i, fi[i]) for i in range(5)]
    K["fscat"] = [V("fscat%d" % i, fs[i]) for i in range(5)]
    K["poles"] = []
    for i in range(5):
        K["poles"] += [V("g%d%d" % (i, j), g[i][j]) for j in range(5)] + [V("pmass%d" % i, pm[i])]
    return K


class KMatrix3:
    """D0 -> pi0 pi+ pi- with the pi+ pi- S-wave as a K-matrix and a rho+ (oracle/ref_driver.cpp case_kmatrix)"""

    def __init__(self, grid=90):
        k = self.keep = Keep()
        V = gf.Variable
        self.m12, self.m13 = gf.Observable("m12", 0, 3), gf.Observable("m13", 0, 3)
        self.m12.numbins = self.m13.numbins = grid
        self.evt = gf.EventNumber("eventNumber")
        info = self.info = gf.DecayInfo3()
        info.motherMass, info.daug1Mass, info.daug2Mass, info.daug3Mass = 1.86484, 0.1349766, 0.13957018, 0.13957018
        info.meson_radius, info.mother_meson_radius = 1.5, 5.0
        K = self.K = kmatrix_pars()
        R = gf.Resonances
        self.res = [k(R.kMatrix("kmat", V("km_r", 0.7), V("km_i", -0.4), K["sA0"], K["sA"], K["s0_prod"], K["s0_scatt"], K["beta_r"], K["beta_i"],
                                K["fprod_r"], K["fprod_i"], K["fscat"], K["poles"], 0, 2)),
                    k(R.RBW("rhop", V("rhop_r", 1.0), V("rhop_i", 0.0), V("rhop_m", 0.7758), V("rhop_w", 0.1503), 1, 0))]
        info.resonances = self.res
        self.eff = k(gf.PolynomialPdf("constantEff", [self.m12, self.m13], [V("one", 1)], [V("zero", 0)] * 2, 0))
        self.signal = gf.Amp3Body("kmatrix3", self.m12, self.m13, self.evt, info, self.eff)

    def dataset(self, m12, m13):
        return dataset_from_columns([self.m12, self.m13, self.evt], [m12, m13, np.arange(len(m12), dtype=np.float64)])


class KMatrix4:
    """D0 -> K- pi+ pi+ pi-: a (pi pi)_S K-matrix x K*(892) amplitude and a rho x K* one (oracle/ref_driver.cpp case_kmatrix)"""

    def __init__(self):
        k = self.keep = Keep()
        O, V = gf.Observable, gf.Variable
        self.m12, self.m34 = O("m12", 0, 3), O("m34", 0, 3)
        self.cos12, self.cos34, self.phi = O("cos12", -1, 1), O("cos34", -1, 1), O("phi", -3.5, 3.5)
        self.evt = gf.EventNumber("eventNumber", 0, 2147483647)
        mD0, piP, Km = 1.8645, 0.13957018, 0.493677
        info = self.info = gf.DecayInfo4()
        info.meson_radius = 1.5
        info.particle_masses = [mD0, piP, piP, Km, piP]
        K = self.K = kmatrix_pars()
        L, P, S = gf.Lineshapes, gf.DP4Pair, gf.SF_4Body
        ksM, ksW = V("Kstar_m", 0.89581), V("Kstar_w", 0.0474)
        L1 = [k(L.kMatrix("kmat", V("km_r", 1), V("km_i", 0), K["sA0"], K["sA"], K["s0_prod"], K["s0_scatt"], K["beta_r"], K["beta_i"],
                          K["fprod_r"], K["fprod_i"], K["fscat"], K["poles"], 0, int(P.M_12))),
              k(L.RBW("Kstar", ksM, ksW, 1, int(P.M_34)))]
        S1 = [k(gf.SpinFactor("SF", S.DtoVS_VtoP1P2_StoP3P4, mD0, 2, 3, 0, 1))]
        L2 = [k(L.RBW("rho", V("rho_m", 0.77526), V("rho_w", 0.1478), 1, int(P.M_12))), k(L.RBW("Kstar", ksM, ksW, 1, int(P.M_34)))]
        S2 = [k(gf.SpinFactor("SF", S.DtoV1V2_V1toP1P2_V2toP3P4_S, mD0, 0, 1, 2, 3))]
        for a in (k(gf.Amplitude("KS", V("a1r", 0.02), V("a1i", 0.01), L1, S1, 1)), k(gf.Amplitude("RK", V("a2r", 1.0), V("a2i", 0.0), L2, S2, 1))):
            info.add_amplitude(a)
        self.observables = [self.m12, self.m34, self.cos12, self.cos34, self.phi, self.evt]
        self.eff = k(gf.PolynomialPdf("constantEff", self.observables, [V("constantOne", 1)], [V("constantZero", 0)] * 2, 0))
        self.dp = gf.Amp4Body("kmatrix4", self.observables, info, self.eff, 0)

    def dataset(self, cols5):
        n = len(cols5[0])
        return dataset_from_columns(self.observables, list(cols5) + [np.arange(n, dtype=np.float64)])


class TD3:
    """Amp3Body_TD on D0 -> pi0 pi+ pi- (oracle/ref_driver.cpp td3_one): non-resonant term, three rho(770) charge states, f0(980)"""

    def __init__(self, three_gauss=False, grid=120):
        k = self.keep = Keep()
        O, V = gf.Observable, gf.Variable
        self.dtime, self.sigmat = O("dtime", 0, 10), O("sigmat", -3, 3)
        self.m12, self.m13 = O("m12", 0, 3), O("m13", 0, 3)
        self.mistag, self.charmtag = O("mistag", 0, 1), O("charmtag", -2, 2)
        self.evt = gf.EventNumber("eventNumber")
        self.m12.numbins = self.m13.numbins = grid
        self.tau, self.xm, self.ym = V("tau", 0.4101, 0.001, 0.3, 0.5), V("xmixing", 0.0049, 0.001, -0.1, 0.1), V("ymixing", 0.0061, 0.001, -0.1, 0.1)
        self.dx, self.dy = V("deltax", 0.0011, 0.001, -0.1, 0.1), V("deltay", -0.0007, 0.001, -0.1, 0.1)
        info = self.info = gf.DecayInfo3t(self.tau, self.xm, self.ym, self.dx, self.dy)
        info.motherMass, info.daug1Mass, info.daug2Mass, info.daug3Mass = 1.86484, 0.1349766, 0.13957018, 0.13957018
        info.meson_radius, info.mother_meson_radius = 1.5, 5.0
        MD02 = 1.86484 ** 2
        self.rhoM, self.rhoW = V("rho_mass", 0.7758, 0.01, 0.7, 0.8), V("rho_width", 0.1503, 0.01, 0.1, 0.2)
        self.f0M, self.f0W = V("f0_980_mass", 0.980, 0.01, 0.8, 1.2), V("f0_980_width", 0.044, 0.001, 0.001, 0.08)
        self.c4r = V("rhom_amp_real", 0.714, 0.001, 0, 0)
        R = gf.Resonances
        self.res = [k(R.NonRes("nonr", V("nonr_amp_real", -0.5595), V("nonr_amp_imag", 0.108761))),
                    k(R.RBW("rhop", V("rhop_amp_real", 1.0), V("rhop_amp_imag", 0.0), self.rhoM, self.rhoW, 1, 0)),
                    k(R.RBW("rho0", V("rho0_amp_real", 0.565), V("rho0_amp_imag", 0.164), self.rhoM, self.rhoW, 1, 2)),
                    k(R.RBW("rhom", self.c4r, V("rhom_amp_imag", -0.025), self.rhoM, self.rhoW, 1, 1)),
                    k(R.RBW("f0_980", V("f0_amp_real", 0.008 * (-MD02)), V("f0_amp_imag", -0.013 * (-MD02)), self.f0M, self.f0W, 0, 2))]
        info.resonances = self.res
        if three_gauss:
            vals = [("coreFrac", 0.8), ("tailFrac", 0.15), ("coreBias", 0.05), ("coreScale", 1.0), ("tailBias", 0.1), ("tailScale", 2.0),
                    ("outlBias", -0.05), ("outlScale", 5.0), ("selBias", 0.02), ("cbo", 0.01), ("cws", 1.1), ("cbo2", -0.02), ("cws2", 0.9)]
            self.resol = k(gf.ThreeGaussResolution(*[V(n, v) for n, v in vals]))
        else:
            self.resol = k(gf.TruthResolution())
        self.eff = k(gf.PolynomialPdf("constantEff", [self.m12, self.m13], [V("constantOne", 1)], [V("constantZero", 0)] * 2, 0))
        self.dp = gf.Amp3Body_TD("tddalitz", self.dtime, self.sigmat, self.m12, self.m13, self.evt, info, self.resol, self.eff, self.mistag, self.charmtag)
        self.observables = [self.dtime, self.sigmat, self.m12, self.m13, self.evt, self.mistag, self.charmtag]

    def dataset(self, dtime, sigmat, m12, m13, mistag_value, charmtag):
        n = len(m12)
        self.mistag.value = mistag_value  # the normalisation reads the Observable's own value (Amp3Body_TD.cu:735-744)
        return dataset_from_columns(self.observables, [dtime, sigmat, m12, m13, np.arange(n, dtype=np.float64), np.full(n, mistag_value), charmtag])
