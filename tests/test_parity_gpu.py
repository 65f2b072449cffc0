"""Parity of the CUDA path against the oracle (the unmodified reference built with its OpenMP backend,
oracle/ref_driver.cpp). Every test calls the product through the host API -> C ABI -> sm_100a kernels.
Tolerances: per-event values 1e-12 relative (libm vs CUDA libm differ at the ulp level); NLL sums 1e-12
relative (north star); fitted parameters 0.01 sigma (north star)."""
import math
import os

import numpy as np
import pytest

import goofit_b200 as gf
from models import Dalitz, Zoo3, dataset_from_columns

pytestmark = pytest.mark.gpu

RTOL_EVT = 1e-12
RTOL_NLL = 1e-12


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


def fast_vs_interpreter(pdf, expect):
    """The shape-specialised likelihood kernel (csrc/gf_fast.cu) and the interpreter on the same program and data: `expect`
    is the copy that must have been launched; returns (fast NLL, interpreter NLL)."""
    fast = pdf.calculateNLL()
    assert gf.last_kernel_name() == expect, gf.last_kernel_name()
    was = gf.fast_enable(False)
    try:
        slow = pdf.calculateNLL()
        assert gf.last_kernel_name() == "k_eval_sum"
    finally:
        gf.fast_enable(was)
    assert rel(fast, slow) < RTOL_NLL, (fast, slow)  # two compiled copies of the same arithmetic: FMA contraction differs
    return fast, slow


def kstar_move_fused_vs_separate(sig, dp, var, want_nll, want_integral, rtol_int):
    """Move the K*(892) mass away and back with the one-pass update (k_amp4_update: DMMA Gram sum) switched off, i.e. through
    k_amp4_cache + k_amp4_amps + the 4x4-tile k_amp4_gram, then once more with it on: same NLL and integral both ways."""
    for fused in (False, True):
        was = gf.fused_enable(fused)
        try:
            for value in (0.8931, 0.8975):
                var.changed = True
                sig.updateVariable(var, value)
                sig.updateParameters()
                nll = sig.calculateNLL()
        finally:
            gf.fused_enable(was)
        assert rel(nll, want_nll) < RTOL_NLL, (fused, nll, want_nll)
        assert rel(1.0 / dp.getNormalization(), want_integral) < rtol_int, fused


def check_pdf(golden, tag, pdf, data, gridvar=None, comps=True, rtol_evt=RTOL_EVT):
    pdf.setData(data)
    nll = pdf.calculateNLL()
    assert rel(nll, golden[tag + ".nll"]) < RTOL_NLL, (tag, nll, golden[tag + ".nll"])
    if comps:
        probs = pdf.getCompProbsAtDataPoints()
        np.testing.assert_allclose(probs[0], golden.arr(tag + ".prob"), rtol=rtol_evt, atol=0)
        k = 0
        while golden.has(tag + ".comp%d" % k):
            np.testing.assert_allclose(probs[1 + k], golden.arr(tag + ".comp%d" % k), rtol=rtol_evt, atol=0)
            k += 1
        # the NLL must be the compensated sum of the per-event terms
        terms = -np.log(np.asarray(probs[0])[np.asarray(probs[0]) > 0])
        import math
        if tag != "addext":
            assert rel(nll, math.fsum(terms)) < RTOL_NLL
    pdf.setFitControl(gf.UnbinnedNllFit())
    pdf.setData(data)
    again = pdf.calculateNLL()
    assert again == nll, "NLL must be bit-reproducible"
    if gridvar is not None and golden.has(tag + ".grid"):
        g = pdf.evaluateAtPoints(gridvar)
        np.testing.assert_allclose(g, golden.arr(tag + ".grid"), rtol=rtol_evt, atol=1e-300)
        pdf.setFitControl(gf.UnbinnedNllFit())


def test_exponential(golden):
    x = gf.Observable("x", 0, 10)
    d = dataset_from_columns([x], [golden.arr("exp.x")])
    alpha = gf.Variable("alpha", -2, 0.1, -10, 10)
    pdf = gf.ExpPdf("exppdf", x, alpha)
    pdf.setData(d)
    assert rel(pdf.integrate(0, 10), golden["exp.integrate_0_10"]) < 1e-15
    assert rel(pdf.normalize(), golden["exp.normalize"]) < 1e-15
    check_pdf(golden, "exp", pdf, d, x)
    pdf.setData(d)
    fast, slow = fast_vs_interpreter(pdf, "k_eval_logleaf")  # -(a x + log norm) against -log(exp(a x) norm)
    assert rel(fast, golden["exp.nll"]) < RTOL_NLL and rel(slow, golden["exp.nll"]) < RTOL_NLL
    pdf.setData(d)
    fm = gf.FitManager(pdf)
    fm.setVerbosity(0)
    fm.fit()
    assert bool(fm)
    sigma = golden["exp.fit_alpha_err"]
    assert abs(alpha.value - golden["exp.fit_alpha"]) < 0.01 * sigma
    assert rel(alpha.error, sigma) < 1e-3
    assert rel(pdf.calculateNLL(), golden["exp.fit_nll"]) < 1e-8


def test_gaussian(golden):
    x = gf.Observable("x", -10, 10)
    d = dataset_from_columns([x], [golden.arr("gauss.x")])
    m, s = gf.Variable("m", 1.5, 0.1, -10, 10), gf.Variable("s", 2.5, 0.1, 0.5, 10)
    pdf = gf.GaussianPdf("gauss", x, m, s)
    pdf.setData(d)
    assert rel(pdf.normalize(), golden["gauss.normalize"]) < 1e-15
    assert abs(pdf.normalize() - 6.26657) < 1e-5  # tests/simple/NormalizeTest.cpp:194
    check_pdf(golden, "gauss", pdf, d, x)
    pdf.setData(d)
    fast, slow = fast_vs_interpreter(pdf, "k_eval_logleaf")
    assert rel(fast, golden["gauss.nll"]) < RTOL_NLL and rel(slow, golden["gauss.nll"]) < RTOL_NLL


def test_normalize_test_known_answers():
    # tests/simple/NormalizeTest.cpp:68-99 — Gaussian(1.5, 2.5) probabilities on the 100-bin grid over [-5, 10]
    x = gf.Observable("xvar", -5, 10)
    m, s = gf.Variable("mean", 1.5), gf.Variable("sigma", 2.5)
    pdf = gf.GaussianPdf("gausspdf", x, m, s)
    grid = pdf.makeGrid()
    pdf.setData(grid)
    for idx, val in [(0, -4.925), (20, -1.925), (40, 1.075), (60, 4.075)]:
        assert grid.getValue(x, idx) == pytest.approx(val)
    vv = pdf.getCompProbsAtDataPoints()
    assert len(vv) == 1 and len(vv[0]) == 100
    for idx, val in [(0, 0.00587129964482), (20, 0.0624318782242), (40, 0.157287605852), (60, 0.0938855055588)]:
        assert vv[0][idx] == pytest.approx(val, rel=1e-10)
    v = pdf.evaluateAtPoints(x)  # NormalizeTest.cpp:189-194
    for idx, val in [(0, 0.00587129964482), (20, 0.0624318782242), (40, 0.157287605852), (60, 0.0938855055588)]:
        assert v[idx] == pytest.approx(val, rel=1e-10)
    assert pdf.normalize() == pytest.approx(6.26657, rel=1e-5)


def test_add_norm_known_answer():
    # tests/simple/AddNormTest.cpp:16-44: f*A + (1-f)*B must normalise every component, including the last
    x = gf.Observable("xvar", -5, 10)
    mean = gf.Variable("mean", 1.5)
    ga = gf.GaussianPdf("gaussA", x, mean, gf.Variable("sigmaA", 2.5))
    gb = gf.GaussianPdf("gaussB", x, mean, gf.Variable("sigmaB", 0.8))
    total = gf.AddPdf("sum", gf.Variable("frac", 0.6), ga, gb)
    grid = total.makeGrid()
    total.setData(grid)
    vv = total.getCompProbsAtDataPoints()
    assert len(vv) == 3 and len(vv[0]) == len(vv[1]) == len(vv[2])
    for i in range(0, len(vv[0]), 7):
        assert vv[0][i] == pytest.approx(0.6 * vv[1][i] + 0.4 * vv[2][i], rel=1e-12)


def test_shapes(golden):
    V = gf.Variable
    x = gf.Observable("x", 0.5, 4)
    d = dataset_from_columns([x], [golden.arr("shapes.x")])
    a, x0 = V("a", -0.7, 0.1, -5, 5), V("x0", 0.25, 0.1, -1, 1)
    check_pdf(golden, "expoff", gf.ExpPdf("expoff", x, a, x0), d, x)
    w = [V("w0", 0.3, 0.1, -5, 5), V("w1", -0.8, 0.1, -5, 5), V("w2", 0.05, 0.1, -5, 5)]
    check_pdf(golden, "exppoly", gf.ExpPdf("exppoly", x, w), d, x)
    check_pdf(golden, "exppolyoff", gf.ExpPdf("exppolyoff", x, w, x0), d, x)
    c = [V("c0", 1.0, 0.1, -5, 5), V("c1", 0.4, 0.1, -5, 5), V("c2", 0.15, 0.1, -5, 5)]
    check_pdf(golden, "poly", gf.PolynomialPdf("poly", x, c), d, x)
    check_pdf(golden, "poly_low1", gf.PolynomialPdf("poly_low1", x, c, 1), d, x)
    jo = gf.JohnsonSUPdf("johnson", x, V("jm", 2.0, 0.1, 0, 4), V("js", 0.6, 0.1, 0.1, 4), V("jg", -0.4, 0.1, -3, 3), V("jd", 1.3, 0.1, 0.1, 5))
    check_pdf(golden, "johnson", jo, d, x)
    sg = gf.ScaledGaussianPdf("scaledgauss", x, V("sm", 2.1, 0.1, 0, 4), V("ss", 0.7, 0.1, 0.1, 4), V("sd", 0.1, 0.1, -1, 1), V("se", 0.2, 0.1, -1, 1))
    check_pdf(golden, "scaledgauss", sg, d, x)
    asl, apw = V("asl", -1.5, 0.1, -10, 10), V("apw", 0.5, 0.1, 0.1, 3)
    check_pdf(golden, "argus_upper", gf.ArgusPdf("argus_upper", x, V("am0", 4.2, 0.1, 4, 5), asl, True, apw), d, x)
    check_pdf(golden, "argus_lower", gf.ArgusPdf("argus_lower", x, V("bm0", 0.3, 0.1, 0, 0.5), V("asl2", -0.02, 0.01, -1, 1), False, apw), d, x)


def make_zach(golden):
    from models import zach_model
    return zach_model(golden.arr("dm.x"))


def test_zachfit_pieces(golden):
    o = make_zach(golden)
    check_pdf(golden, "klbw", o["kl"], o["d"], o["dm"])
    check_pdf(golden, "argus_dm", o["bkg"], o["d"], o["dm"])
    # convolution sums 900 products per table entry; same i-order as the reference, CUDA FMA contraction differs
    check_pdf(golden, "conv", o["conv"], o["d"], o["dm"], comps=False, rtol_evt=1e-11)
    check_pdf(golden, "convadd", o["total"], o["d"], o["dm"], rtol_evt=1e-11)
    # cache state machine: only the resolution moves, then only the model (ConvolutionPdf.cu:330-353)
    total, rs, mean, width = o["total"], o["rs"], o["mean"], o["width"]
    total.setData(o["d"])
    total.calculateNLL()
    rs.changed, mean.changed, width.changed = True, False, False
    total.updateVariable(rs, 0.0003)
    total.updateParameters()
    assert rel(total.calculateNLL(), golden["convadd.nll_res_moved"]) < 1e-11
    rs.changed, width.changed = False, True
    total.updateVariable(width, 0.0005)
    total.updateParameters()
    assert rel(total.calculateNLL(), golden["convadd.nll_model_moved"]) < 1e-11


def test_add_prod(golden):
    V = gf.Variable
    x = gf.Observable("x", 0, 10)
    d = dataset_from_columns([x], [golden.arr("add.x")])
    g = gf.GaussianPdf("ag", x, V("am", 4.0, 0.1, 0, 10), V("as", 1.2, 0.1, 0.1, 5))
    e = gf.ExpPdf("ae", x, V("aalpha", -0.5, 0.1, -5, 5))
    add = gf.AddPdf("add", V("afrac", 0.35, 0.1, 0, 1), g, e)
    check_pdf(golden, "add", add, d, x)
    ext = gf.AddPdf("addext", [V("n1", 500, 10, 0, 5000), V("n2", 1100, 10, 0, 5000)], [g, e])
    check_pdf(golden, "addext", ext, d, x)
    y = gf.Observable("y", -5, 5)
    d2 = dataset_from_columns([x, y], [golden.arr("prod.x"), golden.arr("prod.y")])
    gy = gf.GaussianPdf("gy", y, V("my", 0.5, 0.1, -5, 5), V("sy", 1.7, 0.1, 0.1, 5))
    prod = gf.ProdPdf("prod", [g, gy])
    check_pdf(golden, "prod", prod, d2)


def test_event_weighted_add_and_mapped(golden):
    """combine/EventWeightedAddPdf.cu (one weight column per component but the last; or one per component = extended, divided by
    the total weight) and combine/MappedPdf.cu (the rounded value of the mapping function picks the shape) against the
    reference on the same 700 events (oracle/ref_driver.cpp case_combine2)."""
    V = gf.Variable
    x = gf.Observable("x", -5, 5)
    w1 = gf.Observable("w1", 0, 1)
    w2 = gf.Observable("w2", 0, 5)
    sel = gf.Observable("sel", 0, 2)
    X, W1, W2, S = (golden.arr("combine2." + k) for k in ("x", "w1", "w2", "sel"))
    m1, s1 = V("m1", 0.2, 0.1, -3, 3), V("s1", 1.1, 0.1, 0.3, 4)
    m2, s2 = V("m2", -0.8, 0.1, -3, 3), V("s2", 2.0, 0.1, 0.3, 4)
    al = V("alpha", -0.3, 0.1, -2, 2)

    def check(pdf, data, tag):
        pdf.setData(data)
        nll = pdf.calculateNLL()
        assert gf.last_kernel_name() == "k_eval_sum"  # no shape-specialised copy for these: the interpreter runs them
        assert rel(nll, golden["combine2.%s_nll" % tag]) < RTOL_NLL, (tag, nll)
        probs = np.asarray(pdf.getCompProbsAtDataPoints()[0])
        np.testing.assert_allclose(probs, golden.arr("combine2.%s_prob" % tag), rtol=RTOL_EVT, atol=0)
        assert rel(nll, math.fsum(-np.log(probs))) < RTOL_NLL
        pdf.setFitControl(gf.UnbinnedNllFit())
        pdf.setData(data)
        assert pdf.calculateNLL() == nll

    ga = gf.GaussianPdf("ga", x, m1, s1)
    gb = gf.GaussianPdf("gb", x, m2, s2)
    ex = gf.ExpPdf("ex", x, al)
    check(gf.EventWeightedAddPdf("ew", [w1], [ga, ex]), dataset_from_columns([w1, x], [W1, X]), "ew")
    check(gf.EventWeightedAddPdf("ewx", [w1, w2], [ga, gb]), dataset_from_columns([w1, w2, x], [W1, W2, X]), "ewx")
    mapping = gf.PolynomialPdf("map", sel, [V("zero", 0), V("one", 1)])
    mp = gf.MappedPdf("mapped", mapping, [ga, gb, ex])
    check(mp, dataset_from_columns([sel, x], [S, X]), "map")
    # a parameter of one of the mapped shapes moves: only that third of the events changes
    before = np.asarray(mp.getCompProbsAtDataPoints()[0])
    mp.updateVariable(m2, -0.5)
    mp.updateParameters()
    mp.setFitControl(gf.UnbinnedNllFit())
    mp.calculateNLL()
    after = np.asarray(mp.getCompProbsAtDataPoints()[0])
    moved = before != after
    assert moved[S == 1].all() and not moved[S != 1].any()


def test_basic2_remaining_closed_form_shapes(golden):
    """The closed-form shapes of PDFs/basic beyond the five configured fits: BW, BifurGauss, CrystalBall (both tails), ExpGaus,
    Novosibirsk (with and without tail), Step (both directions, inside an AddPdf), TrigThreshold (all four forms), Bernstein,
    CorrGaussian, Landau (narrow and wide: every range of the approximation), BinTransform as the mapping function of a MappedPdf
    (oracle/ref_driver.cpp case_basic2)."""
    V = gf.Variable
    x = gf.Observable("x", -4, 6)
    X = golden.arr("basic2.x")
    d = dataset_from_columns([x], [X])
    m, w = V("b2m", 0.8, 0.1, -4, 6), V("b2w", 1.3, 0.1, 0.1, 5)
    sl, sr = V("b2sl", 0.9, 0.1, 0.1, 5), V("b2sr", 1.7, 0.1, 0.1, 5)
    ap, an, pw = V("b2ap", 1.1, 0.1, 0.1, 5), V("b2an", -0.9, 0.1, -5, -0.1), V("b2pw", 2.5, 0.1, 0.5, 10)
    tau = V("b2tau", 0.7, 0.1, 0.05, 5)
    tail, notail = V("b2tail", 0.35, 0.05, -2, 2), V("b2notail", 0.0)
    check_pdf(golden, "bw", gf.BWPdf("bw", x, m, w), d, x)
    check_pdf(golden, "bifur", gf.BifurGaussPdf("bifur", x, m, sl, sr), d, x)
    check_pdf(golden, "cbpos", gf.CrystalBallPdf("cbpos", x, m, w, ap, pw), d, x)
    check_pdf(golden, "cbneg", gf.CrystalBallPdf("cbneg", x, m, w, an), d, x)
    check_pdf(golden, "expgaus", gf.ExpGausPdf("expgaus", x, m, sl, tau), d, x)
    check_pdf(golden, "novo", gf.NovosibirskPdf("novo", x, m, w, tail), d, x)
    check_pdf(golden, "novo0", gf.NovosibirskPdf("novo0", x, m, w, notail), d, x)
    x0 = V("b2x0", 0.4, 0.1, -4, 6)
    g = gf.GaussianPdf("b2g", x, m, sr)
    f = V("b2f", 0.3, 0.1, 0, 1)
    up, dn = gf.StepPdf("stepup", x, x0, 1), gf.StepPdf("stepdn", x, x0, 0)
    assert up.integrate(-4, 6) == golden["stepup.integral"] and dn.integrate(-4, 6) == golden["stepdn.integral"]
    check_pdf(golden, "stepup_add", gf.AddPdf("stepup_add", f, up, g), d, x)
    check_pdf(golden, "stepdn_add", gf.AddPdf("stepdn_add", f, dn, g), d, x)
    th, tc, lc = V("b2th", 1.0, 0.1, -4, 6), V("b2tc", 0.12, 0.01, 0.01, 2), V("b2lc", 0.6, 0.01, 0, 1)
    check_pdf(golden, "trigup", gf.TrigThresholdPdf("trigup", x, th, tc, lc, True), d, x)
    check_pdf(golden, "trigdn", gf.TrigThresholdPdf("trigdn", x, th, tc, lc, False), d, x)
    bc = [V("bc0", 1.0, 0.1, 0, 5), V("bc1", 0.3, 0.1, 0, 5), V("bc2", 1.6, 0.1, 0, 5), V("bc3", 0.7, 0.1, 0, 5)]
    check_pdf(golden, "bern", gf.BernsteinPdf("bern", x, bc), d, x)
    check_pdf(golden, "bern1", gf.BernsteinPdf("bern1", x, bc, 1), d, x)
    mpv = V("b2mpv", -0.5, 0.1, -4, 6)
    # the narrow Landau is 0 for the events far below its peak (as upstream): their terms are left out of the sum on both sides
    # (far below the peak the density is exp(-1/u) with u = exp(v + 1): a relative error eps of u becomes eps / u of the value, so the
    # values around 1e-298 carry 1e-12 where libm's exp differs by an ulp; the NLL is still held to 1e-12)
    check_pdf(golden, "landau_narrow", gf.LandauPdf("landau_narrow", x, mpv, V("b2ls1", 0.02, 0.01, 0.001, 5)), d, x, rtol_evt=1e-10)
    check_pdf(golden, "landau_wide", gf.LandauPdf("landau_wide", x, mpv, V("b2ls2", 0.6, 0.1, 0.001, 5)), d, x)
    y = gf.Observable("y", -3, 3)
    y.setNumBins(60)
    x.setNumBins(80)
    d2 = dataset_from_columns([x, y], [X, golden.arr("basic2.y")])
    m2, s2, corr = V("b2m2", 0.1, 0.1, -3, 3), V("b2s2", 0.8, 0.1, 0.1, 5), V("b2corr", 0.25, 0.05, -1, 1)
    check_pdf(golden, "corrgauss", gf.CorrGaussianPdf("corrgauss", x, y, m, w, m2, s2, corr), d2)
    mc = V("b2mc", 3.0, 0.1, 0, 10)
    check_pdf(golden, "trig2up", gf.TrigThresholdPdf("trig2up", x, y, th, tc, lc, mc, True), d2)
    check_pdf(golden, "trig2dn", gf.TrigThresholdPdf("trig2dn", x, y, th, tc, lc, mc, False), d2)
    bt = gf.BinTransformPdf("bintr", [x, y], [-4, -3], [5, 3], [2, 2])
    comps = [gf.GaussianPdf("bt0", x, m, sl), gf.GaussianPdf("bt1", x, m, sr), gf.GaussianPdf("bt2", y, m2, s2), gf.ExpPdf("bt3", y, corr)]
    check_pdf(golden, "bintr_mapped", gf.MappedPdf("bintr_mapped", bt, comps), d2)
    x.setNumBins(100)


def test_smooth_histogram_pdf(golden):
    """basic/SmoothHistogramPdf.cu: the bin content of a 1-D (50 bins) and a 2-D (20 x 12) histogram, each bin averaged with its
    neighbours; the smoothed table is built by the host class and attached to the record with gf_table_set
    (oracle/ref_driver.cpp case_smoothhist)."""
    V = gf.Variable
    x, y = gf.Observable("x", -4, 6), gf.Observable("y", -3, 3)
    x.setNumBins(50)
    y.setNumBins(12)
    h1 = gf.BinnedDataSet(x)
    for i, v in enumerate(golden.arr("smoothhist.h1")):
        h1.setBinContent(i, v)
    d = dataset_from_columns([x], [golden.arr("smoothhist.x")])
    s1 = gf.SmoothHistogramPdf("smooth1", h1, V("sm1", 0.4, 0.05, 0, 2))
    check_pdf(golden, "smooth1", s1, d, x)
    assert rel(s1.normalize(), golden["smooth1.normalize"]) < RTOL_NLL
    check_pdf(golden, "smooth0", gf.SmoothHistogramPdf("smooth0", h1, V("sm0", 0.0)), d, x)
    x2 = gf.Observable("x", -4, 6)
    x2.setNumBins(20)
    h2 = gf.BinnedDataSet(x2, y)
    for i, v in enumerate(golden.arr("smoothhist.h2")):
        h2.setBinContent(i, v)
    d2 = dataset_from_columns([x2, y], [golden.arr("smoothhist.x"), golden.arr("smoothhist.y")])
    s2 = gf.SmoothHistogramPdf("smooth2", h2, V("sm2", 0.25, 0.05, 0, 2))
    check_pdf(golden, "smooth2", s2, d2)
    assert rel(s2.normalize(), golden["smooth2.normalize"]) < RTOL_NLL


def test_voigtian_pdf(golden):
    """basic/VoigtianPdf.cu: Re w(z) by the Matta-Reichel series. Wide; narrow (events near the real axis, where the h = 0.53 node
    grid takes over around half-integer Re z); very narrow with events on the peak (|z|^2 < 1e-7: the Pade form); and the two
    degenerate forms. The constructor's (m, s, w) reach the device function as (m, width, sigma), as upstream
    (oracle/ref_driver.cpp case_voigt)."""
    V = gf.Variable
    x = gf.Observable("x", -4, 6)
    d = dataset_from_columns([x], [golden.arr("voigt.x")])
    m = V("vm", 0.8, 0.1, -4, 6)
    a1, b1 = V("va1", 0.9, 0.1, 0, 5), V("vb1", 0.7, 0.1, 0, 5)
    a2, b2, a3, zero = V("va2", 0.001, 0.0001, 0, 5), V("vb2", 0.5, 0.1, 0, 5), V("va3", 0.0002, 0.0001, 0, 5), V("vzero", 0.0)
    check_pdf(golden, "voigt_wide", gf.VoigtianPdf("voigt_wide", x, m, a1, b1), d, x)
    check_pdf(golden, "voigt_narrow", gf.VoigtianPdf("voigt_narrow", x, m, a2, b2), d, x, rtol_evt=1e-11)
    check_pdf(golden, "voigt_tiny", gf.VoigtianPdf("voigt_tiny", x, m, a3, b2), d, x, rtol_evt=1e-11)
    check_pdf(golden, "voigt_bw", gf.VoigtianPdf("voigt_bw", x, m, a1, zero), d, x)
    check_pdf(golden, "voigt_gauss", gf.VoigtianPdf("voigt_gauss", x, m, zero, b1), d, x)


def test_variable_bin_transform():
    """utilities/VariableBinTransform1DPdf.cu: the number of ascending limits |x| has reached. Parity unpinned: upstream's class
    cannot be run (its constructor never calls initialize() and its device function never advances the table cursor, :8-43; on
    its own it segfaults in the reference build), so this is held to its documented meaning: on its own the values are
    proportional to the bin number, and as the mapping function of a MappedPdf it picks the component of that bin."""
    V = gf.Variable
    x = gf.Observable("x", -4, 6)
    X = np.random.default_rng(41).normal(0.5, 1.5, 4000)
    X = X[(X > -4) & (X < 6)][:500]
    d = dataset_from_columns([x], [X])
    limits = [0.5, 1.25, 2.0, 3.5]
    which = np.searchsorted(np.asarray(limits), np.abs(X), side="right")
    vb = gf.VariableBinTransform1DPdf("varbin", x, limits)
    vb.setData(d)
    vb.calculateNLL()
    p = np.asarray(vb.getCompProbsAtDataPoints()[0])
    unit = p[which == 1][0]
    np.testing.assert_allclose(p, which * unit, rtol=1e-14, atol=0)
    m = V("vbm", 0.5, 0.1, -4, 6)
    comps = [gf.GaussianPdf("vb%d" % k, x, m, V("vbs%d" % k, 0.6 + 0.4 * k, 0.1, 0.1, 5)) for k in range(5)]
    mp = gf.MappedPdf("varbin_mapped", gf.VariableBinTransform1DPdf("varbin_map", x, limits), comps)
    mp.setData(d)
    mp.calculateNLL()
    got = np.asarray(mp.getCompProbsAtDataPoints()[0])
    for k, c in enumerate(comps):
        c.setData(d)
        c.calculateNLL()
        want = np.asarray(c.getCompProbsAtDataPoints()[0])
        sel = which == k
        assert sel.any()
        np.testing.assert_allclose(got[sel], want[sel], rtol=RTOL_EVT, atol=0)


@pytest.mark.parametrize("kind,tag", [(1, "dalitzeff"), (2, "dalitzsq")])
def test_dalitz_efficiency_helpers(golden, kind, tag):
    """physics/DalitzVetoPdf.cu (kind 1: two veto bands) and physics/SquareDalitzEffPdf.cu (kind 2) as the efficiency of the
    16-resonance Amp3Body: NLL and per-event values against the reference, where neither helper can be normalised on its own: the grid integral
    writes every dimension's bin centre into the first observable's slot (MetricTaker.cu:73-88, kept here), so m13 stays 0 and no
    grid point is inside the Dalitz plot (oracle/ref_driver.cpp case_dalitzeff)."""
    M = Dalitz(float_masses=False, grid=120, eff_kind=kind)
    m12, m13 = golden.arr("dalitzeff.m12"), golden.arr("dalitzeff.m13")
    d = M.dataset(m12, m13)
    M.signal.setData(d)
    M.signal.setDataSize(len(d))
    nll = M.signal.calculateNLL()
    assert rel(nll, golden[tag + ".nll"]) < RTOL_NLL, (nll, golden[tag + ".nll"])
    wrap = gf.ProdPdf("wrap", [M.signal])
    wrap.setData(d)
    M.signal.setDataSize(len(d))
    probs = np.asarray(wrap.getCompProbsAtDataPoints()[0])
    want = golden.arr(tag + ".prob")
    assert ((probs == 0) == (want == 0)).all() and (want == 0).sum() > 20
    np.testing.assert_allclose(probs, want, rtol=1e-11, atol=0)


def test_empty_and_single_event_datasets():
    """Edge sizes: no event at all (uploading it must not crash; the empty sum is refused like upstream's), one event (a partial first block), and 33 events (one past a warp): the NLL equals -sum(log(prob))."""
    V = gf.Variable
    x = gf.Observable("x", -5, 5)
    g = gf.GaussianPdf("edge_g", x, V("edge_m", 0.4, 0.1, -5, 5), V("edge_s", 1.3, 0.1, 0.1, 5))
    for n in (0, 1, 33):
        X = np.linspace(-2.0, 2.5, n) if n else np.empty(0)
        d = dataset_from_columns([x], [X])
        g.setData(d)
        if n == 0:  # the empty sum is 0, which calculateNLL refuses as the reference does (GooPdf.cu:268-270 "zero NLL")
            with pytest.raises(Exception, match="zero NLL"):
                g.calculateNLL()
            continue
        nll = g.calculateNLL()
        want = -np.log(np.exp(-0.5 * ((X - 0.4) / 1.3) ** 2) / (1.3 * math.sqrt(2 * math.pi)))
        probs = np.asarray(g.getCompProbsAtDataPoints()[0])
        assert len(probs) == n
        np.testing.assert_allclose(-np.log(probs), want, rtol=1e-9, atol=0)  # analytic integral over all x vs the range: tails 1e-4
        assert rel(nll, math.fsum(-np.log(probs))) < RTOL_NLL


def test_misuse_raises_instead_of_crashing(golden):
    """Calls out of order or with the wrong sizes end in an exception with a message, never in a crash or a silent wrong number."""
    V = gf.Variable
    x = gf.Observable("x", -5, 5)
    g = gf.GaussianPdf("mis_g", x, V("mis_m", 0.4, 0.1, -5, 5), V("mis_s", 1.3, 0.1, 0.1, 5))
    with pytest.raises(Exception):
        g.calculateNLL()  # no data yet
    # the amplitude caches follow the dataset: a larger, then a smaller dataset on the same model give what a fresh model gives
    m12, m13 = golden.arr("dalitz.m12"), golden.arr("dalitz.m13")
    M = Dalitz(float_masses=True, grid=120)
    out = []
    for n in (200, len(m12), 50):
        d = M.dataset(m12[:n], m13[:n])
        M.signal.setData(d)
        M.signal.setDataSize(len(d))
        out.append(M.signal.calculateNLL())
    assert rel(out[1], golden["dalitz.nll"]) < RTOL_NLL
    F = Dalitz(float_masses=True, grid=120)
    d = F.dataset(m12[:50], m13[:50])
    F.signal.setData(d)
    F.signal.setDataSize(len(d))
    assert F.signal.calculateNLL() == out[2]
    # the cache is sized from the uploaded events, not from setDataSize's argument: a wrong argument cannot make it too small
    d = M.dataset(m12[:300], m13[:300])
    M.signal.setData(d)
    M.signal.setDataSize(100)
    wrong = M.signal.calculateNLL()
    M.signal.setDataSize(len(d))
    assert M.signal.calculateNLL() == wrong and math.isfinite(wrong)


def test_composite_pdf(golden):
    """combine/CompositePdf.cu: f = h(g(event)); the shell h reads the core's value where the composite's first observable sits in
    the event, neither is normalised inside, the composite is (grid integral). Alone, inside an AddPdf, and with a core of two
    observables (oracle/ref_driver.cpp case_composite)."""
    V = gf.Variable
    x = gf.Observable("x", -5, 5)
    y = gf.Observable("y", -2, 2)
    X, Y = golden.arr("composite.x"), golden.arr("composite.y")
    c0, c1 = V("c0", 0.5), V("c1", 0.8)
    mean, sigma = V("cmean", 0.3, 0.1, -3, 3), V("csigma", 1.5, 0.1, 0.3, 4)

    def check(pdf, data, tag):
        pdf.setData(data)
        nll = pdf.calculateNLL()
        assert gf.last_kernel_name() == "k_eval_sum"
        assert rel(nll, golden["composite.%snll" % tag]) < RTOL_NLL, (tag, nll, golden["composite.%snll" % tag])
        probs = np.asarray(pdf.getCompProbsAtDataPoints()[0])
        np.testing.assert_allclose(probs, golden.arr("composite.%sprob" % tag), rtol=RTOL_EVT, atol=0)
        pdf.setFitControl(gf.UnbinnedNllFit())
        pdf.setData(data)
        assert pdf.calculateNLL() == nll

    core = gf.PolynomialPdf("core", x, [c0, c1])
    shell = gf.GaussianPdf("shell", x, mean, sigma)
    comp = gf.CompositePdf("comp", core, shell)
    d = dataset_from_columns([x], [X])
    check(comp, d, "")
    ex = gf.ExpPdf("cex", x, V("calpha", -0.2, 0.1, -2, 2))
    check(gf.AddPdf("csum", V("cfrac", 0.6, 0.1, 0, 1), comp, ex), d, "sum_")
    k0, zero = V("k0", 0.25), V("kzero", 0)
    core2 = gf.PolynomialPdf("core2", [x, y], [k0, c1, c1, k0], [zero, zero], 1)
    shell2 = gf.GaussianPdf("shell2", x, mean, sigma)
    check(gf.CompositePdf("comp2", core2, shell2), dataset_from_columns([x, y], [X, Y]), "two_")
    with pytest.raises(Exception):
        gf.CompositePdf("nested", core, gf.CompositePdf("inner", core, shell))  # one substitution at a time: rejected on the host


def test_dalitz_values(golden):
    M = Dalitz(float_masses=True, grid=120)
    m12, m13 = golden.arr("dalitz.m12"), golden.arr("dalitz.m13")
    d = M.dataset(m12, m13)
    s = M.signal
    s.setData(d)
    s.setDataSize(len(d))
    nll = s.calculateNLL()
    assert rel(1.0 / s.getNormalization(), golden["dalitz.integral"]) < 1e-12
    assert rel(nll, golden["dalitz.nll"]) < RTOL_NLL
    ref = golden.arr("dalitz.waves")
    for i in range(16):
        w = np.asarray(s.getCachedWave(i))
        r = ref[i, 0::2] + 1j * ref[i, 1::2]
        scale = np.maximum(np.abs(r), 1e-300)
        assert np.max(np.abs(w - r) / scale) < 1e-11, i
    wrap = gf.ProdPdf("wrap", [s])
    wrap.setData(d)
    s.setDataSize(len(d))
    probs = wrap.getCompProbsAtDataPoints()
    np.testing.assert_allclose(probs[0], golden.arr("dalitz.prob"), rtol=1e-11, atol=0)
    assert np.all(np.asarray(probs[0])[-3:] == 0)  # the three events outside the Dalitz boundary
    s.setFitControl(gf.UnbinnedNllFit())
    s.setData(d)
    s.setDataSize(len(d))
    assert s.calculateNLL() == nll
    # coefficients move, masses stay: caches reused
    for v in M.masses + M.widths:
        v.changed = False
    before = gf.launch_count()
    s.updateVariable(M.coefs[4], 0.6)
    s.updateVariable(M.coefs[7], -0.1)
    s.updateParameters()
    assert rel(s.calculateNLL(), golden["dalitz.nll_coef_moved"]) < RTOL_NLL
    assert gf.launch_count() - before == 1, "steady-state call must be exactly one kernel launch"
    assert rel(1.0 / s.getNormalization(), golden["dalitz.integral_coef_moved"]) < 1e-12
    # one mass and one width move: only the affected waves / integrals are redone
    M.masses[0].changed = True
    M.widths[9].changed = True
    s.updateVariable(M.masses[0], 0.77)
    s.updateVariable(M.widths[9], 0.05)
    s.updateParameters()
    assert rel(s.calculateNLL(), golden["dalitz.nll_mass_moved"]) < RTOL_NLL
    assert rel(1.0 / s.getNormalization(), golden["dalitz.integral_mass_moved"]) < 1e-12
    # fit fractions from the cached waves at this point: W_ij summed on the device (gf_amp3_wave_gram), coefficients applied on the
    # host; the reference sums Re(c_i conj(c_j) w_i conj(w_j)) event by event for every ordered pair (Amp3Body.cu:376-428)
    ff = np.asarray(s.fit_fractions())
    want = golden.arr("dalitz.fitfrac")
    assert ff.shape == want.shape == (16, 16)
    np.testing.assert_allclose(ff, want, rtol=1e-11, atol=1e-13 * np.abs(want).max())
    assert abs(ff.sum() - 1.0) < 1e-12


def test_zoo3_resonance_shapes(golden):
    """LASS / FLATTE / GS / Gauss / RhoOmegaMix / symmetrised and unnormalised RBW against the reference's resonances/*.cu"""
    M = Zoo3(grid=90)
    d = M.dataset(golden.arr("zoo3.m12"), golden.arr("zoo3.m13"))
    s = M.signal
    s.setData(d)
    s.setDataSize(len(d))
    nll = s.calculateNLL()
    ref = golden.arr("zoo3.waves")
    for i in range(len(M.resonances)):
        w = np.asarray(s.getCachedWave(i))
        r = ref[i, 0::2] + 1j * ref[i, 1::2]
        scale = np.maximum(np.abs(r), 1e-300)
        assert np.max(np.abs(w - r) / scale) < 1e-11, (i, np.max(np.abs(w - r) / scale))
    assert rel(1.0 / s.getNormalization(), golden["zoo3.integral"]) < 1e-12
    assert rel(nll, golden["zoo3.nll"]) < RTOL_NLL
    for v in M.pars:
        v.changed = False
    byname = {v.name: v for v in M.pars}
    for n, v in (("lass_r_", 3.1), ("f0_g1", 0.18), ("gs_m", 0.79), ("mix_im", 0.004), ("sp_re3", 0.9)):
        # LASS range, FLATTE coupling, GS mass, rho-omega mixing, one spline knot (its second derivatives stay frozen, as upstream)
        byname[n].changed = True
        s.updateVariable(byname[n], v)
    s.updateParameters()
    assert rel(s.calculateNLL(), golden["zoo3.nll_moved"]) < RTOL_NLL
    assert rel(1.0 / s.getNormalization(), golden["zoo3.integral_moved"]) < 1e-12


@pytest.mark.parametrize("tag,three_gauss,mistag", [("td3", False, 0.0), ("td3mis", False, 0.05), ("td3g3", True, 0.05)])
def test_amp3body_td(golden, tag, three_gauss, mistag):
    """Amp3Body_TD (Amp3Body_TD.cu:40-760, SpecialWaveCalculator.cu, SpecialDalitzIntegrator.cu): cached (A_i, B_i) planes, six grid
    integrals per ordered pair, D0 / D0bar mixing shifts by charm tag, mistagged fraction, Truth and ThreeGauss resolution."""
    from models import TD3
    M = TD3(three_gauss=three_gauss)
    d = M.dataset(golden.arr("td3.dtime"), golden.arr("td3.sigmat"), golden.arr("zoo3.m12"), golden.arr("zoo3.m13"), mistag, golden.arr("td3.charmtag"))
    dp = M.dp
    dp.setData(d)
    dp.setDataSize(len(d), 7)
    nll = dp.calculateNLL()
    ref = golden.arr("td3.waves")  # [resonance, 4 * event] = (ai_re, ai_im, bi_re, bi_im)
    for i in range(5):
        a, b = np.asarray(dp.getCachedWaveA(i)), np.asarray(dp.getCachedWaveB(i))
        ra, rb = ref[i][0::4] + 1j * ref[i][1::4], ref[i][2::4] + 1j * ref[i][3::4]
        assert np.max(np.abs(a - ra) / np.maximum(np.abs(ra), 1e-300)) < 1e-11, i
        assert np.max(np.abs(b - rb) / np.maximum(np.abs(rb), 1e-300)) < 1e-11, i
    assert rel(1.0 / dp.getNormalization(), golden[tag + ".integral"]) < 1e-11
    assert rel(nll, golden[tag + ".nll"]) < RTOL_NLL
    wrap = gf.ProdPdf("wrap", [dp])
    wrap.setData(d)
    dp.setDataSize(len(d), 7)
    probs = wrap.getCompProbsAtDataPoints()
    want = golden.arr(tag + ".prob")
    np.testing.assert_allclose(probs[0], want, rtol=1e-10, atol=1e-15 * float(np.median(want)) * 16)
    dp.setFitControl(gf.UnbinnedNllFit())
    dp.setData(d)
    dp.setDataSize(len(d), 7)
    assert dp.calculateNLL() == nll
    for v in (M.rhoM, M.rhoW, M.f0M, M.f0W):
        v.changed = False
    dp.updateVariable(M.xm, 0.0123)
    dp.updateVariable(M.ym, -0.0071)
    dp.updateVariable(M.dx, 0.002)
    dp.updateVariable(M.tau, 0.4153)
    dp.updateVariable(M.c4r, 0.75)
    dp.updateParameters()
    before = gf.launch_count()
    assert rel(dp.calculateNLL(), golden[tag + ".nll_mix_moved"]) < RTOL_NLL
    assert gf.launch_count() - before == 1, "mixing parameters and coefficients move: the NLL kernel only"
    assert rel(1.0 / dp.getNormalization(), golden[tag + ".integral_mix_moved"]) < 1e-11
    M.rhoM.changed = True
    dp.updateVariable(M.rhoM, 0.77)
    dp.updateParameters()
    assert rel(dp.calculateNLL(), golden[tag + ".nll_mass_moved"]) < RTOL_NLL
    assert rel(1.0 / dp.getNormalization(), golden[tag + ".integral_mass_moved"]) < 1e-11


def test_kmatrix_three_and_four_body(golden):
    """The pi pi S-wave K-matrix (resonances/kMatrix.cu:19-121, lineshapes/kMatrix.cu:18-119, kMatrixUtils.cu): five channels, five
    poles, LU inverse of 1 - i adler K rho. Same events and MC sample as the zoo3 / zoo4 cases."""
    from models import KMatrix3, KMatrix4
    M = KMatrix3()
    d = M.dataset(golden.arr("zoo3.m12"), golden.arr("zoo3.m13"))
    s = M.signal
    s.setData(d)
    s.setDataSize(len(d))
    nll = s.calculateNLL()
    w = np.asarray(s.getCachedWave(0))
    ref = golden.arr("kmatrix3.wave")[0]
    r = ref[0::2] + 1j * ref[1::2]
    assert np.max(np.abs(w - r) / np.maximum(np.abs(r), 1e-300)) < 1e-10
    assert rel(nll, golden["kmatrix3.nll"]) < RTOL_NLL
    assert rel(1.0 / s.getNormalization(), golden["kmatrix3.integral"]) < 1e-11
    K = M.K
    K["beta_r"][1].changed = True
    K["s0_prod"].changed = True
    s.updateVariable(K["beta_r"][1], 11.0)
    s.updateVariable(K["s0_prod"], -0.2)
    s.updateParameters()
    assert rel(s.calculateNLL(), golden["kmatrix3.nll_moved"]) < RTOL_NLL
    assert rel(1.0 / s.getNormalization(), golden["kmatrix3.integral_moved"]) < 1e-11
    for mode in ("Gram", "PerEvent"):
        M = KMatrix4()
        dp = M.dp
        dp.setNormMode(getattr(gf.Amp4BodyBase.NormMode, mode))
        names = ("m12", "m34", "cos12", "cos34", "phi")
        mc = np.vstack([golden.arr("zoo4.mc_" + n) for n in names])
        assert mc.shape[1] == int(golden["kmatrix4.mc_events"])
        dp.setNormEvents(mc)
        d = M.dataset([golden.arr("zoo4." + n) for n in names])
        dp.setData(d)
        dp.setDataSize(len(d), 6)
        nll = dp.calculateNLL()
        ref = golden.arr("kmatrix4.amps")
        for a in range(2):
            w = np.asarray(dp.getCachedAmp(a))
            r = ref[:, 2 * a] + 1j * ref[:, 2 * a + 1]
            assert np.max(np.abs(w - r) / np.maximum(np.abs(r), 1e-300)) < 1e-10, a
        assert rel(nll, golden["kmatrix4.nll"]) < RTOL_NLL
        assert rel(1.0 / dp.getNormalization(), golden["kmatrix4.integral"]) < 1e-11
        K = M.K
        K["beta_r"][1].changed = True
        K["s0_prod"].changed = True
        dp.updateVariable(K["beta_r"][1], 11.0)
        dp.updateVariable(K["s0_prod"], -0.2)
        dp.updateParameters()
        assert rel(dp.calculateNLL(), golden["kmatrix4.nll_moved"]) < RTOL_NLL
        assert rel(1.0 / dp.getNormalization(), golden["kmatrix4.integral_moved"]) < 1e-11


def test_gspline_lineshape_self_consistent():
    """Lineshapes::GSpline (lineshapes/GSpline.cu:31-161). Upstream's version cannot be pinned: make_spline_curvatures fills an
    uninitialised Eigen matrix and getSpline reads past the constants it registered (see csrc/gf_amp4.cuh). The device function is
    checked against a numpy restatement of the same formula on the same constants (parity with upstream: unpinned)."""
    from models import Keep
    k = Keep()
    O, V = gf.Observable, gf.Variable
    obs = [O("m12", 0, 3), O("m34", 0, 3), O("cos12", -1, 1), O("cos34", -1, 1), O("phi", -3.5, 3.5), gf.EventNumber("eventNumber", 0, 2147483647)]
    mD0, piP, Km = 1.8645, 0.13957018, 0.493677
    info = gf.DecayInfo4()
    info.meson_radius = 1.5
    info.particle_masses = [mD0, piP, piP, Km, piP]
    knots = [V("k%d" % i, 0.3 + 0.2 * math.sin(0.7 * i) + 0.05 * i) for i in range(8)]
    smin, smax, nb = 0.08, 1.6, 8
    mass, width, radius = V("gs_m", 0.9), V("gs_w", 0.25), 1.5
    L, P, S = gf.Lineshapes, gf.DP4Pair, gf.SF_4Body
    ls = [k(L.GSpline("gspline", mass, width, 0, int(P.M_12), gf.FF.BL2, radius, knots, (smin, smax, nb))), k(L.One("one", mass, width, 0, int(P.M_34)))]
    sf = [k(gf.SpinFactor("SF", S.ONE, mD0, 0, 1, 2, 3))]
    info.add_amplitude(k(gf.Amplitude("A", V("ar", 1.0), V("ai", 0.0), ls, sf, 1)))
    eff = k(gf.PolynomialPdf("constantEff", obs, [V("constantOne", 1)], [V("constantZero", 0)] * 2, 0))
    dp = gf.Amp4Body("gs", obs, info, eff, 0)
    n = 4000
    cols, _ = gf.generatePhaseSpace5(info.particle_masses, n, 3)
    dp.setNormEvents(np.asarray(cols)[:, :500])
    dp.setData(dataset_from_columns(obs, list(np.asarray(cols)) + [np.arange(n, dtype=np.float64)]))
    dp.setDataSize(n, 6)
    dp.calculateNLL()
    got = np.asarray(dp.getCachedAmp(0))
    curv = np.asarray(L.GSpline.curvatures(knots, (smin, smax, nb)))
    c = np.concatenate([curv, np.zeros(2 * nb + 2)])  # reads past the registered block yield zero

    def spline(x, continued):
        spacing = (smax - smin) / (nb - 1.0)
        out = np.empty_like(x)
        lo, hi = x <= smin, x >= smax
        out[lo] = c[0] if continued else 0.0
        out[hi] = c[nb - 1] if continued else 0.0
        mid = ~(lo | hi)
        dx = np.fmod(x[mid] - smin, spacing)
        b = ((x[mid] - smin) / spacing).astype(int)
        x0, x1, f0, f1 = c[b], c[b + 1], c[b + nb], c[b + nb + 1]
        out[mid] = x0 + dx * ((x1 - x0) / spacing - (f1 + 2 * f0) * spacing / 6) + dx * dx * f0 + dx ** 3 * (f1 - f0) / (6 * spacing)
        return out

    m12 = np.asarray(cols)[0]
    s_, m, w = m12 ** 2, mass.value, width.value
    q2 = np.abs(s_ / 4.0 - (piP ** 2 + piP ** 2) / 2.0)
    BF = np.exp(-q2 * radius ** 2 / 2)
    ws, wn = w * spline(s_, True), w * spline(np.array([m * m]), False)[0]
    gamma = m * math.sqrt(m * m + w * w)
    kf = math.sqrt(2 * math.sqrt(2.0) * m * w * gamma / (math.pi * math.sqrt(m * m + gamma)))
    rw = np.zeros_like(ws) if wn == 0 else w * ws / wn
    want = kf * BF / ((m * m - s_) - 1j * m * rw)
    np.testing.assert_allclose(got, want, rtol=1e-11)


def test_dalitz_fit_matches_reference(golden):
    M = Dalitz(float_masses=False, grid=240)
    d = M.dataset(golden.arr("dalitzfit.m12"), golden.arr("dalitzfit.m13"))
    s = M.signal
    s.setData(d)
    s.setDataSize(len(d))
    assert rel(s.calculateNLL(), golden["dalitzfit.nll_start"]) < RTOL_NLL
    fm = gf.FitManager(s)
    fm.setVerbosity(0)
    res = fm.fit()
    assert bool(fm)
    val, err = np.array(golden["dalitzfit.values"]), np.array(golden["dalitzfit.errors"])
    mine = np.array([c.value for c in M.coefs])
    free = err > 0
    pulls = np.abs(mine[free] - val[free]) / err[free]
    assert pulls.max() < 0.01, pulls
    assert rel(res.Fval(), golden["dalitzfit.fval"]) < 1e-9


def test_run_to_run_bit_reproducible():
    from models import dalitz_uniform_events
    a, b = dalitz_uniform_events(200003, 5)
    vals = []
    for _ in range(3):
        M = Dalitz()
        d = M.dataset(a, b)
        M.signal.setData(d)
        M.signal.setDataSize(len(d))
        vals.append(M.signal.calculateNLL())
    assert vals[0] == vals[1] == vals[2]


@pytest.mark.parametrize("mode", ["Gram", "PerEvent"])
def test_dp4_values(golden, mode):
    """mode = how the MC integral is summed: Gram (default; G_ab rebuilt only when a lineshape moved, O(nAmp^2) host sum
    per call) or PerEvent (the reference's association, one pass over the cached MC amplitudes per call)"""
    from models import DP4
    M = DP4(mc_events=0, float_ls=True)
    M.dp.setNormMode(getattr(gf.Amp4BodyBase.NormMode, mode))
    cols = [golden.arr("dp4." + n) for n in ("m12", "m34", "cos12", "cos34", "phi")]
    mc = np.vstack([golden.arr("dp4.mc_" + n) for n in ("m12", "m34", "cos12", "cos34", "phi")])
    assert mc.shape[1] == int(golden["dp4.mc_events"])
    M.dp.setNormEvents(mc)
    d = M.dataset(cols)
    sig, dp = M.signal, M.dp
    assert dp.numLineshapes() == 16 and dp.numSpinFactors() == 16
    sig.setData(d)
    dp.setDataSize(len(d), 6)
    nll = sig.calculateNLL()
    ref = golden.arr("dp4.amps")  # [event, 2*amp]
    for a in range(8):
        w = np.asarray(dp.getCachedAmp(a))
        r = ref[:, 2 * a] + 1j * ref[:, 2 * a + 1]
        assert np.max(np.abs(w - r) / np.maximum(np.abs(r), 1e-300)) < 1e-10, a
    assert rel(1.0 / dp.getNormalization(), golden["dp4.integral"]) < 1e-12
    assert rel(nll, golden["dp4.nll"]) < RTOL_NLL
    fast, slow = fast_vs_interpreter(sig, "k_eval_fast_amp4")
    assert fast == nll and rel(slow, golden["dp4.nll"]) < RTOL_NLL
    probs = sig.getCompProbsAtDataPoints()
    np.testing.assert_allclose(probs[0], golden.arr("dp4.prob"), rtol=1e-10, atol=0)
    sig.setFitControl(gf.UnbinnedNllFit())
    sig.setData(d)
    dp.setDataSize(len(d), 6)
    assert sig.calculateNLL() == nll
    for v in M.lsvars:
        v.changed = False
    sig.updateVariable(M.coefs[0], -0.12)
    sig.updateVariable(M.coefs[5], 0.45)
    sig.updateParameters()
    before = gf.launch_count()
    assert rel(sig.calculateNLL(), golden["dp4.nll_coef_moved"]) < RTOL_NLL
    assert rel(1.0 / dp.getNormalization(), golden["dp4.integral_coef_moved"]) < 1e-12
    # steady state: NLL kernel only (+ the per-event MC integral kernel in PerEvent mode); the background's grid
    # integral is not re-summed because none of its inputs moved
    assert gf.launch_count() - before == (1 if mode == "Gram" else 2)
    M.lsvars[2].changed = True
    sig.updateVariable(M.lsvars[2], 0.8975)
    sig.updateParameters()
    before = gf.launch_count()
    assert rel(sig.calculateNLL(), golden["dp4.nll_kstar_moved"]) < RTOL_NLL
    assert rel(1.0 / dp.getNormalization(), golden["dp4.integral_kstar_moved"]) < 1e-12
    if mode == "Gram":
        # a mass moved: ONE pass over the MC sample (lineshapes -> amplitudes -> G_ab on FP64 MMAs), one over the data, the NLL
        assert gf.launch_count() - before == 3
        kstar_move_fused_vs_separate(sig, dp, M.lsvars[2], golden["dp4.nll_kstar_moved"], golden["dp4.integral_kstar_moved"], 1e-12)


@pytest.mark.parametrize("mode", ["Gram", "PerEvent"])
def test_zoo4_lineshapes_and_spin_factors(golden, mode):
    """One / NonRes / Flatte / GLASS / FOCUS lineshapes and the spin factors DP4 leaves out, against lineshapes/*.cu and SpinFactors.cu"""
    from models import Zoo4
    M = Zoo4()
    dp = M.dp
    dp.setNormMode(getattr(gf.Amp4BodyBase.NormMode, mode))
    cols = [golden.arr("zoo4." + n) for n in ("m12", "m34", "cos12", "cos34", "phi")]
    mc = np.vstack([golden.arr("zoo4.mc_" + n) for n in ("m12", "m34", "cos12", "cos34", "phi")])
    assert mc.shape[1] == int(golden["zoo4.mc_events"])
    dp.setNormEvents(mc)
    d = M.dataset(cols)
    dp.setData(d)
    dp.setDataSize(len(d), 6)
    nll = dp.calculateNLL()
    ref = golden.arr("zoo4.amps")
    for a in range(5):
        w = np.asarray(dp.getCachedAmp(a))
        r = ref[:, 2 * a] + 1j * ref[:, 2 * a + 1]
        err = np.max(np.abs(w - r) / np.maximum(np.abs(r), 1e-300))
        assert err < 1e-10, (a, err)
    assert rel(1.0 / dp.getNormalization(), golden["zoo4.integral"]) < 1e-12
    assert rel(nll, golden["zoo4.nll"]) < RTOL_NLL
    for v in M.pars.values():
        v.changed = False
    for n, v in (("glass_phiF", 0.35), ("f0_m", 0.98)):
        M.pars[n].changed = True
        dp.updateVariable(M.pars[n], v)
    dp.updateParameters()
    assert rel(dp.calculateNLL(), golden["zoo4.nll_moved"]) < RTOL_NLL
    assert rel(1.0 / dp.getNormalization(), golden["zoo4.integral_moved"]) < 1e-12


def test_dp4_fit_matches_reference(golden):
    from models import DP4
    M = DP4(mc_events=0, float_ls=False)
    names = ("m12", "m34", "cos12", "cos34", "phi")
    M.dp.setNormEvents(np.vstack([golden.arr("dp4fit.mc_" + n) for n in names]))
    d = M.dataset([golden.arr("dp4fit." + n) for n in names])
    sig, dp = M.signal, M.dp
    sig.setData(d)
    dp.setDataSize(len(d), 6)
    fm = gf.FitManager(sig)
    fm.setVerbosity(0)
    assert rel(fm.getFCN()(), golden["dp4fit.nll_start"]) < RTOL_NLL
    res = fm.fit()
    assert bool(fm)
    val, err = np.array(golden["dp4fit.values"]), np.array(golden["dp4fit.errors"])
    mine = np.array([c.value for c in M.coefs[:6]])
    pulls = np.abs(mine - val) / err
    assert pulls.max() < 0.01, pulls
    assert rel(res.Fval(), golden["dp4fit.fval"]) < 1e-9

TD_DATA = ("m12", "m34", "cos12", "cos34", "phi", "dtime", "sigmat")
TD_MC = ("m12", "m34", "cos12", "cos34", "phi")


@pytest.mark.parametrize("mode", ["Gram", "PerEvent"])
@pytest.mark.parametrize("tag,three_gauss,shared", [("tddp4", False, False), ("tddp4g3", True, False), ("tddp4sh", False, True)])
def test_tddp4_values(golden, tag, three_gauss, shared, mode):
    """Amp4Body_TD (Amp4Body_TD.cu:167-261,744-799) with TruthResolution / ThreeGaussResolution, and with one amplitude
    object shared by the A and B lists (flag 2). All three cases run on the same data and MC sample."""
    from models import TDDP4
    M = TDDP4(three_gauss=three_gauss, float_ls=True, float_mix=True, shared=shared)
    M.dp.setNormMode(getattr(gf.Amp4BodyBase.NormMode, mode))
    mc = np.vstack([golden.arr("tddp4.mc_" + n) for n in TD_MC])
    assert mc.shape[1] == int(golden["tddp4.mc_events"])
    M.dp.setNormEvents(mc)
    d = M.dataset([golden.arr("tddp4." + n) for n in TD_DATA])
    sig, dp = M.signal, M.dp
    namp = 5 if shared else 6
    assert dp.numLineshapes() == 4 and dp.numSpinFactors() == 6 and dp.numAmplitudes() == namp
    sig.setData(d)
    dp.setDataSize(len(d), 8)
    nll = sig.calculateNLL()
    ref = golden.arr(tag + ".amps")  # [event, 2*amp]
    for a in range(namp):
        w = np.asarray(dp.getCachedAmp(a))
        r = ref[:, 2 * a] + 1j * ref[:, 2 * a + 1]
        assert np.max(np.abs(w - r) / np.maximum(np.abs(r), 1e-300)) < 1e-10, a
    # The time integral tau/(1-y^2) (|A|^2+|B|^2) + tau/(1+x^2) (|A|^2-|B|^2) - ... cancels the dominant |B|^2 sum to
    # O(x^2+y^2) on the host (TruthResolution.cu:56-70), which amplifies the ~1e-16 association differences of the four MC
    # sums by ~|B|^2/|result| ~ 1e3-1e4: the integral is compared at 1e-11, the NLL (which sees N log(integral)) at 1e-12.
    RTOL_INT = 1e-11
    assert rel(1.0 / dp.getNormalization(), golden[tag + ".integral"]) < RTOL_INT
    assert rel(nll, golden[tag + ".nll"]) < RTOL_NLL
    fast, slow = fast_vs_interpreter(sig, "k_eval_fast_amp4_td")
    assert fast == nll and rel(slow, golden[tag + ".nll"]) < RTOL_NLL
    probs = sig.getCompProbsAtDataPoints()
    # per-event values: 1e-10 relative; events deep in the resolution tail (p ~ 1e-6 of the typical value) are sums of
    # cancelling erfc/exp terms, so they additionally get an absolute allowance of a few ulp of the typical value
    want = golden.arr(tag + ".prob")
    np.testing.assert_allclose(probs[0], want, rtol=1e-10, atol=1e-15 * float(np.median(want)) * 16)
    sig.setFitControl(gf.UnbinnedNllFit())
    sig.setData(d)
    dp.setDataSize(len(d), 8)
    assert sig.calculateNLL() == nll
    sig.updateParameters()
    assert rel(sig.calculateNLL(), golden[tag + ".nll_after_update"]) < RTOL_NLL
    for v in M.lsvars:
        v.changed = False
    sig.updateVariable(M.mix[0], 0.4153)
    sig.updateVariable(M.mix[1], 0.0123)
    sig.updateVariable(M.mix[2], -0.0071)
    sig.updateParameters()
    before = gf.launch_count()
    assert rel(sig.calculateNLL(), golden[tag + ".nll_mix_moved"]) < RTOL_NLL
    assert rel(1.0 / dp.getNormalization(), golden[tag + ".integral_mix_moved"]) < RTOL_INT
    # steady state: NLL kernel only (+ the four-sum MC integral kernel in PerEvent mode)
    assert gf.launch_count() - before == (1 if mode == "Gram" else 2)
    M.lsvars[2].changed = True
    sig.updateVariable(M.lsvars[2], 0.8975)
    sig.updateParameters()
    before = gf.launch_count()
    assert rel(sig.calculateNLL(), golden[tag + ".nll_kstar_moved"]) < RTOL_NLL
    assert rel(1.0 / dp.getNormalization(), golden[tag + ".integral_kstar_moved"]) < RTOL_INT
    if mode == "Gram":
        assert gf.launch_count() - before == 3
        kstar_move_fused_vs_separate(sig, dp, M.lsvars[2], golden[tag + ".nll_kstar_moved"], golden[tag + ".integral_kstar_moved"], RTOL_INT)


@pytest.mark.parametrize("tag,three_gauss", [("tddp4w", False), ("tddp4wg3", True)])
def test_tddp4_weighted_norm_events(golden, tag, three_gauss):
    """NormEvents_4Body_Weighted (detail/NormEvents_4Body_Weighted.cu, NormIntegrator_TD_Weighted.cu:15-122): every MC event has
    its own decay time, resolution and weight; the integral is sum_m w_m pdf(x_m, t_m) over the sum of the initial weights -
    which upstream forms with an integer accumulator (sum of floor(w)), reproduced and pinned here."""
    from models import TDDP4
    mc = [golden.arr("tddp4.mc_" + n) for n in TD_MC]
    t, sg, w = golden.arr("tddp4w.mc_dtime"), golden.arr("tddp4w.mc_sigma"), golden.arr("tddp4w.mc_weight")
    ne = [gf.NormEvents_4Body_Weighted(*mc, t, sg, w)]
    M = TDDP4(three_gauss=three_gauss, float_ls=True, float_mix=True, norm_events=ne)
    d = M.dataset([golden.arr("tddp4." + n) for n in TD_DATA])
    sig, dp = M.signal, M.dp
    assert dp.hasWeightedNormEvents() and dp.getNumAccNormEvents() == len(w)
    assert dp.getSumInitNormEventWeights() == golden[tag + ".sum_init_weights"] == np.floor(w).sum()
    sig.setData(d)
    dp.setDataSize(len(d), 8)
    assert rel(sig.calculateNLL(), golden[tag + ".nll"]) < RTOL_NLL
    assert rel(1.0 / dp.getNormalization(), golden[tag + ".integral"]) < 1e-11
    for v in M.lsvars:
        v.changed = False
    sig.updateVariable(M.mix[0], 0.4153)
    sig.updateVariable(M.mix[1], 0.0123)
    sig.updateVariable(M.mix[2], -0.0071)
    sig.updateParameters()
    before = gf.launch_count()
    assert rel(sig.calculateNLL(), golden[tag + ".nll_mix_moved"]) < RTOL_NLL
    assert rel(1.0 / dp.getNormalization(), golden[tag + ".integral_mix_moved"]) < 1e-11
    assert gf.launch_count() - before == 2  # the weighted MC integral (it depends on tau, x, y event by event) and the NLL
    M.lsvars[2].changed = True
    sig.updateVariable(M.lsvars[2], 0.8975)
    sig.updateParameters()
    assert rel(sig.calculateNLL(), golden[tag + ".nll_kstar_moved"]) < RTOL_NLL
    assert rel(1.0 / dp.getNormalization(), golden[tag + ".integral_kstar_moved"]) < 1e-11


def test_tddp4_fit_matches_reference(golden):
    from models import TDDP4
    M = TDDP4(three_gauss=False, float_ls=False, float_mix=True)
    M.dp.setNormEvents(np.vstack([golden.arr("tddp4fit.mc_" + n) for n in TD_MC]))
    d = M.dataset([golden.arr("tddp4fit." + n) for n in TD_DATA])
    sig, dp = M.signal, M.dp
    M.mix[0].value, M.mix[1].value, M.mix[2].value = 0.40, 0.01, 0.0
    sig.setData(d)
    dp.setDataSize(len(d), 8)
    fm = gf.FitManager(sig)
    fm.setVerbosity(0)
    assert rel(fm.getFCN()(), golden["tddp4fit.nll_start"]) < RTOL_NLL
    res = fm.fit()
    assert bool(fm)
    val, err = np.array(golden["tddp4fit.values"]), np.array(golden["tddp4fit.errors"])
    mine = np.array([c.value for c in M.mix])
    pulls = np.abs(mine - val) / err
    assert pulls.max() < 0.01, pulls
    assert rel(res.Fval(), golden["tddp4fit.fval"]) < 1e-9


def test_tddp4_generated_norm_sample():
    """the constructor's own normalisation sample: phase space -> accept-reject -> compacted accepted events
    (NormEvents_4Body_Base.cu:24-78), deterministic in the seed"""
    from models import TDDP4
    A, B, C = TDDP4(mc_to_gen=20000, norm_seed=7), TDDP4(mc_to_gen=20000, norm_seed=7), TDDP4(mc_to_gen=20000, norm_seed=8)
    a, b, c = A.dp.getNormEvents(), B.dp.getNormEvents(), C.dp.getNormEvents()
    assert a.shape[0] == 5 and 0 < a.shape[1] < 20000 and A.dp.getNumAccNormEvents() == a.shape[1]
    assert np.array_equal(a, b) and (a.shape != c.shape or not np.array_equal(a, c))
    assert A.dp.getSumInitNormEventWeights() == a.shape[1]


def _binned(x, values):
    b = gf.BinnedDataSet(x)
    for v in values:
        x.value = float(v)
        b.addEvent()
    return b


def test_binned_fits(golden):
    """BinnedDataSet upload (PdfBase.cu:439-527) and the binned metrics BinAvg / BinWithError / Chisq
    (MetricPointer.cu:23-58) against the reference, plus binned fits (tests/simple/BinningTest.cpp)."""
    V = gf.Variable
    x = gf.Observable("x", 0, 10)
    x.numbins = 100
    b = _binned(x, golden.arr("binned.exp_x"))
    alpha = V("alpha", -2, 0.1, -10, 10)
    pdf = gf.ExpPdf("exppdf", x, alpha)
    pdf.setFitControl(gf.BinnedNllFit())
    pdf.setData(b)
    assert rel(pdf.calculateNLL(), golden["binned.exp_nll"]) < RTOL_NLL
    pdf.setFitControl(gf.BinnedChisqFit())
    pdf.setData(b)
    assert rel(pdf.calculateNLL(), golden["binned.exp_chisq"]) < RTOL_NLL
    pdf.setFitControl(gf.BinnedErrorFit())
    for i in range(b.getNumBins()):
        b.setBinError(i, np.sqrt(max(1.0, b.getBinContent(i))))
    pdf.setData(b)
    assert rel(pdf.calculateNLL(), golden["binned.exp_binerr"]) < RTOL_NLL
    pdf.setFitControl(gf.BinnedNllFit())
    pdf.setData(b)
    fm = gf.FitManager(pdf)
    fm.setVerbosity(0)
    fm.fit()
    assert bool(fm) == bool(golden["binned.exp_fit_valid"])
    assert abs(alpha.value - golden["binned.exp_fit_alpha"]) < 0.01 * golden["binned.exp_fit_alpha_err"]
    x = gf.Observable("x", -10, 10)
    x.numbins = 200
    b = _binned(x, golden.arr("binned.gauss_x"))
    m, sg = V("m", 1.0, 0.1, -10, 10), V("s", 2.0, 0.1, 0.5, 10)
    pdf = gf.GaussianPdf("gauss", x, m, sg)
    pdf.setFitControl(gf.BinnedNllFit())
    pdf.setData(b)
    assert rel(pdf.calculateNLL(), golden["binned.gauss_nll"]) < RTOL_NLL
    pdf.setFitControl(gf.BinnedChisqFit())
    pdf.setData(b)
    assert rel(pdf.calculateNLL(), golden["binned.gauss_chisq"]) < RTOL_NLL
    fm = gf.FitManager(pdf)
    fm.setVerbosity(0)
    fm.fit()
    val, err = np.array(golden["binned.gauss_fit"]), np.array(golden["binned.gauss_fit_err"])
    assert np.max(np.abs(np.array([m.value, sg.value]) - val) / err) < 0.01


def test_two_gpu_sharded_parity():
    """events and MC normalisation sample sharded over two GPUs, sums all-reduced over NCCL inside the library: every
    golden NLL / integral is reproduced (tests/mgpu_worker.py). Needs two visible GPUs."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    here = os.path.dirname(os.path.abspath(__file__))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29541", os.path.join(here, "mgpu_worker.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MGPU_OK" in r.stdout, r.stdout[-1500:] + r.stderr[-3000:]


def test_cpp_example_exponential():
    # the reference's examples/exponential flow compiled against the drop-in C++ headers (examples/exponential.cpp)
    import os
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples", "bin", "exponential")
    if not os.path.exists(exe):
        pytest.skip("examples/bin/exponential not built")
    r = subprocess.run([exe, "--nosplash", "-q"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("name", ["ref_exponential", "ref_product"])
def test_reference_example_sources_unmodified(name):
    """The reference's own examples/exponential/exponential.cpp and examples/product/product.cpp, compiled UNMODIFIED (where they
    lie in the reference checkout, `make examples`) against this repo's headers and linked to its libraries: the program's own exit
    status is the check (exponential.cpp returns 1 unless the fitted alpha is within 1 % of -1, then the Minuit status)."""
    import os
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples", "bin", name)
    if not os.path.exists(exe):
        pytest.skip("examples/bin/%s not built (needs the reference checkout at build time)" % name)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


def test_generate_sig_three_and_four_body():
    """Amp3Body::GenerateSig / Amp4Body::GenerateSig (Amp3Body.cu:430-547, Amp4Body.cu:560-677): the returned weight is the
    phase-space weight times the un-normalised PDF value at the generated point; checked against the cached waves /
    amplitudes of those very events, which the call leaves installed as the PDF's dataset (as upstream does)."""
    from models import DP4
    M = Dalitz(float_masses=False, grid=120)
    s = M.signal
    n = 20000
    P, V, w, flags = s.GenerateSig(n, 11)
    assert P.shape == (3, n, 3) and V.shape == (3, n) and w.shape == (n,) and flags.shape == (n,)
    masses = [1.86484, 0.1349766, 0.13957018, 0.13957018]
    P4, w_ps = gf.generatePhaseSpaceParticles(masses, n, 11, 0)
    np.testing.assert_allclose(P, P4[:, :, :3], rtol=1e-10, atol=1e-12)  # GPU stream == host restatement of the stream (libm aside)
    m12 = (P4[0, :, 0] + P4[1, :, 0]) ** 2 - ((P4[0, :, 1:] + P4[1, :, 1:]) ** 2).sum(axis=1)
    np.testing.assert_allclose(V[0], m12, rtol=1e-10)
    np.testing.assert_allclose(V.sum(axis=0), sum(m * m for m in masses), rtol=1e-12)  # m12^2 + m23^2 + m13^2
    coef = np.array([M.coefs[2 * i].value + 1j * M.coefs[2 * i + 1].value for i in range(16)])
    tot = sum(coef[i] * np.asarray(s.getCachedWave(i)) for i in range(16))
    np.testing.assert_allclose(w, w_ps * np.abs(tot) ** 2, rtol=1e-10)
    frac = flags.mean()
    assert abs(frac - w.mean() / w.max()) < 5 * math.sqrt(frac * (1 - frac) / n) + 1e-3
    # a second batch continues the stream
    s.setGenerationOffset(n)
    P2, _, _, _ = s.GenerateSig(100, 11)
    # the GPU generator draws the host generator's stream (same counter-based numbers; libm differences only)
    np.testing.assert_allclose(P2, gf.generatePhaseSpaceParticles(masses, 100, 11, n)[0][:, :, :3], rtol=1e-11, atol=1e-13)
    # the model can still be fitted to a dataset afterwards
    acc = np.flatnonzero(flags)
    d = M.dataset(V[0][acc], V[2][acc])
    s.setData(d)
    s.setDataSize(len(d))
    assert np.isfinite(s.calculateNLL())

    D = DP4(mc_events=20000, float_ls=False)
    n4 = 5000
    P, V, w, flags = D.dp.GenerateSig(n4, 5)
    assert P.shape == (4, n4, 4) and V.shape == (5, n4)
    masses4 = [1.8645, 0.13957018, 0.13957018, 0.493677, 0.13957018]
    _, w_ps = gf.generatePhaseSpaceParticles(masses4, n4, 5, 0)
    cols, _ = gf.generatePhaseSpace5(masses4, n4, 5)
    np.testing.assert_allclose(V, np.asarray(cols), rtol=1e-9, atol=1e-9)
    # info.amplitudes order is (KF, AS, AP, AD, KK, K1R, A1R, A1RD); coefs were created AS, AP, AD, KF, ...
    order = [3, 0, 1, 2, 4, 5, 6, 7]
    coef = np.array([D.coefs[2 * k].value + 1j * D.coefs[2 * k + 1].value for k in order])
    tot = sum(coef[a] * np.asarray(D.dp.getCachedAmp(a)) for a in range(8))
    np.testing.assert_allclose(w, w_ps * np.abs(tot) ** 2, rtol=1e-10)
    assert 0 < flags.mean() < 1


def test_device_phase_space_matches_host_stream():
    """k_genbod (csrc/gf_gen.cu) against the host restatement of the same GENBOD stream: four-vectors, weights and the
    five-variable form (detail/Dim5.h:19-111), for three and four daughters and with an offset into the stream."""
    for masses in ([1.8645, 0.13957018, 0.13957018, 0.493677, 0.13957018], [1.86484, 0.13957018, 0.13957018, 0.1349766]):
        n = 20000
        Pd, wd, v5 = gf.devicePhaseSpace(masses, n, 17, 0)
        Ph, wh = gf.generatePhaseSpaceParticles(masses, n, 17, 0)
        np.testing.assert_allclose(Pd, Ph, rtol=1e-10, atol=1e-12)
        np.testing.assert_allclose(wd, wh, rtol=1e-11)
        if len(masses) == 5:
            cols, w5 = gf.generatePhaseSpace5(masses, n, 17)
            np.testing.assert_allclose(v5, np.asarray(cols), rtol=1e-9, atol=1e-9)
        Po, wo, _ = gf.devicePhaseSpace(masses, 500, 17, 1500)
        assert np.array_equal(Po, Pd[:, 1500:2000]) and np.array_equal(wo, wd[1500:2000])  # counter-based: batch-independent


def test_generate_sig_time_dependent():
    """Amp4Body_TD::GenerateSig (Amp4Body_TD.cu:825-1018): unweighted phase space, decay times under exp(-gammamin t), the PDF at
    t = 0 as `weights`, accept/reject against the envelope. Properties: deterministic in (seed, offset); the returned weights are
    the un-normalised PDF at t = 0; the accepted decay times follow the model's own time dependence."""
    from models import TDDP4
    M = TDDP4(three_gauss=False, float_ls=False, float_mix=True, mc_to_gen=20000)
    n = 2000000
    P, V, w, flags = M.dp.GenerateSig(n, 23)
    nacc = V.shape[1]
    assert 0.01 * n < nacc < n and P.shape == (4, nacc, 4) and V.shape == (6, nacc) and w.shape == (nacc,) and flags.shape == (nacc,)
    assert np.all(V[5] > 0) and np.all(w >= 0)
    M.dp.setMaxWeight(0)
    P2, V2, w2, f2 = M.dp.GenerateSig(n, 23)
    assert np.array_equal(V, V2) and np.array_equal(w, w2) and np.array_equal(flags, f2)
    # the PDF at t = 0, evaluated independently through the ordinary path
    d = M.dataset([V[k] for k in range(5)] + [np.zeros(nacc), np.zeros(nacc)])
    M.signal.setData(d)
    M.dp.setDataSize(nacc, 8)
    M.dp.setFitControl(gf.EvalFit())
    M.dp.normalize()
    direct = np.asarray(M.dp.evaluate_with_metric())
    M.dp.setFitControl(gf.UnbinnedNllFit())
    np.testing.assert_allclose(w, direct, rtol=1e-10)
    # the accept/reject decision, recomputed from its definition (exp_functor, Amp4Body_TD.cu:76-117): u x envelope < pdf(t), with
    # the envelope exp(-gammamin t) x 1.1 x max pdf(t = 0) and the PDF at the drawn times from the ordinary evaluation path
    tau, y = M.mix[0].value, M.mix[2].value
    gammamin = 1.0 / tau - abs(y) / tau
    assert abs(V[5].mean() * gammamin - 1.0) < 4.0 / math.sqrt(nacc)  # candidates follow exp(-gammamin t)
    d = M.dataset([V[k] for k in range(5)] + [V[5], np.zeros(nacc)])
    M.signal.setData(d)
    M.dp.setDataSize(nacc, 8)
    M.dp.setFitControl(gf.EvalFit())
    M.dp.normalize()
    at_t = np.asarray(M.dp.evaluate_with_metric())
    M.dp.setFitControl(gf.UnbinnedNllFit())
    u = np.asarray(gf.counterUniform(23, np.arange(nacc), 4000))
    want = u * np.exp(-V[5] * gammamin) * 1.1 * w.max() < at_t
    assert (want != flags).sum() <= 2 and 0.003 < flags.mean() < 0.95
    assert M.dp.getMaxWeight() == 1.1 * w.max()


def test_generate_sig_with_an_empty_normalisation_sample():
    """python/examples/generate_sig_example.py of the reference builds Amp4Body_TD with ONE normalisation candidate because it only
    generates: when the accept-reject step keeps none, GenerateSig still works (it never integrates: Amp4Body_TD.cu:946-950) and a
    later normalize() says why it cannot."""
    from models import TDDP4
    M = None
    for seed in range(1, 40):  # one candidate is accepted with its phase-space weight as probability: find a seed that rejects it
        M = TDDP4(three_gauss=False, float_ls=False, float_mix=True, mc_to_gen=1, norm_seed=seed)
        P, V, w, flags = M.dp.GenerateSig(20000, 5)
        assert V.shape[0] == 6 and V.shape[1] == w.shape[0] > 100 and np.all(np.isfinite(w))
        d = M.dataset([V[k] for k in range(5)] + [V[5], np.zeros(V.shape[1])])
        M.signal.setData(d)
        M.dp.setDataSize(V.shape[1], 8)
        try:
            M.dp.normalize()
        except gf.GeneralError as e:
            assert "no normalisation event was accepted" in str(e)
            return
    pytest.fail("every seed accepted its single normalisation candidate")
