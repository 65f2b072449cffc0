"""The flows of the reference's own Python tests (python/tests/test_exp.py, test_min_conv.py, test_comb_conv.py — SURVEY 7
step 10) run against goofit_b200, plus the goofit.minuit2 module driving a GooFit likelihood. Written against the same public
names (`from goofit_b200 import *` stands for `from goofit import *`); nothing is copied from the reference's files."""
import numpy as np
import pytest

import goofit_b200 as gf
from goofit_b200.minuit2 import MnMigrad, MnUserParameters

pytestmark = pytest.mark.gpu


def _exp_sample(n=100000, seed=5):
    return np.random.default_rng(seed).exponential(size=n)


def test_exp_event_by_event_and_fit_manager():
    xdata = _exp_sample(20000)
    xvar = gf.Observable("xvar", 0, np.max(xdata) + 1)
    data = gf.UnbinnedDataSet(xvar)
    for v in xdata:
        xvar.value = v
        data.addEvent()
    alpha = gf.Variable("alpha", -2, 0.1, -10, 10)
    exppdf = gf.ExpPdf("exppdf", xvar, alpha)
    exppdf.setData(data)
    fitter = gf.FitManager(exppdf)
    fitter.setVerbosity(0)
    fitter.fit()
    assert abs(alpha.value + 1) < 0.03
    assert alpha.error < 0.05


def test_exp_from_numpy_and_fit_to():
    xdata = _exp_sample()
    xvar = gf.Observable("xvar", 0, np.max(xdata) + 1)
    data = gf.UnbinnedDataSet(xvar)
    data.from_numpy(xdata.reshape(1, -1))
    alpha = gf.Variable("alpha", -2, 0.1, -10, 10)
    exppdf = gf.ExpPdf("exppdf", xvar, alpha)
    exppdf.fitTo(data, 0)
    assert abs(alpha.value + 1) < 0.01
    assert alpha.error < 0.05


def test_exp_matrix_round_trip_and_getitem():
    xdata = _exp_sample().reshape([1, -1])
    xvar = gf.Observable("xvar", 0, np.max(xdata) + 1)
    data = gf.UnbinnedDataSet(xvar)
    data.from_matrix(xdata, False)
    np.testing.assert_array_equal(xdata, data.to_matrix())
    np.testing.assert_array_equal(xdata, data.to_numpy())
    alpha = gf.Variable("alpha", -2, 0.1, -10, 10)
    exppdf = gf.ExpPdf("exppdf", xvar, alpha)
    exppdf.fitTo(data, 0)
    assert abs(alpha.value + 1) < 0.01
    small = np.array([[1.0, 3.0, 4.0, 5.0, 9.0]])
    xv = gf.Observable("xvar", 0, 10)
    d2 = gf.UnbinnedDataSet(xv)
    d2.from_matrix(small, False)
    assert d2[0, 0] == 1.0 and d2[0, 1] == 3.0 and d2[0, -1] == 9.0
    np.testing.assert_array_equal(small[0], d2[0, :])


def test_minimal_convolution_probabilities_on_a_grid():
    xvar = gf.Observable("xvar", 1690, 1890)
    xvar.setNumBins(200)
    pureExp = gf.ExpPdf("pureExp", xvar, gf.Variable("expAlpha", 0.4676863526561826), gf.Variable("expX0", 1860.0))
    gauss = gf.GaussianPdf("gauss", xvar, gf.Variable("zero", 0.0), gf.Variable("sigma", 3.4, 0.1, 0.1, 5))
    convolution = gf.ConvolutionPdf("convolution", xvar, pureExp, gauss)
    convolution.setIntegrationConstants(1695, 1885, 0.1)
    grid = convolution.makeGrid()
    assert grid.to_matrix().flatten().shape == (200,)
    convolution.setData(grid)
    probs = np.asarray(convolution.getCompProbsAtDataPoints())
    assert probs.shape == (3, 200)
    assert np.sum(probs, axis=1) == pytest.approx(np.array([1, 1, 0.0]))


def test_convolution_of_an_added_model():
    xvar = gf.Observable("xvar", 1690, 1890)
    xvar.setNumBins(200)
    Exp_A = gf.ExpPdf("Exp_A", xvar, gf.Variable("expAlpha_A", 0.4676863526561826), gf.Variable("expX0_A", 1860.0))
    Exp_B = gf.ExpPdf("Exp_B", xvar, gf.Variable("expAlpha_B", 0.02429336729138952), gf.Variable("expX0_B", 1860.0))
    added = gf.AddPdf("photosShape", [gf.Variable("fracA", 0.5), gf.Variable("fracB", 0.5)], [Exp_A, Exp_B])
    grid = added.makeGrid()
    added.setData(grid)
    assert len(added.getCompProbsAtDataPoints()) == 3
    gauss = gf.GaussianPdf("gauss", xvar, gf.Variable("zero", 0.0), gf.Variable("sigma", 3.4, 0.1, 0.1, 5))
    convolution = gf.ConvolutionPdf("convolution", xvar, added, gauss)
    grid = convolution.makeGrid()
    convolution.setIntegrationConstants(1695, 1885, 0.1)
    convolution.setData(grid)
    probs = convolution.getCompProbsAtDataPoints()
    assert len(probs) == 3 and len(probs[0]) == 200
    assert np.all(np.isfinite(np.asarray(probs)))


def test_minuit2_module_drives_a_goofit_likelihood():
    """goofit.minuit2 (python/Minuit2/*.cpp): MnMigrad over GooFit's own FCN gives what FitManager gives"""
    xdata = _exp_sample(50000, 9)
    xvar = gf.Observable("xvar", 0, np.max(xdata) + 1)
    data = gf.UnbinnedDataSet(xvar)
    data.from_numpy(xdata.reshape(1, -1))
    alpha = gf.Variable("alpha", -2, 0.1, -10, 10)
    pdf = gf.ExpPdf("exppdf", xvar, alpha)
    pdf.setData(data)
    fm = gf.FitManager(pdf)
    fcn = fm.getFCN()
    upar = MnUserParameters()
    upar.Add("alpha", -2.0, 0.1, -10.0, 10.0)
    migrad = MnMigrad(fcn, upar)
    res = migrad()
    assert res.IsValid()
    a1, e1 = res.UserState().Value(0), res.UserState().Error(0)
    alpha.value = -2.0
    fm2 = gf.FitManager(pdf)
    fm2.setVerbosity(0)
    r2 = fm2.fit()
    assert abs(a1 - alpha.value) < 1e-6 and abs(e1 - alpha.error) < 1e-6 * max(1.0, e1)
    assert abs(res.Fval() - r2.Fval()) < 1e-6 * abs(r2.Fval())
    assert "Minimum" in str(res) or "minimum" in str(res).lower()


def test_module_level_info_functions(capsys):
    """`from goofit import *; print_goofit_info()` — the first two lines of every script under the reference's python/examples
    (python/goofit/Application.cpp, Version.cpp): the functions and constants exist at module level under both package names."""
    import goofit
    ns = {}
    exec("from goofit import *", ns)
    ns["print_goofit_info"]()
    out = capsys.readouterr().out
    assert "GooFit" in out and "SMs" in out, out
    info = goofit.goofit_info(0)
    assert "sm_100a" in info and goofit.GOOFIT_VERSION in info
    assert goofit.__version__ == goofit.GOOFIT_VERSION_TUPLE == (2, 3, 0) and goofit.GOOFIT_DEVICE == "CUDA" and goofit.ROOT_FOUND == 0
    assert ns["DalitzPlotPdf"] is gf.Amp3Body and ns["TddpPdf"] is gf.Amp3Body_TD
    goofit.set_floating_exceptions()
    goofit.print_splash()
