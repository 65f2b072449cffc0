"""Parity AT BASELINE SCALE: the CUDA path against the unmodified reference (oracle/_ref) run on this very
box, on identical inputs of 10^6 - 10^7 events. The truth value at these sizes is a compensated (Neumaier) sum of the
reference's OWN per-event terms (SURVEY 8(c) "oracle summation order"): both the reference's plain OpenMP reduce and the
CUDA tree reduction are reported against it; the CUDA value must agree within 1e-12 relative, fitted parameters within
0.01 sigma of the reference fit. Needs a GPU and the prebuilt oracle/_ref (which travels to the GPU box)."""
import shutil

import numpy as np
import pytest

import goofit_b200 as gf

import scale
from models import DP4, TDDP4, Dalitz, ZachFull, dalitz_uniform_events, dataset_from_columns

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not scale.have_reference(), reason="oracle/_ref not built")]

TOL = 1e-12


def _movers(fm, variables):
    params = fm.getParams()
    return np.array(params.make_minuit_vector()), [v.fitter_index for v in variables]


def _check_schedule(fcn, v0, idx, ref, steps, warmup):
    """evaluate the likelihood at the points the reference arm visited and compare with its values"""
    out = {}
    last = None
    for k in range(warmup + steps):
        last = fcn(list(scale.mover_vector(v0, idx, k)))
    out["last_vs_neumaier"] = scale.rel(last, ref["neumaier_last_nll"])
    out["last_vs_ref_reduce"] = scale.rel(last, ref["last_nll"])
    out["ref_reduce_vs_neumaier"] = scale.rel(ref["last_nll"], ref["neumaier_last_nll"])
    return out, last


@pytest.mark.parametrize("n", [1_000_000, 10_000_000])
def test_dalitz_nll_at_scale(n):
    m12, m13 = dalitz_uniform_events(n, 2024)
    d = scale.write_columns(scale.scratch_dir("dalitz"), {"m12": m12, "m13": m13})
    try:
        ref = scale.run_reference("dalitz", 0, 2, 1, data=d, neumaier=True)
    finally:
        shutil.rmtree(d, ignore_errors=True)
    assert ref["n_events"] == n
    M = Dalitz(float_masses=False, grid=240)
    data = M.dataset(m12, m13)
    M.signal.setData(data)
    M.signal.setDataSize(len(data))
    first = M.signal.calculateNLL()
    assert scale.rel(first, ref["first_nll"]) < TOL, (first, ref["first_nll"])
    fm = gf.FitManager(M.signal)
    v0, idx = _movers(fm, M.coefs[2:])
    r, last = _check_schedule(fm.getFCN(), v0, idx, ref, 2, 1)
    print("dalitz n=%d: %s (reference on %d threads, %.1f ms/call)" % (n, r, ref["threads"], ref["ms_per_step"]))
    assert r["last_vs_neumaier"] < TOL, r
    # run-to-run bit identity at this size (grid-stride accumulation + 592-partial last-block tree)
    again = fm.getFCN()(list(scale.mover_vector(v0, idx, 2)))
    assert again == last


@pytest.mark.parametrize("mode", ["Gram", "PerEvent"])
def test_dp4_nll_at_scale(mode):
    n = n_mc = 1_000_000
    data_cols, _ = scale.phase_space(gf, n, 4242)
    mc = scale.unweighted_phase_space(gf, n_mc, 99)
    cols = {nm: data_cols[k] for k, nm in enumerate(scale.NAMES5)}
    cols.update({"mc_" + nm: mc[k] for k, nm in enumerate(scale.NAMES5)})
    d = scale.write_columns(scale.scratch_dir("dp4"), cols)
    try:
        ref = scale.run_reference("dp4", 0, 2, 1, data=d, neumaier=True, lsmove=True)
    finally:
        shutil.rmtree(d, ignore_errors=True)
    assert ref["n_events"] == n and ref["n_mc"] == n_mc
    M = DP4(mc_events=0, float_ls=True)
    M.dp.setNormMode(getattr(gf.Amp4BodyBase.NormMode, mode))
    M.dp.setNormEvents(mc)
    data = M.dataset(list(data_cols))
    M.signal.setData(data)
    M.dp.setDataSize(len(data), 6)
    first = M.signal.calculateNLL()
    assert scale.rel(first, ref["first_nll"]) < TOL, (first, ref["first_nll"])
    fm = gf.FitManager(M.signal)
    v0, idx = _movers(fm, M.coefs[:6])
    fcn = fm.getFCN()
    for k in range(3):
        last = fcn(list(scale.mover_vector(v0, idx, k)))
    # the reference then moved the K*(892) mass four times and came back to the last steady-state point
    li = M.lsvars[2].fitter_index
    v = scale.mover_vector(v0, idx, 2)
    for k in range(4):
        v[li] = v0[li] * (1.0 + 1e-4 * (k + 1))
        ls_last = fcn(list(v))
    v[li] = v0[li]
    back = fcn(list(v))
    res = {"last": scale.rel(back, ref["neumaier_last_nll"]), "lsmove": scale.rel(ls_last, ref["lsmove_last_nll"]),
           "ref_reduce_vs_neumaier": scale.rel(ref["last_nll"], ref["neumaier_last_nll"])}
    print("dp4 %s: %s (reference %.1f ms/call, %.1f ms when the K* mass moves, %d threads)" % (mode, res, ref["ms_per_step"],
                                                                                          ref["lsmove_ms_per_step"], ref["threads"]))
    assert res["last"] < TOL and res["lsmove"] < 2e-12 and scale.rel(last, ref["last_nll"]) < 2e-12, res


@pytest.mark.parametrize("mode", ["Gram", "PerEvent"])
def test_tddp4_nll_at_scale(mode):
    n = n_mc = 1_000_000
    rng = np.random.default_rng(5)
    data_cols, _ = scale.phase_space(gf, n, 31337)
    mc = scale.unweighted_phase_space(gf, n_mc, 7)
    t = rng.exponential(0.4101, n).clip(0, 9.99)
    s = rng.uniform(0.02, 0.08, n)
    cols = {nm: data_cols[k] for k, nm in enumerate(scale.NAMES5)}
    cols.update({"mc_" + nm: mc[k] for k, nm in enumerate(scale.NAMES5)})
    cols.update({"dtime": t, "sigmat": s})
    d = scale.write_columns(scale.scratch_dir("tddp4"), cols)
    try:
        ref = scale.run_reference("tddp4", 0, 2, 1, data=d, neumaier=True, lsmove=True)
    finally:
        shutil.rmtree(d, ignore_errors=True)
    assert ref["n_events"] == n and ref["n_mc"] == n_mc
    M = TDDP4(three_gauss=False, float_ls=True, float_mix=True)
    M.dp.setNormMode(getattr(gf.Amp4BodyBase.NormMode, mode))
    M.dp.setNormEvents(mc)
    data = M.dataset(list(data_cols) + [t, s])
    M.signal.setData(data)
    M.dp.setDataSize(len(data), 8)
    first = M.signal.calculateNLL()
    assert scale.rel(first, ref["first_nll"]) < TOL, (first, ref["first_nll"])
    fm = gf.FitManager(M.signal)
    v0, idx = _movers(fm, M.mix)
    fcn = fm.getFCN()
    for k in range(3):
        last = fcn(list(scale.mover_vector(v0, idx, k)))
    li = M.lsvars[2].fitter_index
    v = scale.mover_vector(v0, idx, 2)
    for k in range(4):
        v[li] = v0[li] * (1.0 + 1e-4 * (k + 1))
        ls_last = fcn(list(v))
    v[li] = v0[li]
    back = fcn(list(v))
    res = {"last": scale.rel(back, ref["neumaier_last_nll"]), "lsmove": scale.rel(ls_last, ref["lsmove_last_nll"]),
           "ref_reduce_vs_neumaier": scale.rel(ref["last_nll"], ref["neumaier_last_nll"])}
    print("tddp4 %s: %s (reference %.1f ms/call, %.1f ms when the K* mass moves)" % (mode, res, ref["ms_per_step"], ref["lsmove_ms_per_step"]))
    assert res["last"] < TOL and res["lsmove"] < 2e-12 and scale.rel(last, ref["last_nll"]) < 2e-12, res


def test_exponential_nll_at_scale():
    n = 20_000_000
    rng = np.random.default_rng(137)
    x = rng.exponential(1.0, n)
    x = x[x < 10.0]
    d = scale.write_columns(scale.scratch_dir("exp"), {"x": x})
    try:
        ref = scale.run_reference("exponential", 0, 2, 1, data=d, neumaier=True)
    finally:
        shutil.rmtree(d, ignore_errors=True)
    xo = gf.Observable("x", 0, 10)
    alpha = gf.Variable("alpha", -2, 0.1, -10, 10)
    pdf = gf.ExpPdf("exppdf", xo, alpha)
    pdf.setData(dataset_from_columns([xo], [x]))
    first = pdf.calculateNLL()
    assert scale.rel(first, ref["first_nll"]) < TOL
    fm = gf.FitManager(pdf)
    v0, idx = _movers(fm, [alpha])
    r, _ = _check_schedule(fm.getFCN(), v0, idx, ref, 2, 1)
    print("exponential n=%d: %s" % (len(x), r))
    assert r["last_vs_neumaier"] < TOL, r


@pytest.mark.parametrize("johnson", [False, True])
def test_zachfit_full_tree_at_scale(johnson):
    """BASELINE configs[1] as configured: three ConvolutionPdfs inside nested AddPdfs, 10 000-point workspaces, 2 700 norm bins.
    The reference convolves per event (6e4 flop/event), so it is run on 10^5 events; the fit is compared too."""
    n = 100_000
    rng = np.random.default_rng(11)
    core = rng.normal(0.14543, 0.0004, n)
    tail = 0.1395 + (0.1665 - 0.1395) * rng.beta(1.5, 4.0, n)
    dm = np.where(rng.uniform(size=n) < 0.9, core, tail)
    dm = dm[(dm > 0.1396) & (dm < 0.1664)]
    cfg = "zachfit_jsu" if johnson else "zachfit"
    d = scale.write_columns(scale.scratch_dir("zach"), {"dm": dm})
    try:
        ref = scale.run_reference(cfg, 0, 2, 1, data=d, neumaier=True, fit=True)
    finally:
        shutil.rmtree(d, ignore_errors=True)
    Z = ZachFull(johnson=johnson)
    Z.total.setData(Z.dataset(dm))
    first = Z.total.calculateNLL()
    assert scale.rel(first, ref["first_nll"]) < 1e-11, (first, ref["first_nll"])
    fm = gf.FitManager(Z.total)
    v0, idx = _movers(fm, Z.free)
    r, _ = _check_schedule(fm.getFCN(), v0, idx, ref, 2, 1)
    print("%s n=%d: %s" % (cfg, len(dm), r))
    assert r["last_vs_neumaier"] < 1e-11, r  # ~10^4-term convolution sums per event: 1e-11 as in test_zachfit_pieces
    # fit from the constructor values on both sides
    Z = ZachFull(johnson=johnson)
    Z.total.setData(Z.dataset(dm))
    fm, res = scale.fit_with(gf, Z.total)
    rf = ref["fit"]
    vals = [v.value for v in Z.free]
    p = scale.pulls(vals, rf["values"], rf["errors"])
    print("%s fit: nfcn %d vs %d, fval %.6f vs %.6f, max pull %.2e, wall %.2f s vs %.2f s" % (cfg, res.NFcn(), rf["nfcn"], res.Fval(), rf["fval"],
                                                                                         p.max(), fm.lastFitSeconds(), rf["wall_s"]))
    assert bool(fm) == rf["valid"]
    assert p.max() < 0.01, (vals, rf)


def test_dalitz_fit_at_scale():
    """10^6 events drawn from the model, fitted on both sides from the same start values (30 free coefficients)."""
    n = 1_000_000
    M = Dalitz(float_masses=False, grid=240)
    # Whether Migrad ends below its EDM goal with 30 free coefficients depends on the sample (tools/diag_fit.py: valid at 10^5
    # and 10^6 events of this draw, "above max EDM" at 3x10^5); this draw converges, and both arms must agree on that.
    m12, m13 = scale.dalitz_model_events(gf, M, n, 31, chunk=4_000_000)
    d = scale.write_columns(scale.scratch_dir("dalitzfit"), {"m12": m12, "m13": m13})
    try:
        ref = scale.run_reference("dalitz", 0, 1, 0, data=d, fit=True)
    finally:
        shutil.rmtree(d, ignore_errors=True)
    M = Dalitz(float_masses=False, grid=240)
    data = M.dataset(m12, m13)
    M.signal.setData(data)
    M.signal.setDataSize(len(data))
    assert scale.rel(M.signal.calculateNLL(), ref["first_nll"]) < TOL
    fm, res = scale.fit_with(gf, M.signal)
    rf = ref["fit"]
    vals = [c.value for c in M.coefs]
    p = scale.pulls(vals, rf["values"], rf["errors"])
    print("dalitz fit 1e6: nfcn %d vs %d, fval %.6f vs %.6f, max pull %.2e, wall %.2f s vs reference %.1f s (%d threads)"
          % (res.NFcn(), rf["nfcn"], res.Fval(), rf["fval"], p.max(), fm.lastFitSeconds(), rf["wall_s"], ref["threads"]))
    assert bool(fm) == rf["valid"], (bool(fm), rf["valid"], res.Edm(), res.IsAboveMaxEdm())
    assert p.max() < 0.01, p
    assert scale.rel(res.Fval(), rf["fval"]) < 1e-10
    assert rf["valid"] and bool(fm)


def test_dp4_fit_at_scale():
    """10^6 model-drawn data events + 10^6 MC normalisation events, six free coefficients (examples/DP4/DP4.cu:140-177)."""
    n = n_mc = 1_000_000
    mc = scale.unweighted_phase_space(gf, n_mc, 99)
    M = DP4(mc_events=0, float_ls=False)
    ev = scale.dp4_model_events(gf, M, n, 8, mc)
    cols = {nm: ev[k] for k, nm in enumerate(scale.NAMES5)}
    cols.update({"mc_" + nm: mc[k] for k, nm in enumerate(scale.NAMES5)})
    d = scale.write_columns(scale.scratch_dir("dp4fit"), cols)
    try:
        ref = scale.run_reference("dp4", 0, 1, 0, data=d, fit=True)
    finally:
        shutil.rmtree(d, ignore_errors=True)
    M = DP4(mc_events=0, float_ls=False)
    M.dp.setNormEvents(mc)
    data = M.dataset(list(ev))
    M.signal.setData(data)
    M.dp.setDataSize(len(data), 6)
    assert scale.rel(M.signal.calculateNLL(), ref["first_nll"]) < TOL
    fm, res = scale.fit_with(gf, M.signal)
    rf = ref["fit"]
    vals = [c.value for c in M.coefs[:6]]
    p = scale.pulls(vals, rf["values"], rf["errors"])
    print("dp4 fit 1e6+1e6: nfcn %d vs %d, fval %.6f vs %.6f, max pull %.2e, wall %.2f s vs reference %.1f s"
          % (res.NFcn(), rf["nfcn"], res.Fval(), rf["fval"], p.max(), fm.lastFitSeconds(), rf["wall_s"]))
    assert rf["valid"] and bool(fm)
    assert p.max() < 0.01, p


def test_tddp4_fit_floating_lineshapes_at_scale():
    """BASELINE configs[4] as configured: tau, x, y AND the rho / K*(892) masses and widths float (TDDP4.cu:50-53),
    so most Minuit calls move a lineshape and go through the one-pass update (k_amp4_update, DMMA Gram). 3x10^5 model-drawn data
    events + 3x10^5 MC normalisation events on both arms, seven free parameters."""
    n = n_mc = 300_000
    mc = scale.unweighted_phase_space(gf, n_mc, 199)
    M = TDDP4(three_gauss=False, float_ls=True, float_mix=True)
    ev = scale.tddp4_model_events(gf, M, n, 9, mc)
    names = list(scale.NAMES5) + ["dtime", "sigmat"]
    cols = {nm: ev[k] for k, nm in enumerate(names)}
    cols.update({"mc_" + nm: mc[k] for k, nm in enumerate(scale.NAMES5)})
    d = scale.write_columns(scale.scratch_dir("tddp4fit"), cols)
    try:
        ref = scale.run_reference("tddp4", 0, 1, 0, data=d, fit=True, lsmove=True)
    finally:
        shutil.rmtree(d, ignore_errors=True)
    M = TDDP4(three_gauss=False, float_ls=True, float_mix=True)
    M.dp.setNormEvents(mc)
    data = M.dataset(list(ev))
    M.signal.setData(data)
    M.dp.setDataSize(len(data), 8)
    assert scale.rel(M.signal.calculateNLL(), ref["first_nll"]) < TOL
    l0 = gf.launch_count()
    fm, res = scale.fit_with(gf, M.signal)
    rf = ref["fit"]
    assert rf["names"] == ["tau", "xmixing", "ymixing", "rho_mass", "rho_width", "KstarM", "KstarW"], rf["names"]
    vals = [v.value for v in list(M.mix) + list(M.lsvars)]
    p = scale.pulls(vals, rf["values"], rf["errors"])
    print("tddp4 fit (7 free, lineshapes floating) 3e5+3e5: nfcn %d vs %d, fval %.6f vs %.6f, max pull %.2e, wall %.2f s vs reference %.1f s, %.1f launches/call"
          % (res.NFcn(), rf["nfcn"], res.Fval(), rf["fval"], p.max(), fm.lastFitSeconds(), rf["wall_s"], (gf.launch_count() - l0) / max(res.NFcn(), 1)))
    assert bool(fm) == rf["valid"]
    assert p.max() < 0.01, p
    assert scale.rel(res.Fval(), rf["fval"]) < 1e-9
