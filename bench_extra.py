"""bench_extra.py — secondary measurements bench.py attaches to its JSON line at N=1 (key "workloads"):

  fp64_peak    DFMA micro-benchmark (gf_fp64_peak): the measured denominator of the FP64-bound kernels
  dalitz_cache k_amp3_waves over all 16 resonances (what reruns when masses/widths float), CUDA-event timed
  dp4          BASELINE configs[3]: Amp4Body D0->K-pi+pi+pi-, steady-state FCN calls (MC integral + NLL kernel)
  tddp4        BASELINE configs[4] shape at one GPU: Amp4Body_TD with TruthResolution

Nothing here touches oracle/. Every number is taken through the host API or the C ABI on the ctx stream.
"""
import ctypes as C
import json
import os
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))


def _profile_counts():
    """per-launch FP64 instruction counts taken once with ncu (profiles/traffic.json), or {}"""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def fp64_peak(lib, ctx):
    v = C.c_double(0)
    if lib.gf_fp64_peak(ctx, C.byref(v)) != 0:
        return None
    return v.value


def timed(lib, ctx, fn, reps):
    ms = C.c_float(0)
    lib.gf_sync(ctx)
    assert lib.gf_timer_start(ctx) == 0
    for _ in range(reps):
        fn()
    assert lib.gf_timer_stop(ctx, C.byref(ms)) == 0
    return float(ms.value) / reps


def dalitz_cache(gf, lib, ctx, M, n_events, peak_tf):
    """all 16 waves recomputed: the per-call cost when every mass/width floats"""
    node = M.signal.getFunctionIndex()
    nres = 16
    mask = (C.c_uint8 * nres)(*([1] * nres))

    def run():
        assert lib.gf_amp3_cache_waves(ctx, C.c_int(node), mask) == 0

    run()
    ms = timed(lib, ctx, run, 5)
    out = {"kernel": "k_amp3_waves", "kernel_ms": ms, "resonance_evals_per_s": nres * n_events / (ms * 1e-3),
           "hbm_write_gbs": 16.0 * nres * n_events / (ms * 1e-3) / 1e9}
    flops = _profile_counts().get("k_amp3_waves_fp64_flops_per_launch")
    if flops and peak_tf:
        ach = flops / (ms * 1e-3) / 1e12
        out["roofline"] = {"bound": "fp64", "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s", "frac": ach / peak_tf,
                           "flops_per_launch": flops, "flops_source": "ncu sass dfma*2+dmul+dadd (profiles/)",
                           "peak_source": "gf_fp64_peak DFMA micro-benchmark, this run"}
    return out


def dalitz_fit(gf, lib, ctx, M, n_events):
    """fit wall time: FitManager(Minuit2 Migrad, strategy 1) over the 30 free coefficients on the resident dataset"""
    fm = gf.FitManager(M.signal)
    fm.setVerbosity(0)
    l0 = gf.launch_count()
    t0 = time.perf_counter()
    res = fm.fit()
    wall = time.perf_counter() - t0
    nfcn = int(res.NFcn())
    return {"wall_s": wall, "nfcn": nfcn, "valid": bool(fm), "fval": res.Fval(), "ms_per_call": 1e3 * wall / max(nfcn, 1),
            "events_x_evals_per_s": n_events * nfcn / wall, "kernel_launches": int(gf.launch_count() - l0),
            "note": "synthetic events are uniform in the Dalitz plot, so the minimum is wherever Migrad finds it; the number is the "
                    "cost of a full Minuit2 minimisation with Hesse on 10^7 events. Fits of model-drawn events, valid on both arms and "
                    "timed beside the reference: workloads.reference_side"}


def _tile(cols, n):
    reps = (n + cols.shape[1] - 1) // cols.shape[1]
    return np.ascontiguousarray(np.tile(cols, (1, reps))[:, :n])


def four_body(gf, lib, ctx, kind, n_data, n_mc, steps, peak_hbm):
    """steady-state likelihood calls of the 4-body fits: coefficients (DP4) or tau/x/y (TDDP4) move, lineshapes fixed.
    Per call: MC-integral kernel over n_mc cached-amplitude rows + fused NLL kernel over n_data rows."""
    from models import DP4, TDDP4
    unique = 2_000_000
    masses = [1.8645, 0.13957018, 0.13957018, 0.493677, 0.13957018]
    t0 = time.perf_counter()
    ph, _ = gf.devicePhaseSpace5(masses, unique, 4242)  # GENBOD on the GPU (k_genbod)
    t_gen = time.perf_counter() - t0
    rng = np.random.default_rng(99)
    if kind == "dp4":
        M = DP4(mc_events=0, float_ls=True)
        namp, bytes_evt = 8, 8 * 16  # cached amplitudes only: the flat background and constant efficiency read no column
        d = M.dataset(list(_tile(ph, n_data)))
        movers = M.coefs[:6]
    else:
        M = TDDP4(three_gauss=False, float_ls=True, float_mix=True)
        namp, bytes_evt = 6, 6 * 16 + 16  # cached amplitudes + decay time and its error
        cols = list(_tile(ph, n_data))
        cols.append(rng.exponential(0.4101, n_data).clip(0, 9.99))
        cols.append(rng.uniform(0.02, 0.08, n_data))
        d = M.dataset(cols)
        movers = M.mix
    M.dp.setNormEvents(_tile(ph[:, ::-1].copy(), n_mc))
    sig, dp = M.signal, M.dp
    t0 = time.perf_counter()
    sig.setData(d)
    dp.setDataSize(len(d), 6 if kind == "dp4" else 8)
    t_setdata = time.perf_counter() - t0
    t0 = time.perf_counter()
    sig.calculateNLL()
    t_first = time.perf_counter() - t0
    fm = gf.FitManager(sig)
    fcn, params = fm.getFCN(), fm.getParams()
    v0 = np.array(params.make_minuit_vector())
    idx = [c.fitter_index for c in movers]

    def vec(k):
        v = v0.copy()
        for j, i in enumerate(idx):
            v[i] = v0[i] * (1.0 + 1e-3 * ((k + j) % 3 - 1))
        return list(v)

    for k in range(3):
        fcn(vec(k))
    lib.gf_sync(ctx)
    l0 = gf.launch_count()
    vecs = [vec(3 + k) for k in range(steps)]
    t0 = time.perf_counter()
    for k in range(steps):
        last = fcn(vecs[k])
    lib.gf_sync(ctx)
    t_e2e = (time.perf_counter() - t0) / steps
    launches = (gf.launch_count() - l0) / steps
    # the two streaming kernels alone, queued back to back on the ctx stream (CUDA events)
    node, root = dp.getFunctionIndex(), sig.getFunctionIndex()

    def pair():
        assert lib.gf_amp4_norm_async(ctx, C.c_int(node)) == 0
        assert lib.gf_eval_sum_async(ctx, C.c_int(root), C.c_int(1)) == 0

    pair()
    ms_pair = timed(lib, ctx, pair, 10)
    ms_norm = timed(lib, ctx, lambda: lib.gf_amp4_norm_async(ctx, C.c_int(node)), 10)
    ms_eval = timed(lib, ctx, lambda: lib.gf_eval_sum_async(ctx, C.c_int(root), C.c_int(1)), 10)
    kernel_name = lib.gf_last_kernel_name(ctx).decode()
    # A/B: the same call on the interpreter copy of the kernel
    was = lib.gf_set_option(ctx, b"fast", C.c_int(0))
    lib.gf_eval_sum_async(ctx, C.c_int(root), C.c_int(1))
    ms_interp = timed(lib, ctx, lambda: lib.gf_eval_sum_async(ctx, C.c_int(root), C.c_int(1)), 10)
    lib.gf_set_option(ctx, b"fast", C.c_int(was))
    lib.gf_eval_sum_async(ctx, C.c_int(root), C.c_int(1))
    # default (Gram) mode: with the lineshapes fixed the MC integral is a host sum over G_ab, so the step is k_eval_sum
    ms = ms_eval
    bytes_step = bytes_evt * n_data
    ach = bytes_step / (ms * 1e-3) / 1e9
    # lineshape move: K*(892) mass changes -> its lineshapes, the amplitudes using them and the Gram matrix are redone
    def ls_move(reps=5):
        for k in range(2):  # warm-up (first use of a path allocates its scratch)
            M.lsvars[2].changed = True
            sig.updateVariable(M.lsvars[2], 0.8958 + 1e-4 * (k + 7))
            sig.updateParameters()
            sig.calculateNLL()
        lib.gf_sync(ctx)
        l0 = gf.launch_count()
        t0 = time.perf_counter()
        for k in range(reps):
            M.lsvars[2].changed = True
            sig.updateVariable(M.lsvars[2], 0.8958 + 1e-4 * (k + 1))
            sig.updateParameters()
            sig.calculateNLL()
        lib.gf_sync(ctx)
        return (time.perf_counter() - t0) / reps, (gf.launch_count() - l0) / reps

    t_ls, launches_ls = ls_move()
    was = lib.gf_set_option(ctx, b"fused", C.c_int(0))
    t_ls_sep, launches_ls_sep = ls_move()
    lib.gf_set_option(ctx, b"fused", C.c_int(was))
    ls_move(1)
    return {"workload": "examples/DP4: Amp4Body, 8 amplitudes" if kind == "dp4" else "examples/TDDP4: Amp4Body_TD, 6 amplitudes, TruthResolution",
            "data_events": n_data, "mc_norm_events": n_mc, "unique_generated": unique,
            "value": n_data / (ms * 1e-3), "unit": "events*evals/s", "ms_per_step": ms,
            "norm_mode": "Gram (default): per-call MC pass only when a lineshape moved",
            "lineshape_move": {"ms_per_call": 1e3 * t_ls, "launches_per_call": launches_ls,
                               "separate_kernels_ms_per_call": 1e3 * t_ls_sep, "separate_kernels_launches": launches_ls_sep,
                               "note": "K*(892) mass moves. One pass per sample (k_amp4_update): flagged lineshapes -> amplitudes -> G_ab on FP64 "
                                       "tensor-core MMAs (MC), then the NLL kernel; 'separate' = k_amp4_cache + k_amp4_amps per sample + 4x4-tile k_amp4_gram"},
            "per_event_mode": {"ms_per_step": ms_pair, "value": n_data / (ms_pair * 1e-3),
                               "note": "reference association: MC-integral kernel every call"},
            "e2e": {"value": n_data / t_e2e, "ms_per_step": 1e3 * t_e2e, "launches_per_step": launches,
                    "setdata_s": t_setdata, "first_call_s": t_first, "phase_space_gen_s": t_gen},
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peak_hbm, "unit": "GB/s", "frac": ach / peak_hbm,
                         "kernel": kernel_name, "interpreter_kernel_ms": ms_interp,
                         "algorithmic_bytes_per_step": bytes_step,
                         "norm_kernel_ms": ms_norm, "norm_gbs": namp * 16 * n_mc / (ms_norm * 1e-3) / 1e9,
                         "eval_kernel_ms": ms_eval, "eval_gbs": bytes_evt * n_data / (ms_eval * 1e-3) / 1e9},
            "nll": last}


def simple_fits(gf, lib, ctx, peak_hbm, steps=20):
    """BASELINE configs[0] and [1]: the event-streaming fits. exponential = ExpPdf on 10^6 events (8 MB: L2-resident, so
    latency- not HBM-bound; also run at 10^8 events where it streams). zachFit = AddPdf(Argus, KinLimitBW (x) resolution)
    on 10^7 delta-M values with the reference's 2.7e-6 integration step (10 000-point model workspace)."""
    out = {}
    rng = np.random.default_rng(137)
    for tag, n in (("exponential_1e6", 1_000_000), ("exponential_1e8", 100_000_000)):
        x = gf.Observable("x", 0, 10)
        xs = rng.exponential(1.0, n)
        xs = xs[xs < 10.0]
        from models import dataset_from_columns
        d = dataset_from_columns([x], [xs])
        alpha = gf.Variable("alpha", -2, 0.1, -10, 10)
        pdf = gf.ExpPdf("exppdf", x, alpha)
        pdf.setData(d)
        pdf.calculateNLL()
        fm = gf.FitManager(pdf)
        fcn = fm.getFCN()
        vecs = [[-2.0 + 0.01 * (k % 5)] for k in range(steps + 3)]
        for k in range(3):
            fcn(vecs[k])
        t0 = time.perf_counter()
        for k in range(steps):
            fcn(vecs[3 + k])
        lib.gf_sync(ctx)
        t = (time.perf_counter() - t0) / steps
        root = pdf.getFunctionIndex()
        ms = timed(lib, ctx, lambda: lib.gf_eval_sum_async(ctx, C.c_int(root), C.c_int(1)), 10)
        kname = lib.gf_last_kernel_name(ctx).decode()
        was = lib.gf_set_option(ctx, b"fast", C.c_int(0))
        lib.gf_eval_sum_async(ctx, C.c_int(root), C.c_int(1))
        ms_interp = timed(lib, ctx, lambda: lib.gf_eval_sum_async(ctx, C.c_int(root), C.c_int(1)), 10)
        lib.gf_set_option(ctx, b"fast", C.c_int(was))
        ach = 8.0 * len(xs) / (ms * 1e-3) / 1e9
        out[tag] = {"events": int(len(xs)), "value": len(xs) / (ms * 1e-3), "ms_per_step": ms,
                    "e2e": {"value": len(xs) / t, "ms_per_step": 1e3 * t},
                    "roofline": {"bound": "hbm", "achieved": ach, "peak": peak_hbm, "unit": "GB/s", "frac": ach / peak_hbm,
                                 "kernel": kname, "interpreter_kernel_ms": ms_interp, "algorithmic_bytes_per_step": 8 * len(xs)}}
        del pdf, d
    from models import zach_model
    n = 10_000_000
    dmv = 0.1395 + (0.1665 - 0.1395) * rng.beta(2.0, 5.0, n)
    o = zach_model(dmv, step=2.7e-6, numbins=2700)
    total = o["total"]
    t0 = time.perf_counter()
    total.setData(o["d"])
    first = total.calculateNLL()
    t_first = time.perf_counter() - t0
    fm = gf.FitManager(total)
    fcn, params = fm.getFCN(), fm.getParams()
    v0 = np.array(params.make_minuit_vector())

    def run(var, label):
        i = var.fitter_index
        vecs = []
        for k in range(steps + 2):
            v = v0.copy()
            v[i] = v0[i] * (1.0 + 1e-3 * (k % 3 - 1))
            vecs.append(list(v))
        for k in range(2):
            fcn(vecs[k])
        lib.gf_sync(ctx)
        l0 = gf.launch_count()
        t0 = time.perf_counter()
        for k in range(steps):
            fcn(vecs[2 + k])
        lib.gf_sync(ctx)
        t = (time.perf_counter() - t0) / steps
        return {"moved": label, "ms_per_call": 1e3 * t, "value": n / t, "launches_per_call": (gf.launch_count() - l0) / steps}

    out["zachfit"] = {"events": n, "workspace_points": 10000, "norm_bins": 2700, "first_call_s": t_first, "nll": first,
                      "unit": "events*evals/s", "calls": [run(o["frac"], "bkg_frac (no workspace refill)"),
                                                           run(o["rs"], "res_sigma (resolution workspace + table)"),
                                                           run(o["width"], "kl_width (model workspace + table)")]}
    return out


def call_latency(gf, lib, ctx, reps=400):
    """the per-call floor: a Dalitz likelihood on 2 000 events (kernel ~ 10 us), so what is left is launch + parameter upload +
    completion wait + host code. host_api = FCN::operator() from Python; c_abi = gf_set_params_async + gf_eval_sum."""
    from models import Dalitz, dalitz_uniform_events
    M = Dalitz(float_masses=False, grid=60)
    a, b = dalitz_uniform_events(2000, 3)
    d = M.dataset(a, b)
    M.signal.setData(d)
    M.signal.setDataSize(len(d))
    M.signal.calculateNLL()
    fm = gf.FitManager(M.signal)
    fcn, params = fm.getFCN(), fm.getParams()
    v = list(params.make_minuit_vector())
    for _ in range(20):
        fcn(v)
    t0 = time.perf_counter()
    for _ in range(reps):
        fcn(v)
    t_host = (time.perf_counter() - t0) / reps
    root = M.signal.getFunctionIndex()
    p = np.array(gf.program_params(), dtype=np.float64)
    n = np.array(gf.program_norms(), dtype=np.float64)
    dptr = C.POINTER(C.c_double)
    pp, np_ = p.ctypes.data_as(dptr), n.ctypes.data_as(dptr)
    out = C.c_double(0)
    t0 = time.perf_counter()
    for _ in range(reps):
        lib.gf_set_params_async(ctx, pp, C.c_int(len(p)), np_, C.c_int(len(n)))
        lib.gf_eval_sum(ctx, C.c_int(root), C.c_int(1), C.byref(out))
    t_abi = (time.perf_counter() - t0) / reps
    t0 = time.perf_counter()
    for _ in range(reps):
        lib.gf_eval_sum(ctx, C.c_int(root), C.c_int(1), C.byref(out))
    t_eval = (time.perf_counter() - t0) / reps
    t0 = time.perf_counter()
    for _ in range(reps):
        params.from_minuit_vector(v)
    t_params = (time.perf_counter() - t0) / reps
    ms = timed(lib, ctx, lambda: lib.gf_eval_sum_async(ctx, C.c_int(root), C.c_int(1)), 50)
    return {"events": 2000, "host_api_us_per_call": 1e6 * t_host, "c_abi_upload_plus_eval_us": 1e6 * t_abi, "c_abi_eval_only_us": 1e6 * t_eval,
            "from_minuit_vector_us": 1e6 * t_params, "kernel_us_queued": 1e3 * ms}


def reference_side(gf, lib, ctx, n=1_000_000):
    """The other half of BASELINE's metric: fit wall time, Dalitz and DP4, on BOTH arms from the same model-drawn events and
    start values (the reference prints its own at FitManagerMinuit2.cpp:43,52,86-87), plus the reference's per-call cost of
    the DP4 / TDDP4 / zachFit likelihoods (steady state and with the K*(892) mass moving) on a bounded sample: n events
    (+ n MC normalisation events) on all host cores. Inputs go to the reference as .npy columns (tests/scale.py)."""
    import shutil
    import scale
    from models import DP4, Dalitz
    out = {"sample": "%d events (+ %d MC normalisation events for the four-body fits), identical on both arms" % (n, n)}

    def both(tag, cfg, cols, fit_model, fit_vars, ndims=None, ref_kwargs=None):
        d = scale.write_columns(scale.scratch_dir(tag), cols)
        try:
            ref = scale.run_reference(cfg, 0, 1, 0, data=d, fit=True, **(ref_kwargs or {}))
        finally:
            shutil.rmtree(d, ignore_errors=True)
        t0 = time.perf_counter()
        fm, res = scale.fit_with(gf, fit_model)
        wall = time.perf_counter() - t0
        rf = ref["fit"]
        pulls = scale.pulls([v.value for v in fit_vars], rf["values"][:len(fit_vars)], rf["errors"][:len(fit_vars)])
        return {"events": n, "fit_wall_s": wall, "nfcn": int(res.NFcn()), "valid": bool(fm), "fval": res.Fval(),
                "reference": {"fit_wall_s": rf["wall_s"], "nfcn": rf["nfcn"], "valid": rf["valid"], "fval": rf["fval"], "cores": ref["threads"],
                              "ms_per_call": 1e3 * rf["wall_s"] / max(rf["nfcn"], 1)},
                "max_pull_vs_reference_sigma": float(pulls.max()), "speedup_fit_wall": rf["wall_s"] / wall}

    try:  # ---- Dalitz: 30 free coefficients
        M = Dalitz(float_masses=False, grid=240)
        m12, m13 = scale.dalitz_model_events(gf, M, n, 31, chunk=4_000_000)
        M = Dalitz(float_masses=False, grid=240)
        data = M.dataset(m12, m13)
        M.signal.setData(data)
        M.signal.setDataSize(len(data))
        M.signal.calculateNLL()
        out["dalitz_fit"] = both("dalitzfit", "dalitz", {"m12": m12, "m13": m13}, M.signal, M.coefs)
    except Exception as e:
        out["dalitz_fit"] = {"error": str(e)[:300]}
    try:  # ---- DP4: six free coefficients, 10^6 MC normalisation events; and the reference's per-call cost
        mc = scale.unweighted_phase_space(gf, n, 99)
        M = DP4(mc_events=0, float_ls=False)
        ev = scale.dp4_model_events(gf, M, n, 8, mc)
        cols = {nm: ev[k] for k, nm in enumerate(scale.NAMES5)}
        cols.update({"mc_" + nm: mc[k] for k, nm in enumerate(scale.NAMES5)})
        M = DP4(mc_events=0, float_ls=False)
        M.dp.setNormEvents(mc)
        data = M.dataset(list(ev))
        M.signal.setData(data)
        M.dp.setDataSize(len(data), 6)
        M.signal.calculateNLL()
        out["dp4_fit"] = both("dp4fit", "dp4", cols, M.signal, M.coefs[:6])
        d = scale.write_columns(scale.scratch_dir("dp4cpu"), cols)
        try:
            r = scale.run_reference("dp4", 0, 3, 1, data=d, lsmove=True)
        finally:
            shutil.rmtree(d, ignore_errors=True)
        out["dp4_cpu_baseline"] = {"value": r["events_x_evals_per_s"], "unit": "events*evals/s", "cores": r["threads"], "kind": "reference",
                                   "ms_per_step": r["ms_per_step"], "lineshape_move_ms_per_step": r.get("lsmove_ms_per_step"),
                                   "sample": "%d data + %d MC events x 3 FCN calls, unmodified reference (OMP backend, all host cores)" % (n, n)}
    except Exception as e:
        out.setdefault("dp4_fit", {"error": str(e)[:300]})
    try:  # ---- TDDP4 as configured (C5): tau, x, y and the rho / K* masses and widths free; most calls move a lineshape
        from models import TDDP4
        m = n // 2
        mc = scale.unweighted_phase_space(gf, m, 199)
        M = TDDP4(three_gauss=False, float_ls=True, float_mix=True)
        ev = scale.tddp4_model_events(gf, M, m, 9, mc)
        names = list(scale.NAMES5) + ["dtime", "sigmat"]
        cols = {nm: ev[k] for k, nm in enumerate(names)}
        cols.update({"mc_" + nm: mc[k] for k, nm in enumerate(scale.NAMES5)})
        M = TDDP4(three_gauss=False, float_ls=True, float_mix=True)
        M.dp.setNormEvents(mc)
        data = M.dataset(list(ev))
        M.signal.setData(data)
        M.dp.setDataSize(len(data), 8)
        M.signal.calculateNLL()
        n_keep, n = n, m
        out["tddp4_fit"] = both("tddp4fit", "tddp4", cols, M.signal, list(M.mix) + list(M.lsvars), ref_kwargs={"lsmove": True})
        n = n_keep
        out["tddp4_fit"]["events"] = m
        out["tddp4_fit"]["note"] = "%d data + %d MC events; 7 free parameters incl. lineshape masses and widths (TDDP4.cu:50-53)" % (m, m)
    except Exception as e:
        out["tddp4_fit"] = {"error": str(e)[:300]}
    try:  # ---- TDDP4 per-call cost of the reference
        rng = np.random.default_rng(5)
        mc = scale.unweighted_phase_space(gf, n, 7)
        data_cols, _ = scale.phase_space(gf, n, 31337)
        cols = {nm: data_cols[k] for k, nm in enumerate(scale.NAMES5)}
        cols.update({"mc_" + nm: mc[k] for k, nm in enumerate(scale.NAMES5)})
        cols.update({"dtime": rng.exponential(0.4101, n).clip(0, 9.99), "sigmat": rng.uniform(0.02, 0.08, n)})
        d = scale.write_columns(scale.scratch_dir("tdcpu"), cols)
        try:
            r = scale.run_reference("tddp4", 0, 3, 1, data=d, lsmove=True)
        finally:
            shutil.rmtree(d, ignore_errors=True)
        out["tddp4_cpu_baseline"] = {"value": r["events_x_evals_per_s"], "unit": "events*evals/s", "cores": r["threads"], "kind": "reference",
                                     "ms_per_step": r["ms_per_step"], "lineshape_move_ms_per_step": r.get("lsmove_ms_per_step"),
                                     "sample": "%d data + %d MC events x 3 FCN calls, unmodified reference (OMP backend, all host cores)" % (n, n)}
    except Exception as e:
        out["tddp4_cpu_baseline"] = {"error": str(e)[:300]}
    try:  # ---- zachFit as configured (BASELINE configs[1]): three ConvolutionPdfs inside nested AddPdfs, 10 000-point workspaces
        from models import ZachFull

        def dm_sample(m, seed):
            rng = np.random.default_rng(seed)
            core = rng.normal(0.14543, 0.0004, m)
            tail = 0.1395 + (0.1665 - 0.1395) * rng.beta(1.5, 4.0, m)
            dm = np.where(rng.uniform(size=m) < 0.9, core, tail)
            return dm[(dm > 0.1396) & (dm < 0.1664)]

        zres = {}
        for cfg, johnson in (("zachfit", False), ("zachfit_jsu", True)):
            dm_small = dm_sample(100_000, 11)  # the reference convolves per event (6e4 flop/event): bounded to 10^5 events
            d = scale.write_columns(scale.scratch_dir("zach"), {"dm": dm_small})
            try:
                r = scale.run_reference(cfg, 0, 2, 1, data=d)
            finally:
                shutil.rmtree(d, ignore_errors=True)
            dm = dm_sample(10_000_000, 12)
            Z = ZachFull(johnson=johnson)
            Z.total.setData(Z.dataset(dm))
            t0 = time.perf_counter()
            Z.total.calculateNLL()
            t_first = time.perf_counter() - t0
            fm = gf.FitManager(Z.total)
            fcn = fm.getFCN()
            v0 = np.array(fm.getParams().make_minuit_vector())
            idx = [v.fitter_index for v in Z.free]
            for k in range(3):
                fcn(list(scale.mover_vector(v0, idx, k)))
            lib.gf_sync(ctx)
            l0 = gf.launch_count()
            t0 = time.perf_counter()
            for k in range(10):
                fcn(list(scale.mover_vector(v0, idx, 3 + k)))
            lib.gf_sync(ctx)
            t = (time.perf_counter() - t0) / 10
            zres[cfg] = {"events": int(len(dm)), "ms_per_call": 1e3 * t, "value": len(dm) / t, "unit": "events*evals/s",
                         "launches_per_call": (gf.launch_count() - l0) / 10, "first_call_s": t_first,
                         "moved_per_call": "width, delta, epsilon, slope, bkg_frac (all three convolution tables refilled)",
                         "cpu_baseline": {"value": r["events_x_evals_per_s"], "unit": "events*evals/s", "cores": r["threads"], "kind": "reference",
                                          "ms_per_step": r["ms_per_step"],
                                          "sample": "%d events x 2 FCN calls (the reference convolves per event), same tree and schedule" % r["n_events"]}}
        out["zachfit_full"] = zres
    except Exception as e:
        out["zachfit_full"] = {"error": str(e)[:300]}
    return out


def run_all(gf, lib, ctx, M_dalitz, n_events, peak_hbm, steps=20, with_reference=True):
    out = {}
    peak_tf = fp64_peak(lib, ctx)
    out["fp64_peak_tflops_measured"] = peak_tf
    try:
        out["dalitz_fit"] = dalitz_fit(gf, lib, ctx, M_dalitz, n_events)
    except Exception as e:
        out["dalitz_fit"] = {"error": str(e)[:200]}
    try:
        out["dalitz_cache"] = dalitz_cache(gf, lib, ctx, M_dalitz, n_events, peak_tf)
    except Exception as e:  # secondary numbers never break the headline line
        out["dalitz_cache"] = {"error": str(e)[:200]}
    for kind in ("dp4", "tddp4"):
        try:
            out[kind] = four_body(gf, lib, ctx, kind, 10_000_000, 10_000_000, steps, peak_hbm)
        except Exception as e:
            out[kind] = {"error": str(e)[:200]}
    try:
        out.update(simple_fits(gf, lib, ctx, peak_hbm, steps))
    except Exception as e:
        out["simple_fits"] = {"error": str(e)[:300]}
    if with_reference:
        try:
            import scale
            if scale.have_reference():
                out["reference_side"] = reference_side(gf, lib, ctx)
        except Exception as e:
            out["reference_side"] = {"error": str(e)[:300]}
    try:  # installs its own small program: after everything that uses the caller's model
        out["call_latency"] = call_latency(gf, lib, ctx)
    except Exception as e:
        out["call_latency"] = {"error": str(e)[:200]}
    return out
