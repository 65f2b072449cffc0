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
                    "cost of a full Minuit2 minimisation with Hesse on 10^7 events"}


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
    ph, _ = gf.generatePhaseSpace5(masses, unique, 4242)
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
    try:  # installs its own small program: after everything that uses the caller's model
        out["call_latency"] = call_latency(gf, lib, ctx)
    except Exception as e:
        out["call_latency"] = {"error": str(e)[:200]}
    return out
