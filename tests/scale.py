"""Shared helpers for everything that runs at BASELINE scale (10^6 - 10^7 events): tests/test_scale_gpu.py and bench.py.

The product and the reference must see IDENTICAL inputs. Inputs are generated on the product side (numpy + the product's
own phase-space generator and, for model-drawn toys, its own PDF evaluation on the GPU), written as .npy columns into a
scratch directory, and handed to `oracle/_ref/ref_driver bench <config> ... --data DIR`, which loads them into the
unmodified reference (data columns through UnbinnedDataSet, the MC normalisation sample into Amp4Body's / NormEvents'
own containers). The oracle is only ever the checker / the timed CPU baseline here — nothing under oracle/ is imported
by the product path.
"""
import json
import os
import shutil
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
MASSES4 = [1.8645, 0.13957018, 0.13957018, 0.493677, 0.13957018]
NAMES5 = ("m12", "m34", "cos12", "cos34", "phi")


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def have_reference():
    return os.path.exists(REF_DRIVER)


def scratch_dir(tag):
    """a directory for the exchanged columns: RAM-backed when /dev/shm has room (10^7 events x 12 columns ~ 1 GB)"""
    base = "/dev/shm" if os.path.isdir("/dev/shm") and shutil.disk_usage("/dev/shm").free > (4 << 30) else None
    return tempfile.mkdtemp(prefix="goofit_b200_%s_" % tag, dir=base)


def write_columns(path, cols):
    os.makedirs(path, exist_ok=True)
    for name, v in cols.items():
        np.save(os.path.join(path, name + ".npy"), np.ascontiguousarray(v, dtype=np.float64))
    return path


def run_reference(config, n_events=0, steps=3, warmup=1, data=None, neumaier=False, fit=False, lsmove=False, timeout=3000):
    """run the unmodified reference (oracle/_ref, OpenMP Thrust backend, all host cores) and return its JSON line.
    torch.distributed.run exports OMP_NUM_THREADS=1 to its workers: never inherit that."""
    if not have_reference():
        return None
    env = dict(os.environ)
    env["OMP_NUM_THREADS"] = str(host_cores())
    env.pop("OMP_PROC_BIND", None)
    cmd = [REF_DRIVER, "bench", config, str(n_events), str(steps), str(warmup)]
    if data:
        cmd += ["--data", data]
    if neumaier:
        cmd.append("--neumaier")
    if fit:
        cmd.append("--fit")
    if lsmove:
        cmd.append("--lsmove")
    out = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout)
    for line in out.stdout.splitlines()[::-1]:
        if line.startswith("{"):
            r = json.loads(line)
            if r.get("threads", 0) == 1 and host_cores() > 1:
                raise RuntimeError("reference arm ran on one thread on a %d-core box" % host_cores())
            return r
    raise RuntimeError("ref_driver failed (rc %d): %s" % (out.returncode, (out.stderr or out.stdout)[-800:]))


def mover_vector(v0, idx, k):
    """the parameter schedule both arms use: mover j at step k is start_j * (1 + 1e-3 * ((k + j) % 3 - 1))
    (oracle/ref_driver.cpp bench())"""
    v = np.array(v0, dtype=np.float64)
    for j, i in enumerate(idx):
        v[i] = v0[i] * (1.0 + 1e-3 * ((k + j) % 3 - 1))
    return v


def rel(a, b):
    return abs(a - b) / abs(b)


# ---- toy generation on the product side ------------------------------------------------------------------------------------
def eval_raw(gf, pdf):
    """un-normalised PDF value per uploaded event (EvalFit metric), leaving the PDF in NLL mode"""
    pdf.setFitControl(gf.EvalFit())
    pdf.normalize()
    v = np.asarray(pdf.evaluate_with_metric())
    pdf.setFitControl(gf.UnbinnedNllFit())
    return v


def accept_reject(weights, rng, wmax):
    return rng.uniform(0.0, wmax, size=len(weights)) < weights


def dalitz_model_events(gf, M, n, seed, chunk=8_000_000):
    """n (m12, m13) points drawn from the Amp3Body model at its current parameter values: uniform candidates inside the
    Dalitz boundary, accept-reject against the PDF evaluated on the GPU (what DalitzPlotter::fillDataSetMC /
    Amp3Body::GenerateSig are for upstream, DalitzPlotter.h:64-95)."""
    from models import dalitz_uniform_events
    rng = np.random.default_rng(seed)
    out12, out13, have, wmax, k = [], [], 0, None, 0
    while have < n:
        a, b = dalitz_uniform_events(chunk, seed * 1000 + k)
        k += 1
        d = M.dataset(a, b)
        M.signal.setData(d)
        M.signal.setDataSize(len(d))
        w = eval_raw(gf, M.signal)
        if wmax is None:
            wmax = 1.25 * float(w.max())
        acc = accept_reject(np.minimum(w, wmax), rng, wmax)
        out12.append(a[acc])
        out13.append(b[acc])
        have += int(acc.sum())
    return np.concatenate(out12)[:n].copy(), np.concatenate(out13)[:n].copy()


def phase_space(gf, n, seed):
    """weighted four-body phase space in the five variables, generated on the GPU (k_genbod)"""
    cols, w = gf.devicePhaseSpace5(MASSES4, n, seed)
    return np.asarray(cols), np.asarray(w)


def unweighted_phase_space(gf, n, seed, chunk=2_000_000):
    """n accepted phase-space points (accept-reject on the GENBOD weight): the MC normalisation sample"""
    rng = np.random.default_rng(seed + 17)
    parts, have, k = [], 0, 0
    while have < n:
        cols, w = phase_space(gf, chunk, seed * 100 + k)
        k += 1
        acc = accept_reject(w, rng, 1.0)  # generatePhaseSpace5 weights are already divided by the maximum weight
        parts.append(cols[:, acc])
        have += int(acc.sum())
    return np.ascontiguousarray(np.concatenate(parts, axis=1)[:, :n])


def dp4_model_events(gf, M, n, seed, mc_cols, chunk=2_000_000):
    """n five-variable points drawn from the Amp4Body model (accept-reject of weighted phase space against the PDF value,
    as examples/DP4/DP4.cu:140-177 does through GenerateSig)"""
    rng = np.random.default_rng(seed)
    M.dp.setNormEvents(mc_cols)
    parts, have, wmax, k = [], 0, None, 0
    while have < n:
        cols, w = phase_space(gf, chunk, seed * 100 + k)
        k += 1
        d = M.dataset(list(cols))
        M.signal.setData(d)
        M.dp.setDataSize(len(d), 6)
        v = eval_raw(gf, M.dp) * w
        if wmax is None:
            wmax = 1.5 * float(v.max())
        acc = accept_reject(np.minimum(v, wmax), rng, wmax)
        parts.append(cols[:, acc])
        have += int(acc.sum())
    return np.ascontiguousarray(np.concatenate(parts, axis=1)[:, :n])


def tddp4_model_events(gf, M, n, seed, mc_cols, tau=0.4101, chunk=2_000_000):
    """n seven-variable points (kinematics, decay time, its resolution) drawn from the Amp4Body_TD model: candidates are
    weighted phase space with t ~ Exp(tau) (t < 10) and sigma_t ~ U(0.02, 0.08); the accept weight is the PDF value divided
    by the exponential the times were drawn from (oracle/ref_driver.cpp case_tddp4_fit does the same)"""
    rng = np.random.default_rng(seed)
    M.dp.setNormEvents(mc_cols)
    parts, have, wmax, k = [], 0, None, 0
    while have < n:
        cols, w = phase_space(gf, chunk, seed * 100 + k)
        k += 1
        t = rng.exponential(tau, cols.shape[1])
        t = np.where(t < 9.99, t, rng.uniform(0, 9.99, cols.shape[1]))
        s = rng.uniform(0.02, 0.08, cols.shape[1])
        d = M.dataset(list(cols) + [t, s])
        M.signal.setData(d)
        M.dp.setDataSize(len(d), 8)
        v = eval_raw(gf, M.dp) * w / np.exp(-t / tau)
        if wmax is None:
            wmax = 1.5 * float(v.max())
        acc = accept_reject(np.minimum(v, wmax), rng, wmax)
        parts.append(np.vstack([cols, t, s])[:, acc])
        have += int(acc.sum())
    return np.ascontiguousarray(np.concatenate(parts, axis=1)[:, :n])


def fit_with(gf, pdf, verbosity=0):
    fm = gf.FitManager(pdf)
    fm.setVerbosity(verbosity)
    res = fm.fit()
    return fm, res


def pulls(values, ref_values, ref_errors):
    v, r, e = np.asarray(values), np.asarray(ref_values), np.asarray(ref_errors)
    free = e > 0
    return np.abs(v[free] - r[free]) / e[free]
