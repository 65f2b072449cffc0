"""tools/profile_once.py — launches every hot kernel a small, fixed number of times at BASELINE sizes so that one
`ncu --set full -k regex:"k_amp|k_eval_sum"` capture holds all of them (Dalitz: k_amp3_waves/grid/gram + k_eval_sum;
DP4 and TDDP4: k_amp4_sf, k_amp4_update (first call and K*(892)-mass move), k_eval_fast_amp4; exp: k_eval_logleaf and the
interpreter copy of k_eval_sum on 10^8 events). Not a benchmark."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import goofit_b200 as gf  # noqa: E402
from models import DP4, TDDP4, Dalitz, dalitz_uniform_events  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
which = sys.argv[2] if len(sys.argv) > 2 else "dalitz,dp4,tddp4,exp"
if "dalitz" in which:
    M = Dalitz(float_masses=False, grid=240)
    m12, m13 = dalitz_uniform_events(N, 12345)
    d = M.dataset(m12, m13)
    M.signal.setData(d)
    M.signal.setDataSize(len(d))
    print("dalitz", M.signal.calculateNLL(), M.signal.calculateNLL())
masses = [1.8645, 0.13957018, 0.13957018, 0.493677, 0.13957018]
ph, _ = gf.devicePhaseSpace5(masses, min(N, 2_000_000), 4242)
reps = (N + ph.shape[1] - 1) // ph.shape[1]
cols = np.ascontiguousarray(np.tile(ph, (1, reps))[:, :N])
mc = np.ascontiguousarray(np.tile(ph[:, ::-1], (1, reps))[:, :N])
rng = np.random.default_rng(99)
for kind in ("dp4", "tddp4"):
    if kind not in which.split(","):
        continue
    if kind == "dp4":
        M = DP4(mc_events=0, float_ls=True)
        d = M.dataset(list(cols))
    else:
        M = TDDP4(three_gauss=(len(sys.argv) > 3 and sys.argv[3] == "g3"), float_ls=True, float_mix=True)
        d = M.dataset(list(cols) + [rng.exponential(0.4101, N).clip(0, 9.99), rng.uniform(0.02, 0.08, N)])
    M.dp.setNormEvents(mc)
    M.signal.setData(d)
    M.dp.setDataSize(len(d), 6 if kind == "dp4" else 8)
    print(kind, M.signal.calculateNLL())
    # the K*(892) mass moves: only its lineshapes and the amplitudes using them are redone
    for v in M.lsvars:
        v.changed = False
    M.lsvars[2].changed = True
    M.signal.updateVariable(M.lsvars[2], 0.8975)
    M.signal.updateParameters()
    print(kind, M.signal.calculateNLL())
if "exp" in which:
    from models import dataset_from_columns
    x = gf.Observable("x", 0, 10)
    xs = np.random.default_rng(137).exponential(1.0, 10 * N)
    xs = xs[xs < 10.0]
    pdf = gf.ExpPdf("exppdf", x, gf.Variable("alpha", -2, 0.1, -10, 10))
    pdf.setData(dataset_from_columns([x], [xs]))
    print("exp", len(xs), pdf.calculateNLL(), gf.last_kernel_name())
    gf.fast_enable(False)
    print("exp (interpreter)", pdf.calculateNLL(), gf.last_kernel_name())
    gf.fast_enable(True)
if "zach" in which:
    from models import ZachFull
    rz = np.random.default_rng(12)
    m = N
    core = rz.normal(0.14543, 0.0004, m)
    tail = 0.1395 + (0.1665 - 0.1395) * rz.beta(1.5, 4.0, m)
    dm = np.where(rz.uniform(size=m) < 0.9, core, tail)
    dm = dm[(dm > 0.1396) & (dm < 0.1664)]
    Z = ZachFull()
    Z.total.setData(Z.dataset(dm))
    print("zach", len(dm), Z.total.calculateNLL(), Z.total.calculateNLL(), gf.last_kernel_name())
print("launches", gf.launch_count())
