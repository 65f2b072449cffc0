"""GPU-side diagnostic: Minuit2 status flags of the model-drawn Dalitz fit at several sample sizes (no reference run)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import goofit_b200 as gf
import scale
from models import Dalitz

for n in (100_000, 300_000, 1_000_000):
    M = Dalitz(float_masses=False, grid=240)
    m12, m13 = scale.dalitz_model_events(gf, M, n, 31, chunk=max(4 * n, 1_000_000))
    M = Dalitz(float_masses=False, grid=240)
    data = M.dataset(m12, m13)
    M.signal.setData(data)
    M.signal.setDataSize(len(data))
    fm = gf.FitManager(M.signal)
    fm.setVerbosity(0)
    res = fm.fit()
    print(n, "nfcn", res.NFcn(), "fval", res.Fval(), "edm", res.Edm(), "IsValid", res.IsValid(), "HasCov", res.HasCovariance(),
          "ValidCov", res.HasValidCovariance(), "Accurate", res.HasAccurateCovar(), "CallLimit", res.HasReachedCallLimit(),
          "AboveMaxEdm", res.IsAboveMaxEdm(), "fm", bool(fm), flush=True)
