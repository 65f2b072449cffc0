"""Worker for test_parity_gpu.py::test_two_gpu_sharded_parity — run under torch.distributed.run with one process per GPU.
Events and the MC normalisation sample are sharded over the ranks (PdfBase.cu:364-378 split, remainder on the last
rank); the partial NLL, the MC integral sums and the Gram matrix are all-reduced over NCCL inside the library. Every rank
must reproduce the single-process reference values of tests/golden."""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import goofit_b200 as gf  # noqa: E402
from goofit_b200 import capi  # noqa: E402
from models import DP4, TDDP4, Dalitz  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
gf.set_device(local)
gf.set_shard(rank, world)
dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world)
lib = capi.lib()
ctx = C.c_void_p(gf.ctx_handle())
uid = (C.c_char * 128)()
if rank == 0:
    assert lib.gf_comm_unique_id(uid) == 0, capi.last_error()
box = [bytes(uid)]
dist.broadcast_object_list(box, src=0)
uid = (C.c_char * 128).from_buffer_copy(box[0])
assert lib.gf_comm_init(ctx, uid, rank, world) == 0, capi.last_error()
path = capi.comm_path(ctx)
assert gf.comm_attached() == world

G = os.path.join(ROOT, "tests", "golden")
gold = {}
for f in os.listdir(G):
    if f.startswith("manifest") and f.endswith(".json"):
        gold.update(json.load(open(os.path.join(G, f))))
arr = lambda n: np.load(os.path.join(G, n + ".npy"))
rel = lambda a, b: abs(a - b) / abs(b)
res = {}

M = Dalitz(float_masses=True, grid=120)
d = M.dataset(arr("dalitz.m12"), arr("dalitz.m13"))
M.signal.setData(d)
M.signal.setDataSize(len(d))
res["dalitz"] = rel(M.signal.calculateNLL(), gold["dalitz.nll"])
assert gf.local_events() < len(d)
# per-event values under sharding: every rank gets ALL events back in event order (the reference's MPI_Allgatherv, GooPdf.cu:139-172)
wrap = gf.ProdPdf("wrap", [M.signal])  # as DalitzPlotter.h:60-63 and tests/test_parity_gpu.py::test_dalitz_values do
wrap.setData(d)
M.signal.setDataSize(len(d))
probs = np.asarray(wrap.getCompProbsAtDataPoints()[0])
want = arr("dalitz.prob")
assert probs.shape == want.shape, (probs.shape, want.shape)
res["dalitz.prob"] = float(np.max(np.abs(probs - want) / np.maximum(np.abs(want), 1e-300)))
M.signal.setFitControl(gf.UnbinnedNllFit())
M.signal.setData(d)
M.signal.setDataSize(len(d))
M.signal.calculateNLL()
# fit fractions: the wave Gram matrix is a sum over ranks
for v in M.masses + M.widths:
    v.changed = False
ff = np.asarray(M.signal.fit_fractions())
res["dalitz.fitfrac_sum"] = abs(float(ff.sum()) - 1.0)

names = ("m12", "m34", "cos12", "cos34", "phi")
for mode in ("Gram", "PerEvent"):
    M = DP4(mc_events=0, float_ls=True)
    M.dp.setNormMode(getattr(gf.Amp4BodyBase.NormMode, mode))
    M.dp.setNormEvents(np.vstack([arr("dp4.mc_" + n) for n in names]))
    d = M.dataset([arr("dp4." + n) for n in names])
    M.signal.setData(d)
    M.dp.setDataSize(len(d), 6)
    res["dp4." + mode] = rel(M.signal.calculateNLL(), gold["dp4.nll"])
    res["dp4.int." + mode] = rel(1.0 / M.dp.getNormalization(), gold["dp4.integral"])
    M = TDDP4(three_gauss=False, float_ls=True, float_mix=True)
    M.dp.setNormMode(getattr(gf.Amp4BodyBase.NormMode, mode))
    M.dp.setNormEvents(np.vstack([arr("tddp4.mc_" + n) for n in names]))
    d = M.dataset([arr("tddp4." + n) for n in names + ("dtime", "sigmat")])
    M.signal.setData(d)
    M.dp.setDataSize(len(d), 8)
    res["tddp4." + mode] = rel(M.signal.calculateNLL(), gold["tddp4.nll"])
    res["tddp4.int." + mode] = rel(1.0 / M.dp.getNormalization(), gold["tddp4.integral"])

bad = {k: v for k, v in res.items() if v > (1e-11 if k.startswith("tddp4.int") or k == "dalitz.prob" else 1e-12)}
assert not bad, (rank, bad)
dist.barrier()
lib.gf_comm_destroy(ctx)
if rank == 0:
    print("MGPU_OK", path, "|", json.dumps(res))
dist.destroy_process_group()
