#!/usr/bin/env python
"""bench.py — events x NLL-evaluations / s of the Amp3Body Dalitz fit (BASELINE.json configs[2]).

Workload (config.workload): examples/dalitz — Amp3Body D0 -> pi+ pi- pi0, 16 resonances, constant efficiency,
240x240 normalisation grid, 10^7 synthetic events PER GPU (weak scaling; events uniform inside the Dalitz
boundary, seeded). One "step" = one Minuit-style likelihood call: the 30 free amplitude coefficients move
(masses/widths fixed as the shipped example does, dalitz.cpp:298-305), the normalisation integral is re-summed
from the cached I_ij on the host, and the fused evaluate-log-reduce kernel streams the cached waves.

  value : K steps queued back to back through the C ABI with everything resident in HBM, timed with CUDA events
          on the library's stream (parameter + normalisation upload, kernel, NCCL all-reduce when N > 1).
  e2e   : the same K steps through the reference-facing host API (FCN::operator() = Params::from_minuit_vector +
          GooPdf::calculateNLL): host parameter vector in, host double out, one synchronisation per call.
          Dataset upload (setData) and the first call (cache fill + integrals) are reported beside it.
  --impl reference : the unmodified reference (oracle/_ref, OpenMP Thrust backend) on the host cores.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "events_x_nll_evals_per_s"
UNIT = "events*evals/s"
WORKLOAD = "examples/dalitz: Amp3Body D0->pi+pi-pi0, 16 resonances, 240x240 norm grid"
BYTES_PER_EVENT = 272  # m12, m13 (16 B) + 16 cached complex waves (256 B); evtNum is the row index
REF_DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver")


import scale  # tests/scale.py: shared input generation + the reference-arm runner (identical inputs for both arms)  # noqa: E402
from scale import host_cores, run_reference  # noqa: E402


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dalitz_events(n, rank):
    """the synthetic workload: n points uniform inside the Dalitz boundary, numpy Generator seeded 12345 + rank. Rank 0's
    shard is also the whole fixed-size (strong-scaling) dataset and what the reference arm is timed on."""
    from models import dalitz_uniform_events
    return dalitz_uniform_events(n, 12345 + rank)


def reference_on_events(m12, m13, steps, warmup, neumaier):
    d = scale.write_columns(scale.scratch_dir("bench_dalitz"), {"m12": m12, "m13": m13})
    try:
        return run_reference("dalitz", 0, steps, warmup, data=d, neumaier=neumaier)
    finally:
        import shutil
        shutil.rmtree(d, ignore_errors=True)


def reference_arm(args, rank):
    """--impl reference: the reference's own CPU path (oracle/_ref: the unmodified sources, OpenMP Thrust backend) on this
    box's host cores, all of them, on the SAME 10^7 events the B200 arm's rank 0 streams; rank 0 only. At N > 1 the B200
    arm's weak-scaling workload is N such shards; the reference is timed on one of them (bounded sample, per-event metric)."""
    if rank != 0:
        return 0
    if not scale.have_reference():
        emit({"impl": "reference", "unavailable": "oracle/_ref/ref_driver not built (needs /root/reference at build time)"})
        return 0
    n = args.events
    m12, m13 = dalitz_events(n, 0)
    r = reference_on_events(m12, m13, args.steps, args.warmup, False)
    v = r["events_x_evals_per_s"]
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "events_per_gpu": n, "total_events": n, "free_params": 30,
                   "note": "the same %d events (numpy seed 12345) and the same parameter schedule as the B200 arm's rank 0; at N > 1 "
                           "this is one shard of the weak-scaling workload (bounded sample; the metric is per event)" % n},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": r["threads"], "kind": "reference",
                         "sample": "%d events x %d FCN calls, unmodified reference, GOOFIT_DEVICE=OMP, -O3 -march=x86-64-v3" % (r["n_events"], args.steps)},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "setdata_s": r["setdata_s"], "first_call_s": r["first_call_s"], "nll": r["last_nll"],
    }
    emit(line)
    return 0


_REAL_STDOUT = None


def claim_stdout():
    """stdout carries exactly ONE JSON line: anything a library prints there (NCCL's version banner, for one) is sent to
    stderr by pointing fd 1 at fd 2 for the duration of the run; emit() writes the line to the saved descriptor."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--events", type=int, default=10_000_000, help="events per GPU (weak scaling) = total events of the fixed-size run")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the DP4 / TDDP4 / zachFit / fit / cache-kernel extras (N=1 only)")
    ap.add_argument("--no-strong", action="store_true", help="skip the fixed-size (strong-scaling) pass at N > 1")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else max(args.warmup, 1)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return reference_arm(args, rank)

    import torch
    import torch.distributed as dist

    import goofit_b200 as gf
    from goofit_b200 import capi
    from models import Dalitz

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    gf.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world)

    N = args.events
    K, W = args.steps, args.warmup
    lib = capi.lib()
    ctx = C.c_void_p(gf.ctx_handle())
    t_comm = 0.0
    if world > 1:  # attach the communicator: gf_eval_sum then returns the all-rank sum (warm-up exchange included)
        t0 = time.perf_counter()
        uid = (C.c_char * 128)()
        if rank == 0:
            assert lib.gf_comm_unique_id(uid) == 0, capi.last_error()
        box = [bytes(uid)]
        dist.broadcast_object_list(box, src=0)
        uid = (C.c_char * 128).from_buffer_copy(box[0])
        assert lib.gf_comm_init(ctx, uid, rank, world) == 0, capi.last_error()
        t_comm = time.perf_counter() - t0

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        lib.gf_sync(ctx)

    dptr = C.POINTER(C.c_double)

    def measure(m12, m13, sampler=None):
        """one pass of the benchmark on the dataset (m12, m13) as currently sharded: e2e through the host API, then the
        same K steps queued through the C ABI and timed with CUDA events on the library's stream"""
        M = Dalitz(float_masses=False, grid=240)
        data = M.dataset(m12, m13)
        t0 = time.perf_counter()
        M.signal.setData(data)
        M.signal.setDataSize(len(data))
        t_setdata = time.perf_counter() - t0
        t0 = time.perf_counter()
        M.signal.calculateNLL()
        t_first = time.perf_counter() - t0
        fm = gf.FitManager(M.signal)
        fcn, params = fm.getFCN(), fm.getParams()
        v0 = np.array(params.make_minuit_vector())
        idx = [c.fitter_index for c in M.coefs[2:]]
        vec = lambda k: scale.mover_vector(v0, idx, k)
        # ---- e2e: host API, host vector in -> host double out ----
        for k in range(W):
            fcn(list(vec(k)))
        if sampler:
            sampler.start()
        barrier()
        launches0 = gf.launch_count()
        vecs = [list(vec(W + k)) for k in range(K)]  # Minuit hands over a ready vector: build them outside the timed region
        t0 = time.perf_counter()
        for k in range(K):
            last = fcn(vecs[k])
        barrier()
        t_e2e = time.perf_counter() - t0
        launches_e2e = gf.launch_count() - launches0
        # ---- value: resident inputs, K steps queued back to back through the C ABI, CUDA-event timed ----
        root = M.signal.getFunctionIndex()
        poff = M.signal.getParameterIndex()
        nslot = M.signal.getNormalizationIndex()
        base_p = np.array(gf.program_params(), dtype=np.float64)
        base_n = np.array(gf.program_norms(), dtype=np.float64)
        I = np.asarray(M.signal.getIntegrals()).reshape(16, 16)
        bin_area = M.m12.bin_size * M.m13.bin_size
        coef_fit_index = [c.fitter_index for c in M.coefs]

        def blocks(k):  # parameter + normalisation block of step k (what FCN::operator() would push)
            v = vec(k)
            p = base_p.copy()
            cv = v[coef_fit_index]
            p[poff:poff + 32] = cv
            c = cv[0::2] + 1j * cv[1::2]
            integral = float(np.real(c @ I @ np.conj(c))) * bin_area  # Amp3Body.cu:334-357
            n = base_n.copy()
            n[nslot] = 1.0 / integral
            return p, n

        def step(pn):
            p, n = pn
            rc = lib.gf_set_params_async(ctx, p.ctypes.data_as(dptr), C.c_int(len(p)), n.ctypes.data_as(dptr), C.c_int(len(n)))
            rc |= lib.gf_eval_sum_async(ctx, C.c_int(root), C.c_int(1))
            assert rc == 0, capi.last_error()

        pre = [blocks(k) for k in range(W + K)]
        for k in range(W):
            step(pre[k])
        barrier()
        launches0 = gf.launch_count()
        assert lib.gf_timer_start(ctx) == 0
        for k in range(K):
            step(pre[W + k])
        ms = C.c_float(0)
        assert lib.gf_timer_stop(ctx, C.byref(ms)) == 0, capi.last_error()
        barrier()
        launches_value = gf.launch_count() - launches0
        # the dominant kernel alone: CUDA events around each of K more launches of the same queued steps (opt-in profiling:
        # the timed regions above run without per-launch events)
        assert lib.gf_profile_begin(ctx) == 0
        for k in range(K):
            step(pre[W + k])
        kms, kn = C.c_float(0), C.c_int(0)
        assert lib.gf_profile_end(ctx, C.byref(kms), C.byref(kn)) == 0 and kn.value == K, capi.last_error()
        barrier()
        clocks = sampler.stop() if sampler else None
        # the queued path must reproduce the host-API value of the same parameter point
        out = C.c_double(0)
        lib.gf_eval_sum(ctx, C.c_int(root), C.c_int(1), C.byref(out))
        check = fcn(list(vec(W + K - 1)))
        if abs(out.value - check) > 1e-12 * abs(check):
            raise SystemExit("bench.py: queued C-ABI path and host API disagree: %r vs %r" % (out.value, check))
        nll_check = fcn(list(vec(3)))  # the point the reference's cpu_baseline run ends on (steps 3, warmup 1)
        stats = torch.tensor([float(ms.value) * 1e-3, t_e2e, float(kms.value), t_setdata, t_first], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(stats, op=dist.ReduceOp.MAX)
        t_dev, t_e2e, k_ms, t_setdata, t_first = [float(x) for x in stats.cpu()]
        return dict(M=M, t_dev=t_dev, t_e2e=t_e2e, k_ms=k_ms, t_setdata=t_setdata, t_first=t_first, last=last, nll_check=nll_check,
                    launches_value=int(launches_value), launches_e2e=int(launches_e2e), clocks=clocks,
                    h2d=int(8 * (len(base_p) + len(base_n))), local_events=int(gf.local_events()))

    # ---- weak scaling: every rank owns its own 10^7-event shard ----
    m12, m13 = dalitz_events(N, rank)
    R = measure(m12, m13, ClockSampler(local_rank))
    # ---- strong scaling: the SAME 10^7 events (rank 0's) sharded over the ranks by the library (PdfBase.cu:364-378 split) ----
    S = None
    if world > 1 and not args.no_strong:
        del R["M"]
        g12, g13 = dalitz_events(N, 0)
        gf.set_shard(rank, world)
        S = measure(g12, g13)
        gf.set_shard(0, 1)

    if rank == 0:
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        total_events = N * world
        t_dev, t_e2e, k_ms = R["t_dev"], R["t_e2e"], R["k_ms"]
        achieved = BYTES_PER_EVENT * N / (k_ms * 1e-3) / 1e9
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            with open(tp) as f:
                traffic = json.load(f).get("k_eval_sum_dram_bytes_per_launch")
        comm = "none (one GPU)" if world == 1 else capi.comm_path(ctx)
        line = {
            "metric": METRIC, "value": total_events * K / t_dev, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": 1e3 * t_dev / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "events_per_gpu": N, "total_events": total_events, "free_params": 30,
                       "parallelism": "event-sharded x%d; partial NLL exchanged by: %s" % (world, comm),
                       "l2": "inputs (%.2f GB of cached waves + columns per GPU) exceed the 126 MB L2; no flush needed" % (BYTES_PER_EVENT * N / 1e9)},
            "e2e": {"value": total_events * K / t_e2e, "unit": UNIT, "h2d_bytes_per_step": R["h2d"],
                    "d2h_bytes_per_step": 8, "ms_per_step": 1e3 * t_e2e / K,
                    "note": "FCN::operator() through the host API; dataset resident after setData",
                    "setdata_s": R["t_setdata"], "setdata_h2d_bytes": int(16 * N), "first_call_s": R["t_first"], "comm_init_s": t_comm},
            "gpu_launches": R["launches_value"], "gpu_launches_e2e": R["launches_e2e"],
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "kernel": "k_eval_sum", "kernel_ms": k_ms,
                         "algorithmic_bytes_per_launch": BYTES_PER_EVENT * N,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)"},
            "clocks": R["clocks"], "nll": R["last"],
            "overhead_us_per_call": {"device_timed": 1e3 * (1e3 * t_dev / K - k_ms), "e2e": 1e3 * (1e3 * t_e2e / K - k_ms)},
        }
        line["e2e"]["including_upload_and_first_call"] = total_events * K / (t_e2e + R["t_setdata"] + R["t_first"])
        if S is not None:
            line["strong"] = {
                "scaling": "strong", "total_events": N, "events_per_gpu": S["local_events"], "value": N * K / S["t_dev"], "unit": UNIT,
                "ms_per_step": 1e3 * S["t_dev"] / K, "kernel_ms": S["k_ms"],
                "e2e": {"value": N * K / S["t_e2e"], "ms_per_step": 1e3 * S["t_e2e"] / K, "setdata_s": S["t_setdata"], "first_call_s": S["t_first"]},
                "overhead_us_per_call": {"device_timed": 1e3 * (1e3 * S["t_dev"] / K - S["k_ms"]), "e2e": 1e3 * (1e3 * S["t_e2e"] / K - S["k_ms"])},
                "note": "the fixed-size workload BASELINE.json quotes: rank 0's %d events sharded over %d GPUs" % (N, world)}
        # ---- CPU baseline + parity at scale: the unmodified reference on the SAME events (rank 0's shard == the fixed-size set) ----
        if not args.no_cpu_baseline and scale.have_reference():
            try:
                g12, g13 = dalitz_events(N, 0)
                r = reference_on_events(g12, g13, 3, 1, True)
                line["cpu_baseline"] = {"value": r["events_x_evals_per_s"], "unit": UNIT, "cores": r["threads"], "kind": "reference",
                                        "sample": "the same %d events x 3 FCN calls, unmodified reference (OMP backend, all host cores)" % r["n_events"],
                                        "ms_per_step": r["ms_per_step"], "setdata_s": r["setdata_s"], "first_call_s": r["first_call_s"]}
                mine = (S if S is not None else R)["nll_check"]
                line["parity_at_scale"] = {
                    "rel": scale.rel(mine, r["neumaier_last_nll"]), "n_events": r["n_events"], "n_gpus": world,
                    "nll": mine, "reference_neumaier_nll": r["neumaier_last_nll"], "reference_reduce_nll": r["last_nll"],
                    "reference_reduce_rel": scale.rel(r["last_nll"], r["neumaier_last_nll"]),
                    "what": ("all-reduced NLL of the %d events sharded over %d GPUs" % (N, world) if world > 1 else "NLL of the %d events" % N)
                            + " vs a Neumaier sum of the reference's own per-event terms on the same events at the same parameter point "
                              "(step 3 of the shared schedule)"}
            except Exception as e:  # the baseline is reported, never required
                line["cpu_baseline"] = {"error": str(e)[:300]}
        if world == 1 and not args.no_secondary:
            import bench_extra
            line["workloads"] = bench_extra.run_all(gf, lib, ctx, R["M"], N, peak, with_reference=not args.no_cpu_baseline)
        emit(line)
    if world > 1:
        lib.gf_comm_destroy(ctx)
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
