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


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def run_reference(config, n_events, steps, warmup):
    """time the reference's own CPU implementation (oracle/_ref/ref_driver bench ...)"""
    if not os.path.exists(REF_DRIVER):
        return None
    env = dict(os.environ)
    # torch.distributed.run exports OMP_NUM_THREADS=1 to its workers: the reference arm must not inherit that
    env["OMP_NUM_THREADS"] = str(host_cores())
    env.pop("OMP_PROC_BIND", None)
    out = subprocess.run([REF_DRIVER, "bench", config, str(n_events), str(steps), str(warmup)], env=env,
                         capture_output=True, text=True, timeout=1500)
    for line in out.stdout.splitlines()[::-1]:
        if line.startswith("{"):
            return json.loads(line)
    raise RuntimeError("ref_driver failed: " + out.stderr[-500:])


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


def reference_arm(args, rank):
    """--impl reference: the reference's own CPU path on this box's host cores; rank 0 only."""
    if rank != 0:
        return 0
    sample = args.ref_events
    r = run_reference("dalitz", sample, args.steps, args.warmup)
    if r is None:
        emit({"impl": "reference", "unavailable": "oracle/_ref/ref_driver not built (needs /root/reference at build time)"})
        return 0
    v = r["events_x_evals_per_s"]
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "events_per_step": sample, "note": "bounded sample of the 10^7-event workload; metric is per event"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": r["threads"], "kind": "reference",
                         "sample": "%d events x %d FCN calls, unmodified reference, GOOFIT_DEVICE=OMP, -O3 -march=x86-64-v3" % (sample, args.steps)},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "setdata_s": r["setdata_s"], "first_call_s": r["first_call_s"],
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
    ap.add_argument("--events", type=int, default=10_000_000, help="events per GPU")
    ap.add_argument("--ref-events", type=int, default=1_000_000, help="sample size of the CPU reference arm")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the DP4 / TDDP4 / cache-kernel / FP64-peak extras (N=1 only)")
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
    from models import Dalitz, dalitz_uniform_events

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

    # ---- build the model and the synthetic dataset (each rank owns its own 10^7-event shard) ----
    M = Dalitz(float_masses=False, grid=240)
    m12, m13 = dalitz_uniform_events(N, 12345 + rank)
    data = M.dataset(m12, m13)
    if world > 1:  # attach the NCCL communicator: gf_eval_sum then returns the all-rank sum
        uid = (C.c_char * 128)()
        if rank == 0:
            assert lib.gf_comm_unique_id(uid) == 0, capi.last_error()
        box = [bytes(uid)]
        dist.broadcast_object_list(box, src=0)
        uid = (C.c_char * 128).from_buffer_copy(box[0])
        assert lib.gf_comm_init(ctx, uid, rank, world) == 0, capi.last_error()

    t0 = time.perf_counter()
    M.signal.setData(data)
    M.signal.setDataSize(len(data))
    t_setdata = time.perf_counter() - t0
    t0 = time.perf_counter()
    nll0 = M.signal.calculateNLL()
    t_first = time.perf_counter() - t0

    fm = gf.FitManager(M.signal)
    fcn, params = fm.getFCN(), fm.getParams()
    v0 = np.array(params.make_minuit_vector())
    idx = [c.fitter_index for c in M.coefs[2:]]

    def vec(k):
        v = v0.copy()
        for j, i in enumerate(idx):
            v[i] = v0[i] * (1.0 + 1e-3 * ((k + j) % 3 - 1))
        return v

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        lib.gf_sync(ctx)

    # ---- e2e: host API, host vector in -> host double out ----
    for k in range(W):
        fcn(list(vec(k)))
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    kernel_ms = []
    launches0 = gf.launch_count()
    vecs = [list(vec(W + k)) for k in range(K)]  # Minuit hands over a ready vector: build them outside the timed region
    t0 = time.perf_counter()
    for k in range(K):
        last = fcn(vecs[k])
        kernel_ms.append(gf.last_kernel_ms())
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
    dptr = C.POINTER(C.c_double)

    def step(k):
        v = vec(k)
        p = base_p.copy()
        cv = v[coef_fit_index]
        p[poff:poff + 32] = cv
        c = cv[0::2] + 1j * cv[1::2]
        integral = float(np.real(c @ I @ np.conj(c))) * bin_area  # Amp3Body.cu:334-357
        n = base_n.copy()
        n[nslot] = 1.0 / integral
        rc = lib.gf_set_params_async(ctx, p.ctypes.data_as(dptr), C.c_int(len(p)), n.ctypes.data_as(dptr), C.c_int(len(n)))
        rc |= lib.gf_eval_sum_async(ctx, C.c_int(root), C.c_int(1))
        assert rc == 0, capi.last_error()

    for k in range(W):
        step(k)
    barrier()
    launches0 = gf.launch_count()
    assert lib.gf_timer_start(ctx) == 0
    for k in range(K):
        step(W + k)
    ms = C.c_float(0)
    assert lib.gf_timer_stop(ctx, C.byref(ms)) == 0, capi.last_error()
    barrier()
    launches_value = gf.launch_count() - launches0
    clocks = sampler.stop()
    # the queued path must reproduce the host-API value of the same parameter point
    out = C.c_double(0)
    lib.gf_eval_sum(ctx, C.c_int(root), C.c_int(1), C.byref(out))
    check = fcn(list(vec(W + K - 1)))
    if abs(out.value - check) > 1e-12 * abs(check):
        raise SystemExit("bench.py: queued C-ABI path and host API disagree: %r vs %r" % (out.value, check))

    t_dev = float(ms.value) * 1e-3
    stats = torch.tensor([t_dev, t_e2e, float(np.mean(kernel_ms))], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
    t_dev, t_e2e, k_ms = [float(x) for x in stats.cpu()]

    if rank == 0:
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        total_events = N * world
        achieved = BYTES_PER_EVENT * N / (k_ms * 1e-3) / 1e9
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            with open(tp) as f:
                traffic = json.load(f).get("k_eval_sum_dram_bytes_per_launch")
        line = {
            "metric": METRIC, "value": total_events * K / t_dev, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": 1e3 * t_dev / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "events_per_gpu": N, "total_events": total_events, "free_params": 30,
                       "parallelism": "event-sharded x%d, NCCL all-reduce of the partial NLL" % world,
                       "l2": "inputs (%.2f GB of cached waves + columns per GPU) exceed the 126 MB L2; no flush needed" % (BYTES_PER_EVENT * N / 1e9)},
            "e2e": {"value": total_events * K / t_e2e, "unit": UNIT, "h2d_bytes_per_step": int(8 * (len(base_p) + len(base_n))),
                    "d2h_bytes_per_step": 8, "ms_per_step": 1e3 * t_e2e / K,
                    "note": "FCN::operator() through the host API; dataset resident after setData",
                    "setdata_s": t_setdata, "setdata_h2d_bytes": int(16 * N), "first_call_s": t_first},
            "gpu_launches": int(launches_value), "gpu_launches_e2e": int(launches_e2e),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "kernel": "k_eval_sum", "kernel_ms": k_ms,
                         "algorithmic_bytes_per_launch": BYTES_PER_EVENT * N,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)"},
            "clocks": clocks, "nll": last,
        }
        line["e2e"]["including_upload_and_first_call"] = total_events * K / (t_e2e + t_setdata + t_first)
        if world == 1 and not args.no_secondary:
            import bench_extra
            line["workloads"] = bench_extra.run_all(gf, lib, ctx, M, N, peak)
        if not args.no_cpu_baseline:
            try:
                r = run_reference("dalitz", args.ref_events, 5, 1)
                if r:
                    line["cpu_baseline"] = {"value": r["events_x_evals_per_s"], "unit": UNIT, "cores": r["threads"], "kind": "reference",
                                            "sample": "%d of the %d events x 5 FCN calls, unmodified reference (OMP backend)" % (args.ref_events, N),
                                            "ms_per_step": r["ms_per_step"], "setdata_s": r["setdata_s"], "first_call_s": r["first_call_s"]}
            except Exception as e:  # the baseline is reported, never required
                line["cpu_baseline"] = {"error": str(e)[:200]}
        emit(line)
    if world > 1:
        lib.gf_comm_destroy(ctx)
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
