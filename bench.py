#!/usr/bin/env python
"""bench.py -- DM x acceleration trials per second of the per-DM acceleration search.

Workload (BASELINE.json configs[1]/[3]): 8-bit DM trials of 2^23 (+4096) samples, tsamp 64 us,
101 accelerations (-50..50 m/s^2 in steps of 1), 4 folds (16 harmonics), min S/N 9, the
reference's default zaplist; the 1024-DM job is processed in steps of --ndm DM trials per GPU
(DM trials are independent units, so a step is a slice of that job, and ranks each take their
own slice: weak scaling, no data-path collective, one gather of candidate records per step).

  value : trials/s with the u8 trials already resident in HBM (timed: per DM whitening, 101 x
          (resample+R2C+interbin+normalise+harmonic sums+peaks), peak D2H, host distillation)
  e2e   : same through ps_pipeline_search with the trials in pinned HOST memory (H2D inside)
  roofline     : the acceleration-trial kernel chain timed with CUDA events on its own stream
  cpu_baseline : the CPU oracle port on a bounded sample (rank 0, N = 1)
--impl reference runs the reference's own implementation: its unmodified CUDA build
(oracle/_ref: kernels.cu + transform headers + cuFFT, replaying Worker::start) -- the reference
has no CPU search path -- on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

SIZE = 1 << 23
NSAMPS = SIZE + 4096
TSAMP = 64e-6
ACCS = np.arange(-50, 51, 1, dtype=np.float32)
NFOLDS = 4
# dram__bytes_read.sum + dram__bytes_write.sum of the four chain kernels, per acceleration trial, from the
# committed ncu launch list of `tools/profile_chain.py 26` (26 trials per launch): pass 1 40.4 MB,
# pass 2 65.1 MB, spectrum 49.0 MB, harmonic sums + peaks 16.8 MB
NCU_DRAM_BYTES_PER_TRIAL = 171_300_000
NCU_DRAM_SOURCE = "profiles/r01q_launches_ncu.csv (ncu, 26 trials per launch, divided by 26)"
A_INJ, P0 = 31.4, 7.5e-3
ZAP_F = np.array([50.0, 100.0, 150.0, 200.0, 250.0], np.float32)  # misc/default_zaplist.txt
ZAP_W = np.array([0.1, 0.15, 0.15, 0.15, 0.15], np.float32)
ALGO_BYTES_PER_TRIAL = 16 * SIZE  # 16*N bytes per acceleration trial


def synth_trials(ndm, device, seed):
    """Synthetic u8 DM trials generated on the device (Gaussian baseline mean 128 sigma 16, red
    ramp, 50/100 Hz birdies, one accelerated 5 %-duty pulsar brightest in the middle trial)."""
    import torch
    g = torch.Generator(device=device).manual_seed(20140213 + seed)
    j = torch.arange(NSAMPS, device=device, dtype=torch.float64)
    af = float(np.float32(A_INJ) * np.float32(TSAMP)) / (2 * 299792458.0)
    phase = torch.remainder((j - af * j * (j - SIZE)) * (TSAMP / P0), 1.0)
    pulse = (phase < 0.05).float()
    t = (j * TSAMP)
    birdie = (3.0 * torch.sin(2 * np.pi * 50.0 * t) + 2.0 * torch.sin(2 * np.pi * 100.0 * t + 1.0)).float()
    ramp = (8.0 * torch.sin(2 * np.pi * t / 301.0)).float()
    out = torch.empty((ndm, NSAMPS), dtype=torch.uint8, device=device)
    for d in range(ndm):
        w = max(0.0, 1.0 - abs(d - ndm // 2) / 8.0)
        x = 128.0 + 16.0 * torch.randn(NSAMPS, device=device, generator=g) + birdie + ramp + (3.0 * w) * pulse
        out[d] = x.round_().clamp_(0, 255).to(torch.uint8)
    return out


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._halt = index, [], threading.Event()

    def run(self):
        try:
            p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                  "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            return
        while not self._halt.is_set():
            line = p.stdout.readline()
            if not line:
                break
            self.rows.append([c.strip() for c in line.split(",")])
        p.terminate()

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def cpu_baseline(n_acc=12):
    """The CPU oracle port on a bounded sample: one DM trial whitened + n_acc acceleration trials
    of the same 2^23 workload, one thread."""
    from oracle import oracle as O
    sys.path.insert(0, os.path.join(REPO, "tests"))
    from synth import make_trials
    trial = make_trials(1, NSAMPS, SIZE, tsamp=TSAMP, a_inj=A_INJ, p0=P0, amp=3.0, seed=1)
    cfg = O.SearchConfig(SIZE, TSAMP, nharmonics=NFOLDS, zap=(ZAP_F, ZAP_W))
    accs = ACCS[::max(1, len(ACCS) // n_acc)][:n_acc]
    t0 = time.perf_counter()
    O.search(trial, [0.0], accs, cfg)
    dt = time.perf_counter() - t0
    return {"value": len(accs) / dt, "unit": "trials/s", "cores": 1, "kind": "port",
            "sample": f"1 DM trial (2^23 samples) whitened + {len(accs)} acceleration trials, oracle/oracle.py "
                      f"(numpy pocketfft + oracle.cpp), {dt:.1f} s"}


def load_ref():
    path = os.path.join(REPO, "oracle", "_ref", "libpeasoup_ref.so")
    if not os.path.exists(path):
        return None
    r = C.CDLL(path)
    r.ref_last_error.restype = C.c_char_p
    r.ref_worker_search.restype = C.c_long
    return r


def run_reference(args):
    """--impl reference: the reference's own code on a bounded sample, rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    base = {"metric": "DMxaccel trials/sec (2^23-pt)", "unit": "trials/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "impl": "reference", "config": workload_config(args, args.ref_ndm)}
    ref = load_ref()
    if ref is None or not torch.cuda.is_available():
        cb = cpu_baseline(8)
        base.update({"value": cb["value"], "ms_per_step": None, "cpu_baseline": cb,
                     "e2e": {"value": cb["value"], "unit": "trials/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                     "note": "oracle/_ref not available: CPU oracle port timed instead"})
        print(json.dumps(base))
        return
    torch.cuda.set_device(0)
    ndm = args.ref_ndm
    trials = synth_trials(ndm, torch.device("cuda:0"), 0).cpu().numpy()

    class P(C.Structure):
        _fields_ = [("size", C.c_uint), ("nsamps", C.c_uint), ("tsamp", C.c_float), ("nharmonics", C.c_int),
                    ("min_snr", C.c_float), ("min_freq", C.c_float), ("max_freq", C.c_float), ("freq_tol", C.c_float),
                    ("max_harm", C.c_int), ("zapfile", C.c_char_p)]
    zp = "/tmp/peasoup_b200_bench_zaplist.txt"
    with open(zp, "w") as f:
        for a, b in zip(ZAP_F, ZAP_W):
            f.write(f"{a} {b}\n")
    p = P(SIZE, NSAMPS, TSAMP, NFOLDS, 9.0, 0.1, 1100.0, 1e-4, 16, zp.encode())
    cap = 1 << 20
    o = [np.zeros(cap, t) for t in ("f4", "i4", "f4", "i4", "f4", "f4", "i4")]
    dms = np.linspace(0, 100, ndm).astype(np.float32)
    vp = C.c_void_p
    ref.ref_worker_search.argtypes = [C.POINTER(P), vp, vp, C.c_int, vp, C.c_int, C.c_long] + [vp] * 7 + [C.POINTER(C.c_double)]
    secs = C.c_double()
    times = []
    for it in range(args.warmup + args.steps):
        n = ref.ref_worker_search(C.byref(p), trials.ctypes.data, dms.ctypes.data, ndm, ACCS.ctypes.data, len(ACCS), cap,
                                  *[a.ctypes.data for a in o], C.byref(secs))
        if n < 0:
            raise RuntimeError(ref.ref_last_error().decode())
        if it >= args.warmup:
            times.append(secs.value)
    tot = sum(times)
    val = ndm * len(ACCS) * args.steps / tot
    sample = f"{ndm} DM trials x {len(ACCS)} accelerations per step through the reference's unmodified CUDA build " \
             f"(kernels.cu + cuFFT, Worker::start replay, 1 host thread, 1 B200; the reference has no CPU search path)"
    base.update({"value": val, "ms_per_step": 1e3 * tot / args.steps,
                 "cpu_baseline": {"value": val, "unit": "trials/s", "cores": 1, "kind": "reference", "sample": sample},
                 "e2e": {"value": val, "unit": "trials/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "host_cores": os.cpu_count()})
    print(json.dumps(base))


def workload_config(args, ndm):
    return {"workload": "BASELINE configs[1]: 2^23-sample 8-bit DM trials x 101 accel (-50..50 m/s^2), 4 folds (16 harmonics), "
                        "min S/N 9, default zaplist; 1024-DM job processed in steps of ndm_per_step_per_gpu DM trials",
            "fft_size": SIZE, "nsamps": NSAMPS, "tsamp": TSAMP, "n_accel": int(len(ACCS)), "nfolds": NFOLDS,
            "ndm_per_step_per_gpu": ndm, "trials_per_step_per_gpu": ndm * int(len(ACCS)),
            "parallelism": f"dm-shard x{args.gpus}",
            "l2": "per-step input (ndm x 8 MiB u8) and per-trial working set (>=96 MiB) exceed the 126 MB L2"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--ndm", type=int, default=16, help="DM trials per step per GPU")
    ap.add_argument("--ref-ndm", type=int, default=2, help="DM trials per step of the reference arm")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import torch.distributed as dist

    import peasoup_b200 as pb
    from peasoup_b200.host import make_search_params
    from peasoup_b200.shard import gather_candidates

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    lib = pb.load()

    ndm = args.ndm
    d_trials = synth_trials(ndm, device, rank)
    h_trials = torch.empty((ndm, NSAMPS), dtype=torch.uint8).pin_memory()
    h_trials.copy_(d_trials)
    torch.cuda.synchronize()
    dms = np.linspace(0, 100, ndm).astype(np.float32)
    prm = pb.PipelineParams(make_search_params(SIZE, NSAMPS, TSAMP, NFOLDS, 9.0), 1e-4, 16, 0.0, 0.0, 1.1, 64.0, 1382.0, -0.390625)
    offset = rank * ndm

    pipe = pb.Pipeline(prm, zap=(ZAP_F, ZAP_W), device0=local, nworkers=1)

    def step(src):
        recs, done = pipe.run(src, dms, accs=ACCS, dm_idx_offset=offset)
        merged = gather_candidates(recs, device=device) if world > 1 else recs
        return done, merged

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def timed(src, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        done = 0
        merged = None
        for _ in range(steps):
            d, merged = step(src)
            done += d
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, done, merged

    for _ in range(args.warmup):
        step(d_trials)
    sampler = ClockSampler(local)
    sampler.start()
    l0 = lib.ps_launch_count()
    ms, done, merged = timed(d_trials, args.steps)
    launches = lib.ps_launch_count() - l0
    # end to end: trials in pinned host memory, H2D inside the timed region
    h0, d0 = C.c_ulonglong(), C.c_ulonglong()
    lib.ps_transfer_bytes(C.byref(h0), C.byref(d0))
    ms_e2e, done_e2e, _ = timed(h_trials, args.steps)
    h1, d1 = C.c_ulonglong(), C.c_ulonglong()
    lib.ps_transfer_bytes(C.byref(h1), C.byref(d1))
    prof = (C.c_double * 4)()
    lib.ps_pipeline_profile(prof)
    clocks = sampler.stop()

    # roofline: the acceleration-trial kernel chain alone, CUDA events on its own stream
    st = torch.cuda.Stream(device=device)
    s = pb.Search(make_search_params(SIZE, NSAMPS, TSAMP, NFOLDS, 9.0))
    s.set_zaplist(ZAP_F, ZAP_W)
    s.whiten(d_trials[ndm // 2], stream=st)
    s.reset_peaks(stream=st)
    s.accel_batch(ACCS[:8], stream=st)
    st.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    s.accel_batch(ACCS, stream=st)
    e1.record(st)
    st.synchronize()
    chain_ms = e0.elapsed_time(e1) / len(ACCS)
    s.close()
    peaks = json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(REPO, "MEASURED_PEAKS.json")) else {}
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = ALGO_BYTES_PER_TRIAL / (chain_ms * 1e-3) / 1e9

    if rank == 0:
        total_trials = done * world
        out = {
            "metric": "DMxaccel trials/sec (2^23-pt)", "value": total_trials / (ms * 1e-3), "unit": "trials/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, ndm),
            "e2e": {"value": done_e2e * world / (ms_e2e * 1e-3), "unit": "trials/s",
                    "h2d_bytes_per_step": (h1.value - h0.value) // args.steps, "d2h_bytes_per_step": (d1.value - d0.value) // args.steps},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": NCU_DRAM_BYTES_PER_TRIAL, "traffic_source": NCU_DRAM_SOURCE, "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650",
                         "kernel": "acceleration-trial chain: fft_pass1(resample gather) + fft_pass2 + spectrum + harmonic_peaks",
                         "algorithmic_bytes_per_launch": ALGO_BYTES_PER_TRIAL, "ms_per_launch": chain_ms},
            "candidates_last_step": int(len(merged)) if merged is not None else None,
            "host_profile_last_step": {"submit_s": prof[0], "wait_device_s": prof[1], "host_distill_s": prof[2],
                                       "raw_peaks": int(prof[3])},
        }
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline()
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
