#!/usr/bin/env python
"""bench.py -- DM x acceleration trials per second of the per-DM acceleration search.

--config selects a BASELINE.json configuration (numbered as BASELINE.md lists them, 1-based):

  2 (default)  configs[1]: 8-bit DM trials of 2^23 (+4096) samples, tsamp 64 us, 101 accelerations
               (-50..50 m/s^2 in steps of 1), 4 folds (16 harmonics), min S/N 9, the reference's default
               zaplist.  The 1024-DM job is processed in steps of --ndm DM trials per GPU; ranks each take
               their own slice: weak scaling, no data-path collective, one gather of records per step.
  3            configs[2]: harmonic-sum + peak-find microbench on a batch of 2^22-bin spectra, 4 folds
               (16 harmonics); HBM GB/s on 4 B/bin, batch sweep in `config.sweep`.  One GPU.
  4            configs[3]: the FIXED 1024-DM x 101-acceleration job sharded over the ranks (strong
               scaling); a step = the whole job: every rank searches its shard, ONE gather of candidate
               records at the end, and rank 0 runs the reference's post-join DMDistiller +
               HarmonicDistiller (pipeline_multi.cu:353-365) inside the timed region.
  5            configs[4]: 2^24 samples, 201 accelerations (-500..500 m/s^2 in steps of 5), 4096 DM
               trials on 8 GPUs = 512 per GPU, generated per rank on the device; with fewer GPUs the
               job is the sub-sample of 512 DM trials per GPU that is present (stated in `config`).

  value : trials/s with the u8 trials already resident in HBM (timed: per DM whitening, n_acc x
          (resample+R2C+interbin+normalise+harmonic sums+peaks), peak D2H, host distillation)
  e2e   : same through the pipeline call with the trials in pinned HOST memory (H2D inside)
  roofline     : the acceleration-trial kernel chain timed with CUDA events on its own stream
  cpu_baseline : the CPU oracle port on a bounded sample (rank 0, N = 1)
--impl reference runs the reference's own implementation: its unmodified CUDA build
(oracle/_ref: kernels.cu + transform headers + cuFFT, replaying Worker::start) -- the reference
has no CPU search path -- on the same synthetic DM trials, one worker per GPU like the
reference's `-t N` (pipeline_multi.cu:342-359): every rank runs it on its own device, rank 0 adds up.
"""
from __future__ import annotations

import argparse
import ctypes as C
import glob
import json
import os
import re
import subprocess
import sys
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

TSAMP = 64e-6
NFOLDS = 4
A_INJ, P0 = 31.4, 7.5e-3
ZAP_F = np.array([50.0, 100.0, 150.0, 200.0, 250.0], np.float32)  # misc/default_zaplist.txt
ZAP_W = np.array([0.1, 0.15, 0.15, 0.15, 0.15], np.float32)

# module-level workload of the default configuration (tools/*.py import these)
SIZE = 1 << 23
NSAMPS = SIZE + 4096
ACCS = np.arange(-50, 51, 1, dtype=np.float32)
ALGO_BYTES_PER_TRIAL = 16 * SIZE  # 16*N bytes per acceleration trial


class Workload:
    def __init__(self, config):
        self.config = config
        if config == 5:
            self.size = 1 << 24
            self.accs = np.arange(-500, 501, 5, dtype=np.float32)
            self.a_inj = 312.5
        else:
            self.size = SIZE
            self.accs = ACCS
            self.a_inj = A_INJ
        self.nsamps = self.size + 4096
        self.algo_bytes = 16 * self.size
        lg = self.size.bit_length() - 1
        self.metric = "DMxaccel trials/sec (2^%d-pt)" % lg


def synth_trials(ndm, device, seed, size=SIZE, nsamps=NSAMPS, a_inj=A_INJ, centre=None):
    """Synthetic u8 DM trials generated on the device (Gaussian baseline mean 128 sigma 16, red
    ramp, 50/100 Hz birdies, one accelerated 5 %-duty pulsar brightest in trial `centre` (default the
    middle one) and falling off over +-8 trials)."""
    import torch
    g = torch.Generator(device=device).manual_seed(20140213 + seed)
    j = torch.arange(nsamps, device=device, dtype=torch.float64)
    af = float(np.float32(a_inj) * np.float32(TSAMP)) / (2 * 299792458.0)
    phase = torch.remainder((j - af * j * (j - size)) * (TSAMP / P0), 1.0)
    pulse = (phase < 0.05).float()
    t = (j * TSAMP)
    birdie = (3.0 * torch.sin(2 * np.pi * 50.0 * t) + 2.0 * torch.sin(2 * np.pi * 100.0 * t + 1.0)).float()
    ramp = (8.0 * torch.sin(2 * np.pi * t / 301.0)).float()
    del j, phase, t
    out = torch.empty((ndm, nsamps), dtype=torch.uint8, device=device)
    c = ndm // 2 if centre is None else centre
    for d in range(ndm):
        w = max(0.0, 1.0 - abs(d - c) / 8.0)
        x = 128.0 + 16.0 * torch.randn(nsamps, device=device, generator=g) + birdie + ramp + (3.0 * w) * pulse
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


def ncu_traffic():
    """DRAM bytes per acceleration trial from the newest committed ncu launch list of the chain
    (profiles/r*_chain<N>_ncu.csv: `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,
    dram__bytes_write.sum` around ONE N-trial ps_search_accel_batch of tools/profile_chain.py, captured
    between cudaProfilerStart/Stop).  Returns (bytes per trial, per-kernel MB per trial, source) or
    (None, None, reason)."""
    files = sorted(glob.glob(os.path.join(REPO, "profiles", "r*_chain*_ncu.csv")))
    if not files:
        return None, None, "no profiles/r*_chain<N>_ncu.csv"
    path = files[-1]
    m = re.search(r"_chain(\d+)_ncu\.csv$", path)
    ntrials = int(m.group(1))
    import csv
    rows = [r for r in csv.reader(open(path, newline="")) if r]
    hdr = next((r for r in rows if "Kernel Name" in r and "Metric Name" in r), None)
    if hdr is None:
        return None, None, f"{os.path.basename(path)}: no ncu CSV header"
    ik, im, iu, iv = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Unit"), hdr.index("Metric Value")
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    per_kernel, total = {}, 0.0
    for r in rows[rows.index(hdr) + 1:]:
        if len(r) <= iv or not r[im].startswith("dram__bytes_"):
            continue
        b = float(r[iv].replace(",", "")) * scale.get(r[iu], 1.0)
        name = re.sub(r"^void\s+|ps::|\(.*$", "", r[ik]).split("<")[0]
        per_kernel[name] = per_kernel.get(name, 0.0) + b
        total += b
    if not total:
        return None, None, f"{os.path.basename(path)}: no dram__bytes rows"
    return total / ntrials, {k: round(v / ntrials / 1e6, 1) for k, v in per_kernel.items()}, \
        f"profiles/{os.path.basename(path)} (ncu, one {ntrials}-trial launch of each chain kernel, divided by {ntrials})"


def load_ref():
    path = os.path.join(REPO, "oracle", "_ref", "libpeasoup_ref.so")
    if not os.path.exists(path):
        return None
    r = C.CDLL(path)
    r.ref_last_error.restype = C.c_char_p
    r.ref_worker_search.restype = C.c_long
    return r


def dist_env():
    return int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))


def workload_config(args, wl, ndm, world):
    lg = wl.size.bit_length() - 1
    acc_hi = int(abs(wl.accs[0]))
    base = {"fft_size": wl.size, "nsamps": wl.nsamps, "tsamp": TSAMP, "n_accel": int(len(wl.accs)), "nfolds": NFOLDS,
            "parallelism": f"dm-shard x{world}",
            "l2": "per-step input (ndm x %d MiB u8) and per-trial working set (>= %d MiB) exceed the 126 MB L2"
                  % (wl.size >> 20, 12 * wl.size >> 20)}
    if args.config == 4:
        base.update({"workload": "BASELINE configs[3]: fixed job of %d DM trials (2^23 samples) x 101 accel (-50..50 m/s^2), 4 folds, "
                                 "rednoise + default zaplist, DM-sharded over the ranks; one gather at the end, rank-0 DM + harmonic "
                                 "distillation of the merged list inside the timed region" % args.ndm_total,
                     "ndm_total": args.ndm_total, "trials_per_step": args.ndm_total * int(len(wl.accs))})
    elif args.config == 5:
        base.update({"workload": "BASELINE configs[4]: 2^24-sample 8-bit DM trials x 201 accel (-500..500 m/s^2), 4 folds, default "
                                 "zaplist; the full job is 4096 DM trials on 8 GPUs = 512 per GPU; this run searches %d DM trials per "
                                 "GPU per step on %d GPU(s) (%s)" % (ndm, world, "the full job" if ndm * world == 4096 else
                                                                     "a sub-sample of the 4096"),
                     "ndm_per_step_per_gpu": ndm, "trials_per_step_per_gpu": ndm * int(len(wl.accs))})
    else:
        base.update({"workload": "BASELINE configs[1]: 2^%d-sample 8-bit DM trials x %d accel (-%d..%d m/s^2), 4 folds (16 harmonics), "
                                 "min S/N 9, default zaplist; 1024-DM job processed in steps of ndm_per_step_per_gpu DM trials"
                                 % (lg, len(wl.accs), acc_hi, acc_hi),
                     "ndm_per_step_per_gpu": ndm, "trials_per_step_per_gpu": ndm * int(len(wl.accs))})
    return base


def base_line(args, wl, world, ndm, scaling):
    return {"metric": wl.metric, "unit": "trials/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, wl, ndm, world)}


def run_reference(args, wl):
    """--impl reference: the reference's own code, one worker per GPU (every rank on its own device,
    like the reference's -t N), on the same synthetic DM trials as the repo arm."""
    import torch
    world, rank, local = dist_env()
    if args.config == 3:
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": "config 3 is a kernel microbench of this repo's fused stage; "
                              "the reference's materialised equivalent is timed by tools/harmonic_bench.py"}))
        return
    ref = load_ref()
    if ref is None or not torch.cuda.is_available():
        if rank == 0:
            cb = cpu_baseline(8)
            out = base_line(args, wl, 1, 1, "weak")
            out.update({"impl": "reference", "value": cb["value"], "ms_per_step": None, "cpu_baseline": cb,
                        "e2e": {"value": cb["value"], "unit": "trials/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                        "note": "oracle/_ref not available: CPU oracle port timed instead"})
            print(json.dumps(out))
        return
    import torch.distributed as dist
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    # per-rank sample: the repo arm's own step (config 2), or a bounded share of the bigger jobs
    if args.config == 4:
        ndm, note = max(1, args.ref_ndm_total // world), f"bounded sample: {args.ref_ndm_total} of the {args.ndm_total} DM trials, split over the ranks"
        trials = synth_trials(ndm, device, 1000 + rank, wl.size, wl.nsamps, wl.a_inj, centre=ndm // 2 if rank == 0 else -100)
    elif args.config == 5:
        ndm, note = args.ref_ndm, f"bounded sample: {args.ref_ndm} DM trials per GPU per step"
        trials = synth_trials(ndm, device, rank, wl.size, wl.nsamps, wl.a_inj)
    else:
        ndm, note = args.ndm, "the repo arm's step: same DM trials (same generator and seed per rank), same acceleration list"
        trials = synth_trials(ndm, device, rank, wl.size, wl.nsamps, wl.a_inj)
    trials = trials.cpu().numpy()

    class P(C.Structure):
        _fields_ = [("size", C.c_uint), ("nsamps", C.c_uint), ("tsamp", C.c_float), ("nharmonics", C.c_int),
                    ("min_snr", C.c_float), ("min_freq", C.c_float), ("max_freq", C.c_float), ("freq_tol", C.c_float),
                    ("max_harm", C.c_int), ("zapfile", C.c_char_p)]
    zp = "/tmp/peasoup_b200_bench_zaplist_%d.txt" % rank
    with open(zp, "w") as f:
        for a, b in zip(ZAP_F, ZAP_W):
            f.write(f"{a} {b}\n")
    p = P(wl.size, wl.nsamps, TSAMP, NFOLDS, 9.0, 0.1, 1100.0, 1e-4, 16, zp.encode())
    cap = 1 << 20
    o = [np.zeros(cap, t) for t in ("f4", "i4", "f4", "i4", "f4", "f4", "i4")]
    dms = np.linspace(0, 100, ndm).astype(np.float32)
    vp = C.c_void_p
    ref.ref_worker_search.argtypes = [C.POINTER(P), vp, vp, C.c_int, vp, C.c_int, C.c_long] + [vp] * 7 + [C.POINTER(C.c_double)]
    secs = C.c_double()
    accs = np.ascontiguousarray(wl.accs)
    tot = 0.0
    for it in range(args.warmup + args.steps):
        if world > 1:
            dist.barrier()
        n = ref.ref_worker_search(C.byref(p), trials.ctypes.data, dms.ctypes.data, ndm, accs.ctypes.data, len(accs), cap,
                                  *[a.ctypes.data for a in o], C.byref(secs))
        if n < 0:
            raise RuntimeError(ref.ref_last_error().decode())
        if it >= args.warmup:
            tot += secs.value
    if world > 1:  # slowest worker decides, as after the reference's pthread_join
        t = torch.tensor([tot], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        tot = float(t.item())
        dist.destroy_process_group()
    if rank != 0:
        return
    val = world * ndm * len(accs) * args.steps / tot
    sample = f"{ndm} DM trials x {len(accs)} accelerations per step per GPU through the reference's unmodified CUDA build " \
             f"(kernels.cu + cuFFT, Worker::start replay), one host thread and one B200 per worker, {world} worker(s); " \
             f"{note}; the reference has no CPU search path"
    out = base_line(args, wl, world, ndm, "weak")
    out.update({"impl": "reference", "value": val, "ms_per_step": 1e3 * tot / args.steps,
                "cpu_baseline": {"value": val, "unit": "trials/s", "cores": world, "kind": "reference", "sample": sample},
                "e2e": {"value": val, "unit": "trials/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "host_cores": os.cpu_count()})
    print(json.dumps(out))


def measured_peak():
    path = os.path.join(REPO, "MEASURED_PEAKS.json")
    peaks = json.load(open(path)) if os.path.exists(path) else {}
    return float(peaks.get("hbm_gbs", 6650.0)), ("measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650")


def run_config3(args):
    """BASELINE configs[2]: harmonic sums (2/4/8/16 harmonics) + threshold + compaction of a batch of
    2^22-bin normalised spectra through ps_search_harmonic_peaks; algorithmic bytes 4 B per bin."""
    import torch
    import peasoup_b200 as pb
    from peasoup_b200.host import make_search_params
    world, rank, local = dist_env()
    if rank != 0:
        return
    torch.cuda.set_device(local)
    lib = pb.load()
    size = SIZE
    nb = size // 2 + 1
    stride = (nb + 63) & ~63   # padded, 16-byte aligned spectra: the TMA-staged kernel
    s = pb.Search(make_search_params(size, size, TSAMP, NFOLDS, 9.0))
    g = torch.Generator(device="cuda").manual_seed(20140213)
    st = torch.cuda.Stream()

    def spectra(n):
        x = torch.randn((n, stride), generator=g, device="cuda", dtype=torch.float32)
        x[:, ::32] += 1.0                                       # src/harmonic_sum_test.cpp's pattern on top of the noise
        for b in range(n):
            x[b, (5000 + 11 * b) * torch.arange(1, 17, device="cuda")] += 8.0   # a 16-harmonic comb
        return x

    def time_batch(x, n, reps, src_host=None):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(st)
        for _ in range(reps):
            if src_host is not None:
                with torch.cuda.stream(st):
                    x.copy_(src_host, non_blocking=True)
            s.reset_peaks(stream=st)
            s.harmonic_peaks(x, stride, n, stream=st)
            if src_host is not None:
                s.fetch_peaks(stream=st)
        e1.record(st)
        st.synchronize()
        return e0.elapsed_time(e1) * 1e-3 / reps

    sweep = {}
    for n in (1, 4, 16, 64, 256):
        x = spectra(n)
        time_batch(x, n, 2)
        sweep[str(n)] = round(4.0 * nb * n / time_batch(x, n, 3) / 1e9, 1)   # 256 spectra = 4.3 GB: inputs exceed the L2
        del x
    n = args.batch
    x = spectra(n)
    for _ in range(args.warmup):
        time_batch(x, n, 1)
    sampler = ClockSampler(local)
    sampler.start()
    l0 = lib.ps_launch_count()
    t = time_batch(x, n, args.steps)
    launches = lib.ps_launch_count() - l0
    npk = len(s.fetch_peaks())
    host = torch.empty((n, stride), dtype=torch.float32).pin_memory()
    host.copy_(x)
    time_batch(x, n, 1, host)
    t_e2e = time_batch(x, n, args.steps, host)
    clocks = sampler.stop()
    peak, peak_src = measured_peak()
    gbs = 4.0 * nb * n / t / 1e9
    out = {"metric": "harmonic-sum + peak-find HBM GB/s (2^22-bin spectra, 16 harmonics, 4 B/bin)", "value": gbs, "unit": "GB/s",
           "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": "BASELINE configs[2]: batch of %d normalised spectra of 2^22+1 bins (N(0,1) + 1 every 32 bins + a "
                                  "16-harmonic comb), 4 folds = 16 harmonics, threshold 9, fused sums + threshold + compaction "
                                  "(sums never materialised)" % n,
                      "bins": nb, "batch": n, "nfolds": NFOLDS, "us_per_spectrum": 1e6 * t / n,
                      "sweep": {"GB/s by batch size": sweep},
                      "l2": "batch of %d x 16.8 MB exceeds the 126 MB L2" % n},
           "e2e": {"value": 4.0 * nb * n / t_e2e / 1e9, "unit": "GB/s", "h2d_bytes_per_step": int(n * stride * 4),
                   "d2h_bytes_per_step": int(4 + 16 * npk)},
           "gpu_launches": int(launches), "clocks": clocks, "peaks_per_step": npk,
           "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak, "traffic": None,
                        "peak_source": peak_src, "kernel": "harmonic_peaks_rows_kernel<4>",
                        "algorithmic_bytes_per_launch": int(4 * nb * n), "ms_per_launch": 1e3 * t,
                        "note": "the stage pulls 8.5 words per bin out of the L2 (the operand ranges of the 15 fractions): "
                                "L2 -> SM traffic, not HBM, bounds it"}}
    s.close()
    print(json.dumps(out))


def synth_shard(start, count, ndm_total, device, wl):
    """DM trials [start, start+count) of the fixed ndm_total-trial job: noise seeded per GLOBAL DM index,
    the pulsar brightest at trial ndm_total/2 -- the job's input does not depend on how it is sharded."""
    import torch
    out = torch.empty((count, wl.nsamps), dtype=torch.uint8, device=device)
    for d in range(count):
        out[d] = synth_trials(1, device, 7919 * (start + d), wl.size, wl.nsamps, wl.a_inj, centre=ndm_total // 2 - (start + d))[0]
    return out


def inprocess_workers_check(pb, make_search_params):
    """The reference's own multi-GPU model -- one host thread per GPU inside ONE process
    (pipeline_multi.cu:342-351) -- on real hardware: ps_pipeline_search with nworkers = 2 on devices 0 and 1
    must return the records of nworkers = 1.  Small job, outside every timed region."""
    sys.path.insert(0, os.path.join(REPO, "tests"))
    from synth import make_trials
    size, tsamp = 1 << 18, 64e-6
    trials = make_trials(12, size + 64, size, tsamp=tsamp, a_inj=20.0, amp=8.0)
    dms = np.linspace(0, 30, 12).astype(np.float32)
    accs = np.array([0.0, 20.0, -20.0], np.float32)
    prm = pb.PipelineParams(make_search_params(size, size + 64, tsamp), 1e-4, 16, 0.0, 0.0, 1.1, 64.0, 1382.0, -0.390625)
    one, d1 = pb.pipeline_search(trials, dms, prm, accs=accs, device0=0, nworkers=1)
    two, d2 = pb.pipeline_search(trials, dms, prm, accs=accs, device0=0, nworkers=2)
    return bool(d1 == d2 == 36 and len(one) == len(two) > 0 and all(np.array_equal(one[k], two[k]) for k in one.dtype.names))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--config", type=int, default=2, choices=[2, 3, 4, 5], help="BASELINE configuration (1-based)")
    ap.add_argument("--ndm", type=int, default=0, help="DM trials per step per GPU (config 2: 16, config 5: 512)")
    ap.add_argument("--ndm-total", type=int, default=1024, help="config 4: DM trials of the fixed job")
    ap.add_argument("--ref-ndm", type=int, default=2, help="config 5: DM trials per GPU per step of the reference arm")
    ap.add_argument("--ref-ndm-total", type=int, default=64, help="config 4: DM trials of the reference arm's bounded sample")
    ap.add_argument("--batch", type=int, default=64, help="config 3: spectra per launch")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    wl = Workload(args.config)
    if not args.ndm:
        args.ndm = 512 if args.config == 5 else 16
    if args.impl == "reference":
        run_reference(args, wl)
        return
    if args.config == 3:
        run_config3(args)
        return

    import torch
    import torch.distributed as dist

    import peasoup_b200 as pb
    from peasoup_b200.host import distill_merged, make_search_params
    from peasoup_b200.shard import gather_candidates, shard_range

    world, rank, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    lib = pb.load()

    strong = args.config == 4
    if strong:
        offset, ndm = shard_range(args.ndm_total, rank, world)
        d_trials = synth_shard(offset, ndm, args.ndm_total, device, wl)
        dms = np.linspace(0, 100, args.ndm_total).astype(np.float32)[offset:offset + ndm]
    else:
        ndm = args.ndm
        offset = rank * ndm
        d_trials = synth_trials(ndm, device, rank, wl.size, wl.nsamps, wl.a_inj)
        dms = np.linspace(0, 100, ndm).astype(np.float32)
    h_trials = torch.empty((ndm, wl.nsamps), dtype=torch.uint8).pin_memory()
    h_trials.copy_(d_trials)
    torch.cuda.synchronize()
    sp = make_search_params(wl.size, wl.nsamps, TSAMP, NFOLDS, 9.0, max_peaks=(1 << 22) if args.config == 5 else 0)
    prm = pb.PipelineParams(sp, 1e-4, 16, 0.0, 0.0, 1.1, 64.0, 1382.0, -0.390625)
    pipe = pb.Pipeline(prm, zap=(ZAP_F, ZAP_W), device0=local, nworkers=1, max_out=1 << 21)

    def step(src):
        recs, done = pipe.run(src, dms, accs=wl.accs, dm_idx_offset=offset)
        merged = gather_candidates(recs, device=device) if world > 1 else recs
        if strong and rank == 0:  # the post-join distillations on the merged list (pipeline_multi.cu:353-365)
            merged = distill_merged(merged, 1e-4, 16)
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
            t = torch.tensor([ms, float(done)], device=device, dtype=torch.float64)
            mx = t.clone()
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            ms, done = float(mx[0].item()), int(t[1].item())
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
    s = pb.Search(sp)
    s.set_zaplist(ZAP_F, ZAP_W)
    s.whiten(d_trials[ndm // 2], stream=st)
    s.reset_peaks(stream=st)
    s.accel_batch(wl.accs[:8], stream=st)
    st.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    s.accel_batch(wl.accs, stream=st)
    e1.record(st)
    st.synchronize()
    chain_ms = e0.elapsed_time(e1) / len(wl.accs)
    s.close()
    peak, peak_src = measured_peak()
    achieved = wl.algo_bytes / (chain_ms * 1e-3) / 1e9
    workers_ok = None
    if world > 1:
        barrier()
        if rank == 0 and torch.cuda.device_count() >= 2:
            workers_ok = inprocess_workers_check(pb, make_search_params)
        barrier()

    if rank == 0:
        traffic, per_kernel, traffic_src = ncu_traffic() if args.config in (2, 4) else (None, None, "no ncu capture at this transform size")
        out = base_line(args, wl, world, ndm, "strong" if strong else "weak")
        out.update({
            "value": done / (ms * 1e-3), "ms_per_step": ms / args.steps,
            "e2e": {"value": done_e2e / (ms_e2e * 1e-3), "unit": "trials/s",
                    "h2d_bytes_per_step": (h1.value - h0.value) // args.steps, "d2h_bytes_per_step": (d1.value - d0.value) // args.steps,
                    "bytes_scope": "rank 0"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "frac_uses": "algorithmic bytes (16 N per acceleration trial, SURVEY 8d) / event-timed chain",
                         "traffic": traffic, "traffic_per_kernel_MB": per_kernel, "traffic_source": traffic_src,
                         "achieved_on_traffic": (traffic / (chain_ms * 1e-3) / 1e9) if traffic else None, "peak_source": peak_src,
                         "kernel": "acceleration-trial chain: fft_pass1(resample gather) + fft_pass2 + spectrum + harmonic_peaks",
                         "algorithmic_bytes_per_launch": wl.algo_bytes, "ms_per_launch": chain_ms},
            "candidates_last_step": int(len(merged)) if merged is not None else None,
            "host_profile_last_step": {"submit_s": prof[0], "wait_device_s": prof[1], "host_distill_s": prof[2],
                                       "raw_peaks": int(prof[3])},
        })
        if workers_ok is not None:
            out["inprocess_workers_ok"] = workers_ok
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline()
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
