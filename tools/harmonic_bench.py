"""BASELINE.json configs[2]: harmonic-sum + peak-find microbench.  Batch of 2^22-bin spectra
(size 2^23), 4 folds = 16 harmonics, 9 sigma threshold; sweep of the batch size.

fused        ps_search_harmonic_peaks: sums never materialised; algorithmic bytes 4*B per spectrum
materialised ps_harmonic_sum (writes the four HarmonicSums buffers the reference API exposes) +
             ps_find_peaks x 5: algorithmic bytes 4*B*(1+4) written/read once, + 5 threshold reads
Prints a markdown table (GB/s = algorithmic bytes / CUDA-event time)."""
import ctypes as C
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import peasoup_b200 as pb
from peasoup_b200.host import make_search_params

SIZE, TSAMP, NF = 1 << 23, 64e-6, 4
B = SIZE // 2 + 1
torch.cuda.set_device(0)
lib = pb.load()
peak = json.load(open(os.path.join(pb.REPO, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6546.9) if os.path.exists(
    os.path.join(pb.REPO, "MEASURED_PEAKS.json")) else 6546.9
s = pb.Search(make_search_params(SIZE, SIZE, TSAMP, NF, 9.0))
g = torch.Generator(device="cuda").manual_seed(20140213)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


SB = (B + 63) & ~63   # padded, 16-byte aligned spectra: the TMA-staged kernel; the plain stride B: the gather kernel


def spectra(n):
    x = torch.randn((n, SB), generator=g, device="cuda", dtype=torch.float32)
    x[:, ::32] += 1.0
    for b in range(n):
        x[b, (5000 + 11 * b) * torch.arange(1, 17, device="cuda")] += 8.0
    return x


def timed(fn, reps=5):
    best = 1e9
    for _ in range(reps):
        flush.fill_(1)  # > L2
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e-3)
    return best


print("| batch | staged us/spectrum | staged GB/s (4B per bin) | frac of %.0f GB/s | gather kernel us/spectrum | peaks |" % peak)
print("|---|---|---|---|---|---|")
for n in (1, 2, 4, 8, 16, 32, 64, 128, 256):
    x = spectra(n)
    s.reset_peaks(); s.harmonic_peaks(x, SB, n); npk = len(s.fetch_peaks())
    def run():
        s.reset_peaks(); s.harmonic_peaks(x, SB, n)
    t = timed(run)
    os.environ["PS_HSMEM_OFF"] = "1"
    y = x[:, :B].contiguous()   # odd stride: gather kernel
    def run2():
        s.reset_peaks(); s.harmonic_peaks(y, B, n)
    run2(); t2 = timed(run2)
    gbs = 4.0 * B * n / t / 1e9
    print(f"| {n} | {t / n * 1e6:.1f} | {gbs:.0f} | {gbs / peak:.3f} | {t2 / n * 1e6:.1f} | {npk} |")
    del x, y

# materialised path, the reference's API shape: one spectrum at a time
x = spectra(1)[:, :B].contiguous()
sums = [torch.empty(B, dtype=torch.float32, device="cuda") for _ in range(NF)]
ptrs = (C.c_void_p * NF)(*[t.data_ptr() for t in sums])
ws = C.c_void_p()
pb.check(lib.ps_peak_workspace_create(100000, C.byref(ws)))
idx = np.zeros(100000, np.int32); snr = np.zeros(100000, np.float32)
from peasoup_b200.host import peak_range
bw = 1.0 / float(np.float32(np.float32(SIZE) * np.float32(TSAMP)))
ranges = [peak_range(B, bw, nh, 0.1, 1100.0) for nh in range(NF + 1)]


def staged():
    pb.check(lib.ps_harmonic_sum(x.data_ptr(), ptrs, B, NF, None))
    for nh in range(NF + 1):
        src = x if nh == 0 else sums[nh - 1]
        start, limit, _ = ranges[nh]
        n = lib.ps_find_peaks(limit, start, src.data_ptr(), C.c_float(9.0), idx.ctypes.data, snr.ctypes.data, 100000, ws, None)
        assert n >= 0


staged()
t = timed(staged)
print()
print(f"materialised (ps_harmonic_sum + 5 x ps_find_peaks, batch 1): {t * 1e6:.1f} us/spectrum, "
      f"{4.0 * B * (1 + NF) / t / 1e9:.0f} GB/s on 4B(1+{NF}) bytes")
