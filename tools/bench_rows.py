"""Timings of the stages either side of the hot path on synthetic data.
Prints markdown.  All calls go through the C ABI with HOST buffers (these entry points are
synchronous and copy in/out themselves), so the figures are end-to-end."""
import ctypes as C, os, sys, time, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import peasoup_b200 as pb
from synth import write_sigproc
lib = pb.load()
torch.cuda.set_device(0)
torch.zeros(1, device="cuda")


def best(fn, reps=3):
    t = 1e9
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); t = min(t, time.perf_counter() - t0)
    return t


print("## f1 dedispersion (ps_dedisperse: host filterbank in, host 8-bit trials out)\n")
print("| nchans | nbits | nsamps | ndm | seconds | G channel-samples/s (ndm x nout x nchans / s) |")
print("|---|---|---|---|---|---|")
rng = np.random.default_rng(1)
for nchans, nbits, nsamps, ndm in ((64, 2, 1 << 21, 128), (1024, 8, 1 << 20, 128), (1024, 8, 1 << 21, 256)):
    raw = rng.integers(0, 256, nsamps * nchans * nbits // 8, dtype=np.uint8)
    dms = np.linspace(0, 100, ndm).astype(np.float32)
    tsamp, fch1, foff = 64e-6, 1510.0, -300.0 / nchans
    md = lib.ps_dedisp_max_delay(dms.ctypes.data, ndm, nchans, tsamp, fch1, foff)
    nout = nsamps - md
    out = np.zeros((ndm, nout), np.uint8)
    def run():
        pb.check(lib.ps_dedisperse(raw.ctypes.data, nsamps, nchans, nbits, tsamp, fch1, foff, dms.ctypes.data, ndm, None,
                                   out.ctypes.data, nout))
    run()
    t = best(run)
    print(f"| {nchans} | {nbits} | 2^{int(np.log2(nsamps))} | {ndm} | {t:.3f} | {ndm * nout * nchans / t / 1e9:.0f} |")

print("\n## f3 folding + optimisation (device series in, host result out)\n")
n = 1 << 23
x = torch.randn(n, device="cuda")
fold = torch.zeros(16 * 64, device="cuda")
h = C.c_void_p(); pb.check(lib.ps_fold_optimiser_create(64, 16, C.byref(h)))
res = pb.FoldResult(); of = np.zeros(1024, np.float32); op = np.zeros(64, np.float32)
def fold_run():
    pb.check(lib.ps_fold_timeseries(x.data_ptr(), fold.data_ptr(), n, 16, 0.0123, 64e-6, 64, None))
def opt_run():
    pb.check(lib.ps_fold_optimise(h, fold.data_ptr(), 0.0123, 536.87, of.ctypes.data, op.ctypes.data, C.byref(res), None))
fold_run(); opt_run()
tf, to = best(fold_run, 5), best(opt_run, 5)
print(f"fold 2^23 samples into 16 x 64: {tf*1e6:.0f} us ({4.0*n/tf/1e9:.0f} GB/s of series read); optimise (63 widths x 64 drifts x 64 bins): {to*1e6:.0f} us")
lib.ps_fold_optimiser_destroy(h)

print("\n## f4 coincidencer (ps_coincidence_files: 13 beams of 2^20 samples, 64 channels, 8 bit)\n")
d = tempfile.mkdtemp()
files = []
for b in range(13):
    data = np.clip(np.rint(100 + rng.normal(0, 6, ((1 << 20) + 321, 64))), 0, 255).astype(np.uint8)
    p = os.path.join(d, f"b{b}.fil"); write_sigproc(p, data, 64e-6, 1510.0, -1.0); files.append(p)
arr = (C.c_char_p * len(files))(*[f.encode() for f in files])
prm = pb.CoincidenceParams(4.0, 4, None, None)
def cc():
    assert lib.ps_coincidence_files(arr, len(files), C.byref(prm), None, 0, None, 0) > 0, lib.ps_last_error()
cc()
print(f"{best(cc, 2):.2f} s for 13 beams (file read + DM-0 dedispersion + chirp-z transforms of 1,048,897 points + matching)")
