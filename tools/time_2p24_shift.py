"""Chain time per acceleration trial at 2^24 samples against the acceleration range of the list (profiles/r02_summary.md
r2l): the band of the series the trials of a launch read grows with the range; PS_BAND_MB cuts pass 1 into groups."""
import sys, os
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np, torch
import peasoup_b200 as pb
from peasoup_b200.host import make_search_params
from synth import make_trials
torch.cuda.set_device(0)
for lg, amax, nacc, nfolds in ((24, 5.0, 67, 4), (24, 50.0, 67, 4), (24, 150.0, 67, 4), (24, 500.0, 67, 4), (24, 500.0, 201, 4)):
    size, tsamp = 1 << lg, 64e-6
    trials = make_trials(1, size + 4096, size, tsamp=tsamp, a_inj=amax * 0.6, amp=3.0, seed=lg)
    accs = np.linspace(-amax, amax, nacc).astype(np.float32)
    s = pb.Search(make_search_params(size, size + 4096, tsamp, nfolds, 9.0))
    st = torch.cuda.Stream()
    d = torch.from_numpy(trials[0]).cuda()
    s.whiten(d, stream=st); s.reset_peaks(stream=st); s.accel_batch(accs[:40], stream=st); st.synchronize()
    best = 1e9
    for rep in range(3):
        s.reset_peaks(stream=st)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st); s.accel_batch(accs, stream=st); e1.record(st); st.synchronize()
        best = min(best, e0.elapsed_time(e1) / len(accs))
    print(f"2^{lg} +-{amax:g} x{nacc}: {best*1e3:.1f} us/trial")
    s.close()
