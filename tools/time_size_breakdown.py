"""Chain time per acceleration trial at 2^LG samples with kernels left out (PS_SKIP bit mask: 1 pass 1,
2 pass 2, 4 spectrum, 8 harmonic): python tools/time_size_breakdown.py LG  (run once per mask)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import peasoup_b200 as pb
from peasoup_b200.host import make_search_params
from synth import make_trials
lg = int(sys.argv[1])
size, tsamp, amax, nacc = 1 << lg, 64e-6, 500.0, 64
trials = make_trials(1, size + 4096, size, tsamp=tsamp, a_inj=amax * 0.6, amp=3.0, seed=lg)
accs = np.linspace(-amax, amax, nacc).astype(np.float32)
s = pb.Search(make_search_params(size, size + 4096, tsamp, 4, 9.0))
st = torch.cuda.Stream()
d = torch.from_numpy(trials[0]).cuda()
s.whiten(d, stream=st); s.reset_peaks(stream=st); s.accel_batch(accs[:32], stream=st); st.synchronize()
best = 1e9
for rep in range(3):
    s.reset_peaks(stream=st)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st); s.accel_batch(accs, stream=st); e1.record(st); st.synchronize()
    best = min(best, e0.elapsed_time(e1) / len(accs))
print(f"2^{lg}: {best*1e3:.1f} us/trial  (PS_SKIP={os.environ.get('PS_SKIP','0')})")
