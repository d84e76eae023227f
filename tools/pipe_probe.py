import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import peasoup_b200 as pb
from peasoup_b200.host import make_search_params
import bench
torch.cuda.set_device(0)
dev = torch.device("cuda:0")
ndm = int(sys.argv[1]); snr = float(sys.argv[2])
d = bench.synth_trials(ndm, dev, 0)
dms = np.linspace(0, 100, ndm).astype(np.float32)
sp = make_search_params(bench.SIZE, bench.NSAMPS, bench.TSAMP, bench.NFOLDS, snr)
prm = pb.PipelineParams(sp, 1e-4, 16, 0.0, 0.0, 1.1, 64.0, 1382.0, -0.390625)
pipe = pb.Pipeline(prm, zap=(bench.ZAP_F, bench.ZAP_W), device0=0, nworkers=1)
pipe.run(d, dms, accs=bench.ACCS)
torch.cuda.synchronize()
t0 = time.perf_counter()
recs, done = pipe.run(d, dms, accs=bench.ACCS)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
import ctypes as C
prof = (C.c_double * 4)(); pb.load().ps_pipeline_profile(prof)
print(f"ndm {ndm} snr {snr}: {dt*1e3/ndm:.3f} ms per DM, {done/dt:.0f} trials/s, {dt*1e6/done:.2f} us/trial, host: submit {prof[0]*1e3/ndm:.3f} wait {prof[1]*1e3/ndm:.3f} distill {prof[2]*1e3/ndm:.3f} ms/DM raw {prof[3]/ndm:.0f}/DM cands {len(recs)}")
