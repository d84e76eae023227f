#!/bin/bash
# per-kernel share of the acceleration-trial chain by leaving kernels out (PS_SKIP bit mask);
# needs a profiling build of the library: make clean && make EXTRA=-DPS_PROFILING
for m in 0 1 2 4 8 3 12; do PS_SKIP=$m PS_HP_VARIANT="skip=$m" python tools/time_chain.py; done
