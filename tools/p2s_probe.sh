#!/bin/bash
# direct timing of the pass-2 + spectrum portion (pass 1 and harmonic kernels left out)
export PS_SKIP=9
PS_P2S=0 PS_HP_VARIANT="separate kernels" python tools/time_chain.py
for d in 0 1 3 4 6; do PS_DBG=$d PS_HP_VARIANT="fused dbg=$d" python tools/time_chain.py; done
