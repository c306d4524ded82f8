#!/bin/bash
# persistent prefetching strided pass (FHE_B200_NTT_PF): parity on the NTT tests, then timing against the default
FHE_B200_NTT_PF=2 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "ntt or NTT or chain or switch" 2>&1 | tail -3
for pf in 0 1; do for b in 64 1024; do FHE_B200_NTT_PF=$pf python tools/prof_ntt.py 16 30 $b 3 0 | sed "s/^/pf=$pf /"; done; done
for pf in 0 1; do FHE_B200_NTT_PF=$pf python tools/prof_ntt.py 15 16 512 3 0 | sed "s/^/pf=$pf /"; FHE_B200_NTT_PF=$pf python tools/prof_ntt.py 14 8 2048 3 0 | sed "s/^/pf=$pf /"; done
