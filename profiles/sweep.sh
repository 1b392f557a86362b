#!/bin/bash
# usage: profiles/sweep.sh  -- bench matrix over env knobs (experiments only)
for fr in "0.08" "0.12" "0.05,0.2" "0.08,0.3" "0.04,0.12,0.35"; do
GSPLAT_SLICE_FRACTIONS=$fr profiles/quick.sh fr=$fr
done
for sp in 1 4; do GSPLAT_K5_SPLIT=$sp profiles/quick.sh split=$sp; done
