#!/bin/bash
# usage: profiles/sweep.sh  -- bench matrix over env knobs (experiments only)
for fr in "0.05" "0.065" "0.08" "0.1" "0.13"; do
GSPLAT_SLICE_FRACTIONS=$fr profiles/quick.sh fr=$fr
done
