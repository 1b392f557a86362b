#!/bin/bash
# Copies the round-end measurement pass (gpurun_out/final, written by final_run.sh on the box) into
# profiles/ as the per-round evidence files, summarising the ncu captures on the way.  Run HERE (no GPU).
F=${1:-gpurun_out/final}; R=${2:-r2}
cp $F/bench_6M.json profiles/bench_6M_$R.json
cp $F/stages_6M.json profiles/stages_6M_$R.json
cp $F/bench_ref.json profiles/bench_ref_$R.json
cat $F/bench_cfg0.jsonl > profiles/bench_cfg0_ply_$R.jsonl
cat $F/bench_cfg1.jsonl > profiles/bench_cfg1_orbit_$R.jsonl
cp $F/bench_cfg2_1M.json profiles/bench_cfg2_1M_$R.json
cat $F/bench_cfg3_4k_views.json $F/bench_cfg3_4k_bands_n1.json > profiles/bench_cfg3_4k_n1_$R.jsonl
cp $F/bench_cfg4_multiview_n1.json profiles/bench_cfg4_multiview_n1_$R.json
cp $F/tile_profile_6M.txt profiles/tile_profile_6M_$R.txt
python profiles/frame_launches.py $F/launches.csv > profiles/launches_$R.txt
python profiles/summarize_ncu.py launches $F/launches.csv > profiles/launches_all_$R.txt
python profiles/summarize_ncu.py report $F/prof_frame.ncu-rep > profiles/prof_frame_$R.txt
python profiles/traffic_frame.py $F/prof_frame.ncu-rep > profiles/traffic_$R.json
{ python profiles/stalls.py $F/prof_blend.ncu-rep blend_kernel 0 30; python profiles/stalls.py $F/prof_blend.ncu-rep blend_kernel 1 12;
  python profiles/sass_mix.py $F/prof_blend.ncu-rep blend_kernel 0; } > profiles/prof_blend_$R.txt 2>&1
ls -la profiles/*_$R.*
