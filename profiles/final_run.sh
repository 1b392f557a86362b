#!/bin/bash
# Round-end measurement pass on ONE B200 (run under gpurun): tests, smoke, both bench arms, the other
# BASELINE configs, then the ncu launch list and one full-set capture of ONE frame (each only after
# its plain run exited 0).  Output: gpurun_out/final/ (copied into profiles/ by hand, named per round).
O=gpurun_out/final; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4 > $O/tests.log; cat $O/tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 500 python bench.py --stage-report $O/stages_6M.json > $O/bench_6M.json 2> $O/bench.err || { tail -5 $O/bench.err; exit 1; }
timeout 400 python bench.py --impl reference --steps 4 --warmup 1 > $O/bench_ref.json 2>> $O/bench.err
for w in ply-point_cloud_9-720p-div0.1 ply-point_cloud_9-720p-div1.0; do
  timeout 300 python bench.py --workload $w --steps 100 --warmup 5 >> $O/bench_cfg0.jsonl 2>> $O/bench.err; done
for w in ply-orbit-720p-div0.1 ply-orbit-720p-div1.0; do
  timeout 400 python bench.py --workload $w --steps 32 --warmup 3 >> $O/bench_cfg1.jsonl 2>> $O/bench.err; done
timeout 300 python bench.py --workload synthetic-1M-1080p > $O/bench_cfg2_1M.json 2>> $O/bench.err
timeout 400 python bench.py --workload synthetic-6M-4k --steps 100 > $O/bench_cfg3_4k_views.json 2>> $O/bench.err
timeout 300 python bench.py --mode bands --workload synthetic-6M-4k --steps 120 > $O/bench_cfg3_4k_bands_n1.json 2>> $O/bench.err
timeout 300 python bench.py --mode multiview --workload synthetic-3M-1080p --steps 5 --warmup 3 > $O/bench_cfg4_multiview_n1.json 2>> $O/bench.err
timeout 200 python profiles/tile_profile.py synthetic-6M-1080p 3 > $O/tile_profile_6M.txt 2>&1
# the captures run ONE frame at a time (one lane, fixed slicing): profiles/frames.py
# launch list of the frame as every context runs it (single-sweep sort passes: 33 launches) and, for
# comparison, with the three-kernel passes (GSPLAT_SORT_SWEEP=0: 57 launches)
python profiles/frames.py > $O/frames.log 2>&1 || { tail -5 $O/frames.log; exit 1; }
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_sweep.csv python profiles/frames.py > $O/ncu1.log 2>&1
GSPLAT_SORT_SWEEP=0 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches.csv python profiles/frames.py > $O/ncu1b.log 2>&1
# one frame (33 launches; 2 upload launches + frame 0 are skipped) + the next frame's K1, every section of the full set except the two that
# store per-instruction / per-sample data for each of the 58 launches (SourceCounters, PmSampling):
# gpurun brings back at most 64 MiB.  K5 alone is then captured with the whole --set full and its source.
SECTIONS="--section ComputeWorkloadAnalysis --section InstructionStats --section LaunchStats --section MemoryWorkloadAnalysis --section MemoryWorkloadAnalysis_Tables --section Occupancy --section SchedulerStats --section SpeedOfLight --section WarpStateStats --section WorkloadDistribution"
timeout 900 ncu $SECTIONS --metrics smsp__thread_inst_executed_per_inst_executed.ratio,sm__inst_issued.avg.pct_of_peak_sustained_active --clock-control none --launch-skip 35 --launch-count 34 -f -o $O/prof_frame python profiles/frames.py synthetic-6M-1080p 3 > $O/ncu2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:blend_kernel --launch-skip 2 --launch-count 2 -f -o $O/prof_blend python profiles/frames.py synthetic-6M-1080p 3 > $O/ncu3.log 2>&1
# never lose the pass to the size limit: keep the raw page instead of the report if it is too large
if [ $(du -sm gpurun_out | cut -f1) -gt 58 ]; then
  ncu -i $O/prof_frame.ncu-rep --page raw --csv > $O/prof_frame_raw.csv 2>/dev/null && rm -f $O/prof_frame.ncu-rep
fi
ls -la $O
