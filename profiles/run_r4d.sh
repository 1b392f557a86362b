O=gpurun_out/r4d; mkdir -p $O
python profiles/frames.py synthetic-6M-1080p 3 > $O/frames.log 2>&1 || { tail -5 $O/frames.log; exit 1; }
SECTIONS="--section ComputeWorkloadAnalysis --section InstructionStats --section LaunchStats --section MemoryWorkloadAnalysis --section MemoryWorkloadAnalysis_Tables --section Occupancy --section SchedulerStats --section SpeedOfLight --section WarpStateStats --section WorkloadDistribution"
timeout 600 ncu $SECTIONS --metrics smsp__thread_inst_executed_per_inst_executed.ratio,sm__inst_issued.avg.pct_of_peak_sustained_active --clock-control none --launch-skip 35 --launch-count 34 -f -o $O/prof_frame python profiles/frames.py synthetic-6M-1080p 3 > $O/ncu2.log 2>&1
tail -2 $O/ncu2.log; ls -la $O
