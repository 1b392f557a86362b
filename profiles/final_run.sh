#!/bin/bash
# Round-end measurement pass on one B200 (run under gpurun): tests, smoke, both bench arms, then the
# ncu launch list and one full-set capture of ONE frame (each only after its plain run exited 0).
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/t.log; cat gpurun_out/t.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 400 python bench.py --stage-report gpurun_out/stage_final.json > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err || exit 1
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench_final.err
timeout 300 python bench.py --mode multiview --workload synthetic-3M-1080p --steps 5 --warmup 3 > gpurun_out/bench_multiview.json 2>> gpurun_out/bench_final.err
timeout 300 python bench.py --workload synthetic-1M-1080p --no-cpu-baseline > gpurun_out/bench_1M.json 2>> gpurun_out/bench_final.err
timeout 300 python bench.py --workload synthetic-6M-4k --no-cpu-baseline --steps 100 > gpurun_out/bench_4k.json 2>> gpurun_out/bench_final.err
# the captures run ONE frame at a time (one lane, fixed slicing) so that a frame is 57 consecutive launches
export GSPLAT_SLICE_FRACTIONS=0.08 GSPLAT_CONCURRENT_FRAMES=1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 1768 --launch-count 130 --csv \
  --log-file gpurun_out/launches_final.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --launch-skip 1768 --launch-count 58 -f -o gpurun_out/prof_frame_final \
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out/*.ncu-rep
