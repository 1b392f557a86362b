#!/bin/bash
# The bench half of final_run.sh (no ncu captures): tests, smoke, both bench arms and every other
# BASELINE configuration.  Output: gpurun_out/final_q/.
O=gpurun_out/final_q; mkdir -p $O
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
ls -la $O
