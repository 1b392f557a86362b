O=gpurun_out/r3d; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -6 > $O/tests.log; cat $O/tests.log
python profiles/frames.py synthetic-6M-1080p 3 > $O/frames.log 2>&1 || { tail -5 $O/frames.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:blend_kernel --launch-skip 2 --launch-count 2 -f -o $O/prof_blend python profiles/frames.py synthetic-6M-1080p 3 > $O/ncu3.log 2>&1
tail -2 $O/ncu3.log; ls -la $O
