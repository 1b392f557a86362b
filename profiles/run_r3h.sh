O=gpurun_out/r3h; mkdir -p $O
python profiles/frames.py synthetic-6M-1080p 3 > $O/frames.log 2>&1 || { tail -5 $O/frames.log; exit 1; }
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches.csv python profiles/frames.py > $O/ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --launch-skip 58 --launch-count 58 -f -o $O/prof_frame python profiles/frames.py synthetic-6M-1080p 3 > $O/ncu2.log 2>&1
tail -2 $O/ncu2.log; ls -la $O
