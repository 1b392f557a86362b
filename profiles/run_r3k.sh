O=gpurun_out/r3k; mkdir -p $O
L=gaussian_splat_ipu_b200/libgsplat_b200.so
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "sort or synthetic or benched or scale" 2>&1 | tail -3
bash profiles/ab_k5.sh $O/ab.txt sweep=$L three=$L:GSPLAT_SORT_SWEEP=0 sweepl1=$L:GSPLAT_CONCURRENT_FRAMES=1 threel1=$L:"GSPLAT_SORT_SWEEP=0 GSPLAT_CONCURRENT_FRAMES=1"
