O=gpurun_out/r3s; mkdir -p $O
L=gaussian_splat_ipu_b200
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "synthetic or benched or frames_in_flight or edge or golden or slices" 2>&1 | tail -3
bash profiles/ab_k5.sh $O/ab.txt lb=$L/libgsplat_b200.so nolb=$L/libgsplat_b200_nolb.so lb2=$L/libgsplat_b200.so
