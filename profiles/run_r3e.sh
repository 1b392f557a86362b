O=gpurun_out/r3e; mkdir -p $O
L=gaussian_splat_ipu_b200
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "synthetic or benched or frames_in_flight" 2>&1 | tail -3
bash profiles/ab_k5.sh $O/ab.txt a2=$L/libgsplat_b200.so a1=$L/libgsplat_b200_a1.so a3=$L/libgsplat_b200_a3.so a2s4=$L/libgsplat_b200_a2s4.so a3s6=$L/libgsplat_b200_a3s6.so a3s6b64=$L/libgsplat_b200_a3s6b64.so a4s8b64=$L/libgsplat_b200_a4s8b64.so a1b=$L/libgsplat_b200_a1.so
