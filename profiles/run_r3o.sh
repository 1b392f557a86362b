O=gpurun_out/r3o; mkdir -p $O
L=gaussian_splat_ipu_b200
bash profiles/ab_k5.sh $O/ab2.txt base=$L/libgsplat_b200.so mb3=$L/libgsplat_b200_mb3.so base2=$L/libgsplat_b200.so mb3s1=$L/libgsplat_b200_mb3.so:GSPLAT_K5_SPLIT=1
