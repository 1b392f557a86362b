O=gpurun_out/r3y; mkdir -p $O
L=gaussian_splat_ipu_b200
bash profiles/ab_k5.sh $O/ab.txt base=$L/libgsplat_b200.so k1mb8=$L/libgsplat_b200_k1mb8.so base2=$L/libgsplat_b200.so
