O=gpurun_out/r3u; mkdir -p $O
L=gaussian_splat_ipu_b200
bash profiles/ab_k5.sh $O/ab.txt base=$L/libgsplat_b200.so x4=$L/libgsplat_b200_lb352x4.so x3=$L/libgsplat_b200_lb352x3.so x5=$L/libgsplat_b200_lb352x5.so base2=$L/libgsplat_b200.so
