O=gpurun_out/r3m; mkdir -p $O
L=gaussian_splat_ipu_b200/libgsplat_b200.so
bash profiles/ab_k5.sh $O/ab.txt l3=$L l4=$L:GSPLAT_CONCURRENT_FRAMES=4 l5=$L:GSPLAT_CONCURRENT_FRAMES=5 l6=$L:GSPLAT_CONCURRENT_FRAMES=6 l8=$L:GSPLAT_CONCURRENT_FRAMES=8
