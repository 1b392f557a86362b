O=gpurun_out/r3g; mkdir -p $O
L=gaussian_splat_ipu_b200/libgsplat_b200.so
bash profiles/ab_k5.sh $O/ab.txt pad0=$L pad40k=$L:GSPLAT_K5_SMEM_PAD=40000 pad60k=$L:GSPLAT_K5_SMEM_PAD=60000 "pad40k_l4=$L:GSPLAT_K5_SMEM_PAD=40000 GSPLAT_CONCURRENT_FRAMES=4" l4=$L:GSPLAT_CONCURRENT_FRAMES=4 pad0b=$L
