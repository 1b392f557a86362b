O=gpurun_out/r3n; mkdir -p $O
L=gaussian_splat_ipu_b200/libgsplat_b200.so
bash profiles/ab_k5.sh $O/ab.txt base=$L later2=$L:GSPLAT_K5_SPLIT_LATER=2 later1=$L:GSPLAT_K5_SPLIT_LATER=1 split4=$L:GSPLAT_K5_SPLIT=4 split1=$L:GSPLAT_K5_SPLIT=1 s4l2=$L:"GSPLAT_K5_SPLIT=4 GSPLAT_K5_SPLIT_LATER=2"
