O=gpurun_out/r3r; mkdir -p $O
GSPLAT_DEBUG=1 timeout 200 python profiles/tile_profile.py synthetic-6M-1080p 3 > $O/tile_profile.txt 2> $O/probe.err
grep "K5 probes" $O/probe.err | tail -3; head -4 $O/tile_profile.txt
GSPLAT_DEBUG=1 timeout 200 python profiles/tile_profile.py synthetic-6M-4k 3 2>&1 | grep "K5 probes" | tail -1
GSPLAT_DEBUG=1 timeout 200 python profiles/tile_profile.py ply-point_cloud_9-720p-div1.0 3 2>&1 | grep "K5 probes" | tail -1
