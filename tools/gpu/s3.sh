# ncu --set full of one warm 64-keyframe step of the sorted plan (kernels only), caches left as the run leaves them
ncu --set full --import-source on --clock-control none --cache-control none -k regex:'update_depth|key_prep' -s 10 -c 5 -o gpurun_out/r2_full python tools/kbench.py --keyframes 64 --device-scenes --plans 2 --iters 2 > gpurun_out/r2_full.log 2>&1
ncu -i gpurun_out/r2_full.ncu-rep --page raw --csv > gpurun_out/r2_full_raw.csv 2>/dev/null
ls -la gpurun_out/r2_full*
