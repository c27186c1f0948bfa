ncu --set full --clock-control none --cache-control none -k regex:'prop_|age_|regularize' -s 12 -c 6 -o gpurun_out/r2_comp python tools/companion_bench.py --iters 2 > gpurun_out/r2_comp.log 2>&1
ncu -i gpurun_out/r2_comp.ncu-rep --page raw --csv > gpurun_out/r2_comp_raw.csv 2>/dev/null
ls -la gpurun_out/r2_comp*
