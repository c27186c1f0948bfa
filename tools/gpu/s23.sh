tools/micro/gather_bench
ncu --metrics gpu__time_duration.sum,lts__t_sectors.sum,l1tex__t_sector_hit_rate.pct --clock-control none -k regex:gather -s 3 -c 1 tools/micro/gather_bench 2>&1 | grep -E "gpu__time|lts__t_sectors|hit_rate"
