import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=rows[0]; units=rows[1]
r=rows[2]
keys=[h for h in hdr if any(t in h for t in ("gpu__time_duration.sum","dram__bytes_read.sum","dram__throughput.avg.pct","sm__throughput.avg.pct","warps_active.avg.pct","issue_active.avg.pct","smsp__average_warps_issue_stalled","launch__grid_size","launch__occupancy_limit","sm__cycles_active.avg","smsp__cycles_active.avg","launch__waves","l1tex__t_sector_hit","lts__t_sector_hit","smsp__inst_executed.sum","sm__cycles_elapsed.max","launch__block_size","launch__shared_mem"))]
for k in keys:
    i=hdr.index(k); print(k, r[i], units[i])
