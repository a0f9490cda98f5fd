#!/bin/bash
# ncu --set full captures of the HEAD kernels, exported to CSV on the box (the .ncu-rep files are too big to bring back).
# usage (from the repo root, on the GPU box): bash profiles/ncu_capture.sh gpurun_out
out=${1:-gpurun_out}
cap() {  # name, kernel regex, skip, count, command...
  name=$1; regex=$2; skip=$3; count=$4; shift 4
  ncu --set full --clock-control none --import-source on -k regex:"$regex" -s $skip -c $count -f -o /tmp/$name "$@" > $out/ncu_$name.log 2>&1
  ncu -i /tmp/$name.ncu-rep --page raw --csv > $out/${name}_ncu_raw.csv 2>/dev/null
  rm -f /tmp/$name.ncu-rep
}
MODE=k1_768 cap r2_k1_768 "k1_stream" 3 1 python profiles/once.py
MODE=k1_384 cap r2_k1_384 "k1_stream" 3 1 python profiles/once.py
MODE=k2_1024 cap r2_k2_1024 "k2_gemm|k5_rescore" 4 2 python profiles/once.py
MODE=balanced cap r2_balanced "k1_balanced|k3_filter" 4 2 python profiles/once.py
cap r2_c4_fused "k1_fused" 4 1 python profiles/c4_once.py
CODERAG_B200_K1_FUSED=0 cap r2_c4_k3_k1 "k3_filter|k1_stream" 6 2 python profiles/c4_once.py
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k1_|k2_|k3_|k5_|merge_|signal_" -c 600 --csv --log-file $out/r2_bench_launches.csv python bench.py --steps 20 --warmup 5 --no-configs --cpu-budget 1 > /dev/null 2>&1
ls -la $out/*.csv
