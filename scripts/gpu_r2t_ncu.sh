#!/bin/bash
# Round-2 ncu pass: full captures of the dominant kernels of THIS build (traffic for bench.py's roofline.traffic).
tag=${1:-r2t}
out=gpurun_out
mkdir -p $out
K="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-sub --brief"
timeout 600 $K > $out/${tag}_plain_k512.json 2> $out/${tag}_plain_k512.err || exit 1
N="ncu --set full --clock-control none --import-source on -f"
timeout 900 $N -k regex:k_cn_fast -s 13 -c 1 -o $out/${tag}_cn_fast_k512 $K > $out/${tag}_ncu1.log 2>&1
timeout 900 $N -k regex:k_vn_reg -s 13 -c 1 -o $out/${tag}_vn_reg_k512 $K > $out/${tag}_ncu2.log 2>&1
timeout 900 $N -k regex:k_cn_exact -c 1 -o $out/${tag}_cn_exact_k512 $K > $out/${tag}_ncu3.log 2>&1
H="python bench.py --workload hqc128 --trials 4096 --steps 1 --warmup 1 --no-cpu-baseline --no-sub --brief"
timeout 600 $H > $out/${tag}_plain_hqc128.json 2> $out/${tag}_plain_hqc128.err || exit 1
timeout 900 $N -k regex:k_bp_cn -s 1 -c 1 -o $out/${tag}_bp_cn_hqc128 $H > $out/${tag}_ncu4.log 2>&1
timeout 900 $N -k regex:k_bp_vn -s 10 -c 6 -o $out/${tag}_bp_vn_hqc128 $H > $out/${tag}_ncu5.log 2>&1
for r in $out/${tag}_*.ncu-rep; do python scripts/ncu_summary.py $r > ${r%.ncu-rep}.txt 2>&1; rm -f $r; done
ls -la $out/${tag}_*
