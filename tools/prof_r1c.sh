# final round-1 evidence: launch list of a short bench and one full capture of the BPCG iteration kernels (after the plain run exits 0)
mkdir -p gpurun_out
set -x
B="python bench.py --steps 1 --warmup 1 --iters 8 --lanczos-steps 6 --no-cpu-baseline --minres-iters 6"
$B > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"spmv|bpcg|minres|k_" -c 600 --csv --log-file gpurun_out/r1c_launches.csv $B > gpurun_out/ncu_list.log 2>&1
tail -2 gpurun_out/ncu_list.log
python tools/launch_summary.py gpurun_out/r1c_launches.csv > gpurun_out/r1c_launch_summary.txt; head -24 gpurun_out/r1c_launch_summary.txt
ncu --set full --clock-control none --import-source on -k regex:"spmv_sell_kernel|k_bpcg_b4|k_bpcg_b5" -s 27 -c 5 -o gpurun_out/r1c_bpcg $B > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
ncu -i gpurun_out/r1c_bpcg.ncu-rep --page raw --csv > gpurun_out/r1c_bpcg_raw.csv 2>/dev/null
ls -la gpurun_out | tail -8
set +x
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r1c.json 2> gpurun_out/bench_r1c.err; tail -2 gpurun_out/bench_r1c.err | cut -c1-300
cut -c1-1500 gpurun_out/bench_r1c.json
