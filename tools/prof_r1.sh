# ncu passes for profiles/: launch list of a short bench, then one full capture of the iteration kernels.
set -x
B="python bench.py --steps 1 --warmup 1 --iters 8 --lanczos-steps 6 --no-cpu-baseline"
$B > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"spmv|bpcg|k_" -c 400 --csv --log-file gpurun_out/r1_launches.csv $B > gpurun_out/ncu_list.log 2>&1
tail -2 gpurun_out/ncu_list.log
ncu --set full --clock-control none --import-source on -k regex:"spmv_lx|k_bpcg_b4|k_bpcg_b5" -s 30 -c 5 -o gpurun_out/r1_bpcg_full $B > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
ls -la gpurun_out
