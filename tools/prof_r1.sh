# ncu passes for profiles/: launch list of a short bench, then one full capture of the iteration kernels.
set -x
B="python bench.py --steps 1 --warmup 1 --iters 8 --lanczos-steps 6 --no-cpu-baseline --minres-iters 6 --solve-to-tol 0"
$B > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"spmv|bpcg|minres|k_" -c 600 --csv --log-file gpurun_out/r1_launches_final.csv $B > gpurun_out/ncu_list.log 2>&1
tail -2 gpurun_out/ncu_list.log
# BPCG iteration kernels (B1, B2, B3, b4, b5 of the third iteration of the warm-up step)
ncu --set full --clock-control none --import-source on -k regex:"spmv_sell_kernel|k_bpcg_b4|k_bpcg_b5" -s 27 -c 5 -o gpurun_out/r1_bpcg_final $B > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
ls -la gpurun_out
