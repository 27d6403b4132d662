# What one rank of an 8-GPU run sees: 1/8 of the 20.4M-dof system on one GPU.  Live rate, then the ncu launch list to
# compare the sum of kernel times with the live time per iteration (the difference is launch gaps / tails).
set -x
B="python bench.py --dofs 2.55e6 --steps 5 --warmup 3 --no-cpu-baseline --no-minres --lanczos-steps 50"
$B 2>/dev/null | tail -1 > gpurun_out/small_plain.json
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"spmv|bpcg|k_" -c 900 --csv --log-file gpurun_out/small_launches.csv python bench.py --dofs 2.55e6 --steps 1 --warmup 1 --iters 8 --lanczos-steps 6 --no-cpu-baseline --no-minres > gpurun_out/small_ncu.log 2>&1
tail -2 gpurun_out/small_ncu.log
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-minres --solve-to-tol 1e-10 2>/dev/null | tail -1 > gpurun_out/bench_ttt.json
