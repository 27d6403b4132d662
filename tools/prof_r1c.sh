# final round-1 evidence: GPU parity tests + smoke, ncu launch list of a short bench, one full capture of the BPCG iteration kernels,
# one full capture of the plain A SpMV (roofline.traffic), then the plain bench line (each ncu pass after its command exited 0 without ncu)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
set -x
B="python bench.py --steps 1 --warmup 1 --iters 8 --lanczos-steps 6 --no-cpu-baseline --minres-iters 6 --solve-to-tol 0"
$B > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"spmv|bpcg|minres|k_" -c 600 --csv --log-file gpurun_out/r1c_launches.csv $B > gpurun_out/ncu_list.log 2>&1
tail -2 gpurun_out/ncu_list.log
python tools/launch_summary.py gpurun_out/r1c_launches.csv > gpurun_out/r1c_launch_summary.txt; head -14 gpurun_out/r1c_launch_summary.txt
ncu --set full --clock-control none --import-source on -k regex:"spmv_sell_kernel|k_bpcg_b4|k_bpcg_b5" -s 27 -c 5 -o gpurun_out/r1c_bpcg $B > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
ncu -i gpurun_out/r1c_bpcg.ncu-rep --page raw --csv > gpurun_out/r1c_bpcg_raw.csv 2>/dev/null
P="python tools/spmv_probe.py --dofs 20e6 --which A --reps 6 --warmup 3"
$P > gpurun_out/traffic_plain.json 2> gpurun_out/traffic_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:spmv_sell_kernel -s 5 -c 2 -o gpurun_out/r1c_spmv_A_20M $P > gpurun_out/traffic_ncu.log 2>&1
tail -2 gpurun_out/traffic_ncu.log
ncu -i gpurun_out/r1c_spmv_A_20M.ncu-rep --page raw --csv > gpurun_out/r1c_spmv_A_20M_raw.csv 2>/dev/null
rm -f gpurun_out/r1c_spmv_A_20M.ncu-rep
set +x
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r1c.json 2> gpurun_out/bench_r1c.err; tail -2 gpurun_out/bench_r1c.err | cut -c1-300
cut -c1-700 gpurun_out/bench_r1c.json
timeout 300 python tools/bench_configs.py > gpurun_out/configs_r1c.jsonl 2> gpurun_out/configs_r1c.err; cut -c1-250 gpurun_out/configs_r1c.jsonl
