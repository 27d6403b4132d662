# DRAM bytes per launch of the plain A SpMV at the bench size (roofline.traffic of bench.py): one ncu --set full capture.
set -x
P="python tools/spmv_probe.py --dofs 20e6 --which A --reps 6 --warmup 3"
$P > gpurun_out/traffic_plain.json 2> gpurun_out/traffic_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:spmv_sell_kernel -s 5 -c 2 -o gpurun_out/r1_spmv_A_20M $P > gpurun_out/traffic_ncu.log 2>&1
tail -2 gpurun_out/traffic_ncu.log
ncu -i gpurun_out/r1_spmv_A_20M.ncu-rep --page raw --csv > gpurun_out/r1_spmv_A_20M_raw.csv
ls -la gpurun_out | tail -5
