python -m pytest tests -m gpu -q 2>&1 | tail -8 > gpurun_out/tests.log; tail -4 gpurun_out/tests.log
python tools/spmv_probe.py --dofs 4e6 --which A --reps 5 --warmup 2 > gpurun_out/probe_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:spmv_lx -s 3 -c 1 -o gpurun_out/spmv_r1_lx_a python tools/spmv_probe.py --dofs 4e6 --which A --reps 5 --warmup 2 > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
