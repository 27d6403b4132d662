python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > gpurun_out/tests.log; tail -5 gpurun_out/tests.log
python tools/spmv_probe.py --dofs 20e6 --reps 20 --warmup 5 2>&1 | tail -1 > gpurun_out/probe_lx.json; cat gpurun_out/probe_lx.json
NSB_LX=0 python tools/spmv_probe.py --dofs 20e6 --reps 20 --warmup 5 --which A,B,BT 2>&1 | tail -1 > gpurun_out/probe_csr.json; cat gpurun_out/probe_csr.json
