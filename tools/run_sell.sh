python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > gpurun_out/tests.log; tail -8 gpurun_out/tests.log
NSB_SELL_DEBUG=1 python tools/spmv_probe.py --dofs 20e6 --reps 20 --warmup 5 --which A,B,BT > gpurun_out/probe_sell.json 2> gpurun_out/probe_sell.err; grep built gpurun_out/probe_sell.err; cat gpurun_out/probe_sell.json
