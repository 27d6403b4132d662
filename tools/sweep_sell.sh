# SELL-LX tuning sweep: unroll depth x chunk size (run on the GPU box; rebuilds the library per variant)
P="python tools/spmv_probe.py --dofs 20e6 --reps 20 --warmup 5"
short() { python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print({k:(round(v['ms'],4), d[k+'_format']['rows_per_chunk'], d[k+'_format']['max_distinct']) for k,v in d.items() if isinstance(v,dict) and 'ms' in v})"; }
for U in 8 4; do for MB in 4 6; do
  NSB_NVCC_EXTRA="-DNSB_SELL_UNROLL=$U -DNSB_SELL_MINBLOCKS=$MB" python -m navier_stokes_solver_b200.build > /dev/null 2>&1 || echo BUILD FAIL
  echo "== U=$U MB=$MB default"; $P --which A,B,BT 2>/dev/null | short
  for L in 9 10; do echo "== U=$U MB=$MB A log2R=$L"; NSB_SELL_LOG2R=$L $P --which A 2>/dev/null | short; done
  for L in 5 7; do for T in 2 4 8; do echo "== U=$U MB=$MB B log2R=$L T=$T"; NSB_SELL_T=$T NSB_SELL_LOG2R=$L $P --which B 2>/dev/null | short; done; done
done; done
