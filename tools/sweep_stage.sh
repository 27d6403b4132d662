P="python tools/spmv_probe.py --dofs 20e6 --reps 20 --warmup 5"
short() { python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print({k:(round(v['ms'],4), d[k+'_format']['rows_per_chunk'], d[k+'_format']['max_distinct']) for k,v in d.items() if isinstance(v,dict) and 'ms' in v})"; }
for ST in 4600 7300 9300 14000; do echo "== stage $ST"; NSB_SELL_STAGE=$ST $P --which A,B,BT 2>/dev/null | short; done
