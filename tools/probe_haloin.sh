# the in-kernel-halo variant of the SELL SpMV against the plain one, on one GPU (empty halo, identity order)
mkdir -p gpurun_out
for d in 10e6 2.7e6; do
  for h in 0 1; do
    NSB_BENCH_HALOIN=$h python tools/spmv_probe.py --dofs $d --reps 100 --warmup 20 --which A,B,BT,K > gpurun_out/probe_h${h}_$d.json 2>/dev/null
  done
done
python - <<'PY'
import json
for d in ("10e6","2.7e6"):
    a=json.load(open("gpurun_out/probe_h0_%s.json"%d)); b=json.load(open("gpurun_out/probe_h1_%s.json"%d))
    for k in ("A","B","BT","K"):
        print(d, k, "plain %.1f us  haloin %.1f us  (%+.1f%%)" % (1e3*a[k]["ms"], 1e3*b[k]["ms"], 100*(b[k]["ms"]/a[k]["ms"]-1)))
PY
