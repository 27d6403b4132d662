# round 2, 1 GPU: bank-aware entry order of the long-row SELL panels (NSB_SELL_BANKED) -- parity tests with it on, B / K SpMV timing on/off,
# BPCG + MINRES rates on/off
mkdir -p gpurun_out
echo "== parity tests with the bank-aware layout"; NSB_SELL_BANKED=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "spmv or sell or bpcg_v1 or minres_matches or empty_row" 2>&1 | tail -3
for b in 1 0; do
  NSB_SELL_BANKED=$b python tools/spmv_probe.py --dofs 20e6 --reps 100 --warmup 20 --which A,B,BT,K > gpurun_out/probeB_banked$b.json 2>gpurun_out/probeB_banked$b.err || tail -3 gpurun_out/probeB_banked$b.err
done
python - <<'PY'
import json
a=json.load(open("gpurun_out/probeB_banked1.json")); b=json.load(open("gpurun_out/probeB_banked0.json"))
for k in ("A","B","BT","K"):
    print(k, "banked %.1f us (%.0f GB/s)  ascending %.1f us (%.0f GB/s)" % (1e3*a[k]["ms"], a[k]["GBs"], 1e3*b[k]["ms"], b[k]["GBs"]))
PY
S="--steps 3 --warmup 3 --no-parity --no-simple --no-gs --solve-to-tol 0 --no-cpu-baseline"
for b in 1 0; do NSB_SELL_BANKED=$b timeout 600 python bench.py $S > gpurun_out/r2k_banked$b.json 2>/dev/null; done
python - <<'PY'
import json
for b in (1,0):
    d=json.loads(open("gpurun_out/r2k_banked%d.json"%b).read().strip().splitlines()[-1])
    print("banked", b, "value %.1f loop %.1f minres %.1f  B %.1f us BT %.1f us" % (d["value"], d["iteration_loop"]["iterations_per_s"], d["minres"]["iterations_per_s"], 1e3*d["roofline"]["other"]["spmv_B_ms"], 1e3*d["roofline"]["other"]["spmv_BT_ms"]))
PY
