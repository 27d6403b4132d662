# round 2: GPU parity tests (batched Gram-Schmidt, MypreA golden, MinRes + block smoother), bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 1200 python bench.py --steps 3 --warmup 3 > gpurun_out/r2f_bench1.json 2> gpurun_out/r2f_bench1.err; echo "bench rc=$?"; tail -3 gpurun_out/r2f_bench1.err | cut -c1-400
NSB_GS_BATCHED=0 timeout 600 python bench.py --steps 1 --warmup 3 --no-simple --no-parity --no-minres --solve-to-tol 0 --no-cpu-baseline > gpurun_out/r2f_bench1_gspair.json 2>/dev/null
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2f_bench1.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "minres", d["minres"]["iterations_per_s"])
print("gs", json.dumps(d.get("gram_schmidt")))
print("simple", json.dumps({k:v for k,v in d.get("simple_step",{}).items() if k not in ("formats","nnz","workload")})[:1800])
d=json.loads(open("gpurun_out/r2f_bench1_gspair.json").read().strip().splitlines()[-1])
print("gs pairwise", json.dumps(d.get("gram_schmidt")))
PY
