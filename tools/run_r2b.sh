# round 2: GPU parity tests incl. nsb_simple_step, then the bench line with the SIMPLE time-step leg
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
timeout 1200 python bench.py --steps 3 --warmup 3 > gpurun_out/r2b_bench1.json 2> gpurun_out/r2b_bench1.err; echo "bench rc=$?"; tail -5 gpurun_out/r2b_bench1.err | cut -c1-600
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2b_bench1.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"])
print("simple", json.dumps(d.get("simple_step"))[:3000])
print("gs", json.dumps(d.get("gram_schmidt")))
PY
