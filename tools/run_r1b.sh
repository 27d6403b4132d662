# full 1-GPU check: GPU parity tests, headline bench, configs 1-4
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r1b.json 2> gpurun_out/bench_r1b.err; tail -2 gpurun_out/bench_r1b.err | cut -c1-300
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_r1b.json").read().strip().splitlines()[-1])
print("value %.1f loop %.1f e2e %.1f launches %d ms/step %.2f host %.2f roofline %.3f minres %.1f cpu %.2f"%(d["value"], d["iteration_loop"]["iterations_per_s"], d["e2e"]["value"], d["gpu_launches"], d["ms_per_step"], d["timing"]["host_clock_ms_per_step"], d["roofline"]["frac"], d["minres"]["iterations_per_s"], d["cpu_baseline"]["value"]))
print(d["clocks"])
PY
timeout 600 python tools/bench_configs.py > gpurun_out/configs_r1b.jsonl 2> gpurun_out/configs_r1b.err; cut -c1-330 gpurun_out/configs_r1b.jsonl
