# round-1 check of the programmatic-dependent-launch build: GPU parity tests, smoke, headline bench with NSB_PDL on (default) and off,
# 1/8-size system (one rank's share of the 8-GPU run), configs 1-4, reference arm
mkdir -p gpurun_out
S() { python - "$@" <<PY
import json,sys
for p in sys.argv[1:]:
    try:
        d=json.loads(open(p).read().strip().splitlines()[-1])
        print(p, "value %.1f loop %.1f e2e %.1f launches %d roofline %.3f minres %s cpu %s clocks %s"%(d["value"], d["iteration_loop"]["iterations_per_s"], d["e2e"]["value"], d["gpu_launches"], d["roofline"]["frac"], d.get("minres",{}).get("iterations_per_s"), d.get("cpu_baseline",{}).get("value"), d["clocks"]))
    except Exception as e:
        print(p, "unreadable", e)
PY
}
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_pdl1.log 2>&1; echo "pytest default rc=$?"; tail -4 gpurun_out/pytest_pdl1.log
if ! grep -q " passed" gpurun_out/pytest_pdl1.log || grep -q "failed" gpurun_out/pytest_pdl1.log; then
  NSB_PDL=0 timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_pdl0.log 2>&1; echo "pytest NSB_PDL=0 rc=$?"; tail -4 gpurun_out/pytest_pdl0.log
fi
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_pdl1.json 2> gpurun_out/bench_pdl1.err; tail -2 gpurun_out/bench_pdl1.err | cut -c1-300
NSB_PDL=0 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_pdl0.json 2> gpurun_out/bench_pdl0.err; tail -2 gpurun_out/bench_pdl0.err | cut -c1-300
S gpurun_out/bench_pdl1.json gpurun_out/bench_pdl0.json
for P in 1 0; do
  NSB_PDL=$P timeout 300 python bench.py --dofs 2.55e6 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_small_pdl$P.json 2> gpurun_out/bench_small_pdl$P.err
done
S gpurun_out/bench_small_pdl1.json gpurun_out/bench_small_pdl0.json
for P in 1 0; do
  NSB_PDL=$P timeout 300 python tools/bench_configs.py > gpurun_out/configs_pdl$P.jsonl 2> gpurun_out/configs_pdl$P.err; cut -c1-330 gpurun_out/configs_pdl$P.jsonl
done
timeout 400 python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; cut -c1-400 gpurun_out/bench_reference.json
