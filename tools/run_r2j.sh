# round 2: the allocation-free entry point alone in a fresh process, then the whole GPU suite (it runs last), then the bench line
mkdir -p gpurun_out
echo "== no-alloc test alone"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -3
echo "== full gpu suite"; timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
# (compute-sanitizer is closed on this pool: the memcheck / racecheck passes were refused)
echo "== bench"; timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r2j_bench1.json 2> gpurun_out/r2j_bench1.err; echo "bench rc=$?"; tail -2 gpurun_out/r2j_bench1.err | cut -c1-300
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2j_bench1.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "loop frac", d["iteration_loop"]["frac_of_measured_peak"], "minres", d["minres"]["iterations_per_s"])
print("gs", d["gram_schmidt"]["ms"], d["gram_schmidt"]["launches"], "simple", d["simple_step"].get("ms_per_step"), d["simple_step"].get("frac_of_measured_peak"))
print("parity", d["parity_at_sample"].get("within_contract"), "cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"])
PY
