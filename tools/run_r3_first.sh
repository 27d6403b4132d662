# First GPU call of a next round: everything that was written after the round-2 GPU budget ran out, in the order of what it can break.
#   bash tools/run_r3_first.sh        (1 GPU: steps 1-3)      bash tools/run_r3_first.sh 2   (2 GPUs: adds steps 4-5)
N=${1:-1}
mkdir -p gpurun_out
echo "== 1. verified suite (-x): the adaptor test (tests/test_gpu_y_adaptor.py) has never run on a device"
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
echo "== 2. bench line with the preconditioner leg (own process, last)"
timeout 1200 python bench.py --steps 5 --warmup 3 --precond-dofs 2e6 --precond-timeout 400 > gpurun_out/r3a_bench1.json 2> gpurun_out/r3a_bench1.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r3a_bench1.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "watchdog", d.get("watchdog"))
p = d.get("preconditioners", {})
for k in ("jacobi", "block_jacobi", "block_gs", "aux_gs"):
    print(k, {a: p.get(k, {}).get(a) for a in ("iterations", "loop_seconds", "ms_per_iteration", "launches_per_iteration", "status", "error")})
print({a: p.get(a) for a in ("dofs", "blocks", "colours", "setup_s", "error", "stderr_tail")})
PY
echo "== 3. ncu launch list of the preconditioner leg alone (which kernel of the block sweeps costs what)"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r3a_precond_launches.csv \
    python benchlegs_precond.py --dofs 2e6 --which block_gs --maxsteps 20 > gpurun_out/r3a_precond_ncu.log 2>&1; echo "ncu rc=$?"
echo "== 3b. the drop-in example (never run on a device)"
timeout 300 python examples/stokes_channel.py --csv gpurun_out/r3a_example_errors.csv 2>&1 | tail -4
if [ "$N" -ge 2 ]; then
  echo "== 4. row-partitioned condensed BPCG / block smoother / MypreA / MinRes with the composite: first run on a device"
  NSB_RUN_UNVERIFIED_GPU_TESTS=1 timeout 1500 python -m pytest tests/test_gpu_dist_condensed.py -m gpu -x -q 2>&1 | tail -15
  echo "== 5. the verified 2-rank suite again (same library)"
  timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -x -q 2>&1 | tail -4
fi
