# one rank's share of the N-GPU run on a single GPU (no exchange): what the local kernels alone can do at 1/2, 1/4, 1/8 of the system
mkdir -p gpurun_out
for D in 10.2e6 5.1e6 2.55e6; do
  timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-minres --lanczos-steps 60 --solve-to-tol 0 --dofs $D > gpurun_out/share_$D.json 2> gpurun_out/share_$D.err
  python - gpurun_out/share_$D.json <<PY
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(d["config"]["dofs"], "value %.1f loop %.1f us/it %.1f frac %.3f"%(d["value"], d["iteration_loop"]["iterations_per_s"], 1e6/d["iteration_loop"]["iterations_per_s"], d["iteration_loop"]["frac_of_measured_peak"]))
PY
done
