# SELL kernel variants by NSB_SELL_FLAGS (bit 2: two epilogue rows in flight, bit 3: x staged with cp.async); FLAGS="7 15" selects the list
mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --lanczos-steps 60 --solve-to-tol 0"
run() { # name, dofs, env...
  local name=$1 dofs=$2; shift 2
  env "$@" timeout 300 $B --dofs $dofs > gpurun_out/epi_$name.json 2> gpurun_out/epi_$name.err
  python - gpurun_out/epi_$name.json "$name $*" <<PY
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    o=d["roofline"]["other"]
    print("%-36s value %7.1f loop %7.1f e2e %7.1f minres %7.1f  A %.4f ms  clocks %s %s"%(sys.argv[2], d["value"], d["iteration_loop"]["iterations_per_s"], d["e2e"]["value"], d["minres"]["iterations_per_s"], d["roofline"]["ms"], d["clocks"]["sm_mhz"], d["clocks"]["reasons"]))
except Exception as e:
    print(sys.argv[2], "FAILED", e)
PY
}
for D in 20e6 2.55e6; do
for F in ${FLAGS:-7 15 7 15}; do
run f${F}_$D $D NSB_SELL_FLAGS=$F
done
done
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
