# programmatic-dependent-launch variants of the BPCG iteration (bench.py, short run): attribute mask, trigger / prefetch flags, graphs
mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-minres --lanczos-steps 60 --solve-to-tol 0"
run() { # name, dofs, env...
  local name=$1 dofs=$2; shift 2
  env "$@" timeout 300 $B --dofs $dofs > gpurun_out/pdl_$name.json 2> gpurun_out/pdl_$name.err
  python - gpurun_out/pdl_$name.json "$name $*" <<PY
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    o=d["roofline"]["other"]
    print("%-44s value %7.1f loop %7.1f e2e %7.1f  A %.4f ms  B %.0f BT %.0f GB/s  clocks %s %s"%(sys.argv[2], d["value"], d["iteration_loop"]["iterations_per_s"], d["e2e"]["value"], d["roofline"]["ms"], o["spmv_B_gbs"], o["spmv_BT_gbs"], d["clocks"]["sm_mhz"], d["clocks"]["reasons"]))
except Exception as e:
    print(sys.argv[2], "FAILED", e)
PY
}
for D in 20e6 2.55e6; do
run off_$D $D NSB_PDL=0
run all_$D $D NSB_PDL=3
run spmv_$D $D NSB_PDL=1
run vec_$D $D NSB_PDL=2
run noprefetch_$D $D NSB_PDL=3 NSB_PDL_FLAGS=1
run notrigger_$D $D NSB_PDL=3 NSB_PDL_FLAGS=2
run plainflags_$D $D NSB_PDL=3 NSB_PDL_FLAGS=0
run all_nograph_$D $D NSB_PDL=3 NSB_GRAPHS=0
run off_nograph_$D $D NSB_PDL=0 NSB_GRAPHS=0
done
