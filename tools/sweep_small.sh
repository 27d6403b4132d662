# 1-GPU sweep at the per-rank sizes of the 2- and 8-GPU runs: where does the small-size loss come from, and does a
# smaller chunk target recover it?
mkdir -p gpurun_out
run() {  # tag dofs env...
  tag=$1; dofs=$2; shift 2
  env "$@" NSB_SELL_DEBUG=1 timeout 600 python bench.py --dofs $dofs --steps 3 --warmup 3 --no-cpu-baseline --no-minres > gpurun_out/sw_$tag.json 2> gpurun_out/sw_$tag.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/sw_$tag.json").read().strip().splitlines()[-1])
    r=d["roofline"]; c=d["config"]
    print("$tag dofs %d value %.1f loop %.1f A %.3f ms (%.0f GB/s) B %.0f BT %.0f GB/s iter_frac %.3f"%(c["dofs"], d["value"], d["iteration_loop"]["iterations_per_s"], r["ms"], r["achieved"], r["other"]["spmv_B_gbs"], r["other"]["spmv_BT_gbs"], d["iteration_loop"]["frac_of_measured_peak"]))
except Exception as ex: print("$tag failed", ex)
PY
  grep "built:" gpurun_out/sw_$tag.err | head -3
}
run full 20e6 X=1
run half 10.2e6 X=1
run eighth 2.55e6 X=1
run eighth_c4 2.55e6 NSB_SELL_CHUNKS_PER_BLOCK=4
run eighth_c8 2.55e6 NSB_SELL_CHUNKS_PER_BLOCK=8
run eighth_c16 2.55e6 NSB_SELL_CHUNKS_PER_BLOCK=16
run half_c8 10.2e6 NSB_SELL_CHUNKS_PER_BLOCK=8
