# Where does a row-partitioned iteration wait?  Needs libnsb200.so built with NSB_NVCC_EXTRA=-DNSB_SYNC_STATS (the
# kernels then accumulate the nanoseconds they spend polling peer flags); prints per-rank microseconds per iteration.
mkdir -p gpurun_out
N=${1:-2}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline"
for mode in 1 0; do
  NSB_HALO_FUSED=$mode timeout 600 $T > gpurun_out/stats_n${N}_f$mode.json 2> gpurun_out/stats_n${N}_f$mode.err
  echo "== NSB_HALO_FUSED=$mode"; grep -o "SYNCSTATS.*" gpurun_out/stats_n${N}_f$mode.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/stats_n${N}_f$mode.json").read().strip().splitlines()[-1]); print("value %.1f loop %.1f us/iter %.1f"%(d["value"], d["iteration_loop"]["iterations_per_s"], 1e6/d["iteration_loop"]["iterations_per_s"]))
except Exception as ex: print("failed", ex)
PY
done
