# 2-GPU check of the in-kernel halo wait: parity tests, then the bench with the fused wait and with the wait/copy kernels
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -q -x 2>&1 | tail -5
N=${1:-2}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline"
timeout 600 $T > gpurun_out/bench_n${N}_fused.json 2> gpurun_out/bench_n${N}_fused.err; tail -3 gpurun_out/bench_n${N}_fused.err | cut -c1-300
NSB_HALO_FUSED=0 timeout 600 $T > gpurun_out/bench_n${N}_unfused.json 2> gpurun_out/bench_n${N}_unfused.err
python - <<PY
import json
for k in ("fused","unfused"):
    try:
        d=json.loads(open("gpurun_out/bench_n${N}_%s.json"%k).read().strip().splitlines()[-1]); print(k, "value %.1f loop %.1f e2e %.1f launches %d ms/step %.2f host %.2f"%(d["value"], d["iteration_loop"]["iterations_per_s"], d["e2e"]["value"], d["gpu_launches"], d["ms_per_step"], d["timing"]["host_clock_ms_per_step"]))
    except Exception as ex: print(k, "failed", ex)
PY
