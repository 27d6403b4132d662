# 2-GPU check: parity tests, then the bench with the peer-memory exchanges and with the NCCL fallback
python -m pytest tests/test_gpu_dist.py -m gpu -q -x 2>&1 | tail -5
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --steps 3 --warmup 3 --no-cpu-baseline"
timeout 600 $T > gpurun_out/bench_n$1_p2p.json 2> gpurun_out/bench_n$1_p2p.err; tail -3 gpurun_out/bench_n$1_p2p.err | cut -c1-300
NSB_P2P=0 timeout 600 $T > gpurun_out/bench_n$1_nccl.json 2> gpurun_out/bench_n$1_nccl.err
python - <<PY
import json
for k in ("p2p","nccl"):
    try:
        d=json.loads(open("gpurun_out/bench_n$1_%s.json"%k).read().strip().splitlines()[-1]); print(k, "value %.1f loop %.1f e2e %.1f launches %d"%(d["value"], d["iteration_loop"]["iterations_per_s"], d["e2e"]["value"], d["gpu_launches"]))
    except Exception as ex: print(k, "failed", ex)
PY
