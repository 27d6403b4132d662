# N-GPU bench with the peer-memory exchanges and with the NCCL fallback
N=$1
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline"
timeout 600 $T > gpurun_out/bench_n${N}_p2p.json 2> gpurun_out/bench_n${N}_p2p.err; tail -3 gpurun_out/bench_n${N}_p2p.err | cut -c1-300
NSB_P2P=0 timeout 600 $T > gpurun_out/bench_n${N}_nccl.json 2> gpurun_out/bench_n${N}_nccl.err
python - <<PY
import json
for k in ("p2p","nccl"):
    try:
        d=json.loads(open("gpurun_out/bench_n${N}_%s.json"%k).read().strip().splitlines()[-1]); print(k, "value %.1f loop %.1f e2e %.1f launches %d"%(d["value"], d["iteration_loop"]["iterations_per_s"], d["e2e"]["value"], d["gpu_launches"]))
    except Exception as ex: print(k, "failed", ex)
PY
