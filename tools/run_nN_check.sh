# N-GPU bench (peer-memory path) and its residual-history checkpoints, to be compared with the 1-GPU line
N=$1
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline"
timeout 600 $T > gpurun_out/bench_n${N}_p2p.json 2> gpurun_out/bench_n${N}_p2p.err; tail -3 gpurun_out/bench_n${N}_p2p.err | cut -c1-300
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_n${N}_p2p.json").read().strip().splitlines()[-1])
print("N=${N} value %.1f loop %.1f e2e %.1f launches %d"%(d["value"], d["iteration_loop"]["iterations_per_s"], d["e2e"]["value"], d["gpu_launches"]), d["history_checkpoints"], d["clocks"])
PY
