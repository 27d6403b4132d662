# round 2, 2 GPUs: 2-rank parity tests (BPCG, MINRES, Lanczos, SIMPLE step), bench N=2 with the new objects
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_dist.py -m gpu -x -q 2>&1 | tail -15
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3"
timeout 900 $T > gpurun_out/r2c_bench2.json 2> gpurun_out/r2c_bench2.err; echo "bench rc=$?"; tail -5 gpurun_out/r2c_bench2.err | cut -c1-600
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2c_bench2.json").read().strip().splitlines()[-1])
for k in ("value","iteration_loop","minres","time_to_tolerance","gram_schmidt","simple_step","parity_vs_single","gpu_launches"):
    print(k, json.dumps(d.get(k))[:1500])
PY
