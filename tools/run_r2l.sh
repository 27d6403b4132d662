# round 2, 4 GPUs: 2-rank parity tests (the in-kernel-halo kernel changed since they last ran), the full bench line at N = 4 (ranks with
# two neighbours), and MINRES through push + copy kernels for comparison
N=${1:-4}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dist.py -m gpu -x -q 2>&1 | tail -4
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N"
timeout 600 $T --steps 5 --warmup 3 > gpurun_out/r2l_bench$N.json 2> gpurun_out/r2l_bench$N.err; echo "bench rc=$?"; tail -3 gpurun_out/r2l_bench$N.err | cut -c1-400
NSB_HALO_FUSED=0 timeout 300 $T --steps 2 --warmup 3 --no-parity --no-simple --no-gs --solve-to-tol 0 --no-cpu-baseline > gpurun_out/r2l_n${N}_nohalofused.json 2> gpurun_out/r2l_n${N}_nohalofused.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r2l_bench$N.json").read().strip().splitlines()[-1])
for k in ("value","ms_per_step","iteration_loop","e2e","gpu_launches","minres","time_to_tolerance","gram_schmidt","simple_step","weak_scaling","parity_vs_single","history_checkpoints"):
    v=d.get(k)
    if isinstance(v,dict): v={a:b for a,b in v.items() if a not in ("formats","nnz","workload","spmv_K_format","solution_fingerprint","algorithmic_bytes_per_iteration","algorithmic_bytes_per_step")}
    print(k, json.dumps(v)[:700])
try:
    d=json.loads(open("gpurun_out/r2l_n${N}_nohalofused.json").read().strip().splitlines()[-1])
    print("nohalofused value %.1f loop %.1f launches %d minres %.1f" % (d["value"], d["iteration_loop"]["iterations_per_s"], d["gpu_launches"], d["minres"]["iterations_per_s"]))
except Exception as ex:
    print("nohalofused FAILED", ex)
PY
