# round 2, 2 GPUs: fused push / in-block all-reduce -- 2-rank parity tests (all exchange paths), fault test, bench N=2 (peer + NCCL)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_dist.py -m gpu -x -q 2>&1 | tail -15
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --no-parity"
timeout 900 $T > gpurun_out/r2d_bench2.json 2> gpurun_out/r2d_bench2.err; echo "bench rc=$?"; tail -5 gpurun_out/r2d_bench2.err | cut -c1-600
NSB_PUSH_FUSED=0 timeout 900 $T --no-simple --no-gs --no-minres --solve-to-tol 0 > gpurun_out/r2d_bench2_nopush.json 2> gpurun_out/r2d_bench2_nopush.err
NSB_P2P=0 timeout 900 $T --no-simple --no-gs --solve-to-tol 0 > gpurun_out/r2d_bench2_nccl.json 2> gpurun_out/r2d_bench2_nccl.err
python - <<'PY'
import json
for f in ("r2d_bench2","r2d_bench2_nopush","r2d_bench2_nccl"):
    try:
        d=json.loads(open("gpurun_out/%s.json"%f).read().strip().splitlines()[-1])
    except Exception as ex:
        print(f, "FAILED", ex); continue
    print(f, "value", d["value"], "loop", d["iteration_loop"]["iterations_per_s"], "launches", d["gpu_launches"], "hist", d["history_checkpoints"])
    for k in ("minres","simple_step","time_to_tolerance"):
        if d.get(k): print("   ", k, json.dumps(d[k])[:400])
PY
