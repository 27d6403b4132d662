# round 2, N GPUs (default 2): A/B of the exchange paths on one box -- fused push + in-block all-reduce (default), push kernels,
# all-reduce kernels, both unfused, NCCL
N=${1:-2}
mkdir -p gpurun_out
if [ "$N" = "2" ]; then timeout 1500 python -m pytest tests/test_gpu_dist.py -m gpu -x -q 2>&1 | tail -5; fi
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 --no-parity --no-simple --no-gs --solve-to-tol 0 --no-cpu-baseline"
run() { tag=$1; shift; env "$@" timeout 900 $T > gpurun_out/r2e_n${N}_$tag.json 2> gpurun_out/r2e_n${N}_$tag.err || tail -3 gpurun_out/r2e_n${N}_$tag.err; }
run fused NSB_X=1
run nopush NSB_PUSH_FUSED=0
run noinblock NSB_REDUCE_INBLOCK=0
run neither NSB_PUSH_FUSED=0 NSB_REDUCE_INBLOCK=0
run nccl NSB_P2P=0
run fused2 NSB_X=1
python - <<PY
import json
for f in ("fused","nopush","noinblock","neither","nccl","fused2"):
    try:
        d=json.loads(open("gpurun_out/r2e_n${N}_%s.json"%f).read().strip().splitlines()[-1])
    except Exception as ex:
        print(f, "FAILED", ex); continue
    print("%-10s value %.1f loop %.1f launches %d minres %.1f err100 %r" % (f, d["value"], d["iteration_loop"]["iterations_per_s"], d["gpu_launches"], d["minres"]["iterations_per_s"], d["history_checkpoints"]["err_100"]))
PY
