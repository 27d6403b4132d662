# 2 GPUs: dist tests again + MINRES/BPCG with the in-kernel halo path vs push+copy kernels vs NCCL (after next-chunk staging in HaloIn)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_dist.py -m gpu -x -q 2>&1 | tail -5
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --no-parity --no-simple --solve-to-tol 0 --no-cpu-baseline"
run() { tag=$1; shift; env "$@" timeout 900 $T > gpurun_out/r2h_$tag.json 2> gpurun_out/r2h_$tag.err || tail -3 gpurun_out/r2h_$tag.err; }
run fused NSB_X=1
run nohalofused NSB_HALO_FUSED=0
run nccl NSB_P2P=0
python - <<PY
import json
for f in ("fused","nohalofused","nccl"):
    try:
        d=json.loads(open("gpurun_out/r2h_%s.json"%f).read().strip().splitlines()[-1])
    except Exception as ex:
        print(f, "FAILED", ex); continue
    print("%-12s value %.1f loop %.1f launches %d minres %.1f gs %.2f ms (%d launches)" % (f, d["value"], d["iteration_loop"]["iterations_per_s"], d["gpu_launches"], d["minres"]["iterations_per_s"], d["gram_schmidt"]["ms"], d["gram_schmidt"]["launches"]))
PY
