set -x
python -m pytest tests -m gpu -q 2>&1 | tail -15
for cs in 1 4 16 64; do for tpr in 4 8 16; do
  echo "== cs=$cs tpr=$tpr"; NSB_CHUNK_STEPS=$cs NSB_TPR=$tpr python tools/spmv_probe.py --dofs 20e6 --which A --reps 20 --warmup 5 2>&1 | tail -1
done; done
echo "== default all"; python tools/spmv_probe.py --dofs 20e6 --reps 20 --warmup 5 2>&1 | tail -1
python tools/spmv_probe.py --dofs 2e6 --which A --reps 5 --warmup 2 > gpurun_out/probe_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:spmv_csrvec -s 3 -c 2 -o gpurun_out/spmv_r1_a python tools/spmv_probe.py --dofs 2e6 --which A --reps 5 --warmup 2 > gpurun_out/ncu1.log 2>&1
tail -3 gpurun_out/ncu1.log
