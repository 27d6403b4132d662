# round 2, first pass: GPU parity tests, the reworked bench line (parity_at_sample, fingerprints, MINRES, GS, cpu_baseline on the
# same system) and the reference arm on the full system
mkdir -p gpurun_out
nproc; free -g | head -2
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r2a_bench1.json 2> gpurun_out/r2a_bench1.err; echo "bench rc=$?"; tail -3 gpurun_out/r2a_bench1.err | cut -c1-400
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r2a_ref.json 2> gpurun_out/r2a_ref.err; echo "ref rc=$?"; tail -3 gpurun_out/r2a_ref.err | cut -c1-400
cut -c1-600 gpurun_out/r2a_ref.json
