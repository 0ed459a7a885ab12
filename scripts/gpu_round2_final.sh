# final evidence of the round on one B200: the whole GPU test-suite, the default bench, the CPU reference arm, ncu captures
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2_gputests_final.log 2>&1; echo gputests rc=$?; tail -4 gpurun_out/r2_gputests_final.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err; echo bench rc=$?; tail -c 900 gpurun_out/r2_bench_final.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference.json 2> gpurun_out/r2_bench_reference.err; echo ref rc=$?; tail -c 900 gpurun_out/r2_bench_reference.json
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo smoke rc=$?; tail -2 gpurun_out/r2_smoke.log
timeout 120 python scripts/profile_kernels.py > gpurun_out/r2_profile_kernels_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -o gpurun_out/r2_kernels python scripts/profile_kernels.py > gpurun_out/r2_profile_kernels_ncu.log 2>&1; echo ncu rc=$?
