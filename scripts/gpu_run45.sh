mkdir -p gpurun_out
B=64 NEW=2 REPS=2 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_step_b64_v2.csv python scripts/profile_step.py > gpurun_out/ncu_step.log 2>&1; tail -1 gpurun_out/ncu_step.log
