set -x
export PROFILE_N=2
python tools/profile_target.py all > gpurun_out/profile_target.log 2>&1 || { tail -20 gpurun_out/profile_target.log; exit 1; }
tail -3 gpurun_out/profile_target.log
ncu --set full --clock-control none --import-source on -k regex:"fused|stencil|daxpy" -f -o gpurun_out/prof_all python tools/profile_target.py all > gpurun_out/ncu_all.log 2>&1; tail -5 gpurun_out/ncu_all.log; ls -la gpurun_out/
python benchmarks/sweep.py --out gpurun_out/sweep_n1.jsonl > gpurun_out/sweep_n1.log 2>&1; tail -3 gpurun_out/sweep_n1.log
