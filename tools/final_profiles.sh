# Regenerates the 1-GPU files under profiles/ on the GPU box (outputs into gpurun_out/, copied to profiles/ afterwards).
# Every timed value comes from the plain runs; the ncu passes only provide counters and launch shares.
set -x
python bench.py --steps 20 > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err || exit 1
python tools/profile_kernels.py big > gpurun_out/kernels_big_timings.txt 2>&1
python tools/profile_kernels.py > gpurun_out/kernels_cfg_timings.txt 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu --no-configs > gpurun_out/launches_run.log 2>&1
ncu --set full --clock-control none -k regex:'k_objective|k_gateaux_eval' -o /tmp/big python tools/profile_kernels.py big once > gpurun_out/ncu_big.log 2>&1
python tools/ncu_summary.py /tmp/big.ncu-rep gpurun_out/kernels_big_summary.txt
ncu --set full --clock-control none -k regex:'k_objective|k_gateaux_eval|k_pso_fused' -o /tmp/cfg python tools/profile_kernels.py once > gpurun_out/ncu_cfg.log 2>&1
python tools/ncu_summary.py /tmp/cfg.ncu-rep gpurun_out/kernels_cfg_summary.txt
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'k_objective|k_gateaux_eval' --csv --log-file gpurun_out/traffic.csv \
    python tools/profile_kernels.py big once > gpurun_out/traffic_run.log 2>&1
ls -la gpurun_out/
