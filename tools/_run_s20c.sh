set -x
python -m pytest tests/test_gpu_layers.py -m gpu -x -q > gpurun_out/s20c_layers.log 2>&1; echo "rc=$?" >> gpurun_out/s20c_layers.log
python bench.py --sweep --no-cpu-baseline > gpurun_out/s20c_sweep.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1d_bench_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/s20c_ncu1.log 2>&1
for M in 3 4 5; do
ncu --set full --clock-control none --import-source on -k regex:approx_gemm_lut -c 1 -s 4 -f -o gpurun_out/r1d_lut_M$M python bench.py --mant-width $M --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/s20c_ncu_M$M.log 2>&1
done
ls -la gpurun_out
