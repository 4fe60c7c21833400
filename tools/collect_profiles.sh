# Round evidence, collected on one B200 (run under gpurun from the repo root).  ncu only after the same command has
# exited 0 without it; never under torchrun.
set -x
python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/r01_pytest_gpu.log
python bench.py > gpurun_out/r01_bench.json 2> gpurun_out/bench.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r01_bench_reference_arm.json 2>> gpurun_out/bench.err
python tools/bench_configs.py --configs 3,4,5 > gpurun_out/r01_bench_configs.jsonl 2>> gpurun_out/bench.err
python tools/bench_selective.py > gpurun_out/r01_bench_selective.jsonl 2>> gpurun_out/bench.err
python tools/bench_layers.py > gpurun_out/r01_bench_layers.jsonl 2>> gpurun_out/bench.err
B="python bench.py --steps 2 --warmup 3 --no-cpu"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01_launches.csv $B > gpurun_out/ncu_l.log 2>&1
for k in encrypt_fused weighted_sum decrypt_intt decode_kernel ntt_forward; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -o gpurun_out/full_$k -f $B > gpurun_out/ncu_$k.log 2>&1
done
ls -la gpurun_out | tail -20
tail -n 5 gpurun_out/bench.err
