# Round evidence, collected on one B200 (run under gpurun from the repo root).  ncu only after the same command has
# exited 0 without it; never under torchrun.  The .ncu-rep files stay on the box (gpurun_out/ is capped at 64 MiB):
# only their summaries come back.
set -x
R=/tmp/reps; mkdir -p $R
python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/r01_pytest_gpu.log
python bench.py > gpurun_out/r01_bench.json 2> gpurun_out/bench.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r01_bench_reference_arm.json 2>> gpurun_out/bench.err
python tools/bench_configs.py --configs 3,4,5,6 > gpurun_out/r01_bench_configs.jsonl 2>> gpurun_out/bench.err
python tools/bench_selective.py > gpurun_out/r01_bench_selective.jsonl 2>> gpurun_out/bench.err
python tools/bench_layers.py > gpurun_out/r01_bench_layers.jsonl 2>> gpurun_out/bench.err
python tools/e2e_phases.py > gpurun_out/r01_e2e_phases.json 2>> gpurun_out/bench.err
# (two or more GPUs: torchrun ... tools/peer_gather_check.py and bench.py --gpus N feed r01_peer_gather.jsonl / r01_scaling.jsonl)
B="python bench.py --steps 2 --warmup 3 --no-cpu"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01_launches.csv $B > gpurun_out/ncu_l.log 2>&1
echo "# ncu --set full, one launch per kernel of \`$B\` (tools/collect_profiles.sh), current build" > gpurun_out/r01_ncu_step.md
echo >> gpurun_out/r01_ncu_step.md
first=1
for k in encrypt_fused weighted_sum decrypt_intt decode_kernel ntt_forward; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -o $R/full_$k -f $B > gpurun_out/ncu_$k.log 2>&1
  if [ $first = 1 ]; then python tools/ncu_summary.py $R/full_$k.ncu-rep >> gpurun_out/r01_ncu_step.md; first=0
  else python tools/ncu_summary.py $R/full_$k.ncu-rep | tail -n +3 >> gpurun_out/r01_ncu_step.md; fi
  ncu -i $R/full_$k.ncu-rep --page raw --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum > gpurun_out/traffic_$k.csv 2>/dev/null
done
ncu -i $R/full_encrypt_fused.ncu-rep --page source --csv 2>/dev/null | python tools/ncu_stalls.py 12 > gpurun_out/r01_encrypt_stalls.txt
cat gpurun_out/r01_ncu_step.md
tail -n 5 gpurun_out/bench.err
