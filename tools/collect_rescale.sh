# ncu evidence for the rescale kernels and a fresh launch list of the bench command (run under gpurun from the repo root)
set -x
R=/tmp/reps; mkdir -p $R
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
B="python bench.py --steps 2 --warmup 3 --no-cpu"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r01_launches.csv $B > gpurun_out/ncu_l.log 2>&1
echo "# ncu --set full, one launch of each rescale kernel inside \`$B\` (tools/collect_rescale.sh)" > gpurun_out/r01_ncu_rescale.md
echo >> gpurun_out/r01_ncu_rescale.md
first=1
for k in rescale_intt rescale_ntt; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -o $R/full_$k -f $B > gpurun_out/ncu_$k.log 2>&1
  if [ $first = 1 ]; then python tools/ncu_summary.py $R/full_$k.ncu-rep >> gpurun_out/r01_ncu_rescale.md; first=0
  else python tools/ncu_summary.py $R/full_$k.ncu-rep | tail -n +3 >> gpurun_out/r01_ncu_rescale.md; fi
done
cat gpurun_out/r01_ncu_rescale.md
grep -c rescale gpurun_out/r01_launches.csv
