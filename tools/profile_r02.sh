# Round-2 ncu evidence for the dominant kernel (run under gpurun from the repo root; one GPU).  The plain command runs
# first and must exit 0; ncu wraps the same command.  .ncu-rep files stay on the box, their summaries come back.
set -x
TAG=${1:-r02}
R=/tmp/reps; mkdir -p $R
B="python tools/step_probe.py --steps 2"
$B > gpurun_out/${TAG}_step_plain.json 2> gpurun_out/${TAG}_step_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${TAG}_launches.csv $B > gpurun_out/ncu_l.log 2>&1
echo "# ncu --set full --clock-control none, one launch per kernel of \`$B\` (tools/profile_r02.sh)" > gpurun_out/${TAG}_ncu_step.md
echo >> gpurun_out/${TAG}_ncu_step.md
first=1
for k in encrypt_fused weighted_sum decrypt_fused ntt_forward ntt_inverse; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -o $R/full_$k -f $B > gpurun_out/ncu_$k.log 2>&1
  if [ $first = 1 ]; then python tools/ncu_summary.py $R/full_$k.ncu-rep >> gpurun_out/${TAG}_ncu_step.md; first=0
  else python tools/ncu_summary.py $R/full_$k.ncu-rep | tail -n +3 >> gpurun_out/${TAG}_ncu_step.md; fi
  ncu -i $R/full_$k.ncu-rep --page raw --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum > gpurun_out/traffic_$k.csv 2>/dev/null
done
ncu -i $R/full_encrypt_fused.ncu-rep --page source --csv 2>/dev/null | python tools/ncu_stalls.py 14 > gpurun_out/${TAG}_encrypt_stalls.txt
python tools/ncu_opmix.py $R/full_ntt_forward.ncu-rep 2083840 > gpurun_out/${TAG}_ntt_opmix.txt 2>&1
python tools/ncu_opmix.py $R/full_encrypt_fused.ncu-rep 4070 > gpurun_out/${TAG}_encrypt_opmix.txt 2>&1
cat gpurun_out/${TAG}_ncu_step.md
