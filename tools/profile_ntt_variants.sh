# ncu counters of the two forward-NTT forms on a 59-bit single-limb context (integer pipe): run under gpurun.
# tools/ntt_probe.py launches, per modulus size (52 then 59 bits) and per form, 1 + 3 + 10 transforms.
set -x
R=/tmp/reps; mkdir -p $R
B="python tools/ntt_probe.py"
$B > gpurun_out/r02_ntt_variants.jsonl 2> gpurun_out/ntt_variants.err &&
ncu --set full --clock-control none -k regex:ntt_forward -s 30 -c 1 -o $R/ntt_smem -f $B > gpurun_out/ncu_nttvar.log 2>&1
ncu --set full --clock-control none -k regex:ntt_forward -s 44 -c 1 -o $R/ntt_shfl -f $B >> gpurun_out/ncu_nttvar.log 2>&1
python tools/ncu_summary.py $R/ntt_smem.ncu-rep > gpurun_out/r02_ntt_variants_ncu.md
python tools/ncu_summary.py $R/ntt_shfl.ncu-rep | tail -n +3 >> gpurun_out/r02_ntt_variants_ncu.md
cat gpurun_out/r02_ntt_variants_ncu.md
