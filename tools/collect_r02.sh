# Round-2 evidence, collected on one B200 (run under gpurun from the repo root).  ncu only after the same command has
# exited 0 without it; never under torchrun.  The .ncu-rep files stay on the box; their summaries come back.
set -x
python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/r02_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/bench.err
python bench.py > gpurun_out/r02_bench.json 2>> gpurun_out/bench.err
(python tools/step_probe.py --steps 10; python tools/step_probe.py --steps 10 --per-learner) > gpurun_out/r02_step_probe.json 2>> gpurun_out/bench.err
python tools/latency_probe2.py > gpurun_out/r02_latency.jsonl 2>> gpurun_out/bench.err
(for a in "" "--pageable" "--f64" "--pageable --f64"; do python tools/e2e_probe.py $a; done) > gpurun_out/r02_e2e_probe.jsonl 2>> gpurun_out/bench.err
SHELFI_TIMELINE=1 python tools/e2e_probe.py --steps 5 2> gpurun_out/r02_e2e_timeline.txt > /dev/null
python tools/pcie_probe.py > gpurun_out/r02_pcie_probe.json 2>> gpurun_out/bench.err
(python tools/bench_configs.py --configs 3,4 --shards 8; python tools/bench_configs.py --configs 3 --shards 1) > gpurun_out/r02_bench_configs.jsonl 2>> gpurun_out/bench.err
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-strong"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench.csv $B > gpurun_out/ncu_lb.log 2>&1
bash tools/profile_r02.sh r02 > gpurun_out/profile_r02.log 2>&1
python - <<'PY'
import csv, glob, json
out = {"_source": "ncu --set full --clock-control none, one launch per kernel of `python tools/step_probe.py --steps 2` (tools/profile_r02.sh, tools/collect_r02.sh); dram__bytes_read.sum / dram__bytes_write.sum / gpu__time_duration.sum; final round-2 build"}
mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
tm = {"ns": 1e-3, "us": 1, "ms": 1e3, "usecond": 1, "msecond": 1e3, "nsecond": 1e-3, "second": 1e6}
for f in sorted(glob.glob("gpurun_out/traffic_*.csv")):
    rows = list(csv.reader(open(f)))
    if len(rows) < 3:
        continue
    h, u, r = rows[0], rows[1], rows[2]
    i = {n: k for k, n in enumerate(h)}
    name = r[i["Kernel Name"]].split("(")[0].replace("void ", "").split("<")[0]
    out[name] = {"dram_read_bytes": float(r[i["dram__bytes_read.sum"]]) * mult[u[i["dram__bytes_read.sum"]]],
                 "dram_write_bytes": float(r[i["dram__bytes_write.sum"]]) * mult[u[i["dram__bytes_write.sum"]]],
                 "duration_us": float(r[i["gpu__time_duration.sum"]]) * tm[u[i["gpu__time_duration.sum"]]]}
json.dump(out, open("gpurun_out/r02_traffic.json", "w"), indent=1)
PY
tail -3 gpurun_out/r02_pytest_gpu.log; cat gpurun_out/r02_smoke.log | tail -2; tail -n 5 gpurun_out/bench.err
