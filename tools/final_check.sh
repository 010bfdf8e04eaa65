#!/bin/bash
# Round-end check on a GPU box: tests, smoke, both bench arms, secondary workloads, ncu launch list + full capture.
P=${1:-r2}
python -m pytest tests -m gpu -x -q > gpurun_out/${P}_tests.log 2>&1; echo tests rc=$?; tail -2 gpurun_out/${P}_tests.log
python __graft_entry__.py smoke 2>&1 | tail -1
python bench.py > gpurun_out/${P}_bench.json 2> gpurun_out/${P}_bench.err; echo bench rc=$?
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${P}_ref.json 2>/dev/null
python bench.py --no-cpu-baseline --workload c2 > gpurun_out/${P}_c2.json 2>/dev/null
python bench.py --no-cpu-baseline --workload c1 > gpurun_out/${P}_c1.json 2>/dev/null
python bench.py --no-cpu-baseline --workload c4 --streams 8 --steps 3 > gpurun_out/${P}_c4x8.json 2>/dev/null
python tools/gpu_probe.py 4096 16 | tail -2
python bench.py --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${P}_bench_launches.csv python bench.py --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/ncu1.log 2>&1; echo ncu1 rc=$?
ncu --set full --clock-control none --import-source on -k regex:track_multi_kernel -s 9 -c 1 -o gpurun_out/${P}_track_multi python bench.py --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/ncu2.log 2>&1; echo ncu2 rc=$?
python - <<PY
import json
for f in ["bench","ref","c1","c2","c4x8"]:
    d=json.load(open("gpurun_out/${P}_%s.json"%f)); r=d.get("roofline",{})
    print(f,"value %.4gM e2e %.4gM ms/step %.2f kernel %s frac %s ev/s %.4gM"%(d["value"]/1e6,d["e2e"]["value"]/1e6,d["ms_per_step"],r.get("kernel_ms_per_step"),r.get("frac"),d["events_per_s"]/1e6))
PY
