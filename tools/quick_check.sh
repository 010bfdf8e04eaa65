python -m pytest tests -m gpu -x -q 2>&1 | tail -2; python __graft_entry__.py smoke | tail -1; python bench.py > gpurun_out/quick_bench.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/quick_bench.json')); print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['cpu_baseline']['value'], d['parity']['ok'], d['gpu_launches'], d['clocks'])"
