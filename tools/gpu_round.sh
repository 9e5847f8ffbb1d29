#!/bin/bash
# One GPU-box pass at the state to be judged: parity tests, smoke, bench (all configs), reference arm, ncu launch list
# and the full ncu capture of the top kernel.  Outputs in gpurun_out/, copied to profiles/ by tools/collect_profiles.sh.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log; grep -E "passed|failed" gpurun_out/pytest_gpu.log | tail -2
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
python bench.py --steps 20 --warmup 3 --no-configs > gpurun_out/bench_s20.json 2>> gpurun_out/bench.err; echo "bench s20 rc=$?"
python tools/time_dcn_ref.py > gpurun_out/dcn_comparators.jsonl 2>/dev/null; echo "dcn rc=$?"
python tools/time_ops.py > gpurun_out/time_ops.log 2>&1; echo "time_ops rc=$?"
CMD="python bench.py --steps 20 --warmup 3 --no-configs --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:corr_volume_tc -s 5 -c 1 -f -o gpurun_out/prof_cv $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
python - <<'PY'
import json
j = json.load(open("gpurun_out/bench.json"))
c = j.pop("configs")
print(json.dumps({k: j[k] for k in ("value", "ms_per_step", "roofline", "parity", "e2e", "cpu_baseline", "gpu_eager_reference", "clocks")})[:1800])
for k, v in c.items():
    print(k, {a: (round(b, 3) if isinstance(b, float) else b) for a, b in v.items() if a in ("us", "frac", "parity_max_err", "parity_epe", "reference_extension_us", "torchvision_us", "gpu_eager_reference_us", "error", "pairs_per_s")})
PY
