#!/bin/bash
# Copy the measurements of the last tools/gpu_round.sh pass from gpurun_out/ (scratch) into profiles/ (tracked).
set -u
R=${1:-r02}
cd "$(dirname "$0")/.."
mkdir -p profiles
for f in bench.json bench_ref.json bench_s20.json dcn_comparators.jsonl fullmodel_parity.jsonl exp_precision.jsonl time_ops.log pytest_gpu.log; do
  [ -f gpurun_out/$f ] && cp gpurun_out/$f profiles/${R}_$f
done
for n in 2 4 8; do for v in "" _s20 _nccl _all; do f=gpurun_out/bench_n${n}${v}.json; [ -f $f ] && grep '^{' $f > profiles/${R}_bench_n${n}${v}.json; done; [ -f gpurun_out/check_gather_n$n.log ] && grep "gather to" gpurun_out/check_gather_n$n.log > profiles/${R}_check_gather_n$n.txt; done
[ -f gpurun_out/launches.csv ] && cp gpurun_out/launches.csv profiles/${R}_launches_bench.csv
python - "$R" <<'PY'
import csv, json, sys, collections, subprocess, io
R = sys.argv[1]
# kernel shares from the ncu launch list (serialised, cold-cache times: the SHARE is what carries over)
try:
    rows = [r for r in csv.reader(open("gpurun_out/launches.csv")) if len(r) > 5]
    hdr = next(r for r in rows if "Kernel Name" in r)
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    tot = collections.Counter(); cnt = collections.Counter()
    for r in rows:
        if r is hdr or len(r) <= vi: continue
        try: v = float(r[vi].replace(",", ""))
        except ValueError: continue
        name = r[ki].split("(")[0]
        tot[name] += v; cnt[name] += 1
    s = sum(tot.values())
    with open("profiles/%s_launch_shares.txt" % R, "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none on `python bench.py --steps 20 --warmup 3 --no-configs --no-cpu-baseline`\n")
        f.write("# kernel | launches | total ns | share of all kernel time in the process\n")
        for k, v in tot.most_common(12):
            f.write("%-60s %5d %12.0f %6.3f\n" % (k[:60], cnt[k], v, v / s))
    print(open("profiles/%s_launch_shares.txt" % R).read())
except Exception as e:
    print("launch list:", e)
try:
    out = subprocess.run([sys.executable, "tools/ncu_summary.py", "gpurun_out/prof_cv.ncu-rep"], capture_output=True, text=True).stdout
    j = json.loads(out)
    open("profiles/%s_ncu_full_corr_volume_tc.json" % R, "w").write(json.dumps(j, indent=1))
    def b(k):
        v, u = j[k], j.get(k + "__unit", "byte")
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    t = b("dram__bytes_read.sum") + b("dram__bytes_write.sum")
    json.dump({"corr_volume_bytes_per_launch": int(t), "source": "ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum of one corr_volume_tc_kernel launch at the bench workload (profiles/%s_ncu_full_corr_volume_tc.json)" % R,
               "algorithmic_bytes_per_launch": 139960320}, open("profiles/traffic.json", "w"), indent=1)
    print("traffic", t)
except Exception as e:
    print("ncu full:", e)
PY
ls profiles | grep "^$R" | head -40
