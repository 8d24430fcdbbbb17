#!/bin/bash
# End-of-round validation on the GPU box: the full GPU suite, smoke(), every bench line, and the ncu launch list of the
# bench command.  Everything lands in gpurun_out/final_*.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/final_gpu_tests.txt
cat gpurun_out/final_gpu_tests.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final_smoke.txt 2>&1
tail -1 gpurun_out/final_smoke.txt
timeout 500 python bench.py > gpurun_out/final_bench_default.json 2> gpurun_out/final_bench_default.err
rc=$?
echo "bench default rc=$rc"
timeout 400 python bench.py --workload config4 --no-eager > gpurun_out/final_bench_config4.json 2> gpurun_out/final_bench_config4.err
echo "bench config4 rc=$?"
timeout 500 python bench.py --workload config5 --no-eager > gpurun_out/final_bench_config5.json 2> gpurun_out/final_bench_config5.err
echo "bench config5 rc=$?"
timeout 500 python bench.py --impl reference > gpurun_out/final_bench_reference.json 2> gpurun_out/final_bench_reference.err
echo "bench reference rc=$?"
if [ "$rc" = "0" ]; then
  timeout 600 python bench.py --steps 2 --warmup 1 --no-eager > gpurun_out/final_bench_s2.json 2>/dev/null &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/final_launches.csv \
      python bench.py --steps 2 --warmup 1 --no-eager > gpurun_out/final_ncu_bench.log 2>&1
  echo "launch list rc=$? lines=$(wc -l < gpurun_out/final_launches.csv)"
fi
python - <<'PY'
import json
for n in ("default", "config4", "config5", "reference"):
    try:
        d = json.loads(open("gpurun_out/final_bench_%s.json" % n).read().strip().splitlines()[-1])
        print(n, d.get("value"), d.get("unit"), "ms/step", d.get("ms_per_step"), "e2e", (d.get("e2e") or {}).get("value"),
              "frac", (d.get("roofline") or {}).get("frac"))
    except Exception as e:
        print(n, "unreadable:", e)
PY
