mkdir -p gpurun_out
( timeout 600 python -m pytest tests -q -m gpu -x ) > gpurun_out/final_pytest.log 2>&1; tail -2 gpurun_out/final_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 | tee gpurun_out/final_smoke.log
timeout 600 python bench.py > gpurun_out/final_bench_default.json 2> gpurun_out/final_bench_default.err; cut -c1-300 gpurun_out/final_bench_default.json
timeout 900 python bench.py --config all --no-cpu-baseline > gpurun_out/final_bench_all.jsonl 2> gpurun_out/final_bench_all.err
python - <<'PY'
import json
for l in open("gpurun_out/final_bench_all.jsonl"):
    d=json.loads(l); print(d["config"]["name"], round(d["value"],1), round(d["e2e"]["value"],1), round(d["roofline"]["frac"],3))
PY
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02c_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02c_launches_run.log 2>&1; echo "launch list rc=$?"
timeout 500 bash tools/ncu_capture.sh r02c cornell-box
