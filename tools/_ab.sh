mkdir -p gpurun_out
( timeout 600 python -m pytest tests -q -m gpu -x ) > gpurun_out/ab_pytest.log 2>&1; tail -2 gpurun_out/ab_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py --config cornell-box,room,glass0 --no-cpu-baseline --spp-per-step 256 --steps 6 --warmup 3 2>gpurun_out/ab.err > gpurun_out/ab_bench.jsonl
python - <<'PY'
import json
for l in open("gpurun_out/ab_bench.jsonl"):
    d=json.loads(l); print(d["config"]["name"], round(d["value"],1), round(d["e2e"]["value"],1))
PY
