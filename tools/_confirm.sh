mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -q -m gpu -x ) > gpurun_out/confirm_pytest.log 2>&1; tail -4 gpurun_out/confirm_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/confirm_smoke.log
timeout 600 python bench.py > gpurun_out/confirm_bench.json 2> gpurun_out/confirm_bench.err; cat gpurun_out/confirm_bench.json | cut -c1-1500
