# A/B sweep of two compile-time options of the fast-BVH walk (run on the GPU box; rebuilds libplugpu.so per variant)
OUT=gpurun_out/s23_sstk.log
for V in "0 0" "8 0" "16 0" "0 1" "16 1"; do
  set -- $V
  PLU_SSTK=$1 PLU_ANYHIT=$2 python -m plutracer_b200._build > /dev/null 2>&1 || echo "build fail $V" >> $OUT
  for K in boxes mixed; do
    echo "SSTK=$1 ANYHIT=$2 kind=$K $(python tools/bench_synth.py --kind $K --steps 3 2>&1 | tail -1)" >> $OUT
  done
done
# parity of the most aggressive variant, then back to the default build
PLU_SSTK=16 PLU_ANYHIT=1 python -m plutracer_b200._build > /dev/null 2>&1
python -m pytest tests -m gpu -q -x 2>&1 | tail -3 >> $OUT
python -m plutracer_b200._build > /dev/null 2>&1
python - <<'PY'
import json
for l in open("gpurun_out/s23_sstk.log"):
    if "{" not in l: print(l.strip()); continue
    h, j = l.split("{", 1); d = json.loads("{" + j)
    print(h, round(d["msamples_per_s"], 1), "Msamples/s", d["ms_by_kernel_counted_pass"], round(d["nodes_per_ray"], 2), round(d["prims_per_ray"], 2))
PY
