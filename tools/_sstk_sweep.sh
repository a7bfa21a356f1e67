# A/B sweep of the compile-time options of the fast-BVH walk (run on the GPU box; rebuilds libplugpu.so per variant):
#   PLU_SSTK n      lowest n stack levels in shared memory      PLU_ANYHIT 1   point-light shadow rays stop at the first hit
#   PLU_WIDE_WALK 1 four-child walk (active with PLU_WIDE=1 in the environment at upload)
OUT=gpurun_out/s23_sweep2.log
rm -f $OUT
run() {   # sstk anyhit wide
  PLU_SSTK=$1 PLU_ANYHIT=$2 PLU_WIDE_WALK=$3 python -m plutracer_b200._build > /dev/null 2>&1 || echo "build fail $*" >> $OUT
  for K in boxes mixed; do
    echo "SSTK=$1 ANYHIT=$2 WIDE=$3 kind=$K $(PLU_WIDE=$3 python tools/bench_synth.py --kind $K --steps 3 2>&1 | tail -1)" >> $OUT
  done
}
run 4 1 0
run 12 1 0
run 16 1 0
run 8 1 0
# parity of the last variant (still built), then back to the default build
python -m pytest tests -m gpu -q -x 2>&1 | tail -3 >> $OUT
python -m plutracer_b200._build > /dev/null 2>&1
python - <<'PY'
import json
for l in open("gpurun_out/s23_sweep2.log"):
    if "{" not in l: print(l.strip()); continue
    h, j = l.split("{", 1)
    try: d = json.loads("{" + j)
    except Exception: print(l.strip()[:300]); continue
    print(h, round(d["msamples_per_s"], 1), "Msamples/s", {k: round(v, 1) for k, v in d["ms_by_kernel_counted_pass"].items()}, round(d["nodes_per_ray"], 2), round(d["prims_per_ray"], 2))
PY
