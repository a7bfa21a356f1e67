#!/usr/bin/env bash
# ncu --set full capture of ONE WAVE (2^22 paths) per configuration, after the same command has run without ncu; the raw
# page is exported as CSV on the box (the .ncu-rep files are too large to bring back), tools/ncu_traffic.py reads it.
#   bash tools/ncu_capture.sh <tag> [config ...]    -> gpurun_out/<tag>_<config>_raw.csv (+ _source_<kernel>.csv for the hot kernels)
set -u
tag="${1:-r02}"; shift
configs="${*:-cornell-box synth-boxes synth-mixed synth-spheres}"
out=gpurun_out
export PLU_NO_GRAPH=1   # plain launches: one ncu result per kernel launch
for c in $configs; do
  spp=4
  python tools/one_render.py "$c" --spp $spp > "$out/${tag}_${c}_plain.log" 2>&1 || { echo "plain run of $c failed"; continue; }
  ncu --set full --clock-control none --import-source on -k 'regex:k_extend|k_shade|k_connect' -c 40 -o "/tmp/${tag}_$c" -f python tools/one_render.py "$c" --spp $spp > "$out/${tag}_${c}_ncu.log" 2>&1
  ncu -i "/tmp/${tag}_$c.ncu-rep" --page raw --csv > "$out/${tag}_${c}_raw.csv" 2>/dev/null
  ncu -i "/tmp/${tag}_$c.ncu-rep" --page source --csv --print-source sass > "/tmp/${tag}_${c}_source.csv" 2>/dev/null
  gzip -c "/tmp/${tag}_${c}_source.csv" > "$out/${tag}_${c}_source.csv.gz"
  tail -1 "$out/${tag}_${c}_ncu.log"; ls -la "$out/${tag}_${c}_raw.csv" "$out/${tag}_${c}_source.csv.gz"
done
