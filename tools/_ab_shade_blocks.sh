# A/B of k_shade's resident CTAs per SM (PLU_SHADE_BLOCKS build option): libplugpu_shade<k>.so built beside the default library
mkdir -p gpurun_out
for L in "" shade2 "" shade2; do
  PLU_LIB=$L timeout 200 python bench.py --config cornell-box,room,glass0 --no-cpu-baseline --spp-per-step 256 --steps 6 --warmup 3 2>>gpurun_out/ab.err | sed "s/^{/{\"lib\": \"$L\", /" >> gpurun_out/ab_shade_blocks2.jsonl
done
