# ncu evidence for the shared-memory stack levels: the first wave's eight k_extend_pt launches, local-stack build against the shipped one
M=gpu__time_duration.sum,l1tex__t_sector_hit_rate.pct,smsp__inst_executed_op_local_ld.sum,smsp__inst_executed_op_local_st.sum,smsp__inst_executed_op_shared_ld.sum,smsp__inst_executed_op_shared_st.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,lts__t_sectors_op_write.sum,lts__t_sectors_op_read.sum
for V in "0 0 local" "8 1 shipped"; do
  set -- $V
  PLU_SSTK=$1 PLU_ANYHIT=$2 python -m plutracer_b200._build > /dev/null 2>&1
  ncu --metrics $M --clock-control none -k regex:k_extend_pt -c 8 --csv --log-file gpurun_out/s24_extpt_$3.csv python tools/bench_synth.py --kind boxes --steps 1 > gpurun_out/s24_extpt_$3.log 2>&1
done
python -m plutracer_b200._build > /dev/null 2>&1
tail -n +1 gpurun_out/s24_extpt_local.csv | head -40 | cut -c1-200
