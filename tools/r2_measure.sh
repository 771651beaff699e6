set -x
python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/r2v_tests.log
python bench.py > gpurun_out/r2v_bench.json 2> gpurun_out/r2v_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2v_bench_ref.json 2>> gpurun_out/r2v_bench.err
python tools/slot_sweep.py "1,2,4" "decoder=fused,attention=fused/decoder=split,attention=fused/decoder=split,attention=split" > gpurun_out/r2v_sweep.log 2>&1
python tools/profile_step.py 64 100 "--slots=1 --decoder=fused" > gpurun_out/r2v_plain_fused.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r2v_launches_fused.csv python tools/profile_step.py 64 100 "--slots=1 --decoder=fused" > /dev/null 2>&1
python tools/profile_step.py 64 100 "--slots=1 --decoder=split" > gpurun_out/r2v_plain_split.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r2v_launches_split.csv python tools/profile_step.py 64 100 "--slots=1 --decoder=split" > /dev/null 2>&1
python tools/profile_step.py 64 12 "--slots=1 --decoder=fused" > gpurun_out/r2v_plain12.log 2>&1 && ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"attn_fused_kernel|vocab_tc_kernel|dec_cell_kernel|vocab_final_kernel|dec_out_kernel|attn_merge_kernel" -s 12 -c 6 -o gpurun_out/r2v_full python tools/profile_step.py 64 12 "--slots=1 --decoder=fused" > gpurun_out/r2v_ncu_full.log 2>&1
tail -3 gpurun_out/r2v_tests.log; tail -2 gpurun_out/r2v_bench.err; cat gpurun_out/r2v_sweep.log
