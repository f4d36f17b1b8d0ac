python -m pytest tests -x -q -m gpu > gpurun_out/r2_gputest.log 2>&1; tail -2 gpurun_out/r2_gputest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; tail -c 600 gpurun_out/r2_bench_n1.json
python bench.py --impl reference > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err; tail -c 400 gpurun_out/r2_bench_ref.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2_launches_bench.csv python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; tail -1 gpurun_out/ncu_bench.log | cut -c1-200
ncu --set full --clock-control none --import-source on -k regex:'k_fwd_loss_g8|k_bwd_update|k_update_rows' -s 30 -c 3 -o gpurun_out/r2_train python bench.py --steps 10 --warmup 5 --no-eval --no-cpu-baseline --no-strong > gpurun_out/ncu_train.log 2>&1; tail -1 gpurun_out/ncu_train.log | cut -c1-200
ncu --set full --clock-control none --import-source on -k regex:'k_tc_gemm_topk|k_tc_finish' -s 4 -c 2 -o gpurun_out/r2_tc_final python profiles/eval_tc_diag.py > gpurun_out/ncu_tc.log 2>&1; tail -1 gpurun_out/ncu_tc.log | cut -c1-200
