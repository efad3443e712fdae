set -x
python bench.py --steps 20 --warmup 3 > gpurun_out/r1i_bench_n1.json 2> gpurun_out/r1i_bench_n1.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1i_bench_reference_n1.json 2> gpurun_out/r1i_bench_reference_n1.err
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1i_launches_bench_steps3.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r1i_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_nn_single_pruned -s 5 -c 1 -o gpurun_out/r1i_prof_nn python bench.py --steps 3 --warmup 3 --no-cpu --insert-scans 0 > gpurun_out/r1i_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_insert_rank -s 2 -c 2 -o gpurun_out/r1i_prof_rank python bench.py --only insert_batched > gpurun_out/r1i_ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_insert_commit -s 2 -c 1 -o gpurun_out/r1i_prof_commit python bench.py --only insert_batched > gpurun_out/r1i_ncu4.log 2>&1
cut -c1-300 gpurun_out/r1i_bench_n1.json
