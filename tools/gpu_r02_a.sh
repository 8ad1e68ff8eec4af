#!/bin/bash
# Round 2, GPU call A: the whole GPU tier, the headline bench, the ncu captures the roofline block reads, the other configs.
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv,noheader > $O/a_gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q --durations=15 > $O/a_gpu_tests.log 2>&1; echo "gpu tests exit $?" | tee -a $O/a_gpu_tests.log
tail -25 $O/a_gpu_tests.log
timeout 400 python bench.py > $O/a_bench_config3.json 2> $O/a_bench_config3.err; echo "bench config3 exit $?"
cut -c1-600 $O/a_bench_config3.json
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum
timeout 400 ncu --metrics $M --clock-control none -k regex:'k_trace_(any|closest)' --launch-skip 4 --launch-count 4 --csv --log-file $O/a_traffic_config3.csv \
    python bench.py --workload config3 --steps 1 --warmup 3 --no-cpu-baseline > $O/a_traffic_run.log 2>&1; echo "ncu traffic exit $?"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/a_launches_config3.csv \
    python bench.py --workload config3 --steps 2 --warmup 3 --no-cpu-baseline > $O/a_launches_run.log 2>&1; echo "ncu launches exit $?"
for w in config1 config2 config4; do
  timeout 300 python bench.py --workload $w --steps 10 --warmup 3 > $O/a_bench_$w.json 2> $O/a_bench_$w.err; echo "bench $w exit $?"
done
timeout 400 python bench.py --workload config5 --steps 5 --warmup 3 --no-cpu-baseline > $O/a_bench_config5.json 2> $O/a_bench_config5.err; echo "bench config5 exit $?"
timeout 300 python tools/build_bench.py --flags 0 --sizes 262144,2000000,10000000 > $O/a_build_ploc.jsonl 2>&1; echo "build ploc exit $?"
timeout 300 python tools/build_bench.py --flags 1 > $O/a_build_lbvh.jsonl 2>&1; echo "build lbvh exit $?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/a_bench_reference.json 2> $O/a_bench_reference.err; echo "reference arm exit $?"
ls -la $O | head -40
