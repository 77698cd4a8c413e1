#!/bin/bash
# tools/r2_gpu_session.sh — one GPU-box session: tests, both bench arms, launch list, ncu --set full of the three hot kernels,
# compute-sanitizer of smoke(). Everything lands in gpurun_out/ (scratch); summaries are copied to profiles/ on the CPU box.
O=gpurun_out; mkdir -p $O
NCU="ncu --set full --clock-control none --import-source on"
if [ "$1" != "notests" ]; then timeout 1500 python -m pytest tests -m gpu -x -q > $O/r2_pytest_gpu.log 2>&1; tail -3 $O/r2_pytest_gpu.log; fi
timeout 900 python bench.py > $O/r2_bench_n1.json 2> $O/r2_bench_n1.err; echo "bench rc=$?"; tail -c 600 $O/r2_bench_n1.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/r2_bench_ref.json 2> $O/r2_bench_ref.err; echo "ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_bench_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --adc-rows 0 --flat-rows 0 --shard-rows 0 > $O/ncu_launch.log 2>&1; echo "launch list rc=$?"
python tools/prof.py hnsw --nq 592 --snapshot /tmp/c3.snap > $O/r2_prof_hnsw.json 2>&1
timeout 900 $NCU -k regex:hnsw_search_kernel -c 2 -f -o $O/r2_hnsw python tools/prof.py hnsw --nq 592 --snapshot /tmp/c3.snap > $O/ncu_hnsw.log 2>&1; echo "ncu hnsw rc=$?"
timeout 900 $NCU -k regex:adc_scan_fast_kernel -c 2 -f -o $O/r2_adc python tools/prof.py adc --nq 296 > $O/ncu_adc.log 2>&1; echo "ncu adc rc=$?"
timeout 900 $NCU -k regex:flat_gemm_kernel -c 4 -f -o $O/r2_flat python tools/prof.py flat --n 1000000 --dim 768 --nq 4096 --k 100 > $O/ncu_flat.log 2>&1; echo "ncu flat rc=$?"
timeout 900 compute-sanitizer --tool memcheck --log-file $O/r2_sanitizer_memcheck.log python -c "import __graft_entry__ as g; g.smoke()" > $O/san_mem.out 2>&1; echo "memcheck rc=$?"; tail -3 $O/r2_sanitizer_memcheck.log
timeout 900 compute-sanitizer --tool racecheck --log-file $O/r2_sanitizer_racecheck.log python -c "import __graft_entry__ as g; g.smoke()" > $O/san_race.out 2>&1; echo "racecheck rc=$?"; tail -3 $O/r2_sanitizer_racecheck.log
ls -la $O | head -40
