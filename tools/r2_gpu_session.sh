#!/bin/bash
# tools/r2_gpu_session.sh — one GPU-box session: tests, both bench arms, launch list, ncu --set full of the three hot kernels,
# compute-sanitizer of smoke(). Everything lands in gpurun_out/ (scratch); summaries are copied to profiles/ on the CPU box.
O=gpurun_out; mkdir -p $O
NCU="ncu --set full --clock-control none --import-source on"
if [ "$1" != "notests" ] && [ "$1" != "hnswonly" ]; then timeout 1500 python -m pytest tests -m gpu -x -q > $O/r2_pytest_gpu.log 2>&1; tail -3 $O/r2_pytest_gpu.log; fi
timeout 900 python bench.py > $O/r2_bench_n1.json 2> $O/r2_bench_n1.err; echo "bench rc=$?"; tail -c 600 $O/r2_bench_n1.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/r2_bench_ref.json 2> $O/r2_bench_ref.err; echo "ref rc=$?"
# launch list of the bench steps: the graph is built on the host here (VDB_HOST_GRAPH_BUILD=1) so that the first 400 launches are
# not all graph-construction kernels
VDB_HOST_GRAPH_BUILD=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_bench_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --adc-rows 0 --flat-rows 0 --shard-rows 0 > $O/ncu_launch.log 2>&1; echo "launch list rc=$?"
python tools/prof.py hnsw --nq 592 --snapshot /tmp/c3.snap > $O/r2_prof_hnsw.json 2>&1
timeout 900 $NCU -k regex:hnsw_search_kernel -c 2 -f -o $O/r2_hnsw python tools/prof.py hnsw --nq 592 --snapshot /tmp/c3.snap > $O/ncu_hnsw.log 2>&1; echo "ncu hnsw rc=$?"
[ "$1" = "hnswonly" ] || timeout 900 $NCU -k regex:adc_scan_fast_kernel -c 2 -f -o $O/r2_adc python tools/prof.py adc --nq 296 > $O/ncu_adc.log 2>&1; echo "ncu adc rc=$?"
[ "$1" = "hnswonly" ] || timeout 900 $NCU -k regex:flat_gemm_kernel -c 4 -f -o $O/r2_flat python tools/prof.py flat --n 1000000 --dim 768 --nq 4096 --k 100 > $O/ncu_flat.log 2>&1; echo "ncu flat rc=$?"
# compute-sanitizer is closed on this pool
ls -la $O | head -40
