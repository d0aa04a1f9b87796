# the round-end checks in one call: gpu tests (the no-alloc test alone first, then the whole suite), smoke, default bench, reference arm
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/v_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/v_noalloc_alone.log
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/v_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/v_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/v_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/v_smoke.log
timeout 600 python bench.py > gpurun_out/v_bench.json 2> gpurun_out/v_bench.err; echo "rc=$?" >> gpurun_out/v_bench.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/v_bench_ref.json 2> gpurun_out/v_bench_ref.err
tail -n 3 gpurun_out/v_noalloc_alone.log gpurun_out/v_gpu.log gpurun_out/v_smoke.log gpurun_out/v_bench.err
cut -c1-330 gpurun_out/v_bench.json; cut -c1-200 gpurun_out/v_bench_ref.json
