#!/bin/bash
# final check of the round at HEAD: no-alloc test alone, whole GPU suite, smoke, one default bench line.
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -2 ) | tee gpurun_out/final_pytest_noalloc_alone.log
( timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 ) | tee gpurun_out/final_pytest.log
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 ) | tee gpurun_out/final_smoke.log
timeout 900 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; tail -c 400 gpurun_out/final_bench.json; tail -2 gpurun_out/final_bench.err
