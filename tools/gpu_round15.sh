#!/bin/bash
# the no-allocation entry point alone in a fresh process, then the whole GPU suite (-x, as the round's end runs it), smoke, bench
tag=${1:-r02u}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/${tag}_noalloc_alone.log 2>&1; tail -15 gpurun_out/${tag}_noalloc_alone.log | cut -c1-300
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/${tag}_gputests.log 2>&1; tail -15 gpurun_out/${tag}_gputests.log | cut -c1-300
timeout 600 python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1; tail -2 gpurun_out/${tag}_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -c 3000 gpurun_out/${tag}_bench.json; tail -5 gpurun_out/${tag}_bench.err
