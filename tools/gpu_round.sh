#!/bin/bash
# One GPU-box visit: GPU tests (no -x so every failure shows), smoke, default bench line.  Output under gpurun_out/<tag>_*.
tag=${1:-r02}
mkdir -p gpurun_out
free -g | head -2 > gpurun_out/${tag}_host.txt; nproc >> gpurun_out/${tag}_host.txt
python -m pytest tests -m gpu -q -s > gpurun_out/${tag}_gputests.log 2>&1
tail -25 gpurun_out/${tag}_gputests.log
python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1; tail -2 gpurun_out/${tag}_smoke.log
python bench.py --steps 10 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -c 3000 gpurun_out/${tag}_bench.json; tail -5 gpurun_out/${tag}_bench.err
