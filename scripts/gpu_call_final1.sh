#!/bin/bash
# Final single-GPU call: smoke, the whole GPU suite, the default bench line (all workloads, incumbent, CPU arms), the reference arm.
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 1500 python -m pytest tests -x -q -m gpu --timeout 600 2>&1 | tail -2
timeout 1800 python bench.py > gpurun_out/bench_final.log 2> gpurun_out/bench_final.err; echo "bench rc=$?" >> gpurun_out/bench_final.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_final_ref.log 2> gpurun_out/bench_final_ref.err; echo "ref rc=$?" >> gpurun_out/bench_final_ref.err
tail -c 300 gpurun_out/bench_final.err; tail -c 300 gpurun_out/bench_final_ref.err; cut -c1-300 gpurun_out/bench_final_ref.log
