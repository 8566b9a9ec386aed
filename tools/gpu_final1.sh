#!/bin/bash
# One-GPU end-of-round visit: the whole GPU suite, smoke(), both bench arms, every kbench case, the C++ operator tests.
TAG=${1:-r02}
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/${TAG}_pytest.log; tail -4 gpurun_out/${TAG}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/${TAG}_smoke.log
(timeout 600 tests/host/host_tests gpu 2>&1 | tail -3) | tee gpurun_out/${TAG}_host.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
timeout 900 python bench.py > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err
tail -c 1500 gpurun_out/${TAG}_bench_ref.json; echo; tail -c 3000 gpurun_out/${TAG}_bench_n1.json; tail -3 gpurun_out/${TAG}_bench_n1.err
timeout 600 python tools/kbench.py > gpurun_out/${TAG}_kbench.jsonl 2> gpurun_out/${TAG}_kbench.err
cut -c1-330 gpurun_out/${TAG}_kbench.jsonl; tail -3 gpurun_out/${TAG}_kbench.err
