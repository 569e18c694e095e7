#!/bin/bash
# round-2 verification on one GPU: GPU test suite, smoke, driver-like bench, reference arm
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x --timeout 1500 > gpurun_out/v_tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/v_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/v_smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 20 --warmup 5 > gpurun_out/v_bench.json 2> gpurun_out/v_bench.err; echo "bench rc=$?"
cat gpurun_out/v_bench.json
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/v_ref.json 2> gpurun_out/v_ref.err; echo "ref rc=$?"
cat gpurun_out/v_ref.json
