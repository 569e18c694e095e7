set -u
python -m pytest tests/test_partition.py -m gpu -q -x --timeout 1500 > gpurun_out/n2_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/n2_tests.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/n2_bench.out 2> gpurun_out/n2_bench.err; echo "bench rc=$?"
grep '^{' gpurun_out/n2_bench.out | tail -1 > gpurun_out/n2_bench.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/n2_bench.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['krylov_iterations'], d['checksums'], d['gpu_launches'])
PY
