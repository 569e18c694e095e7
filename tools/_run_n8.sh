set -u
N=8
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/n${N}_bench.out 2> gpurun_out/n${N}_bench.err; echo "bench N=$N rc=$?"
grep '^{' gpurun_out/n${N}_bench.out | tail -1 > gpurun_out/n${N}_bench.json
python - <<PY
import json
d=json.load(open('gpurun_out/n${N}_bench.json'))
print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['krylov_iterations'], d['checksums'], d['gpu_launches'])
print({k:round(v['ms_per_step'],2) for k,v in d['breakdown'].items()})
PY
