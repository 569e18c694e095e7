set -u
run() { # name, env...
  name=$1; shift
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 6 --warmup 5 --no-cpu-baseline > gpurun_out/n8x.out 2> gpurun_out/n8x_$name.err; echo "bench $name rc=$?"
  grep '^{' gpurun_out/n8x.out | tail -1 > gpurun_out/n8x_$name.json
  python - <<PY
import json
d=json.load(open('gpurun_out/n8x_$name.json'))
print('$name', d['value'], d['ms_per_step'], d['krylov_iterations'], d['checksums'], d['config']['parallelism'][:40])
PY
}
run nccl NSB_HALO=nccl
run f32gather NSB_PC_GATHER=f32
