set -u
python bench.py --steps 20 --warmup 5 > gpurun_out/f_n1.json 2> gpurun_out/f_n1.err; echo "n1 rc=$?"
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/f_ref.json 2> gpurun_out/f_ref.err; echo "ref rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/f_n1.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['krylov_iterations'], d['checksums'], d['same_size']['parity_rel_err'], d['same_size']['ratio'], d['gpu_launches'])
print({k:round(v['ms_per_step'],2) for k,v in d['breakdown'].items()})
r=json.load(open('gpurun_out/f_ref.json')); print('ref', r['value'], r['ms_per_step'])
PY
