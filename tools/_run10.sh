set -u
python bench.py --steps 6 --warmup 5 --no-cpu-baseline > gpurun_out/n1d_default.json 2> gpurun_out/n1d.err; echo "rc=$?"
python bench.py --steps 6 --warmup 5 --no-cpu-baseline --solver-option linear_rtol=1e-13 --solver-option correction_rtol=1e-14 > gpurun_out/n1d_tight.json 2> gpurun_out/n1d.err; echo "rc=$?"
python - <<PY
import json
for f in ('default','tight'):
    d=json.load(open('gpurun_out/n1d_%s.json'%f))
    print(f, d['value'], d['ms_per_step'], d['krylov_iterations'], d['checksums'])
PY
python -m pytest tests/test_gpu_scale.py tests/test_gpu_solvers.py -m gpu -q -x 2>&1 | tail -3
