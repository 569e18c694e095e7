set -u
for E in 0 1; do
python bench.py --steps 10 --warmup 5 --no-cpu-baseline --solver-option extrapolate_correction_guess=$E > gpurun_out/e_$E.json 2> gpurun_out/e.err; echo "rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/e_$E.json'))
print('$E', d['value'], d['ms_per_step'], d['krylov_iterations'], d['checksums'], d['breakdown']['spmv_csr']['ms_per_step'])
PY
done
python -m pytest tests/test_gpu_solvers.py tests/test_gpu_scale.py tests/test_gpu_configs.py tests/test_gpu_compat.py -m gpu -q -x 2>&1 | tail -3
