set -u
python -m pytest tests -m gpu -q -x --timeout 1500 > gpurun_out/v_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/v_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/v_smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 20 --warmup 5 > gpurun_out/v_bench.json 2> gpurun_out/v_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/v_bench.json'))
print(d['value'], d['ms_per_step'], d['e2e'], d['newton_iterations'][:4], d['krylov_iterations'], d['checksums'], d['same_size']['parity_rel_err'], d['same_size']['ratio'])
print({k:round(v['ms_per_step'],2) for k,v in d['breakdown'].items()})
print(d['roofline'])
PY
bash tools/profile_r2.sh r2f
