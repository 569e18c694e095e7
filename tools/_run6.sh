set -u
python -m pytest tests/test_gpu_solvers.py -m gpu -q -x --timeout 1500 > gpurun_out/t6.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t6.log
for proj in 8 10 16; do
python - <<PY
import sys, json, time, torch
sys.path.insert(0,'.')
import bench
solver, ts = bench.make_cavity_solver(73, pressure_rhs_projection=$proj)
its=[]
for s in range(16):
    if s == 6:
        torch.cuda.synchronize(); t0=time.time()
    ts.update_coefficients(); solver.solve(); its.append(solver.krylov_iterations[-1]["pressure"]); ts.advance_time(); solver.advance_time()
torch.cuda.synchronize()
print(json.dumps(dict(projection=$proj, ms_per_step=1e3*(time.time()-t0)/10, pressure_iterations=its, checksums=solver.checksums())))
PY
done
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b6.json 2> gpurun_out/b6.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/b6.json'))
print(d['value'], d['ms_per_step'], d['newton_iterations'], d['krylov_iterations'], d['checksums'])
print({k:round(v['ms_per_step'],2) for k,v in d['breakdown'].items()})
PY
