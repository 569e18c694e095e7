set -u
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/v_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/v_smoke.log
timeout 600 python tools/run_configs.py c1 c3 c4 c2 --steps 3 > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"
cut -c1-900 gpurun_out/configs.jsonl
