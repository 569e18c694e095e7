#!/bin/bash
# Round-2 profiling recipe (run on the GPU box through gpurun): launch list of a short bench run, then one
# `ncu --set full` capture per top kernel.  Numbers printed by a run under ncu are never bench values.
set -u
mkdir -p gpurun_out
P=${1:-r2f}
B="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$B > gpurun_out/${P}_plain.json 2> gpurun_out/${P}_plain.err || { echo "plain run failed"; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 3200 --csv --log-file gpurun_out/${P}_launches.csv $B > gpurun_out/${P}_ncu_list.log 2>&1
python tools/summarise_launches.py gpurun_out/${P}_launches.csv > gpurun_out/${P}_launches_summary.md
B1="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
# name:kernel regex:launches to skip.  '^k_spmv$' is the FP64 CSR / Kronecker kernel alone (k_spmv_kron32, k_spmv_bsr_tma have their own lines)
for spec in kron32:k_spmv_kron32:40 pcg:k_pcg_persistent:1 jac:k_jacobian_rowgroups:1 bsr:k_spmv_bsr_tma:10 res:k_residual_cells:2 'masscsr:^k_spmv$:6'; do
  name=${spec%%:*}; rest=${spec#*:}; pat=${rest%%:*}; skip=${rest##*:}
  ncu --set full --clock-control none --import-source on -k "regex:$pat" -s $skip -c 1 -f -o gpurun_out/${P}_$name $B1 > gpurun_out/${P}_ncu_$name.log 2>&1
done
ls -la gpurun_out/${P}_*.ncu-rep
