#!/bin/bash
# ncu evidence for round 2 (run on the GPU box AFTER the same commands have exited 0 without ncu; B200_PROFILING.md):
#   1. launch list of a short default-workload pass (per-launch times: the kernels' SHARE of a step)
#   2. one --set full capture of the class SpMV, the streamed SpMV, a p-multigrid sweep and the Maxwell class stage
# Outputs land in gpurun_out/ ; the raw pages are exported to CSV here so they can be committed under profiles/.
set -u
mkdir -p gpurun_out
NCU="ncu --clock-control none"
$NCU --metrics gpu__time_duration.sum -c 600 --csv --log-file gpurun_out/r02_launches_quad_k3_20m.csv \
    python bench.py --profile-iters 40 > gpurun_out/r02_ncu_launches.log 2>&1
cap() {  # name, kernel regex, command...
    local name=$1 rx=$2; shift 2
    $NCU --set full --import-source on -k "regex:$rx" -s 3 -c 1 -f -o gpurun_out/$name "$@" > gpurun_out/$name.log 2>&1
    ncu -i gpurun_out/$name.ncu-rep --page raw --csv > gpurun_out/${name}_raw.csv 2>/dev/null
}
cap r02_ncu_spmv_class_20m   'dg_spmv_class'          python bench.py --profile-iters 20
YFD_DG_CLASSES=0 cap r02_ncu_spmv_tma_20m 'dg_spmv_tma' python bench.py --profile-iters 20
cap r02_ncu_pmg_sweep_20m    'pmg_sweep'              python bench.py --workload dg_sip_quad_k3_20m_pmg --profile-iters 3
cap r02_ncu_maxwell_class    'maxwell_stage_class'    python bench.py --workload maxwell_tri_k4_r9 --profile-iters 2
cap r02_ncu_dg_sip_cells     'dg_sip_cells'           python bench.py --profile-iters 2
ls -la gpurun_out | tail -20
