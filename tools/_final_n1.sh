set -x
cd $GRAFT_REPO_ROOT
python -m pytest tests -x -q -m gpu 2>&1 | tail -4 > gpurun_out/r02_pytest_gpu_final.txt; cat gpurun_out/r02_pytest_gpu_final.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py > gpurun_out/r02_bench_quad_k3_20m_v5.json 2> gpurun_out/final_err.log; echo "bench exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_reference_arm_v3.json 2>> gpurun_out/final_err.log; echo "ref exit $?"
python bench.py --workload maxwell_tri_k4_r9 > gpurun_out/r02_bench_maxwell_tri_k4_r9_v7.json 2>> gpurun_out/final_err.log; echo "mx exit $?"
python bench.py --workload maxwell_quad_k3_r9 --no-cpu-baseline > gpurun_out/r02_bench_maxwell_quad_k3_r9_v4.json 2>> gpurun_out/final_err.log; echo "mxq exit $?"
python bench.py --workload maxwell_tri_k4_r8_upwind --no-cpu-baseline > gpurun_out/r02_bench_maxwell_tri_k4_r8_upwind.json 2>> gpurun_out/final_err.log; echo "mxu exit $?"
python bench.py --workload dg_sip_quad_k3_20m_pmg --no-cpu-baseline > gpurun_out/r02_bench_quad_k3_20m_pmg_v3.json 2>> gpurun_out/final_err.log; echo "pmg exit $?"
python bench.py --workload cfem_p2_4m --no-cpu-baseline > gpurun_out/r02_bench_cfem_p2_4m_v2.json 2>> gpurun_out/final_err.log; echo "cfem exit $?"
python tools/sanitize_target.py > gpurun_out/r02_sanitize_target_plain.txt 2>&1; echo "selfcheck exit $?"
tail -2 gpurun_out/r02_sanitize_target_plain.txt
tail -5 gpurun_out/final_err.log
