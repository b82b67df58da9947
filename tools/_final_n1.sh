set -x
cd $GRAFT_REPO_ROOT
python -m pytest tests -x -q -m gpu 2>&1 | tail -4 > gpurun_out/r02_pytest_gpu_final.txt; cat gpurun_out/r02_pytest_gpu_final.txt
python bench.py > gpurun_out/r02_bench_quad_k3_20m_v4.json 2> gpurun_out/final_err.log; echo "bench exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_reference_arm_v2.json 2>> gpurun_out/final_err.log; echo "ref exit $?"
python bench.py --workload maxwell_tri_k4_r9 > gpurun_out/r02_bench_maxwell_tri_k4_r9_v6.json 2>> gpurun_out/final_err.log; echo "mx exit $?"
python bench.py --workload maxwell_quad_k3_r9 --no-cpu-baseline > gpurun_out/r02_bench_maxwell_quad_k3_r9_v3.json 2>> gpurun_out/final_err.log; echo "mxq exit $?"
python bench.py --workload dg_sip_tri_k2_r11 --single-shot --pc 1 > gpurun_out/r02_full_tri_k2_r11_n1_jacobi_v2.json 2>> gpurun_out/final_err.log; echo "1e8 exit $?"
python tools/sanitize_target.py > gpurun_out/r02_sanitize_target_plain.txt 2>&1; echo "selfcheck exit $?"
tail -3 gpurun_out/r02_sanitize_target_plain.txt
tail -5 gpurun_out/final_err.log
