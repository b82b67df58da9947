set -x
cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_dg.py -x -q -k owned 2>&1 | tail -2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR tools/multi_gpu_check.py > gpurun_out/r02_multi_gpu_check_n2_v4.txt 2>&1; echo "mgc exit $?"
grep -c "owned rows OK\|owned slices OK" gpurun_out/r02_multi_gpu_check_n2_v4.txt; tail -2 gpurun_out/r02_multi_gpu_check_n2_v4.txt | cut -c1-300
timeout 600 $TR bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r02_bench_quad_k3_20m_n2_v4.json 2> gpurun_out/n2_err.log; echo "bench exit $?"
timeout 600 $TR bench.py --gpus 2 --steps 3 --warmup 3 --workload maxwell_tri_k4_r9 > gpurun_out/r02_bench_maxwell_tri_k4_r9_n2_v3.json 2>> gpurun_out/n2_err.log; echo "mx exit $?"
python tools/show_bench.py gpurun_out/r02_bench_quad_k3_20m_n2_v4.json gpurun_out/r02_bench_maxwell_tri_k4_r9_n2_v3.json 2>/dev/null | cut -c1-330
tail -3 gpurun_out/n2_err.log
