set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541"
timeout 600 $TR tools/multi_gpu_check.py > gpurun_out/r02_multi_gpu_check_n8_v3.txt 2>&1; echo "mgc exit $?"
grep -c "owned rows OK\|owned slices OK" gpurun_out/r02_multi_gpu_check_n8_v3.txt; tail -1 gpurun_out/r02_multi_gpu_check_n8_v3.txt
timeout 600 $TR bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/r02_bench_quad_k3_20m_n8_v3.json 2> gpurun_out/n8_err.log; echo "bench exit $?"
timeout 600 $TR bench.py --gpus 8 --steps 3 --warmup 3 --workload maxwell_tri_k4_r9 > gpurun_out/r02_bench_maxwell_tri_k4_r9_n8_v3.json 2>> gpurun_out/n8_err.log; echo "mx exit $?"
timeout 600 $TR bench.py --gpus 8 --steps 2 --warmup 3 --scaling weak > gpurun_out/r02_weak_n8_v2.json 2>> gpurun_out/n8_err.log; echo "weak exit $?"
python tools/show_bench.py gpurun_out/r02_bench_quad_k3_20m_n8_v3.json gpurun_out/r02_bench_maxwell_tri_k4_r9_n8_v3.json gpurun_out/r02_weak_n8_v2.json 2>/dev/null | cut -c1-300
tail -3 gpurun_out/n8_err.log
