set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR tools/multi_gpu_check.py > gpurun_out/r02_multi_gpu_check_n2_v2.txt 2>&1; echo "mgc exit $?"
tail -5 gpurun_out/r02_multi_gpu_check_n2_v2.txt
timeout 600 $TR bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r02_bench_quad_k3_20m_n2_v2.json 2> gpurun_out/n2_err.log; echo "bench exit $?"
tail -c 1500 gpurun_out/r02_bench_quad_k3_20m_n2_v2.json
timeout 600 $TR bench.py --gpus 2 --steps 3 --warmup 3 --workload maxwell_tri_k4_r9 > gpurun_out/r02_bench_maxwell_tri_k4_r9_n2.json 2>> gpurun_out/n2_err.log; echo "mx exit $?"
tail -c 1200 gpurun_out/r02_bench_maxwell_tri_k4_r9_n2.json
tail -5 gpurun_out/n2_err.log
