set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541"
timeout 300 $TR tools/multi_gpu_check.py > gpurun_out/r02_multi_gpu_check_n8_v4.txt 2>&1; echo "mgc exit $?"
tail -1 gpurun_out/r02_multi_gpu_check_n8_v4.txt
timeout 300 $TR bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/r02_bench_quad_k3_20m_n8_v4.json 2> gpurun_out/n8_err.log; echo "bench exit $?"
python tools/show_bench.py gpurun_out/r02_bench_quad_k3_20m_n8_v4.json 2>/dev/null | cut -c1-260 | head -1
