set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR tools/multi_gpu_check.py > gpurun_out/r02_multi_gpu_check_n2_v5.txt 2>&1; echo "mgc exit $?"
tail -1 gpurun_out/r02_multi_gpu_check_n2_v5.txt | cut -c1-200
YFD_CG_TRACE=1 timeout 600 $TR bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r02_bench_quad_k3_20m_n2_v5.json 2> gpurun_out/n2_err.log; echo "bench exit $?"
grep "yfd trace" gpurun_out/n2_err.log | head -2
python tools/show_bench.py gpurun_out/r02_bench_quad_k3_20m_n2_v5.json 2>/dev/null | cut -c1-250 | head -1
