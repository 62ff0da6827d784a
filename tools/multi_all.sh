# Everything that needs several GPUs, for one gpurun --gpus 8 call: C-ABI sharding checks (both process models), the combine
# micro-benchmark (peer memory vs NCCL), the scaling runs.
NG=${NG:-8}
timeout 300 python tools/multigpu_check.py --gpus $NG --out gpurun_out/${TAG:-r02}_multigpu_check_n${NG}.json; echo "one-process check rc=$?"
timeout 300 python tools/multigpu_check.py --gpus $NG --processes --out gpurun_out/${TAG:-r02}_multigpu_check_procs_n${NG}.json; echo "processes check rc=$?"
timeout 300 python tools/combine_bench.py --gpus $NG --out gpurun_out/${TAG:-r02}_combine_n${NG}.json | cut -c1-250
timeout 300 python -m pytest tests/test_gpu_multi.py -x -q -m gpu 2>&1 | tail -2
bash tools/scale_run.sh
