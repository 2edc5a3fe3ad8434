# compute-sanitizer over the kernels on a small table (memcheck) and over the concurrent-request tests (racecheck).  One GPU.
set -x
mkdir -p gpurun_out
export TT_FULL_ROWS=3000000 TT_CREATOR_ROWS=3000000
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 99 python scripts/all_kernels_small.py > gpurun_out/memcheck_small.log 2>&1; echo "memcheck all_kernels_small rc=$?" | tee -a gpurun_out/memcheck_small.log
tail -4 gpurun_out/memcheck_small.log
timeout 1800 compute-sanitizer --tool memcheck --error-exitcode 99 python -m pytest tests/test_shards_gpu.py -m gpu -x -q --timeout 1500 -k "unsharded_oracle or creator_filter or single_row or exchange_objects" > gpurun_out/memcheck_shards.log 2>&1; echo "memcheck shards rc=$?" | tee -a gpurun_out/memcheck_shards.log
tail -6 gpurun_out/memcheck_shards.log
timeout 1800 compute-sanitizer --tool racecheck --error-exitcode 99 python -m pytest tests/test_threads_gpu.py -m gpu -x -q --timeout 1500 > gpurun_out/racecheck_threads.log 2>&1; echo "racecheck threads rc=$?" | tee -a gpurun_out/racecheck_threads.log
tail -6 gpurun_out/racecheck_threads.log
