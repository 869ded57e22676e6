mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02l_gputest.log 2>&1; echo "pytest exit $?"
timeout 600 python bench.py > gpurun_out/r02l_c4.json 2> gpurun_out/r02l_c4.err; echo "c4 exit $?"
timeout 300 python bench.py --workload c5 --steps 200 --warmup 5 > gpurun_out/r02l_c5.json 2> gpurun_out/r02l_c5.err; echo "c5 exit $?"
timeout 300 python bench.py --workload c3 > gpurun_out/r02l_c3.json 2> gpurun_out/r02l_c3.err; echo "c3 exit $?"
timeout 300 python bench.py --workload c2 > gpurun_out/r02l_c2.json 2> gpurun_out/r02l_c2.err; echo "c2 exit $?"
tail -3 gpurun_out/r02l_gputest.log
