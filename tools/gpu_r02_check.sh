# Round-2 check: the allocation-free entry point alone in a fresh process, the GPU test suite (it again, last), smoke,
# the default bench line (as the driver runs it) and the reference arm.
T=${1:-r02_v1}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/${T}_pytest_noalloc_alone.log 2>&1; echo noalloc-alone rc=$?; tail -5 gpurun_out/${T}_pytest_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo pytest rc=$?; tail -15 gpurun_out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo smoke rc=$?; tail -3 gpurun_out/${T}_smoke.log
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo bench rc=$?; cut -c1-600 gpurun_out/${T}_bench.json; tail -5 gpurun_out/${T}_bench.err
timeout 600 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err; echo ref rc=$?; cut -c1-400 gpurun_out/${T}_bench_reference.json
