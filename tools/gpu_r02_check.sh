# Round-2 check: GPU test suite, smoke, the default bench line (as the driver runs it) and the reference arm.
T=${1:-r02_v1}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo pytest rc=$?; tail -15 gpurun_out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo smoke rc=$?; tail -3 gpurun_out/${T}_smoke.log
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo bench rc=$?; cut -c1-600 gpurun_out/${T}_bench.json; tail -5 gpurun_out/${T}_bench.err
