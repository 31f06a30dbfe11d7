# N-GPU round: the driver's scaling command line (bench.py under torchrun, --steps 20 --warmup 5) and the D2H ceiling probe
N=${1:-2}; T=${2:-r02_multi}
mkdir -p gpurun_out
PORT=29517
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --steps 20 --warmup 5 \
  > gpurun_out/${T}_bench_n$N.json 2> gpurun_out/${T}_bench_n$N.err; echo bench rc=$?
tail -c 3000 gpurun_out/${T}_bench_n$N.json; echo; tail -5 gpurun_out/${T}_bench_n$N.err
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((PORT+1)) tools/d2h_probe.py \
  > gpurun_out/${T}_d2h_n$N.json 2> gpurun_out/${T}_d2h_n$N.err; cat gpurun_out/${T}_d2h_n$N.json
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((PORT+2)) tools/d2h_probe.py --affinity \
  > gpurun_out/${T}_d2h_aff_n$N.json 2> gpurun_out/${T}_d2h_aff_n$N.err; cat gpurun_out/${T}_d2h_aff_n$N.json
nvidia-smi topo -m > gpurun_out/${T}_topo_n$N.txt 2>&1; head -20 gpurun_out/${T}_topo_n$N.txt
