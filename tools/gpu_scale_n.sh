# The two N-GPU lines of the round (run under gpurun --gpus N): configs[4] (8K, 1024 frames accumulated, row bands +
# one NCCL gather, strong scaling) and the bench default (configs[1], frames shard, weak scaling).
N=${1:-8}
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus $N --workload c4 --no-cpu-baseline > gpurun_out/scale_c4_n$N.json 2> gpurun_out/scale_c4_n$N.err
echo "c4 n=$N rc=$?"; cut -c1-300 gpurun_out/scale_c4_n$N.json
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29601 bench.py --gpus $N --no-cpu-baseline > gpurun_out/scale_c1_n$N.json 2> gpurun_out/scale_c1_n$N.err
echo "c1 n=$N rc=$?"; cut -c1-300 gpurun_out/scale_c1_n$N.json
