# 1 -> N GPU scaling of the bench (run under gpurun --gpus N)
NMAX=${1:-8}
mkdir -p gpurun_out
for n in 1 2 4 8; do
  [ $n -gt $NMAX ] && break
  if [ $n -eq 1 ]; then
    timeout 200 python bench.py --no-cpu-baseline > gpurun_out/scale_c1_n$n.json 2> gpurun_out/scale_c1_n$n.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n > gpurun_out/scale_c1_n$n.json 2> gpurun_out/scale_c1_n$n.err
  fi
  echo "c1 n=$n rc=$?"; cut -c1-200 gpurun_out/scale_c1_n$n.json
done
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NMAX --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus $NMAX --workload c4 > gpurun_out/scale_c4_n$NMAX.json 2> gpurun_out/scale_c4_n$NMAX.err
echo "c4 n=$NMAX rc=$?"; cut -c1-200 gpurun_out/scale_c4_n$NMAX.json
