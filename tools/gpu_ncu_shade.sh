# ncu capture of the shading kernel on the bench batch
mkdir -p gpurun_out
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_shade -s 3 -c 1 -o gpurun_out/${1:-shade} -f python tools/shade_time.py > gpurun_out/${1:-shade}.log 2>&1
tail -2 gpurun_out/${1:-shade}.log
