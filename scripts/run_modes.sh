TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 python bench_modes.py --mode c3 --steps 8 > gpurun_out/c3.json 2> gpurun_out/c3.err; tail -2 gpurun_out/c3.err
timeout 300 python bench_modes.py --mode c4 --batch 64 > gpurun_out/c4_n1.json 2> gpurun_out/c4_n1.err; tail -2 gpurun_out/c4_n1.err
timeout 300 $TR --nproc-per-node 2 --master-port 29511 bench_modes.py --mode c4 --batch 64 --gpus 2 > gpurun_out/c4_n2.json 2> gpurun_out/c4_n2.err; tail -2 gpurun_out/c4_n2.err
timeout 400 $TR --nproc-per-node 2 --master-port 29512 bench_modes.py --mode c5 --tile-points 25000000 --gpus 2 --steps 8 > gpurun_out/c5_n2.json 2> gpurun_out/c5_n2.err; tail -2 gpurun_out/c5_n2.err
cat gpurun_out/c3.json gpurun_out/c4_n1.json gpurun_out/c4_n2.json gpurun_out/c5_n2.json
