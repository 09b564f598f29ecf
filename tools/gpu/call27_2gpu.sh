set -x
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $B --master-port 29511 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/bench_n2.log 2>&1; echo "n2 default rc=$?"; tail -1 gpurun_out/bench_n2.log | cut -c1-250
timeout 200 $B --master-port 29512 bench.py --gpus 2 --model tiny --debug_layers 4 --tp 1 --pp 2 --vpp_layers 1 --steps 2 --warmup 1 > gpurun_out/bench_tiny_vpp.log 2>&1; echo "tiny vpp rc=$?"; tail -3 gpurun_out/bench_tiny_vpp.log | cut -c1-600
timeout 200 $B --master-port 29513 bench.py --gpus 2 --model falcon-40b --debug_layers 2 --tp 2 --seq 2048 --steps 2 --warmup 1 > gpurun_out/bench_falcon2l.log 2>&1; echo "falcon 2 layers rc=$?"; tail -3 gpurun_out/bench_falcon2l.log | cut -c1-600
timeout 200 $B --master-port 29514 bench.py --gpus 2 --model llama2-70b --debug_layers 4 --tp 1 --pp 2 --recompute --steps 2 --warmup 1 > gpurun_out/bench_70b4l.log 2>&1; echo "70b 4 layers rc=$?"; tail -3 gpurun_out/bench_70b4l.log | cut -c1-600
