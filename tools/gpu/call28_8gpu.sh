set -x
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 170 $B --master-port 29511 bench.py --gpus 8 --model llama2-13b --tp 4 --pp 2 --vpp_layers 5 --steps 3 --warmup 3 > gpurun_out/bench_13b_tp4pp2_vpp.log 2>&1; echo "13b rc=$?"; tail -1 gpurun_out/bench_13b_tp4pp2_vpp.log | cut -c1-330
timeout 170 $B --master-port 29512 bench.py --gpus 8 --model falcon-40b --tp 8 --seq 2048 --steps 3 --warmup 3 > gpurun_out/bench_falcon40b_tp8.log 2>&1; echo "falcon rc=$?"; tail -1 gpurun_out/bench_falcon40b_tp8.log | cut -c1-330
PYTORCH_CUDA_ALLOC_CONF=expandable_segments:True timeout 200 $B --master-port 29513 bench.py --gpus 8 --model llama2-70b --tp 4 --pp 2 --recompute --steps 2 --warmup 3 > gpurun_out/bench_70b_tp4pp2.log 2>&1; echo "70b rc=$?"; tail -1 gpurun_out/bench_70b_tp4pp2.log | cut -c1-330; grep -E "OutOfMemory|Error" gpurun_out/bench_70b_tp4pp2.log | head -3 | cut -c1-300
nvidia-smi --query-gpu=memory.used --format=csv | head -3
