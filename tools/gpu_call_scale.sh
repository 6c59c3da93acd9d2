#!/bin/bash
# Round 2: bench.py on N GPUs of one box, launched the way the driver launches it (usage: gpurun --gpus N -- 'bash tools/gpu_call_scale.sh N').
cd "$(dirname "$0")/.."
N=${1:-2}
O=gpurun_out
mkdir -p $O
nvidia-smi topo -m > $O/scale_n${N}_topology.txt 2>&1; nproc >> $O/scale_n${N}_topology.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 3 > $O/scale_n${N}_bench.json 2> $O/scale_n${N}_bench.err; echo "bench rc $?" >> $O/scale_n${N}_bench.err
tail -n 5 $O/scale_n${N}_bench.err; cut -c1-400 $O/scale_n${N}_bench.json
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --steps 10 --warmup 3 --workload dielectrics --no-scale-4k > $O/scale_n${N}_bench_dielectrics.json 2> $O/scale_n${N}_bench_dielectrics.err; echo "bench rc $?" >> $O/scale_n${N}_bench_dielectrics.err
tail -n 3 $O/scale_n${N}_bench_dielectrics.err; cut -c1-300 $O/scale_n${N}_bench_dielectrics.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > $O/scale_n${N}_bench_reference.json 2> $O/scale_n${N}_bench_reference.err; echo "bench rc $?" >> $O/scale_n${N}_bench_reference.err
cut -c1-300 $O/scale_n${N}_bench_reference.json
if [ "$N" = "2" ]; then timeout 300 python -m pytest tests/test_peer_exchange.py -m gpu -x -q > $O/scale_n${N}_peer_test.log 2>&1; tail -2 $O/scale_n${N}_peer_test.log; fi
