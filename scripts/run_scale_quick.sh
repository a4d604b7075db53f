#!/bin/bash
# usage: scripts/run_scale_quick.sh N [extra bench args]  - topology + a --quick bench line at N GPUs (GPU box)
N=$1; shift
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
lscpu | grep -i "numa\|socket\|model name\|^CPU(s)" >> gpurun_out/topo_n$N.txt
for d in /sys/bus/pci/devices/*; do if [ -f $d/class ] && grep -q "^0x0302" $d/class; then echo "$d numa=$(cat $d/numa_node) cpus=$(cat $d/local_cpulist)" >> gpurun_out/topo_n$N.txt; fi; done
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5 --quick "$@" 2> gpurun_out/scale_n$N.err
