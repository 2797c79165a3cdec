#!/bin/bash
# Launcher with the reference's environment interface (run.sh:1-16):
#   dnn=resnet50 nworkers=8 compressor=gaussiank density=0.001 ./run.sh
# One process per GPU is started with torchrun on this node (the reference used mpirun over two hosts).
# Extra knobs: nwpernode (GPUs per node, default = nworkers), threshold, master_port, extra (more CLI flags).
set -e
cd "$(dirname "$0")"
dnn="${dnn:-resnet20}"
source exp_configs/$dnn.conf
nworkers="${nworkers:-2}"
density="${density:-1}"
compressor="${compressor:-gaussiank}"
nwpernode="${nwpernode:-$nworkers}"
threshold="${threshold:-524288000}"
master_port="${master_port:-29533}"
PY="${PY:-python}"
GRADSPATH="${GRADSPATH:-./logs/iclr}"

ARGS="--dnn $dnn --dataset $dataset --max-epochs $max_epochs --batch-size $batch_size --nworkers $nworkers \
  --data-dir $data_dir --lr $lr --nsteps-update $nstepsupdate --nwpernode $nwpernode --density $density \
  --compressor $compressor --threshold $threshold --saved-dir $GRADSPATH $extra"

if [ "$nworkers" -gt 1 ]; then
  exec $PY -m torch.distributed.run --nnodes=1 --nproc-per-node "$nworkers" --master-addr 127.0.0.1 \
    --master-port "$master_port" -m gaussiank_sgd_b200.dist_trainer $ARGS
else
  exec $PY -m gaussiank_sgd_b200.dist_trainer $ARGS
fi
