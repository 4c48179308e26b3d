#!/bin/bash
# One B200 box, one process per GPU. The reference ran two single-GPU nodes over TCP
# (REF/scripts/run_node1.sh:3-16: NCCL_SOCKET_IFNAME=ens3, NCCL_IB_DISABLE=1, --nnodes=2); on a
# single NVSwitch box those settings are wrong (the interface does not exist, and NVLink is the
# transport), so both scripts collapse to the same single-node launch. NODE_RANK is kept only so
# that tooling which calls "run_node1.sh" keeps working; run ONE of the two.
set -euo pipefail
export NCCL_DEBUG=${NCCL_DEBUG:-WARN}
NPROC=${NPROC:-$(python -c 'import torch; print(max(torch.cuda.device_count(), 1))')}
cd "$(dirname "$0")/.."
exec python -m torch.distributed.run \
  --nnodes=1 \
  --nproc-per-node="$NPROC" \
  --master-addr 127.0.0.1 \
  --master-port "${MASTER_PORT:-29500}" \
  src/distributed_finetuning.py
