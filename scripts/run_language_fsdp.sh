#!/usr/bin/env bash
# B200 counterpart of the reference's 02_development/run_language_fsdp.sh (:8-23): export a communication preset,
# then launch the 4-GPU FSDP language-model job.  The reference disabled CollNet, raised the NCCL watchdog to 2 h
# and switched async error handling OFF (hangs were tolerated); here NVLS stays on, failures are detected
# (async error handling on, 10-minute watchdog) and there are no RCCL variables.
set -euo pipefail
export NCCL_NVLS_ENABLE=${NCCL_NVLS_ENABLE:-1}
export NCCL_P2P_LEVEL=${NCCL_P2P_LEVEL:-NVL}
export TORCH_NCCL_ASYNC_ERROR_HANDLING=${TORCH_NCCL_ASYNC_ERROR_HANDLING:-1}
export TORCH_NCCL_HEARTBEAT_TIMEOUT_SEC=${TORCH_NCCL_HEARTBEAT_TIMEOUT_SEC:-600}
export HYPERION_NVLS=${HYPERION_NVLS:-1}        # use multimem.* in the fused kernels when a multicast mapping exists
NGPU=${NGPU:-4}
EPOCHS=${EPOCHS:-25}
REPO=$(cd "$(dirname "$0")/.." && pwd)
BASE_DIR=${BASE_DIR:-$REPO}                     # data/processed is read from, data/distributed written under it
python -m torch.distributed.run --standalone --local-addr 127.0.0.1 --nproc-per-node "$NGPU" \
  "$REPO/run_distributed.py" --model language_fsdp --epochs "$EPOCHS" --base_dir "$BASE_DIR" "$@"
