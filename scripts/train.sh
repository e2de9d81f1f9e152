#!/usr/bin/env bash
# Submit `llm-training fit` as a SLURM job (one task per GPU) and follow its output.
#   scripts/train.sh --config config/examples/llama-3.1/llama-3.1-8b_pt_example.yaml [--nodes 2] [--gpus-per-node 8] \
#                    [--partition P] [--time 2-00:00:00] [--job-name NAME] [-- extra llm-training overrides]
# Without sbatch on PATH the command runs locally (the trainer spawns one process per visible GPU).
# Same purpose as the reference's scripts/train.sh:17-77.
set -euo pipefail

CONFIG=""; NODES=1; GPUS_PER_NODE=8; PARTITION=""; TIME=""; JOB_NAME="llm-training"; CKPT=""; EXTRA=()
while [[ $# -gt 0 ]]; do
  case "$1" in
    --config) CONFIG="$2"; shift 2 ;;
    --nodes) NODES="$2"; shift 2 ;;
    --gpus-per-node) GPUS_PER_NODE="$2"; shift 2 ;;
    --partition) PARTITION="$2"; shift 2 ;;
    --time) TIME="$2"; shift 2 ;;
    --job-name) JOB_NAME="$2"; shift 2 ;;
    --ckpt_path) CKPT="$2"; shift 2 ;;
    --) shift; EXTRA=("$@"); break ;;
    *) EXTRA+=("$1"); shift ;;
  esac
done
[[ -n "$CONFIG" ]] || { echo "usage: $0 --config CONFIG [--nodes N] [--gpus-per-node G] [...]" >&2; exit 2; }

REPO="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
CMD=(python -m llm_training_b200.cli.main fit --config "$CONFIG" --trainer.num_nodes "$NODES")
[[ -n "$CKPT" ]] && CMD+=(--ckpt_path "$CKPT")
CMD+=("${EXTRA[@]}")

if ! command -v sbatch >/dev/null 2>&1; then
  echo "sbatch not found: running locally" >&2
  cd "$REPO" && exec "${CMD[@]}" --trainer.devices "$GPUS_PER_NODE"
fi

SBATCH_ARGS=(--job-name "$JOB_NAME" --nodes "$NODES" --ntasks-per-node "$GPUS_PER_NODE" --gpus-per-node "$GPUS_PER_NODE"
             --cpus-per-task 8 --output "slurm-%j.out")
[[ -n "$PARTITION" ]] && SBATCH_ARGS+=(--partition "$PARTITION")
[[ -n "$TIME" ]] && SBATCH_ARGS+=(--time "$TIME")

JOB_ID=$(sbatch --parsable "${SBATCH_ARGS[@]}" --wrap "cd $REPO && MASTER_ADDR=\$(scontrol show hostnames \$SLURM_JOB_NODELIST | head -n1) MASTER_PORT=29500 srun ${CMD[*]}")
echo "Submitted job $JOB_ID"
# wait for the job to start, then follow its first step like `sattach`
while [[ "$(squeue -h -j "$JOB_ID" -o %T 2>/dev/null)" == "PENDING" ]]; do sleep 5; done
if command -v sattach >/dev/null 2>&1; then
  sattach "$JOB_ID.0" || tail -f "slurm-$JOB_ID.out"
else
  tail -f "slurm-$JOB_ID.out"
fi
