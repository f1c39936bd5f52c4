#!/bin/bash
# Node set-up for training with this framework on a B200 / NVSwitch box.
#
# Local counterpart of the reference's IMAGENET/setup.sh (an AWS AMI recipe: pillow-simd, TCP buffer sysctls for the
# 25 Gbit/s EFA-less network its NCCL rings ran over, pip requirements, the validation aspect-ratio index).  Here the
# gradients travel over NVLink peer memory inside the node, so only the rendezvous / multi-node TCP path cares about
# socket buffers; nothing is downloaded (pass a wheelhouse for offline installs).
#
#   tools/setup_node.sh                 # check the box, build the native library
#   tools/setup_node.sh --sysctl        # also raise the TCP buffer limits (needs root; multi-node jobs)
#   tools/setup_node.sh --wheelhouse /opt/wheelhouse   # pip install -r requirements.txt from local wheels only
#   DATA=/data/imagenet tools/setup_node.sh            # also build the validation aspect-ratio index
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
SYSCTL=0
WHEELHOUSE=""
while [ $# -gt 0 ]; do
  case "$1" in
    --sysctl) SYSCTL=1 ;;
    --wheelhouse) WHEELHOUSE="$2"; shift ;;
    -h|--help) sed -n 2,14p "${BASH_SOURCE[0]}"; exit 0 ;;
    *) echo "unknown option $1" >&2; exit 2 ;;
  esac
  shift
done

echo "== GPUs"
if command -v nvidia-smi >/dev/null 2>&1; then
  nvidia-smi --query-gpu=index,name,memory.total,driver_version --format=csv,noheader || true
  echo "== NVLink / NVSwitch topology (every pair should read NV18 on a B200 HGX board)"
  nvidia-smi topo -m 2>/dev/null | head -20 || true
else
  echo "nvidia-smi not found: CPU-only node (the gloo / reference-math engine will be used)"
fi

if [ "$SYSCTL" = 1 ]; then
  echo "== TCP buffers (rendezvous and multi-node traffic)"
  if [ "$(id -u)" != 0 ]; then echo "--sysctl needs root" >&2; exit 1; fi
  sysctl -w net.core.rmem_max=16777216 net.core.wmem_max=16777216 \
            net.core.rmem_default=16777216 net.core.wmem_default=16777216 \
            net.core.netdev_max_backlog=30000 \
            net.ipv4.tcp_rmem='4096 87380 16777216' net.ipv4.tcp_wmem='4096 65536 16777216'
fi

echo "== limits"
ulimit -n 65536 2>/dev/null || true          # many loader workers x many image files
ulimit -l unlimited 2>/dev/null || true      # pinned host batches
echo "open files: $(ulimit -n), locked memory: $(ulimit -l)"

if [ -n "$WHEELHOUSE" ]; then
  echo "== python requirements from $WHEELHOUSE (no index)"
  python -m pip install --no-index --find-links "$WHEELHOUSE" -r "$ROOT/requirements.txt"
fi

echo "== native library (nvcc, sm_100a)"
( cd "$ROOT" && python -c "import __graft_entry__ as g; g.build()" )

if [ -n "${DATA:-}" ]; then
  echo "== validation aspect-ratio index for $DATA"
  python "$ROOT/tools/prepare_imagenet.py" index "$DATA"
fi

cat <<'EOF'
== environment for a run
export NCCL_DEBUG=WARN                 # INFO shows whether NVLS is used for the plumbing collectives
export OMP_NUM_THREADS=8               # loader workers decode JPEGs; keep the trainers' own threads few
python -m layerwise_b200.launch --nproc 8 -- train_imagenet.py $DATA --bf16 --compress layerwise --method Topk -K 0.01 --cuda-graph
EOF
