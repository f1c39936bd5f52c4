"""Real multi-GPU run of the peer-memory engines (skipped with fewer than 2 GPUs)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_fused_and_dense_engines_over_peer_memory():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29571", os.path.join(ROOT, "tools", "check_multigpu.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    out = r.stdout
    d = json.loads(out[out.index('{\n "world"'):out.rindex("}") + 1])
    assert d["ok"] and all(v["ok"] for v in d["results"].values())


def test_cuda_graph_step_over_peer_memory():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29573", os.path.join(ROOT, "tools", "check_graph_multigpu.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    out = r.stdout
    d = json.loads(out[out.index('{\n "world"'):out.rindex("}") + 1])
    assert d["ok"] and all(v["ok"] for v in d["results"].values())
