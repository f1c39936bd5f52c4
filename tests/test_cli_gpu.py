"""Entry points end to end on the GPU (synthetic data, a handful of steps)."""
import json
import os

import pytest
import torch

import layerwise_b200  # noqa: F401

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("network,extra", [("resnet9", ["--compress", "layerwise", "--method", "Topk", "-K", "0.01", "--error_feedback"]),
                                           ("Alexnet", ["--compress", "entiremodel", "--method", "QSGD", "-Q", "255"]),
                                           ("Alexnet1", ["--compress", "layerwise", "--method", "TernGrad"]),
                                           ("vgg16", ["--compress", "layerwise", "--method", "Randomk", "-K", "0.01"]),
                                           ("resnet9", ["--compress", "none"])])
def test_dawn_cli_on_gpu(tmp_path, network, extra):
    import dawn
    from layerwise_b200 import _native
    _native.counters.reset()
    graph = ["--cuda_graph", "--max_batches", "6"] if network == "resnet9" else []      # capture after 3 eager batches
    summary = dawn.main(["--synthetic", "--epochs", "1", "--max_batches", "3", "--batch_size", "128", "--world_size", "1",
                         "--network", network, "--dtype", "bf16", "--momentum", "0.9", "--log_dir", str(tmp_path)] + extra
                        + graph)
    assert summary["epoch"] == 1 and summary["train loss"] == summary["train loss"]
    assert _native.counters.kernels > 0          # the native library really ran


@pytest.mark.parametrize("graph", [False, True])
def test_imagenet_cli_on_gpu(tmp_path, graph):
    from layerwise_b200.train.imagenet_main import main
    phases = json.dumps([{"ep": 0, "sz": 64, "bs": 16 if not graph else 8}, {"ep": [0, 1], "lr": [0.05, 0.1]}])
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        top1, top5 = main(["x", "--synthetic", "64", "--arch", "resnet50", "--phases", phases, "-j", "0", "--bf16",
                           "--logdir", str(tmp_path / "log"), "--compress", "layerwise", "--method", "Topk", "-K", "0.01",
                           "--error_feedback", "--epochs", "1", "--no-bn-wd", "--init-bn0"] + (["--cuda-graph"] if graph else []))
        name = "sz8_checkpoint.path.tar" if graph else "sz16_checkpoint.path.tar"
        assert 0.0 <= top1 <= 100.0 and os.path.exists(tmp_path / name)
        ck = torch.load(tmp_path / name, weights_only=False)
        assert ck["masters"] is not None and ck["ef_residual"] is not None
    finally:
        os.chdir(cwd)


@pytest.mark.parametrize("wrapper", ["ddp", "sparsified"])
def test_trainer_with_ddp_wrappers(wrapper):
    from layerwise_b200.train.imagenet import Trainer
    tr = Trainer(arch="resnet18", num_classes=10, dtype="bf16", lr=0.05, device="cuda:0", wrapper=wrapper, randk=0.1)
    x = torch.randint(0, 256, (8, 3, 64, 64), dtype=torch.uint8, device="cuda")
    y = torch.randint(0, 10, (8,), device="cuda")
    for _ in range(2):
        _, loss = tr.train_step(x, y)
    assert torch.isfinite(loss)


@pytest.mark.parametrize("dtype,compress,method", [("fp16", "entiremodel", "Topk"), ("fp32", "layerwise", "AdaptiveThreshold"),
                                                   ("bf16", "entiremodel", "RandomDithering")])
def test_trainer_precisions_and_modes(dtype, compress, method):
    """--fp16 (loss scale 1024, fp32 masters) as in the reference recipe, fp32, and bf16; entire-model and layer-wise."""
    from layerwise_b200.parallel.exchange import CompressionConfig
    from layerwise_b200.train.imagenet import Trainer
    cfg = CompressionConfig(compress=compress, method=method, ratio=0.05, qstates=255, error_feedback=True)
    tr = Trainer(arch="resnet18", num_classes=10, dtype=dtype, cfg=cfg, lr=0.01, loss_scale=1024.0, device="cuda:0")
    x = torch.randint(0, 256, (8, 3, 64, 64), dtype=torch.uint8, device="cuda")
    y = torch.randint(0, 10, (8,), device="cuda")
    w0 = tr.master_params[0].detach().clone()
    for _ in range(3):
        out, loss = tr.train_step(x, y)
    assert torch.isfinite(loss) and out.dtype == tr.dtype
    assert not torch.equal(w0, tr.master_params[0])
    assert tr.exchanger.engine_name.startswith("fused")
