"""End-to-end smoke of both trainers on CPU with synthetic data."""
import json
import os

import torch

import layerwise_b200  # noqa: F401


def test_dawn_cli_single_process(tmp_path):
    import dawn
    summary = dawn.main(["--synthetic", "--epochs", "1", "--max_batches", "2", "--batch_size", "32", "--world_size", "1",
                         "--network", "Alexnet", "--compress", "entiremodel", "--method", "QSGD", "--qstates", "15",
                         "--log_dir", str(tmp_path)])
    assert summary["epoch"] == 1 and summary["train loss"] == summary["train loss"]
    assert open(tmp_path / "logs.tsv").read().startswith("epoch\thours\ttop1Accuracy")


def test_imagenet_cli_short_epoch_checkpoint_resume(tmp_path):
    from layerwise_b200.train.imagenet_main import main, parse_phases
    phases = json.dumps([{"ep": 0, "sz": 32, "bs": 4}, {"ep": [0, 2], "lr": [0.05, 0.1]}])
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        args = ["x", "--synthetic", "16", "--arch", "resnet18", "--phases", phases, "-j", "0", "--logdir", str(tmp_path / "log"),
                "--compress", "layerwise", "--method", "Topk", "-K", "0.05", "--error_feedback", "--epochs", "1", "--no-bn-wd"]
        main(args)
        ck = torch.load(tmp_path / "sz4_checkpoint.path.tar", weights_only=False)
        assert {"epoch", "state_dict", "best_top5", "optimizer"} <= set(ck) and ck["epoch"] == 1
        assert ck["ef_residual"] is not None
        # resume continues from the stored epoch
        main(args[:-3] + ["--epochs", "1", "--no-bn-wd", "--resume", str(tmp_path / "sz4_checkpoint.path.tar")])
        assert "~~1" in open(tmp_path / "log" / "event.log").read()
    finally:
        os.chdir(cwd)
    assert parse_phases(None)[0]["sz"] == 128


def test_trainer_bf16_masters_cpu():
    from layerwise_b200.parallel.exchange import CompressionConfig
    from layerwise_b200.train.imagenet import Trainer
    tr = Trainer(arch="resnet18", num_classes=10, dtype="bf16", cfg=CompressionConfig(), lr=0.05, device="cpu",
                 overlap=False, channels_last=False)
    x = torch.randint(0, 256, (2, 3, 32, 32), dtype=torch.uint8)
    y = torch.tensor([1, 2])
    w0 = tr.master_params[0].detach().clone()
    out, loss = tr.train_step(x, y)
    assert out.dtype == torch.bfloat16 and torch.isfinite(loss)
    assert not torch.equal(w0, tr.master_params[0])
    assert torch.equal(tr.model_params[0].detach(), tr.master_params[0].detach().to(torch.bfloat16))


def test_graphed_train_step_falls_back_to_eager_on_cpu():
    """``dawn.py --cuda_graph`` on a machine without the fused engine: same loop, stepped eagerly."""
    import torch
    from layerwise_b200.models import build_cifar_model, trainable_params
    from layerwise_b200.ops.sgd import SGD
    from layerwise_b200.parallel.exchange import CompressionConfig, GradientExchanger
    from layerwise_b200.train.cifar import GraphedTrainStep, run_batches
    torch.manual_seed(0)
    model = build_cifar_model("resnet9")
    ex = GradientExchanger(trainable_params(model), CompressionConfig(compress="layerwise", method="Topk", ratio=0.1))
    opt = SGD(trainable_params(model), lr=lambda s: 0.01, momentum=0.9, weight_decay=5e-4, nesterov=True)
    step = GraphedTrainStep(model, ex, opt)
    assert not step.ok                                   # torch engine on CPU: nothing to capture
    batches = [{"input": torch.randn(8, 3, 32, 32), "target": torch.randint(0, 10, (8,))} for _ in range(3)]
    stats = run_batches(model, batches, True, ex, opt.step, graphed=step)
    assert opt.step_number == 3 and step.graph is None
    assert stats.mean("loss") == stats.mean("loss")


def test_apply_bitmask_matches_manual_mask():
    import torch
    from layerwise_b200.ops.conv import apply_bitmask
    torch.manual_seed(0)
    n, c, h, w = 2, 16, 3, 5
    keep = torch.rand(n, h, w, c) > 0.5                  # channels-last pixel order
    bits = keep.reshape(-1, 8).to(torch.uint8)
    mask = sum(bits[:, k] << k for k in range(8)).to(torch.uint8)
    t = torch.randn(n, c, h, w).contiguous(memory_format=torch.channels_last)
    out = apply_bitmask(t, mask)
    assert torch.equal(out, t * keep.permute(0, 3, 1, 2).to(t.dtype))
