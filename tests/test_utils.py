import io
import os

import numpy as np
import pytest
import torch

import layerwise_b200  # noqa: F401
from layerwise_b200 import utils
from layerwise_b200.utils import mixed
from layerwise_b200.utils.misc import format_env, log_environment, run_parallel


def test_piecewise_linear_and_phase_scheduler():
    sched = utils.PiecewiseLinear([0, 5, 24], [0, 0.4, 0])
    assert sched(0) == 0 and abs(sched(5) - 0.4) < 1e-9 and abs(sched(2.5) - 0.2) < 1e-9
    opt = torch.optim.SGD([torch.nn.Parameter(torch.zeros(1))], lr=0.0)
    ps = utils.PhaseScheduler(opt, [{"ep": (0, 2), "lr": (1.0, 2.0)}, {"ep": 2, "lr": 1.0}, {"ep": (3, 5), "lr": (0.1, 0.0)}])
    assert ps.tot_epochs == 5
    assert ps.get_lr(0, 0, 10) == 1.0 and abs(ps.get_lr(1, 5, 10) - 1.75) < 1e-9
    assert ps.get_lr(2, 3, 10) == 1.0
    ps.update_lr(4, 10, 10)
    assert abs(opt.param_groups[0]["lr"] - 0.0) < 1e-9
    assert utils.listify(3) == [3] and utils.listify((1, 2)) == [1, 2] and utils.listify(None) == []


def test_loggers(tmp_path):
    buf = io.StringIO()
    tl = utils.TableLogger(buf)
    tl.append({"epoch": 1, "lr": 0.1, "test acc": np.float32(0.5)})
    tl.append({"epoch": 2, "lr": 0.2, "test acc": np.float32(0.75)})
    lines = buf.getvalue().splitlines()
    assert len(lines) == 3 and "0.7500" in lines[2]
    tsv = utils.TSVLogger()
    tsv.append({"epoch": 1, "total time": 3600.0, "test acc": 0.9412})
    assert str(tsv).splitlines() == ["epoch\thours\ttop1Accuracy", "1\t1.00000000\t94.12"]
    st = utils.StatsLogger(("loss",))
    st.append({"loss": torch.tensor([1.0, 3.0])})
    st.append({"loss": torch.tensor([5.0])})
    assert st.mean("loss") == 3.0
    fl = utils.FileLogger(str(tmp_path), is_master=True, is_rank0=True)
    fl.event("~~epoch\thours")
    fl.verbose("hello")
    assert "~~epoch" in open(tmp_path / "event.log").read() and "hello" in open(tmp_path / "verbose.log").read()
    tb = utils.TensorboardLogger(str(tmp_path), is_master=True, world_size=8)
    tb.log("sizes/lr", 0.5)
    tb.update_step_count(256)
    tb.log_trn_times(0.1, 0.01, 32)
    tb.close()
    import json
    sc = json.load(open(tmp_path / "scalars.json"))
    assert sc["sizes/lr"] == [[0, 0.5]] and sc["times/allgpu_images_per_sec"][0][1] == pytest.approx(2560.0)
    utils.NoOp().anything(1, 2)


def test_meters():
    m = utils.AverageMeter()
    m.update(1.0, 2)
    m.update(4.0, 1)
    assert m.avg == 2.0 and m.val == 4.0
    r, t = utils.network_bytes()
    assert r >= 0 and t >= 0
    nm = utils.NetworkMeter()
    a, b = nm.update_bandwidth()
    assert a >= 0 and b >= 0
    tm = utils.TimeMeter()
    tm.batch_start()
    tm.batch_end()
    assert tm.batch_time.count == 1


def test_checkpoint_reference_format_roundtrip(tmp_path):
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(4, 3), torch.nn.BatchNorm1d(3))
    opt = torch.optim.SGD(net.parameters(), lr=0.1, momentum=0.9)
    net(torch.randn(5, 4)).sum().backward()
    opt.step()
    path = str(tmp_path / "ck.pth.tar")
    state = utils.save_checkpoint(6, net, 93.5, opt, filename=path)
    assert {"epoch", "state_dict", "best_top5", "optimizer"} <= set(state) and state["epoch"] == 7
    net2 = torch.nn.Sequential(torch.nn.Linear(4, 3), torch.nn.BatchNorm1d(3))
    opt2 = torch.optim.SGD(net2.parameters(), lr=0.1, momentum=0.9)
    ep, best = utils.load_checkpoint(path, net2, opt2)
    assert ep == 7 and best == 93.5
    for a, b in zip(net.state_dict().values(), net2.state_dict().values()):
        assert torch.equal(a, b)
    # a file written the reference's way (four keys only, DDP-style "module." prefix) loads too
    ref_style = {"epoch": 3, "state_dict": {"module." + k: v for k, v in net.state_dict().items()},
                 "best_top5": 91.0, "optimizer": opt.state_dict()}
    torch.save(ref_style, path)
    assert utils.load_checkpoint(path, net2, opt2) == (3, 91.0)


def test_mixed_precision_helpers():
    net = torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3), torch.nn.BatchNorm2d(8))
    half = mixed.network_to_half(net, torch.bfloat16)
    assert half[1][0].weight.dtype == torch.bfloat16 and half[1][1].weight.dtype == torch.float32
    model_p, master_p = mixed.prep_param_lists(half)
    assert all(m.dtype == torch.float32 for m in master_p) and len(model_p) == len(master_p)
    for p in model_p:
        p.grad = torch.ones_like(p)
    mixed.model_grads_to_master_grads(model_p, master_p)
    assert all(m.grad.dtype == torch.float32 and float(m.grad.mean()) == 1.0 for m in master_p)
    for m in master_p:
        m.data.fill_(0.5)
    mixed.master_params_to_model_params(model_p, master_p)
    assert all(float(p.float().mean()) == 0.5 for p in model_p)
    groups = mixed.bnwd_optim_params(half, model_p, master_p)
    assert groups[0]["weight_decay"] == 0.0 and len(groups[0]["params"]) == 2 and len(groups[1]["params"]) == 2


def test_misc_helpers():
    obj = [{"ep": 0, "sz": 128, "bs": 64}, {"ep": (0, 5), "lr": (1.0, 2.0)}]
    assert utils.text_unpickle(utils.text_pickle(obj)) == obj
    os.environ["LWC_TEST_FLAG"] = "1"
    assert utils.is_set("LWC_TEST_FLAG") and not utils.is_set("LWC_TEST_UNSET")
    assert len(utils.random_id(7)) == 7
    assert format_env(A=1, B="x") == "A=1 B=x"
    assert "python" in log_environment()
    assert run_parallel(lambda v: v * v, [1, 2, 3]) == [1, 4, 9]
    t = utils.Timer()
    assert t() >= 0 and t.total_time >= 0


def test_minimal_api_example_runs_and_learns():
    import importlib.util
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("ex_min", os.path.join(root, "examples", "compressed_ddp_minimal.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("CPU example test")
    first, last = mod.main(["--steps", "25", "--method", "Topk", "--ratio", "0.1"])
    assert last < first
