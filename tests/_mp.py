"""Helper: run a function in W gloo processes and collect per-rank results."""
import os
import socket
import sys
import tempfile
import traceback

import torch
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _entry(rank, world, port, fn, args, outdir):
    import torch.distributed as dist
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if root not in sys.path:
        sys.path.insert(0, root)
    torch.set_num_threads(1)
    try:
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", world_size=world, rank=rank)
        res = fn(rank, world, *args)
        torch.save(res, os.path.join(outdir, f"r{rank}.pt"))
        dist.barrier()
        dist.destroy_process_group()
    except Exception:
        with open(os.path.join(outdir, f"err{rank}.txt"), "w") as f:
            f.write(traceback.format_exc())
        raise


def run_gloo(fn, world=2, args=()):
    with tempfile.TemporaryDirectory() as d:
        port = _free_port()
        try:
            mp.spawn(_entry, args=(world, port, fn, args, d), nprocs=world, join=True)
        except Exception:
            msgs = []
            for r in range(world):
                p = os.path.join(d, f"err{r}.txt")
                if os.path.exists(p):
                    msgs.append(open(p).read())
            raise RuntimeError("worker failed:\n" + "\n".join(msgs))
        return [torch.load(os.path.join(d, f"r{r}.pt"), weights_only=False) for r in range(world)]
