#!/usr/bin/env python
"""Start TensorBoard on a run directory of this framework and print its URL.

Local counterpart of the reference's ``IMAGENET/tools/launch_tensorboard.py`` (which rents an AWS VM through
``ncluster`` and runs ``tensorboard --logdir`` there): the event files that ``TensorboardLogger`` writes under
``--logdir`` are served from this host instead.

    python tools/launch_tensorboard.py runs/                 # serve every run below runs/
    python tools/launch_tensorboard.py runs/ --port 6007 --dry-run
"""
import argparse
import os
import shutil
import socket
import subprocess
import sys


def command(logdir: str, port: int, host: str):
    exe = shutil.which("tensorboard")
    base = [exe] if exe else [sys.executable, "-m", "tensorboard.main"]
    return base + ["--logdir", logdir, "--port", str(port), "--host", host]


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("logdir", help="run directory (the --logdir given to train_imagenet.py) or a parent of several")
    ap.add_argument("--port", type=int, default=6006)
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--dry-run", action="store_true", help="print the command and URL, start nothing")
    args = ap.parse_args(argv)
    logdir = os.path.abspath(os.path.expanduser(args.logdir))
    if not os.path.isdir(logdir):
        ap.error(f"{logdir} is not a directory")
    cmd = command(logdir, args.port, args.host)
    url = f"http://{socket.gethostname() if args.host == '0.0.0.0' else args.host}:{args.port}"
    print(f"Tensorboard at {url}   ({' '.join(cmd)})")
    if args.dry_run:
        return cmd
    try:
        import tensorboard  # noqa: F401
    except ImportError:
        sys.exit("tensorboard is not installed in this environment")
    return subprocess.call(cmd)


if __name__ == "__main__":
    main()
