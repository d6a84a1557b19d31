"""One-process-per-GPU launcher for a single 8 x B200 box.

The reference starts the DLRM job with ``cloudtik-run --launcher <name> --nproc-per-node N ... script args``
(``RUN/launcher_factory.py:23-54`` picks a ``Launcher`` subclass; ``RUN/launcher.py:11-41`` is the base:
``launch() = resolve(); setup(); run(); finalize()``; distributed flags ``--nnodes --nproc-per-node
--master-addr --master-port`` in ``RUN/distributed_launcher.py:11-48``), RUN = python/cloudtik/runtime/ai/runner.
Its launchers are CPU ones (numactl / MPI / Horovod).  ``GPULauncher`` is the missing ``"gpu"`` entry with the
same shape: it exports the rendezvous variables the training script's ``ext_dist.init_distributed`` reads
(``RANK / WORLD_SIZE / LOCAL_RANK / LOCAL_WORLD_SIZE / MASTER_ADDR / MASTER_PORT``, the names in
``RUN/util/env.py:22-25``) and starts one worker per GPU.  Every worker keeps ALL GPUs visible and selects its
own through ``LOCAL_RANK`` -- the fused embedding exchange maps its peers' HBM, which a per-process
``CUDA_VISIBLE_DEVICES`` mask would forbid (``--isolate-devices`` restores the mask for plain NCCL jobs).

    python -m cloudtik_b200.launcher --nproc-per-node 8 -m cloudtik_b200.train --mini-batch-size=32768 ...
    python -m cloudtik_b200.launcher --launcher gpu --nproc-per-node 2 bench.py --gpus 2

The reference-side patch is one branch in ``create_launcher``:
``elif launcher_name == "gpu": from cloudtik_b200.launcher import GPULauncher; launcher = GPULauncher(args, distributor)``.
"""
from __future__ import annotations

import argparse
import os
import signal
import subprocess
import sys
import time


def add_gpu_launcher_params(parser):
    group = parser.add_argument_group("GPU Launching Parameters")
    group.add_argument("--launcher", default="gpu", type=str, help="launcher name (only 'gpu' lives here)")
    group.add_argument("--nnodes", type=int, default=1, help="must be 1: one NVLink box")
    group.add_argument("--nproc-per-node", "--nproc_per_node", type=int, default=0,
                       help="processes = GPUs to use (0: every visible GPU)")
    group.add_argument("--master-addr", "--master_addr", default="127.0.0.1", type=str)
    group.add_argument("--master-port", "--master_port", default=29500, type=int)
    group.add_argument("--isolate-devices", "--isolate_devices", action="store_true", default=False,
                       help="give every worker CUDA_VISIBLE_DEVICES=<its GPU> (disables the peer-memory exchange)")
    group.add_argument("-m", "--module", action="store_true", default=False,
                       help="run the program as a python module (python -m)")
    group.add_argument("--no-python", "--no_python", action="store_true", default=False,
                       help="the program is not a python script")
    group.add_argument("command", type=str, help="program / module to launch")
    group.add_argument("command_args", nargs=argparse.REMAINDER)
    return parser


def _visible_gpus():
    env = os.environ.get("CUDA_VISIBLE_DEVICES")
    if env is not None:
        return len([x for x in env.split(",") if x.strip() != ""])
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


class GPULauncher:
    """Same life cycle as the reference's ``Launcher`` (``RUN/launcher.py:19-24``)."""

    def __init__(self, args, distributor=None):
        self.args = args
        self.distributor = distributor
        self.environ_set = {}
        self.nproc = 0
        self.procs = []

    def launch(self):
        self.resolve()
        self.setup()
        result = self.run()
        self.finalize()
        return result

    def get_nproc_per_node(self):
        return self.args.nproc_per_node if self.args.nproc_per_node else max(_visible_gpus(), 1)

    def resolve(self):
        if getattr(self.args, "nnodes", 1) not in (0, 1):
            raise ValueError("the gpu launcher drives one node (an 8 x B200 NVLink box)")
        self.nproc = self.get_nproc_per_node()
        if self.distributor is not None:
            self.distributor.resolve(nproc_per_node=self.nproc)

    def set_env(self, env_name, env_value):
        self.environ_set[env_name] = str(env_value)

    def setup(self):
        self.set_env("MASTER_ADDR", self.args.master_addr)
        self.set_env("MASTER_PORT", self.args.master_port)
        self.set_env("WORLD_SIZE", self.nproc)
        self.set_env("LOCAL_WORLD_SIZE", self.nproc)

    def worker_env(self, local_rank):
        env = dict(os.environ)
        env.update(self.environ_set)
        env["RANK"] = env["LOCAL_RANK"] = str(local_rank)
        if self.args.isolate_devices:
            devs = os.environ.get("CUDA_VISIBLE_DEVICES")
            ids = [x for x in devs.split(",") if x.strip() != ""] if devs else [str(i) for i in range(self.nproc)]
            env["CUDA_VISIBLE_DEVICES"] = ids[local_rank]
            env["LOCAL_RANK"] = "0"
        return env

    def command(self):
        a = self.args
        cmd = []
        if not a.no_python and (a.module or a.command.endswith(".py")):
            cmd += [sys.executable, "-u"]
            if a.module:
                cmd.append("-m")
        return cmd + [a.command] + list(a.command_args)

    def run(self):
        cmd = self.command()
        self.procs = [subprocess.Popen(cmd, env=self.worker_env(r)) for r in range(self.nproc)]
        rc = 0
        alive = list(self.procs)
        while alive:  # first failure ends the job: the others would hang in their next collective
            for p in list(alive):
                r = p.poll()
                if r is None:
                    continue
                alive.remove(p)
                if r != 0 and rc == 0:
                    rc = r
                    for q in alive:
                        q.send_signal(signal.SIGTERM)
            time.sleep(0.05)
        return rc

    def finalize(self):
        for p in self.procs:
            if p.poll() is None:
                p.kill()


def create_launcher(launcher_name, args, distributor=None):
    if launcher_name == "gpu":
        return GPULauncher(args, distributor)
    raise ValueError("Launcher type with name {} is not supported.".format(launcher_name))


def main(argv=None):
    parser = argparse.ArgumentParser(description="cloudtik-run style launcher, one process per B200")
    add_gpu_launcher_params(parser)
    args = parser.parse_args(argv)
    return create_launcher(args.launcher, args).launch()


if __name__ == "__main__":
    sys.exit(main())
