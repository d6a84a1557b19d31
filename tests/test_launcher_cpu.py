"""cloudtik_b200.launcher on CPU: two workers get the rendezvous variables
ext_dist.init_distributed reads, rendezvous over gloo and all-reduce; a failing worker fails the job."""
import os
import socket
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(tmp_path, body, nproc=2, extra=()):
    script = tmp_path / "worker.py"
    script.write_text(textwrap.dedent(body))
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    return subprocess.run([sys.executable, "-m", "cloudtik_b200.launcher", "--nproc-per-node", str(nproc),
                           "--master-port", str(_free_port()), *extra, str(script), str(tmp_path)],
                          cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)


def test_two_workers_rendezvous_and_allreduce(tmp_path):
    r = _run(tmp_path, """
        import os, sys, torch
        from cloudtik_b200 import ext_dist
        ext_dist.init_distributed(backend="gloo")
        t = torch.tensor([float(ext_dist.my_rank + 1)])
        torch.distributed.all_reduce(t)
        sl = ext_dist.get_my_slice(26)
        open(os.path.join(sys.argv[1], f"r{ext_dist.my_rank}.txt"), "w").write(
            f"{ext_dist.my_rank} {ext_dist.my_size} {os.environ['LOCAL_RANK']} {os.environ['LOCAL_WORLD_SIZE']} "
            f"{int(t.item())} {sl.start}:{sl.stop} {os.environ.get('CUDA_VISIBLE_DEVICES', 'all') or 'empty'}")
        torch.distributed.destroy_process_group()
    """)
    assert r.returncode == 0, r.stderr[-2000:]
    got = [open(tmp_path / f"r{k}.txt").read().split() for k in range(2)]
    assert got[0][:6] == ["0", "2", "0", "2", "3", "0:13"] and got[1][:6] == ["1", "2", "1", "2", "3", "13:26"]
    assert got[0][6] == got[1][6]  # every worker sees the same devices (peer-memory exchange needs them all)


def test_isolate_devices_and_failure_propagation(tmp_path):
    r = _run(tmp_path, """
        import os, sys
        open(os.path.join(sys.argv[1], f"v{os.environ['RANK']}.txt"), "w").write(
            os.environ["CUDA_VISIBLE_DEVICES"] + " " + os.environ["LOCAL_RANK"])
        sys.exit(3 if os.environ["RANK"] == "1" else 0)
    """, extra=("--isolate-devices",))
    assert r.returncode == 3
    assert open(tmp_path / "v0.txt").read() == "0 0" and open(tmp_path / "v1.txt").read() == "1 0"
