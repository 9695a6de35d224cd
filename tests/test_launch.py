"""Launch layer: the C++ launcher (rank contract, [rank=N] prefixes, kill-a-rank watchdog), the torch_distributed launch module
with a real gloo rendezvous, argument parsing of the launch modules."""
import json
import os
import pathlib
import subprocess
import sys
import time

import pytest

from determined_b200.launch import _common, deepspeed, horovod, torch_distributed

ROOT = pathlib.Path(__file__).resolve().parent.parent
LAUNCHER = _common.NATIVE_LAUNCHER
needs_launcher = pytest.mark.skipif(not LAUNCHER.exists(), reason="native launcher not built")


def _free_port():
    import socket

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@needs_launcher
def test_native_launcher_rank_contract_and_prefixes():
    code = "import os; print('R', os.environ['RANK'], os.environ['LOCAL_RANK'], os.environ['WORLD_SIZE'], os.environ['LOCAL_WORLD_SIZE'], os.environ['MASTER_ADDR'], os.environ['MASTER_PORT'])"
    port = _free_port()
    r = subprocess.run([str(LAUNCHER), "--nproc", "3", "--master-addr", "127.0.0.1", "--master-port", str(port), "--", sys.executable, "-c", code],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=60)
    assert r.returncode == 0, r.stdout
    lines = sorted(l for l in r.stdout.splitlines() if " R " in l or l.startswith("[rank="))
    assert lines == [f"[rank={i}] R {i} {i} 3 3 127.0.0.1 {port}" for i in range(3)], r.stdout


@needs_launcher
def test_native_launcher_kills_the_job_when_a_rank_dies():
    """Kill-a-rank: rank 1 fails, the survivors (sleeping for a minute) are torn down within the grace period and the job exits
    non-zero — the allocation then restarts from its checkpoint under max_restarts (reference pid_server / WorkerError)."""
    code = "import os, sys, time\nif os.environ['RANK'] == '1':\n    time.sleep(0.5); sys.exit(3)\ntime.sleep(60)"
    t0 = time.time()
    r = subprocess.run([str(LAUNCHER), "--nproc", "3", "--grace", "2", "--master-port", str(_free_port()), "--", sys.executable, "-c", code],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=50)
    assert r.returncode != 0 and time.time() - t0 < 20, (r.returncode, time.time() - t0, r.stdout)


def test_torch_distributed_launch_module_runs_a_gloo_job(tmp_path):
    script = tmp_path / "job.py"
    script.write_text(
        "import os, torch, torch.distributed as dist\n"
        "dist.init_process_group('gloo')\n"
        "t = torch.tensor([float(dist.get_rank() + 1)])\n"
        "dist.all_reduce(t)\n"
        "open(os.path.join(os.environ['OUT'], f\"r{dist.get_rank()}\"), 'w').write(str(float(t)))\n"
        "dist.destroy_process_group()\n")
    env = dict(os.environ, PYTHONPATH=str(ROOT), DET_SLOT_IDS=json.dumps([0, 1]), OUT=str(tmp_path), C10D_PORT=str(_free_port()), CUDA_VISIBLE_DEVICES="")
    env.pop("DET_MASTER", None)
    r = subprocess.run([sys.executable, "-m", "determined_b200.launch.torch_distributed", "--", sys.executable, str(script)], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=180)
    assert r.returncode == 0, r.stdout[-2000:]
    assert (tmp_path / "r0").read_text() == "3.0" and (tmp_path / "r1").read_text() == "3.0"


def test_launch_argument_parsing():
    ov, script = torch_distributed.parse_args(["--rdzv_conf", "x=1", "--", "python3", "train.py", "--lr", "1"])
    assert ov == ["--rdzv_conf", "x=1"] and script == ["python3", "train.py", "--lr", "1"]
    ov, script = torch_distributed.parse_args(["--trial", "model_def:MyTrial"])
    assert ov == [] and "model_def:MyTrial" in " ".join(script)
    with pytest.raises(SystemExit):
        torch_distributed.parse_args([])
    assert _common.split_args(["a", "b"]) == ([], ["a", "b"])
    cmd = _common.torchrun_cmd(4, ["python3", "t.py"], ["--foo"], port=1234)
    assert cmd[cmd.index("--nproc_per_node") + 1] == "4" and "--max_restarts" in cmd and cmd[-2:] == ["python3", "t.py"] and "--foo" in cmd
    assert _common.pid_server_wrap(2, ["x"], "/tmp/sock")[-4:] == ["/tmp/sock", "2", "--", "x"]
    for mod in (deepspeed, horovod):
        assert hasattr(mod, "main")


def _pid_server(tmp_path, worker_codes):
    """pid_server supervising one launcher process that starts len(worker_codes) workers under pid_client."""
    addr = str(tmp_path / "pid.sock")
    env = dict(os.environ, PYTHONPATH=str(ROOT))
    workers = " & ".join(f"{sys.executable} -m determined_b200.exec.pid_client {addr} -- {sys.executable} -c \"{code}\"" for code in worker_codes)
    cmd = [sys.executable, "-m", "determined_b200.exec.pid_server", "--on-fail", "SIGTERM", "--on-exit", "WAIT", "--grace-period", "2", "--signal-children", addr,
           str(len(worker_codes)), "--", "/bin/bash", "-c", workers + " & wait"]
    t0 = time.time()
    with open(tmp_path / "out.log", "w") as out:          # a file, not a pipe: an orphaned worker must not keep the test waiting for EOF
        r = subprocess.run(cmd, env=env, stdout=out, stderr=subprocess.STDOUT, timeout=120)
    r.stdout = (tmp_path / "out.log").read_text()
    return r, time.time() - t0


def test_pid_server_passes_when_all_workers_exit_cleanly(tmp_path):
    r, _ = _pid_server(tmp_path, ["print('a')", "print('b')"])
    assert r.returncode == 0, r.stdout[-2000:]


def test_pid_server_fails_fast_when_a_worker_dies(tmp_path):
    """Reference ipc.py:409-502: a worker that vanishes without the graceful-exit byte is a WorkerError; the launcher (and the
    worker still sleeping) are signalled after the grace period."""
    r, dt = _pid_server(tmp_path, ["import time; time.sleep(60)", "import sys, time; time.sleep(0.5); sys.exit(7)"])
    assert r.returncode != 0 and dt < 30, (r.returncode, dt, r.stdout[-2000:])
    assert "without a graceful exit" in r.stdout


@pytest.mark.parametrize("module,flag", [("deepspeed", "USE_DEEPSPEED"), ("horovod", None)])
def test_deepspeed_and_horovod_launch_modules_start_local_ranks(tmp_path, module, flag):
    """Both accept the reference's command lines and start one process per slot with the torch rank contract."""
    script = tmp_path / "job.py"
    script.write_text("import os\nopen(os.path.join(os.environ['OUT'], 'r' + os.environ['RANK']), 'w').write(os.environ['WORLD_SIZE'] + ' ' + os.environ.get('USE_DEEPSPEED', '-'))\n")
    env = dict(os.environ, PYTHONPATH=str(ROOT), DET_SLOT_IDS=json.dumps([0, 1, 2]), OUT=str(tmp_path), C10D_PORT=str(_free_port()), CUDA_VISIBLE_DEVICES="")
    env.pop("DET_MASTER", None)
    r = subprocess.run([sys.executable, "-m", f"determined_b200.launch.{module}", "--", sys.executable, str(script)], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert r.returncode == 0, r.stdout[-2000:]
    got = sorted((p.name, p.read_text()) for p in tmp_path.glob("r[0-9]"))
    assert [g[0] for g in got] == ["r0", "r1", "r2"] and all(g[1].startswith("3 ") for g in got)
    if flag:
        assert all(g[1].endswith(" 1") for g in got)


def test_native_launcher_is_clean_under_address_and_undefined_behaviour_sanitizers(tmp_path):
    """The launcher rebuilt with -fsanitize=address,undefined runs the rank contract and the kill-a-rank path without a
    report (heap misuse in the line splitter, use after close of a pidfd, signed overflow in the grace timer ...)."""
    src = ROOT / "csrc" / "launcher" / "launcher.cpp"
    exe = tmp_path / "launch_asan"
    b = subprocess.run(["g++", "-std=c++17", "-O1", "-g", "-fsanitize=address,undefined", "-fno-omit-frame-pointer", str(src), "-o", str(exe)],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    if b.returncode != 0 and ("asan" in b.stdout.lower() or "ubsan" in b.stdout.lower()):
        pytest.skip("sanitizer runtimes are not installed")
    assert b.returncode == 0, b.stdout[-2000:]
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=1:abort_on_error=0:exitcode=66", UBSAN_OPTIONS="halt_on_error=1:exitcode=66")
    code = "import os; print('x' * 5000); print('R', os.environ['RANK'])"
    r = subprocess.run([str(exe), "--nproc", "3", "--master-port", str(_free_port()), "--", sys.executable, "-c", code], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert r.returncode == 0 and "Sanitizer" not in r.stdout and "runtime error" not in r.stdout, r.stdout[-3000:]
    assert sorted(l for l in r.stdout.splitlines() if " R " in l) == [f"[rank={i}] R {i}" for i in range(3)]
    code = "import os, sys, time\nif os.environ['RANK'] == '1':\n    time.sleep(0.3); sys.exit(3)\ntime.sleep(60)"
    r = subprocess.run([str(exe), "--nproc", "3", "--grace", "1", "--master-port", str(_free_port()), "--", sys.executable, "-c", code], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert r.returncode not in (0, 66) and "Sanitizer" not in r.stdout and "runtime error" not in r.stdout, r.stdout[-3000:]
