"""Container runtimes of the agent: how one task's command line becomes a ``docker`` / ``podman`` / ``apptainer`` invocation.

Parity: reference ``agent/pkg/docker`` (engine API client: pull with registry auth, create with device requests, shm size,
capabilities, bind mounts, host networking), ``agent/pkg/podman`` and ``agent/pkg/singularity`` (CLI drivers: env file,
``--bind``, ``--nv`` / ``--nvccli``, ``--writable-tmpfs``) behind ``agent/pkg/cruntimes``.

Design here: the agent's task machinery (``agent.ProcessRuntime``: process group, output pipe, state file for re-attachment,
TERM-then-KILL) stays the only supervisor.  A driver is a pure function from a :class:`TaskSpec` to the host command that
runs the task in the foreground (``docker run --rm ...`` exits with the task's code and relays its output), plus the command
that delivers a signal to the container when signalling the client process is not enough.  So drivers own no threads and no
state, and can be tested by reading the command lines they produce.

    python -m determined_b200.agent --container-runtime docker --default-image nvcr.io/nvidia/pytorch:25.04-py3

The task's image comes from ``environment.image`` (a string, or ``{cuda: ..., cpu: ...}``), mounts from ``bind_mounts``,
``resources.shm_size`` / ``resources.devices``, ``environment.add_capabilities`` / ``drop_capabilities``,
``environment.force_pull_image`` and ``environment.registry_auth``.
"""
import dataclasses
import json
import os
import shlex
import shutil
import subprocess
from typing import Any, Dict, List, Optional, Sequence, Tuple


@dataclasses.dataclass
class Mount:
    host_path: str
    container_path: str
    read_only: bool = False
    propagation: str = ""          # docker only: rprivate / rshared / ...


@dataclasses.dataclass
class TaskSpec:
    """Everything a driver may need about one task, already resolved for this host."""
    name: str                                   # container name: det-<container id>
    image: str
    argv: List[str]
    env: Dict[str, str]                         # the task's own variables only — never the agent's whole environment
    cwd: str
    mounts: List[Mount] = dataclasses.field(default_factory=list)
    gpu_ids: List[int] = dataclasses.field(default_factory=list)
    gpu_uuids: List[str] = dataclasses.field(default_factory=list)
    shm_size: Optional[int] = None              # bytes
    devices: List[Dict[str, str]] = dataclasses.field(default_factory=list)   # [{host_path, container_path, mode}]
    add_caps: List[str] = dataclasses.field(default_factory=list)
    drop_caps: List[str] = dataclasses.field(default_factory=list)
    user: Optional[str] = None                  # "uid:gid" (agent user group of the workspace / user)
    network: str = "host"
    env_file: str = ""                          # where the driver may write the environment (inside the run directory)
    extra_args: List[str] = dataclasses.field(default_factory=list)          # agent-wide --container-arg


def parse_size(v: Any) -> Optional[int]:
    """``resources.shm_size``: bytes, or a string with a k/m/g/t suffix (reference expconf ``shm_size``)."""
    if v is None or v == "":
        return None
    if isinstance(v, (int, float)):
        return int(v)
    s = str(v).strip().lower().rstrip("b").rstrip("i")
    mult = {"k": 1 << 10, "m": 1 << 20, "g": 1 << 30, "t": 1 << 40}
    if s and s[-1] in mult:
        return int(float(s[:-1]) * mult[s[-1]])
    return int(float(s))


def pick_image(image: Any, has_gpu: bool, default: str) -> str:
    if isinstance(image, dict):
        return str(image.get("cuda" if has_gpu else "cpu") or image.get("gpu") or image.get("cuda") or image.get("cpu") or default)
    return str(image or default)


def write_env_file(path: str, env: Dict[str, str]) -> Tuple[str, Dict[str, str]]:
    """``KEY=value`` lines for ``--env-file``.  Neither docker nor apptainer can carry a newline in an env file: such values
    are returned separately to be passed another way."""
    multi = {k: v for k, v in env.items() if "\n" in v}
    with open(path, "w") as f:
        for k, v in sorted(env.items()):
            if k not in multi:
                f.write(f"{k}={v}\n")
    os.chmod(path, 0o600)
    return path, multi


class Driver:
    """One container runtime.  ``exe`` is the command-line client (overridable: ``--container-exe``)."""
    name = ""
    default_exe = ""

    def __init__(self, exe: Optional[str] = None) -> None:
        self.exe = exe or self.default_exe

    def available(self) -> bool:
        return shutil.which(self.exe) is not None or os.path.isfile(self.exe)

    def pull(self, image: str, force: bool, auth: Optional[Dict[str, str]] = None) -> List[List[str]]:
        """Commands to run before the task so that ``image`` is present (empty: the run command fetches it on demand)."""
        return []

    def run(self, spec: TaskSpec) -> Tuple[List[str], Dict[str, str]]:
        """(host argv, extra host environment) that runs the task in the foreground."""
        raise NotImplementedError

    def signal(self, name: str, sig: str) -> Optional[List[str]]:
        """Command that delivers ``sig`` (TERM / KILL) to the container, or None when signalling the process group does it."""
        return None

    def cleanup(self, name: str) -> Optional[List[str]]:
        return None


class _OCI(Driver):
    """docker and podman share the ``run`` grammar; they differ in how GPUs are named."""

    def _gpu_args(self, spec: TaskSpec) -> List[str]:
        raise NotImplementedError

    def pull(self, image: str, force: bool, auth: Optional[Dict[str, str]] = None) -> List[List[str]]:
        cmds: List[List[str]] = []
        if auth and auth.get("username"):
            server = auth.get("serveraddress") or image.split("/")[0]
            # the password goes on stdin (see agent.ProcessRuntime._prepare), never on the command line
            cmds.append([self.exe, "login", "--username", str(auth["username"]), "--password-stdin", server])
        if force or auth:
            cmds.append([self.exe, "pull", image])
        return cmds

    def run(self, spec: TaskSpec) -> Tuple[List[str], Dict[str, str]]:
        cmd = [self.exe, "run", "--rm", "--name", spec.name, "--init", "--network", spec.network, "--workdir", spec.cwd]
        if spec.network == "host":
            cmd += ["--ipc", "host"]            # NCCL's shared-memory transport between the ranks of a node
        multi: Dict[str, str] = {}
        if spec.env_file:
            _, multi = write_env_file(spec.env_file, spec.env)
            cmd += ["--env-file", spec.env_file]
        else:
            multi = dict(spec.env)
        for k in sorted(multi):
            cmd += ["--env", k]                 # value taken from the client's environment
        for m in spec.mounts:
            opt = f"type=bind,source={m.host_path},target={m.container_path}"
            if m.read_only:
                opt += ",readonly"
            if m.propagation:
                opt += f",bind-propagation={m.propagation}"
            cmd += ["--mount", opt]
        if spec.shm_size:
            cmd += ["--shm-size", str(spec.shm_size)]
        for d in spec.devices:
            cmd += ["--device", ":".join(x for x in (d.get("host_path"), d.get("container_path"), d.get("mode")) if x)]
        for c in spec.add_caps:
            cmd += ["--cap-add", c]
        for c in spec.drop_caps:
            cmd += ["--cap-drop", c]
        if spec.user:
            cmd += ["--user", spec.user]
        cmd += self._gpu_args(spec) + list(spec.extra_args) + [spec.image] + list(spec.argv)
        return cmd, multi

    def signal(self, name: str, sig: str) -> Optional[List[str]]:
        return [self.exe, "kill", "--signal", sig, name]

    def cleanup(self, name: str) -> Optional[List[str]]:
        return [self.exe, "rm", "--force", name]


class Docker(_OCI):
    """``docker run`` with the NVIDIA container toolkit's ``--gpus`` (reference ``docker.go`` DeviceRequests)."""
    name = "docker"
    default_exe = "docker"

    def _gpu_args(self, spec: TaskSpec) -> List[str]:
        ids = spec.gpu_uuids or [str(i) for i in spec.gpu_ids]
        return ["--gpus", '"device=' + ",".join(ids) + '"'] if ids else []


class Podman(_OCI):
    """``podman run`` with CDI device names (``nvidia.com/gpu=<index or uuid>``); rootless, so no ``--user`` remap is needed
    for files written into mounts (``--userns keep-id`` keeps the caller's uid)."""
    name = "podman"
    default_exe = "podman"

    def _gpu_args(self, spec: TaskSpec) -> List[str]:
        ids = spec.gpu_uuids or [str(i) for i in spec.gpu_ids]
        out: List[str] = []
        for i in ids:
            out += ["--device", f"nvidia.com/gpu={i}"]
        if ids:
            out += ["--security-opt", "label=disable"]
        return out + ["--userns", "keep-id"]


class Apptainer(Driver):
    """``apptainer exec`` (Singularity's successor; ``--container-runtime singularity`` selects the old binary name).

    No daemon and no container name: the task is an ordinary child process, so the agent's process-group signals are the
    kill path.  Images are ``docker://`` URIs unless they already name a scheme or a ``.sif`` file."""
    name = "apptainer"
    default_exe = "apptainer"

    @staticmethod
    def image_uri(image: str) -> str:
        if "://" in image or image.endswith(".sif") or os.path.isabs(image):
            return image
        return "docker://" + image

    def pull(self, image: str, force: bool, auth: Optional[Dict[str, str]] = None) -> List[List[str]]:
        return []                               # exec pulls into the user's cache; credentials travel as APPTAINER_DOCKER_*

    def run(self, spec: TaskSpec) -> Tuple[List[str], Dict[str, str]]:
        cmd = [self.exe, "exec", "--no-home", "--writable-tmpfs", "--pwd", spec.cwd, "--cleanenv"]
        host_env: Dict[str, str] = {}
        if spec.env_file:
            _, multi = write_env_file(spec.env_file, spec.env)
            cmd += ["--env-file", spec.env_file]
        else:
            multi = dict(spec.env)
        for k, v in multi.items():              # APPTAINERENV_X=... becomes X inside, whatever the value holds
            host_env[f"{self.name.upper()}ENV_{k}"] = v
        for m in spec.mounts:
            cmd += ["--bind", f"{m.host_path}:{m.container_path}" + (":ro" if m.read_only else "")]
        for d in spec.devices:
            cmd += ["--bind", f"{d.get('host_path')}:{d.get('container_path') or d.get('host_path')}"]
        if spec.gpu_ids or spec.gpu_uuids:
            cmd += ["--nv"]
            ids = spec.gpu_uuids or [str(i) for i in spec.gpu_ids]
            host_env[f"{self.name.upper()}ENV_CUDA_VISIBLE_DEVICES"] = ",".join(ids)
        for c in spec.add_caps:
            cmd += ["--add-caps", c]
        for c in spec.drop_caps:
            cmd += ["--drop-caps", c]
        if spec.network != "host":
            cmd += ["--net", "--network", spec.network]
        cmd += list(spec.extra_args) + [self.image_uri(spec.image)] + list(spec.argv)
        return cmd, host_env


class Singularity(Apptainer):
    name = "singularity"
    default_exe = "singularity"


DRIVERS = {d.name: d for d in (Docker, Podman, Apptainer, Singularity)}


def get(name: str, exe: Optional[str] = None) -> Optional[Driver]:
    """The driver called ``name``; ``process`` (or empty) means no container at all."""
    if not name or name == "process":
        return None
    if name not in DRIVERS:
        raise ValueError(f"unknown container runtime {name!r}; one of: process, {', '.join(sorted(DRIVERS))}")
    return DRIVERS[name](exe)


def build_spec(action: Dict[str, Any], *, cid: str, argv: Sequence[str], env: Dict[str, str], cwd: str, run_dir: str, repo_root: str,
               slots: Sequence[Dict[str, Any]], default_image: str, extra_args: Sequence[str] = ()) -> TaskSpec:
    """The :class:`TaskSpec` of a ``start`` action.  The run directory is mounted at its own path (so every ``DET_*`` path in
    the environment holds inside), this package's tree read-only at its own path (the harness the task imports)."""
    c = action.get("container") or {}
    slot_ids = [int(s) for s in action.get("slot_ids") or []]
    by_index = {int(s.get("index", i)): s for i, s in enumerate(slots)}
    uuids = [str(by_index[i]["uuid"]) for i in slot_ids if i in by_index and str(by_index[i].get("type", "")).endswith("CUDA")
             and str(by_index[i].get("uuid", "")).startswith("GPU-")]
    mounts = [Mount(run_dir, run_dir), Mount(repo_root, repo_root, read_only=True)]
    for b in c.get("bind_mounts") or []:
        cp = str(b.get("container_path") or b.get("host_path"))
        if not os.path.isabs(cp):
            cp = os.path.join(cwd, cp)          # reference: relative container paths hang off the working directory
        mounts.append(Mount(str(b["host_path"]), cp, bool(b.get("read_only")), str(b.get("propagation") or "")))
    ckpt = env.get("DET_CHECKPOINT_HOST_PATH")
    if ckpt:
        mounts.append(Mount(ckpt, ckpt))
    return TaskSpec(name="det-" + "".join(ch if ch.isalnum() or ch in "-_." else "-" for ch in cid), image=pick_image(c.get("image"), bool(slot_ids) and bool(uuids or c.get("gpu")), default_image),
                    argv=list(argv), env=dict(env), cwd=cwd, mounts=mounts, gpu_ids=slot_ids if (uuids or c.get("gpu")) else [], gpu_uuids=uuids,
                    shm_size=parse_size(c.get("shm_size")), devices=list(c.get("devices") or []),
                    add_caps=[str(x) for x in c.get("add_capabilities") or []], drop_caps=[str(x) for x in c.get("drop_capabilities") or []],
                    user=c.get("user"), network=str(c.get("network_mode") or "host"), env_file=os.path.join(run_dir, "task.env"),
                    extra_args=list(extra_args))


def describe(cmd: Sequence[str]) -> str:
    return " ".join(shlex.quote(c) for c in cmd)


def run_prepare(cmds: List[List[str]], password: Optional[str] = None) -> None:
    """Run the pull commands; ``login --password-stdin`` gets the password on stdin."""
    for c in cmds:
        stdin = password if "--password-stdin" in c else None
        r = subprocess.run(c, input=stdin, text=True, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=3600)
        if r.returncode != 0:
            raise RuntimeError(f"{describe(c[:2])} failed ({r.returncode}): {r.stdout.strip()[-500:]}")


def dumps(spec: TaskSpec) -> str:
    return json.dumps(dataclasses.asdict(spec), indent=1, sort_keys=True)
