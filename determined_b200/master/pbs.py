"""PBS (PBS Professional / OpenPBS / Torque) backend: the Slurm backend with PBS's command line.

Parity: reference ``master/internal/rm/dispatcherrm`` serves Slurm and PBS through the same launcher service
(``resource_manager.type: pbs``; expconf ``pbs.pbsbatch_args``).  Here: ``qsub`` to submit (prints the job id), ``qstat`` to watch,
``qdel`` to stop.  Everything else — run directories on the shared work directory, the exit-code file, log shipping, the quota
of slots — is :class:`~determined_b200.master.slurm.SlurmBackend`.

Pool config (``master.yaml``)::

    resource_pools:
      - pool_name: hpc
        provider: {type: pbs, slots: 64, queue: gpu, select: "1:ngpus={slots}", pbsbatch_args: ["-l walltime=04:00:00"], poll_seconds: 2}
"""
import os
import subprocess
from typing import Any, Dict, List

from determined_b200.master import slurm


class PBSBackend(slurm.SlurmBackend):
    NAME = "PBS"
    OUT_FILE = "pbs.out"
    # qstat state letters of jobs that are over: F(inished) / C(ompleted, Torque) / X (sub-job done)
    DONE_STATES = ("F", "C", "X")

    def directives(self, action: Dict[str, Any], run_dir: str, n: int) -> List[str]:
        out = os.path.join(run_dir, self.OUT_FILE)
        lines = [f"#PBS -N det-{action['container_id'][:12]}", f"#PBS -o {out}", "#PBS -j oe"]
        if self.cfg.get("queue"):
            lines.append(f"#PBS -q {self.cfg['queue']}")
        select = str(self.cfg.get("select", "1:ngpus={slots}")) if n else str(self.cfg.get("select_cpu", "1:ncpus=1"))
        if select:
            lines.append("#PBS -l select=" + select.format(slots=n))
        mine = (action.get("hpc") or {}).get("pbs") or {}            # the experiment's own ``pbs:`` section
        for a in list(self.cfg.get("pbsbatch_args") or []) + list(mine.get("pbsbatch_args") or []):
            lines.append(f"#PBS {a}")
        return lines

    def script(self, action: Dict[str, Any], run_dir: str) -> str:
        # PBS only copies the -o file into place when the job ends: redirect inside the script so that logs stream
        body = super().script(action, run_dir)
        head, _, rest = body.partition("\nexport ")
        return head + f"\nexec >> {os.path.join(run_dir, self.OUT_FILE)} 2>&1\nexport " + rest

    def submit(self, path: str) -> str:
        r = subprocess.run([self._cmd("qsub"), path], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=60)
        if r.returncode != 0:
            raise RuntimeError(f"qsub failed ({r.returncode}): {r.stderr.strip() or r.stdout.strip()}")
        return r.stdout.strip().splitlines()[-1].strip()          # "1234.pbs-server"

    def cancel(self, job_id: str) -> None:
        subprocess.run([self._cmd("qdel"), job_id], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=60)

    def _in_queue(self, job_ids: List[str]) -> Dict[str, str]:
        """job id -> state letter, from ``qstat -x`` (``-x`` also lists finished jobs while the server remembers them)."""
        if not job_ids:
            return {}
        r = subprocess.run([self._cmd("qstat"), "-x", *job_ids], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=60)
        out: Dict[str, str] = {}
        for line in r.stdout.splitlines():
            parts = line.split()
            # Job id            Name             User              Time Use S Queue
            if len(parts) >= 6 and parts[0] not in ("Job", "---") and not set(parts[0]) <= {"-"}:
                jid = next((j for j in job_ids if j == parts[0] or j.split(".")[0] == parts[0].split(".")[0]), None)
                if jid is not None:
                    out[jid] = parts[-2]
        return out
