"""``python -m determined_b200.agent``: run an agent until interrupted (reference ``determined-agent run``)."""
import argparse
import logging
import os
import signal
import sys

from determined_b200.agent.agent import Agent


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="determined-agent", description=__doc__)
    ap.add_argument("--master-url", default=os.environ.get("DET_MASTER", "http://127.0.0.1:8080"))
    ap.add_argument("--agent-id", default=os.environ.get("DET_AGENT_ID"))
    ap.add_argument("--slots", type=int, default=None, help="cap the number of slots offered (default: every visible GPU, else 1 CPU slot)")
    ap.add_argument("--slot-type", choices=["auto", "cuda", "cpu"], default="auto")
    ap.add_argument("--resource-pool", default="default")
    ap.add_argument("--work-dir", default=None, help="task run directories and the re-attachment state live here")
    ap.add_argument("--agent-token", default=None, help="shared secret if the master sets DET_AGENT_TOKEN")
    ap.add_argument("--master-cert-file", default=None, help="PEM certificate of an https master to trust, or 'noverify' (default: $DET_MASTER_CERT_FILE)")
    ap.add_argument("--master-cert-name", default=None, help="host name in the master's certificate when it differs from the URL's")
    ap.add_argument("--container-runtime", choices=["process", "docker", "podman", "apptainer", "singularity"],
                    default=os.environ.get("DET_CONTAINER_RUNTIME", "process"), help="run tasks on the host (process) or inside containers of this runtime")
    ap.add_argument("--default-image", default=os.environ.get("DET_DEFAULT_IMAGE", ""), help="image of tasks that name none (environment.image)")
    ap.add_argument("--container-exe", default=None, help="path of the runtime's command-line client when it is not on PATH")
    ap.add_argument("--container-arg", action="append", default=[], help="extra argument for every container run (repeatable)")
    ap.add_argument("--log-level", default="INFO")
    a = ap.parse_args(argv)
    logging.basicConfig(level=getattr(logging, a.log_level.upper(), logging.INFO), format="%(asctime)s %(levelname)s %(name)s: %(message)s")
    agent = Agent(a.master_url, a.agent_id, a.slots, a.slot_type, a.resource_pool, a.work_dir, a.agent_token,
                  a.master_cert_file, a.master_cert_name, a.container_runtime, a.default_image, a.container_exe, a.container_arg)

    def on_signal(signum, frame):  # noqa: ARG001
        # SIGTERM / SIGINT: leave tasks running (they re-attach when the agent comes back), like `systemctl restart`
        agent.stop(kill_tasks=signum == signal.SIGINT)

    signal.signal(signal.SIGTERM, on_signal)
    signal.signal(signal.SIGINT, on_signal)
    agent.run()
    return 0


if __name__ == "__main__":
    sys.exit(main())
