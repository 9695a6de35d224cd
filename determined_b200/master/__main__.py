"""``python -m determined_b200.master`` — run the single-box master (reference
``master/cmd/determined-master`` + ``master.yaml``, ``master/internal/config/config.go``)."""
import argparse
import logging
import os
import signal
import sys
import threading

import yaml

from determined_b200.master import api, config, core


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="determined_b200.master")
    ap.add_argument("--config-file", default=os.environ.get("DET_MASTER_CONFIG"))
    ap.add_argument("--host", default=None)
    ap.add_argument("--port", type=int, default=None)
    ap.add_argument("--db", default=None, help="sqlite file (default ~/.local/share/determined_b200/master.db)")
    ap.add_argument("--scheduler", choices=["fair_share", "priority", "round_robin"], default=None)
    ap.add_argument("--fitting-policy", choices=["best", "worst"], default=None)
    ap.add_argument("--slots", type=int, default=None, help="override detected slot count (CPU testing)")
    ap.add_argument("--agent-token", default=None, help="shared secret agents must present (default: $DET_AGENT_TOKEN, none)")
    ap.add_argument("--otel-endpoint", default=None, help="OTLP/HTTP collector (or file://path) for request traces; config: telemetry.otel_endpoint")
    ap.add_argument("--tls-cert", default=None, help="PEM certificate: serve HTTPS (config: security.tls.cert)")
    ap.add_argument("--tls-key", default=None, help="PEM private key of --tls-cert (config: security.tls.key)")
    ap.add_argument("--root", default=None, help="work directory for task run directories (config: root)")
    ap.add_argument("--print-config", action="store_true", help="print the merged configuration (secrets masked) and exit")
    ap.add_argument("--log-level", default=None)
    args = ap.parse_args(argv)
    try:
        cfg = config.load(args.config_file, flags={
            "host": args.host, "port": args.port, "db.path": args.db, "resource_manager.scheduler.type": args.scheduler,
            "resource_manager.scheduler.fitting_policy": args.fitting_policy, "security.agent_token": args.agent_token,
            "telemetry.otel_endpoint": args.otel_endpoint, "security.tls.cert": args.tls_cert, "security.tls.key": args.tls_key,
            "log.level": args.log_level.lower() if args.log_level else None, "root": args.root})
    except config.MasterConfigError as e:
        print(e, file=sys.stderr)
        return 2
    if args.print_config:
        print(yaml.safe_dump(cfg.printable(), default_flow_style=False, sort_keys=True))
        return 0
    level = {"trace": "DEBUG", "warn": "WARNING", "fatal": "CRITICAL"}.get(cfg["log.level"], cfg["log.level"].upper())
    logging.basicConfig(level=getattr(logging, level), format="%(asctime)s %(levelname)s: %(name)s: %(message)s")
    sch = cfg["resource_manager.scheduler"]
    m = core.Master(db_path=cfg.db_path, host=cfg["host"], port=cfg["port"], scheduler_type=sch["type"], fitting_policy=sch["fitting_policy"],
                    preemption=bool(sch["preemption"]["enabled"]), num_slots=args.slots, work_dir=cfg["root"],
                    cluster_defaults=cfg.cluster_defaults(), resource_pools=cfg["resource_pools"] or None)
    m.apply_config(cfg)
    if cfg["telemetry.otel_endpoint"] and cfg["telemetry.otel_enabled"]:
        from determined_b200.common import tracing

        tracing.configure("determined-master", cfg["telemetry.otel_endpoint"])
    srv = api.APIServer(m, cfg["host"], cfg["port"], tls_cert=cfg["security.tls.cert"], tls_key=cfg["security.tls.key"]).start()
    m.start()
    logging.getLogger("determined.master").info("master %s (%s) listening on %s (%d slots, scheduler=%s)", m.cluster_id, m.cluster_name, m.url,
                                                 m.pool.agents[m.agent_id].num_slots, m.pool.scheduler)
    stop = threading.Event()
    for s in (signal.SIGINT, signal.SIGTERM):
        signal.signal(s, lambda *_: stop.set())
    stop.wait()
    m.stop()
    srv.stop()
    return 0


if __name__ == "__main__":
    sys.exit(main())
