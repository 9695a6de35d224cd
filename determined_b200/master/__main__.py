"""``python -m determined_b200.master`` — run the single-box master (reference
``master/cmd/determined-master`` + ``master.yaml``, ``master/internal/config/config.go``)."""
import argparse
import logging
import os
import signal
import sys
import threading

import yaml

from determined_b200.master import api, core


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
    ap.add_argument("--log-level", default="INFO")
    args = ap.parse_args(argv)
    logging.basicConfig(level=getattr(logging, args.log_level.upper()), format="%(asctime)s %(levelname)s: %(name)s: %(message)s")
    cfg = {}
    if args.config_file:
        with open(args.config_file) as f:
            cfg = yaml.safe_load(f) or {}
    rm = (cfg.get("resource_manager") or {})
    sch = rm.get("scheduler") or {}
    host = args.host or cfg.get("host") or "127.0.0.1"
    port = args.port or int(cfg.get("port") or 8080)
    dbp = args.db or (cfg.get("db") or {}).get("path") or os.path.expanduser("~/.local/share/determined_b200/master.db")
    defaults = {}
    if cfg.get("checkpoint_storage"):
        defaults["checkpoint_storage"] = cfg["checkpoint_storage"]
    m = core.Master(db_path=dbp, host=host, port=port, scheduler_type=args.scheduler or sch.get("type", "fair_share"),
                    fitting_policy=args.fitting_policy or sch.get("fitting_policy", "best"),
                    preemption=bool((sch.get("preemption") or {}).get("enabled", True)), num_slots=args.slots,
                    cluster_defaults=defaults, resource_pools=cfg.get("resource_pools") or None)
    m.agent_token = args.agent_token or (cfg.get("security") or {}).get("agent_token")
    tel = cfg.get("telemetry") or {}
    otel = args.otel_endpoint or (tel.get("otel_endpoint") if tel.get("otel_enabled", True) else None)
    if otel:
        from determined_b200.common import tracing

        tracing.configure("determined-master", otel)
    tls = (cfg.get("security") or {}).get("tls") or {}
    m.tls_cert_name = tls.get("name")
    srv = api.APIServer(m, host, port, tls_cert=args.tls_cert or tls.get("cert"), tls_key=args.tls_key or tls.get("key")).start()
    m.start()
    logging.getLogger("determined.master").info("master %s listening on %s (%d slots, scheduler=%s)", m.cluster_id, m.url,
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
