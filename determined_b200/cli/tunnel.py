"""``python -m determined_b200.cli.tunnel MASTER_URL SERVICE_ID [--listener PORT]``: tunnel a TCP connection to a task through
the master (reference ``harness/determined/cli/tunnel.py``).  Without ``--listener`` stdin / stdout carry the connection
(``ssh -o "ProxyCommand=python -m determined.cli.tunnel %h ..."``); with it, connections to the local port are forwarded."""
import argparse
import time

from determined_b200.cli import proxy
from determined_b200.common import api
from determined_b200.common.api import authentication, certs


def main(argv=None) -> int:  # noqa: ANN001
    ap = argparse.ArgumentParser(description="Tunnel through a Determined master")
    ap.add_argument("master_url", type=api.canonicalize_master_url)
    ap.add_argument("service_uuid")
    ap.add_argument("--cert-file")
    ap.add_argument("--cert-name")
    ap.add_argument("--listener", type=int)
    ap.add_argument("--remote-port", type=int, help="a port the task declared in proxy_ports (default: its service port)")
    ap.add_argument("-u", "--user")
    ap.add_argument("--auth", action="store_true")
    a = ap.parse_args(argv)
    noverify = a.cert_file == "noverify"
    cert = certs.default_load(a.master_url, None if noverify else a.cert_file, a.cert_name, noverify)
    sess = authentication.login_with_cache(a.master_url, a.user, cert=cert) if a.auth else api.UnauthSession(a.master_url, cert)
    if a.listener is not None:
        with proxy.http_tunnel_listener(sess, [proxy.ListenerConfig(a.service_uuid, a.listener, remote_port=a.remote_port)]) as ports:
            print(f"forwarding 127.0.0.1:{ports[0]} to {a.service_uuid}", flush=True)
            while True:
                time.sleep(1)
    proxy.http_connect_tunnel(sess, a.service_uuid, a.remote_port)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
