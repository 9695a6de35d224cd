"""The REST load walk (benchmarks/api_load.py: the read endpoints the reference's k6 suite visits) answers 200 everywhere, and
a request on a kept-alive connection is not held back by Nagle's algorithm."""
import pathlib
import sys
import time

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[1]))


def test_every_walked_endpoint_answers_and_requests_are_not_delayed():
    from benchmarks import api_load

    res = api_load.run(vus=2, seconds=2.0, experiments=3, reports=10)
    bad = {n: r["first_failure"] for n, r in res["endpoints"].items() if r["failures"]}
    assert not bad, bad
    assert all(r["count"] > 0 for r in res["endpoints"].values()) and len(res["endpoints"]) >= 55
    # headers and body leave in two writes; with Nagle on, every request on a kept-alive connection took ~40 ms
    assert res["endpoints"]["get master"]["p50_ms"] < 25, res["endpoints"]["get master"]
