"""SURVEY §8(f) rank 4 — the registration heartbeat towards the upstream PhotoAnalysisServer, re-expressed from
reference src/server/server_connection.py:10-34.  A fake `requests.post` records what would go on the wire:
URL (:15), `api_key` header (:16), JSON body (:17-20), the `connected` flag (:22-26) and that the loop stops when the
server shuts down (:13, :30-33)."""
import threading
import time

import requests

from mlmicroservicetemplate_b200.src.server import dependency as dep
from mlmicroservicetemplate_b200.src.server import server_connection as sc


class _Resp:
    def __init__(self, status):
        self.status_code = status

    def raise_for_status(self):
        if self.status_code >= 400:
            raise requests.exceptions.HTTPError(f"{self.status_code}")


def _run_heartbeat(monkeypatch, post, rounds, api_key="k-123"):
    calls = []
    seen = threading.Event()

    def fake_post(url, headers=None, json=None, timeout=None, **kw):
        calls.append({"url": url, "headers": dict(headers or {}), "json": json, "timeout": timeout})
        try:
            return post(len(calls))
        finally:
            if len(calls) >= rounds:
                seen.set()

    monkeypatch.setattr(sc.requests, "post", fake_post)
    monkeypatch.setattr(sc, "_sleep_unless_shutdown", lambda seconds: time.sleep(0.01))
    monkeypatch.setenv("API_KEY", api_key)
    monkeypatch.setattr(dep, "shutdown", False)
    monkeypatch.setattr(dep, "connected", False)
    t = threading.Thread(target=sc.register_model_to_server, args=("5000", "5005", "example_model"), daemon=True)
    t.start()
    assert seen.wait(5), "the heartbeat never posted"
    return t, calls


def test_heartbeat_posts_the_reference_wire_format_and_tracks_connected(monkeypatch):
    t, calls = _run_heartbeat(monkeypatch, lambda i: _Resp(200), rounds=3)
    assert dep.connected is True
    dep.shutdown = True                      # what the server's shutdown hook sets (reference main.py:88-98)
    t.join(5)
    assert not t.is_alive(), "the registration loop must stop at shutdown"
    c = calls[0]
    assert c["url"] == "http://host.docker.internal:5000/model/register"
    assert c["headers"] == {"api_key": "k-123"}
    assert c["json"] == {"name": "example_model", "socket": "http://host.docker.internal:5005"}
    assert len(calls) >= 3, "it keeps re-registering every WAIT_TIME seconds, as the reference does"


def test_heartbeat_marks_disconnected_on_failure_and_keeps_retrying(monkeypatch):
    def post(i):
        if i == 1:
            raise requests.exceptions.ConnectionError("refused")
        if i == 2:
            return _Resp(503)
        return _Resp(200)

    flags = []
    orig_attempt = sc._attempt

    def spy(url, payload):
        ok = orig_attempt(url, payload)
        flags.append(ok)
        return ok

    monkeypatch.setattr(sc, "_attempt", spy)
    t, calls = _run_heartbeat(monkeypatch, post, rounds=3)
    time.sleep(0.05)
    dep.shutdown = True
    t.join(5)
    assert not t.is_alive()
    assert flags[:3] == [False, False, True]
    assert dep.connected is True


def test_sleep_is_cut_short_by_shutdown(monkeypatch):
    monkeypatch.setattr(dep, "shutdown", True)
    t0 = time.perf_counter()
    sc._sleep_unless_shutdown(dep.WAIT_TIME)
    assert time.perf_counter() - t0 < 0.5
