"""Registration heartbeat towards the upstream PhotoAnalysisServer (control plane; nothing on the
predict path depends on it).

Wire behaviour of reference src/server/server_connection.py:10-34: every WAIT_TIME seconds, until the
server shuts down, POST {"name": <model name>, "socket": "http://host.docker.internal:<model port>"}
with header `api_key` to http://host.docker.internal:<server port>/model/register; `dependency.connected`
records whether the last attempt succeeded.  The key is read from the API_KEY environment variable
rather than from the reference's root-level `secrets.py`, which shadows the standard library module of
the same name.
"""
import os
import time

import requests
from fastapi.logger import logger

from . import dependency as dep

HOST = "http://host.docker.internal"
_RETRYABLE = (requests.exceptions.ConnectionError, requests.exceptions.Timeout, requests.exceptions.HTTPError)


def _attempt(url: str, payload: dict) -> bool:
    try:
        resp = requests.post(url, headers={"api_key": os.getenv("API_KEY", "paste_your_api_key_here")},
                             json=payload, timeout=5)
        resp.raise_for_status()
        return True
    except _RETRYABLE:
        return False


def _sleep_unless_shutdown(seconds: int) -> None:
    for _ in range(seconds):          # one-second slices so shutdown is noticed promptly
        if dep.shutdown:
            return
        time.sleep(1)


def register_model_to_server(server_port, model_port, model_name):
    url = f"{HOST}:{server_port}/model/register"
    payload = {"name": model_name, "socket": f"{HOST}:{model_port}"}
    while not dep.shutdown:
        dep.connected = _attempt(url, payload)
        if not dep.connected:
            logger.debug(f"Registering to server fails. Retry in {dep.WAIT_TIME} seconds")
        _sleep_unless_shutdown(dep.WAIT_TIME)
    logger.debug("[Healthcheck] Server Registration Thread Halted.")
