"""Process-wide server state shared by the routes, the start-up hook and the registration thread.

Same names and meaning as reference src/server/dependency.py (:6-20) — `model_settings.ready_to_predict`,
`PredictionException`, `connected`, `shutdown`, `pool`, `WAIT_TIME` — because src/server/main.py and
server_connection.py address them by name.  The reference builds its settings object on pydantic-v1's
BaseSettings with an un-annotated field, which pydantic 2 rejects; a slotted plain object carries the
one flag here instead.
"""
import concurrent.futures as _futures

WAIT_TIME = 10          # seconds between registration attempts
POOL_THREADS = 10       # init() and the registration loop run on this pool, never on the event loop


class _ReadyFlag:
    """`ready_to_predict` goes False -> True when init() returns and back to False at shutdown."""
    __slots__ = ("ready_to_predict",)

    def __init__(self) -> None:
        self.ready_to_predict = False


class PredictionException(Exception):
    """Raised by a route when the model is not ready; the app maps it to HTTP 503."""


Settings = _ReadyFlag
model_settings = _ReadyFlag()
pool = _futures.ThreadPoolExecutor(max_workers=POOL_THREADS, thread_name_prefix="b2p-server")
shutdown = False        # set at server shutdown; the registration loop polls it
connected = False       # last registration attempt succeeded
