"""HTTP front end — drop-in for reference src/server/main.py; the wire contract is unchanged:

    GET  /         200 ["MLMicroserviceTemplate is Running!"]                            (ref main.py:52-58)
    GET  /status   200 {"status":"success","detail":"Model ready to receive prediction requests."}
                   503 {"status":"failure","detail":"Model is not ready to receive predictions."}
                                                                                          (ref main.py:41-49, 101-116)
    POST /predict  multipart part `image_file` -> 200 {"status":"success","result": <predict() dict>};
                   503 until init() has finished, 422 without the part, 500 on undecodable bytes
                                                                                          (ref main.py:119-140)
    POST /status   the reference README (:143, :154) documents it as "runs init" but the code has no
                   such route (405).  Here it is an idempotent "make sure init() has been started" that
                   answers exactly like GET /status — additive, no existing response changes.
    CORS           same origins / methods / headers as ref main.py:22-38.

Behavioural change on the data path only: the reference awaits nothing and calls predict() inline in
the async handler (ref main.py:135), so requests are served strictly one at a time.  This handler awaits
the plugin's predict_async(): decode runs on a worker thread and the request parks on the
micro-batching queue, so concurrent clients are coalesced into GPU batches.
"""
import contextlib
import io
import logging
import os
import threading

from dotenv import load_dotenv
from fastapi import FastAPI, File, Request, UploadFile
from fastapi.middleware.cors import CORSMiddleware
from starlette.responses import JSONResponse, Response

from ..model import model as plugin
from . import dependency as dep
from .server_connection import register_model_to_server

logger = logging.getLogger("api")

NOT_READY = {"status": "failure", "detail": "Model is not ready to receive predictions."}
READY = {"status": "success", "detail": "Model ready to receive prediction requests."}
ALLOWED_ORIGINS = ["http://localhost"] + [f"http://localhost:{p}" for p in (3000, 5057, 5000, 6379)]

_init_once = threading.Lock()
_init_submitted = False


def _init_then_register() -> None:
    """Pool-thread body: plugin.init(), flip the readiness flag, then start the heartbeat."""
    logger.debug("Beginning Model Initialization Process.")
    plugin.init()
    dep.model_settings.ready_to_predict = True
    logger.debug("Finishing Model Initialization Process.")
    if os.getenv("REGISTER", "1") != "0":
        dep.pool.submit(register_model_to_server, os.getenv("SERVER_PORT"), os.getenv("PORT"), os.getenv("NAME"))


def ensure_init_started() -> None:
    global _init_submitted
    with _init_once:
        if not _init_submitted:
            _init_submitted = True
            dep.pool.submit(_init_then_register)


@contextlib.asynccontextmanager
async def lifespan(_: FastAPI):
    # start-up: answer HTTP immediately, initialise the model in the background
    load_dotenv()
    dep.shutdown = False
    ensure_init_started()
    yield
    # shutdown: stop accepting predictions, stop the heartbeat, drain the batch queue, free the replicas
    dep.model_settings.ready_to_predict = False
    dep.shutdown = True
    plugin.shutdown()
    dep.pool.shutdown(wait=False)


class _Upload:
    """What predict() needs of an upload (ref model.py:23 reads `image_file.file`), without a spooled temp file."""
    __slots__ = ("file", "filename", "content_type")

    def __init__(self, data: bytes, filename: str, content_type: str):
        self.file = io.BytesIO(data)
        self.filename = filename
        self.content_type = content_type


def single_image_part(body: bytes, content_type: str):
    """The request every client of this service sends - a multipart body with ONE part, `image_file`, carrying a file -
    cut out with two substring searches.  -> (file bytes, filename, part content type) or None for anything else
    (other or additional fields, RFC 5987 filenames, malformed framing ...), which takes the framework's parser."""
    if content_type[:19].lower() != "multipart/form-data":
        return None
    i = content_type.find("boundary=")
    if i < 0:
        return None
    boundary = content_type[i + 9:].split(";", 1)[0].strip()
    if len(boundary) >= 2 and boundary[0] == boundary[-1] == '"':
        boundary = boundary[1:-1]
    if not boundary:
        return None
    delim = b"--" + boundary.encode("latin-1")
    h0 = len(delim) + 2
    if body[:h0] != delim + b"\r\n":
        return None
    h1 = body.find(b"\r\n\r\n", h0, h0 + 1024)
    if h1 < 0:
        return None
    filename = part_type = None
    named = False
    for line in body[h0:h1].decode("latin-1").split("\r\n"):
        key, _, value = line.partition(":")
        key = key.strip().lower()
        if key == "content-disposition":
            fields = [f.strip() for f in value.split(";")]
            if not fields or fields[0].lower() != "form-data":
                return None
            for f in fields[1:]:
                k, _, v = f.partition("=")
                if len(v) < 2 or v[0] != '"' or v[-1] != '"' or "\\" in v or "%" in v:
                    return None                     # unquoted / escaped / encoded values: leave them to the real parser
                if k == "name":
                    named = v[1:-1] == "image_file"
                elif k == "filename":
                    filename = v[1:-1]
                else:
                    return None
        elif key == "content-type":
            part_type = value.strip()
        elif key:
            return None
    if not named or filename is None:
        return None
    end = body.find(b"\r\n" + delim, h1 + 4)
    if end < 0 or body[end + 2 + len(delim):] not in (b"--\r\n", b"--"):
        return None                                 # more parts, or no closing delimiter
    return body[h1 + 4:end], filename, part_type or "application/octet-stream"


class FastPredict:
    """ASGI layer in front of the router for POST /predict.  The framework's path costs about a millisecond of Python
    per upload (python-multipart's streaming parser into a spooled temp file, then dependency resolution), more than
    half of this process's time per request; the common request needs neither.  Anything single_image_part() does not
    recognise - and every request while the model is not ready - is REPLAYED to the router unchanged, so validation
    errors, 503s and odd clients get byte-identical answers to before."""

    def __init__(self, app):
        self.app = app

    async def __call__(self, scope, receive, send):
        if scope["type"] != "http" or scope["method"] != "POST" or scope["path"] != "/predict":
            return await self.app(scope, receive, send)
        chunks, more, disconnected = [], True, False
        while more:
            msg = await receive()
            if msg["type"] != "http.request":
                disconnected = True
                break
            chunks.append(msg.get("body", b""))
            more = msg.get("more_body", False)
        body = chunks[0] if len(chunks) == 1 else b"".join(chunks)
        part = None
        if not disconnected and dep.model_settings.ready_to_predict:
            ctype = ""
            for k, v in scope["headers"]:
                if k == b"content-type":
                    ctype = v.decode("latin-1")
                    break
            part = single_image_part(body, ctype)
        if part is None:
            sent = False

            async def replay():
                nonlocal sent
                if disconnected:
                    return {"type": "http.disconnect"}
                if not sent:
                    sent = True
                    return {"type": "http.request", "body": body, "more_body": False}
                return await receive()

            return await self.app(scope, replay, send)
        out = await plugin.predict_async_json(_Upload(*part))
        response = Response(content=b'{"status":"success","result":' + out + b'}', media_type="application/json")
        await response(scope, receive, send)


app = FastAPI(lifespan=lifespan)
if os.getenv("FAST_PREDICT", "1") != "0":
    app.add_middleware(FastPredict)          # inside CORS (added last = outermost), so its answers get the same headers
app.add_middleware(CORSMiddleware, allow_origins=ALLOWED_ORIGINS, allow_credentials=True,
                   allow_methods=["*"], allow_headers=["*"])


@app.exception_handler(dep.PredictionException)
async def _not_ready(request: Request, exc: dep.PredictionException):
    return JSONResponse(status_code=503, content=NOT_READY)


def _require_ready() -> None:
    if not dep.model_settings.ready_to_predict:
        raise dep.PredictionException()


@app.get("/")
async def root():
    return ["MLMicroserviceTemplate is Running!"]


@app.get("/status")
async def check_status():
    _require_ready()
    return READY


@app.post("/status")
async def start_init_and_report_status():
    ensure_init_started()
    _require_ready()
    return READY


@app.post("/predict")
async def create_prediction(image_file: UploadFile = File(...)):
    _require_ready()
    # same JSON text as returning {"status": "success", "result": <predict dict>}, without the per-request encoder walk
    body = await plugin.predict_async_json(image_file)
    return Response(content=b'{"status":"success","result":' + body + b'}', media_type="application/json")
