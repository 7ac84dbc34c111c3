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


app = FastAPI(lifespan=lifespan)
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
