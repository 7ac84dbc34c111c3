"""Host logic on CPU: the micro-batching queue and the round-robin dispatcher, driven with a recording
stand-in for the GPU worker (test-only; the product has no CPU worker)."""
import threading
import time

import numpy as np
import pytest

from mlmicroservicetemplate_b200.runtime import Engine, image_geometry


class RecordingWorker:
    """Same duck type as GpuReplicaWorker.run_batch: returns (probs[n,k], idx[n,k]); idx echoes pixel (0,0,0)."""

    def __init__(self, delay=0.0, fail_on=None):
        self.batches = []
        self.delay = delay
        self.fail_on = fail_on
        self.closed = False

    def run_batch(self, images):
        self.batches.append(len(images))
        if self.delay:
            time.sleep(self.delay)
        if self.fail_on is not None and any(int(im[0, 0, 0]) == self.fail_on for im in images):
            raise RuntimeError("worker exploded")
        n = len(images)
        idx = np.array([[int(im[0, 0, 0])] for im in images], dtype=np.int32)
        return np.ones((n, 1), np.float32), idx

    def close(self):
        self.closed = True


def _img(tag):
    a = np.zeros((4, 4, 3), np.uint8)
    a[0, 0, 0] = tag
    return a


def test_each_request_gets_its_own_result_in_order():
    w = RecordingWorker()
    eng = Engine([w], max_batch=8, max_wait_us=20000)
    out = eng.predict_many([_img(i) for i in range(20)])
    assert [int(idx[0]) for _, idx in out] == list(range(20))
    assert sum(w.batches) == 20 and max(w.batches) <= 8
    eng.close()
    assert w.closed


def test_concurrent_clients_are_coalesced_into_batches():
    w = RecordingWorker(delay=0.02)
    eng = Engine([w], max_batch=16, max_wait_us=50000)
    res = [None] * 64
    def client(i):
        res[i] = eng.submit(_img(i)).result(timeout=10)
    ts = [threading.Thread(target=client, args=(i,)) for i in range(64)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert [int(r[1][0]) for r in res] == list(range(64))
    assert max(w.batches) > 1, "concurrent requests must share batches"
    assert max(w.batches) <= 16 and sum(w.batches) == 64
    eng.close()


def test_round_robin_over_replicas():
    ws = [RecordingWorker(delay=0.005) for _ in range(4)]
    eng = Engine(ws, max_batch=2, max_wait_us=100)
    for i in range(24):                       # one at a time -> 24 single-image batches
        eng.submit(_img(i)).result(timeout=10)
    order = [w for w, _ in eng.dispatch_log]
    assert order == [i % 4 for i in range(len(order))], "whole batches go to replicas in strict rotation"
    assert all(len(w.batches) == len(order) // 4 for w in ws)
    eng.close()


def test_max_wait_bounds_latency_of_a_lone_request():
    eng = Engine([RecordingWorker()], max_batch=64, max_wait_us=5000)
    t0 = time.perf_counter()
    eng.submit(_img(1)).result(timeout=5)
    assert time.perf_counter() - t0 < 0.5
    eng.close()


def test_worker_failure_reaches_every_waiter_and_engine_survives():
    eng = Engine([RecordingWorker(fail_on=7)], max_batch=4, max_wait_us=30000)
    futs = [eng.submit(_img(t)) for t in (5, 6, 7, 8)]
    errs = 0
    for f in futs:
        try:
            f.result(timeout=5)
        except RuntimeError:
            errs += 1
    assert errs >= 1
    assert int(eng.submit(_img(3)).result(timeout=5)[1][0]) == 3      # next batch is fine
    eng.close()


def test_close_drains_queued_requests():
    w = RecordingWorker(delay=0.01)
    eng = Engine([w], max_batch=4, max_wait_us=1000)
    futs = [eng.submit(_img(i)) for i in range(12)]
    eng.close()
    assert all(f.done() for f in futs) and [int(f.result()[1][0]) for f in futs] == list(range(12))
    with pytest.raises(RuntimeError):
        eng.submit(_img(0))


def test_geometry_modes():
    assert image_geometry(480, 640, "direct") == (224, 224, 0, 0)
    assert image_geometry(480, 640, "preset") == (256, 341, 16, 58)
    with pytest.raises(ValueError):
        image_geometry(1, 1, "nope")


class PipelinedWorker(RecordingWorker):
    """Stand-in with the asynchronous duck type of GpuReplicaWorker: submit_batch() returns a ticket, collect()
    waits for it; records how many tickets were outstanding at each submit."""
    DEPTH = 2

    def __init__(self, delay=0.0):
        super().__init__(delay=delay)
        self.outstanding = 0
        self.max_outstanding = 0

    def submit_batch(self, images):
        self.outstanding += 1
        self.max_outstanding = max(self.max_outstanding, self.outstanding)
        assert self.outstanding <= self.DEPTH, "the engine must not exceed the worker's pipeline depth"
        return (time.perf_counter() + self.delay, list(images))

    def collect(self, ticket):
        ready, images = ticket
        time.sleep(max(0.0, ready - time.perf_counter()))
        self.outstanding -= 1
        self.batches.append(len(images))
        idx = np.array([[int(im[0, 0, 0])] for im in images], dtype=np.int32)
        return np.ones((len(images), 1), np.float32), idx, {}


def test_pipelined_worker_gets_the_next_batch_before_the_previous_one_is_collected():
    w = PipelinedWorker(delay=0.02)
    eng = Engine([w], max_batch=4, max_wait_us=500)
    futs = [eng.submit(_img(i)) for i in range(40)]
    assert [int(f.result(timeout=10)[1][0]) for f in futs] == list(range(40))
    assert w.max_outstanding == 2, "batch i+1 must be submitted while batch i is still in flight"
    eng.close()
    assert w.outstanding == 0


def test_least_loaded_dispatch_routes_around_a_slow_replica():
    slow, fast = RecordingWorker(delay=0.05), RecordingWorker(delay=0.001)
    eng = Engine([slow, fast], max_batch=1, max_wait_us=50)
    futs = []
    for i in range(60):
        futs.append(eng.submit(_img(i)))
        time.sleep(0.002)
    assert [int(f.result(timeout=20)[1][0]) for f in futs] == list(range(60))
    assert len(fast.batches) > 2 * len(slow.batches), (len(fast.batches), len(slow.batches))
    eng.close()


def test_a_request_cancelled_while_queued_does_not_poison_its_batch():
    w = RecordingWorker(delay=0.05)
    eng = Engine([w], max_batch=4, max_wait_us=1000)
    first = eng.submit(_img(1))               # occupies the worker
    time.sleep(0.01)
    futs = [eng.submit(_img(t)) for t in (2, 3, 4)]
    assert futs[1].cancel()                   # what asyncio.wrap_future does when the handler task is cancelled
    assert int(first.result(timeout=5)[1][0]) == 1
    assert int(futs[0].result(timeout=5)[1][0]) == 2 and int(futs[2].result(timeout=5)[1][0]) == 4
    assert futs[1].cancelled()
    eng.close()


def test_whole_batches_bypass_the_micro_batcher_and_spread_over_replicas():
    ws = [PipelinedWorker(delay=0.01) for _ in range(4)]
    eng = Engine(ws, max_batch=8, max_wait_us=1000)
    futs = [eng.submit_batch([_img(10 * b + j) for j in range(8)]) for b in range(16)]
    for b, f in enumerate(futs):
        probs, idx, errors = f.result(timeout=10)
        assert [int(v[0]) for v in idx] == [10 * b + j for j in range(8)] and not errors
    assert all(sum(w.batches) == 32 for w in ws), [w.batches for w in ws]
    assert [w for w, _ in eng.dispatch_log[:4]] == [0, 1, 2, 3]
    eng.close()
