"""N>1 plumbing on CPU: two gloo ranks exercise the barrier / max-over-ranks / sharding used by
`bench.py --gpus N` (one process per GPU, independent replicas, no data-path collective)."""
import json
import os
import socket
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import json, os, sys
    sys.path.insert(0, %r)
    from mlmicroservicetemplate_b200.ranks import Ranks
    r = Ranks.from_env(backend="gloo")
    r.barrier()
    elapsed = 10.0 if r.rank == 0 else 25.0            # rank 1 is the slow one
    out = {"rank": r.rank, "world": r.world, "max": r.max_over_ranks(elapsed),
           "value": r.job_throughput(256 * 4, elapsed), "shard": list(r.shard(7)),
           "same": r.same_on_all_ranks(b"identical-digest"), "differ": r.same_on_all_ranks(bytes([r.rank]) * 8)}
    r.barrier()
    r.close()
    print("RESULT " + json.dumps(out))
""" % ROOT)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def test_two_gloo_ranks():
    port = str(_free_port())
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=port)
        procs.append(subprocess.Popen([sys.executable, "-c", WORKER], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        o, e = p.communicate(timeout=120)
        assert p.returncode == 0, e[-2000:]
        outs.append(json.loads([l for l in o.splitlines() if l.startswith("RESULT ")][0][7:]))
    outs.sort(key=lambda d: d["rank"])
    assert [d["world"] for d in outs] == [2, 2]
    assert all(d["max"] == 25.0 for d in outs), "every rank must see the slowest rank's time"
    assert all(abs(d["value"] - 2 * 1024 / 0.025) < 1e-6 for d in outs), "value = all ranks' units / max time"
    assert outs[0]["shard"] + outs[1]["shard"] == list(range(7)), "shards partition the work with no overlap"
    assert all(d["same"] for d in outs) and not any(d["differ"] for d in outs), "digest comparison across ranks"


def test_single_rank_needs_no_process_group():
    from mlmicroservicetemplate_b200.ranks import Ranks
    env = {k: os.environ.pop(k) for k in ("WORLD_SIZE", "RANK", "LOCAL_RANK") if k in os.environ}
    try:
        r = Ranks.from_env(backend="gloo")
        assert (r.world, r.rank) == (1, 0) and r.max_over_ranks(3.5) == 3.5 and list(r.shard(5)) == [0, 1, 2, 3, 4]
        r.barrier(); r.close()
    finally:
        os.environ.update(env)
