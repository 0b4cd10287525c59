#!/usr/bin/env python3
"""Stand-in for the reference's agent loop (app.py:245-318) on boxes where /root/reference does not
exist (the GPU box). TEST INFRASTRUCTURE: it speaks the same wire protocol to a controller —
POST /v1/leases (204 = idle, 200 = {lease_id, tasks}), first task only, op looked up in a two-entry
table whose classify entry is `tpu_ops.map_classify_tpu` called with one positional dict, outcome
posted to /v1/results — so that the op is driven the way the unmodified agent drives it. Where the
reference checkout exists the tests run the real app.py instead.

    CONTROLLER_URL=http://127.0.0.1:PORT python agent_loop_standin.py [max_seconds]
"""
import os
import sys
import time
import traceback

import requests

from tpu_ops import map_classify_tpu

BASE = os.environ["CONTROLLER_URL"].rstrip("/")
TABLE = {"echo": lambda payload: {"ok": True, "echo": payload}, "map_classify_tpu": map_classify_tpu}


def main(max_seconds: float) -> None:
    deadline = time.time() + max_seconds
    handled = 0
    while time.time() < deadline:
        ask = {"agent": os.environ.get("AGENT_NAME", "standin"), "capabilities": {"ops": sorted(TABLE)},
               "max_tasks": 1, "timeout_ms": 3000, "labels": {}, "worker_profile": {}, "metrics": {}}
        reply = requests.post(BASE + "/v1/leases", json=ask, timeout=10)
        if reply.status_code == 204:
            if handled:      # the test controller hands out one task: nothing more will come
                return
            time.sleep(0.05)
            continue
        lease = reply.json()
        task = lease["tasks"][0]
        report = {"lease_id": lease["lease_id"], "job_id": task.get("id") or task.get("job_id"),
                  "job_epoch": task.get("job_epoch"), "status": "succeeded", "result": None, "error": None}
        try:
            report["result"] = TABLE[task["op"]](task.get("payload") or {})
        except Exception as exc:  # what app.py:288-294 turns into status "failed"
            report["status"] = "failed"
            report["error"] = {"type": type(exc).__name__, "message": str(exc),
                               "trace": traceback.format_exc(limit=12)}
        requests.post(BASE + "/v1/results", json=report, timeout=10)
        handled += 1


if __name__ == "__main__":
    main(float(sys.argv[1]) if len(sys.argv) > 1 else 4.0)
