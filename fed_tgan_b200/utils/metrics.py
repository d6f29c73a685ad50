"""Per-round metrics (JSONL) -- the reference computes per-phase timings and drops them
(``Server/dtds/distributed.py:790-825``); here they are kept."""

import json
import time


class MetricsLog:
    def __init__(self, path=None):
        self.path = path
        self.rows = []
        self._extra = {}
        if path:
            open(path, "w").close()

    def add(self, key, value):
        self._extra[key] = value

    def log(self, **kw):
        row = {"t": time.time(), **self._extra, **kw}
        self._extra = {}
        self.rows.append(row)
        if self.path:
            with open(self.path, "a") as fh:
                fh.write(json.dumps(row) + "\n")
        return row
