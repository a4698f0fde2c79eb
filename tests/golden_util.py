"""Helpers shared by the golden replays (CPU oracle and CUDA path)."""
import json
import os

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    with open(os.path.join(GOLDEN, name + ".json")) as f:
        return json.load(f)


def canon_json(c):
    """Canonical dump -> the JSON shape tools/make_golden.py wrote."""
    d = {"nodes": [[list(x) if isinstance(x, tuple) else x for x in row] for row in c["nodes"]],
         "evictable": c["evictable"], "protected": c["protected"]}
    if "dup" in c:
        d["dup"] = sorted([[list(k[0]), k[1], v] for k, v in c["dup"].items()])
    return d
