"""Generates tests/golden/*.json(.gz) from the reference's lowered-graph fixtures (run in the build container,
where /root/reference exists).  The textual files are parsed with moose_b200.textual and re-encoded as compact
JSON operation records, so the GPU box -- which has no /root/reference -- can replay them.

    python tests/golden/make_graph_fixtures.py
"""
import gzip
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from moose_b200 import textual as tx  # noqa: E402

SOURCES = {
    "dotprod_compiled": "/root/reference/tutorials/dotprod-compiled.moose",
    "dotprod_networked": "/root/reference/pymoose/examples/moose-gen/dotprod-networked.moose",
    "rep_computation": "/root/reference/moose/benches/rep_computation.moose",
}

if __name__ == "__main__":
    for name, path in SOURCES.items():
        ops = tx.parse_computation(open(path).read())
        data = tx.dumps_json(ops).encode()
        out = os.path.join(HERE, name + ".json.gz")
        with gzip.GzipFile(out, "wb", mtime=0) as f:
            f.write(data)
        print(name, len(ops), "ops ->", out, os.path.getsize(out), "bytes")
