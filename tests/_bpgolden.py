"""Access to the committed golden vectors of the BP comparison decoder (tests/golden/bp/, made by the reference:
tests/golden/make_bp_golden.py) and to the graph each one belongs to."""
import json
import os

import _alist
import _golden
import pyoracle

BP_DIR = os.path.join(_golden.GOLD, "bp")
BP_MANIFEST = json.load(open(os.path.join(_golden.GOLD, "bp_manifest.json")))
CASES = sorted(BP_MANIFEST)
SMALL = [c for c in CASES if BP_MANIFEST[c]["bytes"] < 300000]


def fixture(name):
    import hashlib
    import lzma
    raw = lzma.open(os.path.join(BP_DIR, name + ".fix.xz")).read()
    assert hashlib.sha256(raw).hexdigest() == BP_MANIFEST[name]["sha256"], "golden file corrupt: " + name
    os.makedirs(_golden.CACHE, exist_ok=True)
    dst = os.path.join(_golden.CACHE, "bp_" + name + ".fix")
    if not os.path.exists(dst):
        tmp = dst + ".tmp%d" % os.getpid()
        open(tmp, "wb").write(raw)
        os.replace(tmp, dst)
    return pyoracle.read_bp_fixture(dst)


def graph(name):
    g = BP_MANIFEST[name]["graph"]
    if g.startswith("synth:"):
        return _alist.graph_from_rows(*_alist.synth_rows(g[6:]))
    return _alist.bundle_graph(g)
