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


def it_file_vars(raw):
    """the int32 / ivec variables of an IT++ it_file (itpp/itpp/base/itfile.h:98-125) as numpy values"""
    import struct
    import numpy as np
    out, off = {}, 5
    while off + 24 <= len(raw):
        hb, db, bb = struct.unpack("<3Q", raw[off:off + 24])
        name, typ = raw[off + 24:off + hb].split(b"\0")[:2]
        if typ == b"ivec":
            out[name.decode()] = np.frombuffer(raw, "<i4", (db - 8) // 4, off + hb + 8).copy()
        elif typ == b"int32":
            out[name.decode()] = int(np.frombuffer(raw, "<i4", 1, off + hb)[0])
        off += bb
    return out


def stored_codec(fname):
    import lzma
    return lzma.open(os.path.join(BP_DIR, fname + ".xz")).read()


def graph(name):
    g = BP_MANIFEST[name]["graph"]
    if g.startswith("file:"):  # a bp_codec.it: edges of variable v in the stored order, jind = v + k * nvar
        import numpy as np
        v = it_file_vars(stored_codec(g[5:]))
        dv, dc, nvar, nchk = v["sumX1"], v["sumX2"], v["nvar"], v["ncheck"]
        off = np.concatenate([[0], np.cumsum(dv)])
        idx = [off[v["jind"][j + nchk * ip] % nvar] + v["jind"][j + nchk * ip] // nvar for j in range(nchk) for ip in range(dc[j])]
        return dict(nvar=nvar, nchk=nchk, num_edges=int(dv.sum()), nchk_lin_indep=v["ncheck_lin_indep"], dv_vec=dv.astype(np.int32),
                    dc_vec=dc.astype(np.int32), cn_msg_idx=np.array(idx, np.int32), codec_file=g[5:])
    if g.startswith("synth:"):
        return _alist.graph_from_rows(*_alist.synth_rows(g[6:]))
    return _alist.bundle_graph(g)
