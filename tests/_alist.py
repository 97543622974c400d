"""TEST HELPER: the parity-check matrix of a golden decoder bundle as an alist file (itpp/itpp/base/gf2mat.cpp:57-150:
1-based, `N M`, max weights, column weights, row weights, column lists, row lists, zero padded).  The bundle holds the
graph as decoder_parameterization flattened it (src/LDPC_Code_LUT.cpp:513-528): edges VN-major, cn_msg_idx = edges of
each check.  /root/reference (where the original alist files live) does not exist on the GPU box."""
import numpy as np

import _golden
import pyoracle


def write_alist(cfg, path, pad=0):
    orc = pyoracle.Oracle(_golden.bundle_path(cfg))
    dv, dc, idx = orc.dv, orc.dc, orc.cn_msg_idx
    edge_vn = np.repeat(np.arange(orc.nvar), dv)
    cs = np.concatenate([[0], np.cumsum(dc)])
    rows = [np.sort(edge_vn[idx[cs[c]:cs[c + 1]]]) for c in range(orc.nchk)]
    cols = [[] for _ in range(orc.nvar)]
    for c, r in enumerate(rows):
        for v in r:
            cols[v].append(c)
    mc, mr = int(dv.max()), int(dc.max())
    with open(path, "w") as f:
        f.write("%d %d\n%d %d\n" % (orc.nvar, orc.nchk, mc, mr))
        f.write(" ".join(str(int(x)) for x in dv) + "\n" + " ".join(str(int(x)) for x in dc) + "\n")
        for cl in cols:
            f.write(" ".join(str(c + 1) for c in cl + [pad - 1] * (mc - len(cl))) + "\n")
        for r in rows:
            f.write(" ".join(str(int(v) + 1) for v in list(r) + [pad - 1] * (mr - len(r))) + "\n")
    return orc
