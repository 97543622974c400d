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


# ---- deterministic synthetic parity-check matrices for the BP comparison decoder: degree mixes no shipped alist has
def synth_rows(kind):
    """(nvar, rows): `mixed` = check degrees 2..9 and 40 over 601 variables (601 is prime, so any stride gives distinct
    columns), variable degrees from 1 up; `wide` = (3,48)-regular, 960 variables, three affine interleavers."""
    if kind == "mixed":
        N, rows = 601, []
        degs = [2, 3, 4, 5, 6, 7, 8, 9, 6, 5, 3, 40]
        for j in range(300):
            d = degs[j % len(degs)]
            start, stride = (j * 37) % N, 1 + (j * 7) % 97
            rows.append(sorted((start + k * stride) % N for k in range(d)))
        rows += [list(range(6 * j, 6 * j + 6)) for j in range(100)] + [[0, 600]]  # every variable is in some check
        return N, rows
    if kind == "wide":
        N, dc, rows = 960, 48, []
        for a, b in ((1, 0), (7, 3), (11, 5)):  # gcd(a, 960) = 1
            layer = [[] for _ in range(N // dc)]
            for v in range(N):
                layer[((a * v + b) % N) // dc].append(v)
            rows += [sorted(r) for r in layer]
        return N, rows
    raise KeyError(kind)


def write_alist_rows(N, rows, path):
    cols = [[] for _ in range(N)]
    for c, r in enumerate(rows):
        assert len(set(r)) == len(r)
        for v in r:
            cols[v].append(c)
    assert min(len(c) for c in cols) >= 1
    mc, mr = max(len(c) for c in cols), max(len(r) for r in rows)
    with open(path, "w") as f:
        f.write("%d %d\n%d %d\n" % (N, len(rows), mc, mr))
        f.write(" ".join(str(len(c)) for c in cols) + "\n" + " ".join(str(len(r)) for r in rows) + "\n")
        for cl in cols:
            f.write(" ".join(str(c + 1) for c in cl + [-1] * (mc - len(cl))) + "\n")
        for r in rows:
            f.write(" ".join(str(v + 1) for v in r + [-1] * (mr - len(r))) + "\n")


def graph_from_rows(N, rows):
    """dv_vec / dc_vec / cn_msg_idx with the reference's edge numbering (src/LDPC_Code_LUT.cpp:513-528)."""
    dv = np.zeros(N, np.int32)
    for r in rows:
        for v in r:
            dv[v] += 1
    off = np.concatenate([[0], np.cumsum(dv)])
    fill = np.zeros(N, np.int64)
    idx = []
    for r in rows:  # checks ascending => within a variable ascending check index
        for v in sorted(r):
            idx.append(off[v] + fill[v])
            fill[v] += 1
    return dict(nvar=N, nchk=len(rows), num_edges=len(idx), nchk_lin_indep=-1, dv_vec=dv,
                dc_vec=np.array([len(r) for r in rows], np.int32), cn_msg_idx=np.array(idx, np.int32))


def bundle_graph(cfg):
    """the same arrays for the code of a golden decoder bundle"""
    orc = pyoracle.Oracle(_golden.bundle_path(cfg))
    return dict(nvar=orc.nvar, nchk=orc.nchk, num_edges=int(orc.dv.sum()), nchk_lin_indep=-1,
                dv_vec=orc.dv.astype(np.int32), dc_vec=orc.dc.astype(np.int32), cn_msg_idx=orc.cn_msg_idx.astype(np.int32))
