"""Access to the committed golden files (outputs of the reference, see tests/golden/make_golden.py).
The .xz files are unpacked once into tests/golden/_cache/ (git-ignored)."""
import hashlib
import json
import lzma
import os

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
CACHE = os.path.join(GOLD, "_cache")
MANIFEST = json.load(open(os.path.join(GOLD, "manifest.json")))


def _unpack(name, want_sha):
    os.makedirs(CACHE, exist_ok=True)
    dst = os.path.join(CACHE, name)
    if not os.path.exists(dst):
        data = lzma.open(os.path.join(GOLD, name + ".xz"), "rb").read()
        assert hashlib.sha256(data).hexdigest() == want_sha, "golden file corrupt: " + name
        tmp = dst + ".tmp%d" % os.getpid()
        with open(tmp, "wb") as f:
            f.write(data)
        os.replace(tmp, dst)
    return dst


def bundle_path(cfg):
    return _unpack(cfg + ".bundle", MANIFEST[cfg]["bundle_sha256"])


def codec_path(cfg):
    """The reference's own lut_codec.it for cfg (input of oracle/_ref/lutldpc_ref sim)."""
    return _unpack(cfg + ".codec.it", MANIFEST[cfg]["codec_sha256"])


def fixture_path(cfg, tag):
    return _unpack("%s.%s.fix" % (cfg, tag), MANIFEST[cfg]["fixtures"][tag]["sha256"])


def fixture_args(cfg, tag):
    """psc/pisc/init-mode the fixture was made with."""
    a = MANIFEST[cfg]["fixtures"][tag]["args"].split()
    kv = {a[i][2:]: a[i + 1] for i in range(0, len(a) - 1) if a[i].startswith("--")}
    return dict(psc=int(kv.get("psc", 1)), pisc=int(kv.get("pisc", 1)),
                init_mode=MANIFEST[cfg]["init_mode"])


def all_fixtures(skip=()):
    return [(c, t) for c in sorted(MANIFEST) for t in sorted(MANIFEST[c]["fixtures"]) if c not in skip]


def sel_llr(cfg):
    """Channel LLRs (float64 [frames][nvar]) of the `sel` fixture of cfg: the committed input the reference decoded
    (tests/golden/<cfg>.sel.llr_i16.xz, cell centres of the 0.01 threshold grid; see make_sel_inputs.py)."""
    import numpy as np
    k = np.frombuffer(lzma.open(os.path.join(GOLD, cfg + ".sel.llr_i16.xz"), "rb").read(), "<i2")
    return ((k.astype(np.float64) + 0.5) * 0.01).reshape(-1, MANIFEST[cfg]["design_report"]["nvar"])


def read_bundle(path):
    """A decoder bundle (layout: oracle/ref_driver.cpp write_bundle) as a dict of numpy arrays / bytes."""
    import struct
    import numpy as np
    b = open(path, "rb").read()
    assert b[:8] == b"LUTLDPC1"
    h = struct.unpack("<10i", b[8:48])
    d = dict(nvar=h[0], nchk=h[1], nchk_lin_indep=h[2], num_edges=h[3], max_iters=h[4], Nq_Cha=h[5], min_lut=h[6])
    off = 48

    def arr(dt, n):
        nonlocal off
        a = np.frombuffer(b, dt, n, off).copy()
        off += a.nbytes
        return a
    d["dv_vec"], d["dc_vec"], d["cn_msg_idx"] = arr("<i4", h[0]), arr("<i4", h[1]), arr("<i4", h[3])
    d["Nq_Msg"], d["reuse_vec"], d["Nq_Cha_2_Nq_Msg_map"] = arr("<i4", h[4]), arr("<i4", h[4]), arr("<i4", h[5])
    d["qb_Cha"], d["qb_Msg"] = arr("<f8", h[5] - 1), arr("<f8", h[7])
    for k in ("var_tree_string", "chk_tree_string"):
        (n,) = struct.unpack_from("<i", b, off)
        d[k] = b[off + 4:off + 4 + n]
        off += 4 + n
    return d
