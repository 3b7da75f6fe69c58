"""ORACLE fixture generator (test infrastructure, not product code).

Runs the REFERENCE itself, in this container, and commits what it produced as small fixtures under
tests/golden/ (the reference tree does not exist on the GPU box):

  * the reference's Python integer golden model  edge-eegnet_wolf/python_utils/golden_model.py:39-113
    (GoldenModel(..., clip_balanced=False, reorder_bn=True|False), per-layer taps included), and
  * the reference's C model  edge-eegnet_wolf/src/cl/net/model.c:42  compiled for the host by
    oracle/build_ref.py (flag sets default / plain / nosimd of edge-eegnet_wolf/Makefile:24-34),
    with net.h/net.c written by the reference's own data/gen_net_header.py:49.

The exported parameter arrays stored in each fixture are parsed back from the reference-generated
net.c / net.h, i.e. they are the reference's conversion (convert_torch_format.py), not ours.

Usage:  python oracle/gen_golden.py          (needs /root/reference)
"""
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import build_ref  # noqa: E402
from mibminet_b200 import synth  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

# (C, T, seed, n_trials, distribution)
CASES = [
    (64, 480, 0, 4, "uniform"),
    (64, 480, 7, 3, "gauss"),
    (22, 1125, 1, 3, "uniform"),
    (38, 480, 2, 3, "uniform"),
    (24, 480, 4, 3, "uniform"),
    (19, 480, 5, 3, "gauss"),
    (16, 480, 6, 3, "uniform"),
    (8, 480, 3, 4, "uniform"),
]

_ARRAYS = ["l1_factor", "l1_offset", "l1_weight_align", "l2_factor", "l2_offset", "l2_weight",
           "l2_weight_reverse", "l3_weight", "l4_factor", "l4_offset", "l4_weight", "l5_bias", "l5_weight"]


def parse_net_c(dirname):
    """net.c / net.h written by the reference's gen_net_header.py -> {name: ndarray}."""
    src = open(os.path.join(dirname, "net.c")).read()
    hdr = open(os.path.join(dirname, "net.h")).read()
    out = {}
    for name in _ARRAYS:
        m = re.search(r"const (int8_t|int32_t) net_%s\[\] = \{([^}]*)\}" % name, src)
        assert m, name
        vals = [int(v) for v in m.group(2).replace("\n", " ").split(",") if v.strip()]
        out[name] = np.asarray(vals, dtype=np.int8 if m.group(1) == "int8_t" else np.int32)
    for name in ("L3_FACTOR", "L5_FACTOR"):
        out[name.lower()] = np.int32(int(re.search(r"#define NET_%s (-?\d+)" % name, hdr).group(1)))
    return out


def main():
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    build_ref._import_reference_python()
    from golden_model import GoldenModel

    for C, T, seed, n, dist in CASES:
        tag = build_ref.tag_for(C, T, seed)
        sd = synth.make_state_dict(C=C, T=T, seed=seed)
        cfg = synth.make_config(C=C, T=T)
        exes = build_ref.build(tag, sd, cfg, flagsets=("default", "plain", "nosimd"))
        d = os.path.join(build_ref.OUT, tag)
        arrays = parse_net_c(d)
        x = synth.make_trials(n, C, T, seed=1000 + seed, dist=dist)

        fx = {"C": C, "T": T, "seed": seed, "x": x}
        fx.update({"arr." + k: v for k, v in arrays.items()})
        fx.update({"sd." + k: v for k, v in sd.items() if k.startswith("batch_norm") or not k.endswith(".weight")})
        for rb in (True, False):
            g = GoldenModel(os.path.join(d, "config.json"), os.path.join(d, "net.npz"),
                            clip_balanced=False, reorder_bn=rb)
            ys, taps = [], [[] for _ in g.layers]
            for i in range(n):
                h = x[i].astype(int)
                for li, layer in enumerate(g.layers):
                    h = layer(h)
                    taps[li].append(np.asarray(h).copy())
                ys.append(np.asarray(h))
            sfx = "reorder" if rb else "plain"
            fx[f"golden_{sfx}"] = np.stack(ys).astype(np.int8)
            for li in range(4):
                t = np.stack(taps[li])
                fx[f"golden_{sfx}_y{li + 1}"] = t.reshape(n, t.shape[1], -1).astype(np.int8)
        y_def, _ = build_ref.run(exes["default"], x)
        y_nos, _ = build_ref.run(exes["nosimd"], x)
        y_pln, _ = build_ref.run(exes["plain"], x)
        fx["c_default"], fx["c_nosimd"], fx["c_plain"] = y_def, y_nos, y_pln
        # the reference's own bit-exactness claim (test/cl/net/model/testcase.py:74-80): C == golden
        assert np.array_equal(y_def, fx["golden_reorder"]), tag
        assert np.array_equal(y_nos, fx["golden_reorder"]), tag
        np.savez_compressed(os.path.join(GOLDEN_DIR, tag + ".npz"), **fx)
        print(tag, "ok; logits[0] =", fx["golden_reorder"][0], " plain C==golden:",
              np.array_equal(y_pln, fx["golden_plain"]))


if __name__ == "__main__":
    main()
