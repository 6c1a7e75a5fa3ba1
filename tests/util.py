"""Helpers shared by the test modules: golden fixtures, regenerated inputs, canonical orders."""
import os
import zlib

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
HEAD_CASES = ["c1_yolov2_dense", "c2_yolov3_dense_b2", "c2_yolov3_sparse_b4", "c3_person_dense_b1", "c4_coco608_b1"]


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def load(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)


def case_heads(g):
    """Regenerate the heads of a golden head case from its seed and verify the crc."""
    from ypp import synth
    heads = synth.make_heads(int(g["batch"]), int(g["nc"]), [int(x) for x in g["grids"]], A=int(g["A"]),
                             g=int(g["g"]), mu_obj=float(g["mu_obj"]))
    got = [crc(h.numpy()) for h in heads]
    assert got == [int(x) for x in g["head_crc"]], "regenerated heads differ from the fixture (torch version?)"
    return heads


def case_anchors(g):
    from ypp import synth
    if str(g["kind"]) == "v2":
        return synth.YOLOV2_SCALED_ANCHORS
    n = int(g["A"]) * len(g["grids"])
    return synth.YOLOV3_ANCHORS[:n] if n == 9 else synth.YOLOV3_ANCHORS[-n:]


def canon(keep, rows):
    """(score desc, index asc)."""
    k = np.asarray(keep, dtype=np.int64)
    order = np.lexsort((k, -rows[k, 4].astype(np.float64)))
    return k[order].astype(np.int32)


def rel_err(a, b):
    with np.errstate(invalid="ignore", divide="ignore"):
        fin = np.isfinite(b)
        return np.where(fin, np.abs(a - b) / np.maximum(np.abs(b), 1e-30), 0.0)
