#!/usr/bin/env python
"""Generates tests/golden/bow_*.npz with the Python restatement of the DBoW2 transform (oracle/py/bow_py.py) on
the synthetic vocabularies of tests/bow_cases.py.  TEST INFRASTRUCTURE.  Run: python oracle/py/make_golden_bow.py
A fixture stores the expected outputs and a CRC of the inputs; the inputs are rebuilt from seeds."""
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [HERE, os.path.join(ROOT, "orb-slam3-noted_b200"), os.path.join(ROOT, "tests")]

import bow_cases as bc  # noqa: E402
import bow_py  # noqa: E402


def crc(*arrays):
    c = 0
    for a in arrays:
        c = zlib.crc32(np.ascontiguousarray(a).tobytes(), c)
    return np.uint32(c & 0xFFFFFFFF)


def main():
    out = os.path.join(ROOT, "tests", "golden")
    for name in bc.BOW_VOC_CASES:
        c = bc.voc_inputs(name)
        v = bow_py.Vocabulary(c["k"], c["L"], c["weighting"], c["norm"])
        for i in range(1, len(c["parent"])):
            v.add_node(int(c["parent"][i]), bool(c["leaf"][i]), c["desc"][i], c["weight"][i])
        bow, fv = v.transform(c["feat"], c["levelsup"])
        words = np.array([w for w, _ in bow], np.uint32)
        values = np.array([x for _, x in bow], np.float64)
        nodes = np.array([n for n, _ in fv], np.uint32)
        offsets = np.cumsum([0] + [len(ix) for _, ix in fv]).astype(np.int32)
        indices = np.array([i for _, ix in fv for i in ix], np.int32)
        np.savez_compressed(os.path.join(out, f"bow_{name}.npz"), in_crc=crc(c["parent"], c["desc"], c["weight"], c["feat"]),
                            words=words, values=values, nodes=nodes, offsets=offsets, indices=indices)
        print(name, len(c["parent"]), "nodes,", len(v.words), "words ->", len(words), "bow entries,", len(nodes), "fv nodes,",
              len(indices), "of", len(c["feat"]), "features kept")


if __name__ == "__main__":
    main()
