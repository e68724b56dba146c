"""ORBVocabulary (DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>) over the C ABI: the BoW transform that
Frame::ComputeBoW (src/Frame.cc:721-728) and KeyFrame::ComputeBoW (src/KeyFrame.cc:105-114) run."""
import ctypes as C

import numpy as np

from ._lib import OrbfeError, lib
from .matcher import FeatureVectorCSR

TF_IDF, TF, IDF, BINARY = 0, 1, 2, 3
NORM_NONE, NORM_L1, NORM_L2 = 0, 1, 2


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class ORBVocabulary:
    def __init__(self, k, L, weighting, norm, parent, is_leaf, descriptors, weights, device=0):
        """Flattened tree: node 0 = root, then the nodes in vocabulary-file order."""
        self._lib = lib()
        parent = np.ascontiguousarray(parent, np.int32)
        is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        descriptors = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        weights = np.ascontiguousarray(weights, np.float64)
        if not (len(parent) == len(is_leaf) == len(descriptors) == len(weights)):
            raise ValueError("parent, is_leaf, descriptors and weights must have one row per node")
        h = C.c_void_p()
        st = self._lib.orbfe_vocabulary_create(k, L, weighting, norm, len(parent), _p(parent), _p(is_leaf), _p(descriptors),
                                               _p(weights), device, C.byref(h))
        if st < 0:
            raise OrbfeError(st, "orbfe_vocabulary_create")
        self._v = h

    @classmethod
    def loadFromTextFile(cls, path, norm=NORM_L1, device=0):
        self = cls.__new__(cls)
        self._lib = lib()
        h = C.c_void_p()
        st = self._lib.orbfe_vocabulary_load_text(str(path).encode(), norm, device, C.byref(h))
        if st < 0:
            raise OrbfeError(st, "orbfe_vocabulary_load_text")
        self._v = h
        return self

    def __del__(self):
        if getattr(self, "_v", None):
            self._lib.orbfe_vocabulary_destroy(self._v)
            self._v = None

    def size(self):
        return self._lib.orbfe_vocabulary_words(self._v)

    def nodes(self):
        return self._lib.orbfe_vocabulary_nodes(self._v)

    def stream(self):
        return self._lib.orbfe_vocabulary_stream(self._v)

    def synchronize(self):
        st = self._lib.orbfe_vocabulary_synchronize(self._v)
        if st < 0:
            raise OrbfeError(st, "orbfe_vocabulary_synchronize")

    def launch_count(self):
        return self._lib.orbfe_vocabulary_launch_count(self._v)

    def transform(self, descriptors, levelsup=4):
        """Returns (BowVector as (word ids, values), FeatureVectorCSR)."""
        d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        n = len(d)
        words, values = np.zeros(max(n, 1), np.uint32), np.zeros(max(n, 1), np.float64)
        nodes, off, idx = np.zeros(max(n, 1), np.uint32), np.zeros(n + 1, np.int32), np.zeros(max(n, 1), np.int32)
        nw, nn = C.c_int(0), C.c_int(0)
        st = self._lib.orbfe_bow_transform(self._v, _p(d), n, levelsup, _p(words), _p(values), C.byref(nw), _p(nodes), _p(off), _p(idx),
                                           C.byref(nn))
        if st < 0:
            raise OrbfeError(st, "orbfe_bow_transform")
        nw, nn = nw.value, nn.value
        fv = FeatureVectorCSR(nodes[:nn].copy(), off[:nn + 1].copy(), idx[:off[nn]].copy())
        return (words[:nw].copy(), values[:nw].copy()), fv

    def transform_batch_device(self, d_desc, frame_stride, d_n, n_max, batch, levelsup, d_words, d_values, d_nwords, d_node_ids,
                               d_offsets, d_indices, d_nnodes):
        st = self._lib.orbfe_bow_transform_batch_device(self._v, d_desc, frame_stride, d_n, n_max, batch, levelsup, d_words,
                                                        d_values, d_nwords, d_node_ids, d_offsets, d_indices, d_nnodes)
        if st < 0:
            raise OrbfeError(st, "orbfe_bow_transform_batch_device")
