"""On-disk graph format of the reference slicer.

Reader mirrors `/root/reference/cslicer/dataset.cpp:8-113` (Dataset: meta.txt
`key=value` lines + raw little-endian arrays); writer mirrors
`/root/reference/python/utils/convert_dgl_dataset.py:99-127`.

  meta.txt                num_nodes, num_edges, feature_dim, csum_features,
                          csum_labels, csum_offsets, csum_edges, num_classes
  indptr.bin              int64[N+1]
  indices.bin             int64[M]      rows sorted, no self loops
  features.bin            float32[N*F]
  labels.bin              int32[N]
  partition_map_opt.bin   int32[N]      (loaded by the reference, never used by
                                         its slicer: owner = id % 4,
                                         pyfrontend.cpp:57)
"""
import os

import numpy as np

META_KEYS = ("num_nodes", "num_edges", "feature_dim", "csum_features", "csum_labels",
             "csum_offsets", "csum_edges", "num_classes")


class Dataset:
    """Host-side graph container (fields named as in cslicer/dataset.h:17-48)."""

    def __init__(self, directory=None, *, indptr=None, indices=None, features=None,
                 labels=None, partition_map=None, verify=True, load_node_data=True):
        if directory is not None:
            self._read(directory, verify, load_node_data)
        else:
            self.BIN_DIR = None
            self.indptr = np.ascontiguousarray(indptr, dtype=np.int64)
            self.indices = np.ascontiguousarray(indices, dtype=np.int64)
            self.num_nodes = int(self.indptr.size - 1)
            self.num_edges = int(self.indices.size)
            self.features = features
            self.labels = labels
            self.partition_map = partition_map
            self.fsize = 0 if features is None else int(features.shape[1])
            self.noClasses = 0
        if self.indptr.size != self.num_nodes + 1 or self.indptr[0] != 0 or \
                int(self.indptr[-1]) != self.num_edges:
            raise ValueError("indptr does not describe num_nodes rows / num_edges edges")

    # dataset.cpp:70-113
    def _read_meta(self, directory):
        meta = {}
        with open(os.path.join(directory, "meta.txt")) as f:
            for line in f:
                line = line.strip()
                if not line or "=" not in line:
                    continue
                k, v = line.split("=", 1)
                meta[k] = int(float(v))  # stoll on the reference side
        for k in ("num_nodes", "num_edges"):
            if k not in meta:
                raise ValueError(f"meta.txt lacks {k}")
        return meta

    def _read(self, directory, verify, load_node_data):
        self.BIN_DIR = directory
        m = self._read_meta(directory)
        self.num_nodes, self.num_edges = m["num_nodes"], m["num_edges"]
        self.fsize = m.get("feature_dim", 0)
        self.noClasses = m.get("num_classes", 0)
        # dataset.cpp:18-36
        self.indptr = np.fromfile(os.path.join(directory, "indptr.bin"), dtype="<i8")
        self.indices = np.fromfile(os.path.join(directory, "indices.bin"), dtype="<i8")
        if self.indptr.size != self.num_nodes + 1 or self.indices.size != self.num_edges:
            raise ValueError("indptr.bin / indices.bin size disagrees with meta.txt")
        if verify:
            # the reference's checksum asserts (dataset.cpp:27,35), compiled out
            # under -DNDEBUG there, enforced here
            if "csum_offsets" in m and int(self.indptr.sum()) != m["csum_offsets"]:
                raise ValueError("indptr checksum mismatch")
            if "csum_edges" in m and int(self.indices.sum()) != m["csum_edges"]:
                raise ValueError("indices checksum mismatch")
        self.features = self.labels = self.partition_map = None
        if load_node_data:
            # dataset.cpp:38-68
            fp = os.path.join(directory, "features.bin")
            if os.path.exists(fp) and self.fsize:
                self.features = np.fromfile(fp, dtype="<f4").reshape(self.num_nodes, self.fsize)
            lp = os.path.join(directory, "labels.bin")
            if os.path.exists(lp):
                self.labels = np.fromfile(lp, dtype="<i4")
            pp = os.path.join(directory, "partition_map_opt.bin")
            if os.path.exists(pp):
                self.partition_map = np.fromfile(pp, dtype="<i4")


def write_dataset(directory, indptr, indices, features=None, labels=None,
                  partition_map=None, num_classes=2):
    """Write a graph in the layout Dataset (and the reference's Dataset) reads.

    Default node data is the cheapest the reference loader accepts:
    feature_dim=1 zeros... features are never read by the slicer.
    """
    os.makedirs(directory, exist_ok=True)
    indptr = np.ascontiguousarray(indptr, dtype="<i8")
    indices = np.ascontiguousarray(indices, dtype="<i8")
    n = indptr.size - 1
    if features is None:
        features = np.ones((n, 1), dtype="<f4")
    features = np.ascontiguousarray(features, dtype="<f4")
    if labels is None:
        labels = (np.arange(n) % num_classes).astype("<i4")
    if partition_map is None:
        partition_map = (np.arange(n) % 4).astype("<i4")
    indptr.tofile(os.path.join(directory, "indptr.bin"))
    indices.tofile(os.path.join(directory, "indices.bin"))
    features.tofile(os.path.join(directory, "features.bin"))
    np.ascontiguousarray(labels, dtype="<i4").tofile(os.path.join(directory, "labels.bin"))
    np.ascontiguousarray(partition_map, dtype="<i4").tofile(
        os.path.join(directory, "partition_map_opt.bin"))
    meta = dict(num_nodes=n, num_edges=int(indices.size), feature_dim=int(features.shape[1]),
                csum_features=int(features.sum(dtype=np.float64)),
                csum_labels=int(np.asarray(labels, dtype=np.int64).sum()),
                csum_offsets=int(indptr.sum()), csum_edges=int(indices.sum()),
                num_classes=int(num_classes))
    with open(os.path.join(directory, "meta.txt"), "w") as f:
        for k in META_KEYS:
            f.write(f"{k}={meta[k]}\n")
    return meta
