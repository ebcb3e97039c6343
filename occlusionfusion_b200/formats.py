"""On-disk formats of the reference's result folders, byte-compatible:

  graph .bin files        /root/reference/utils/utils.py:205-321
      nodes / node_deformations   uint32 num_nodes, then num_nodes*3 float32, row-major
      edges / clusters            uint32 num_nodes, uint32 num_columns, then int32 values, row-major
      edges_weights               uint32 num_nodes, uint32 num_columns, then float32 values, row-major
  volume pickle           TSDFVolume.save_volume / load_volume (tsdf.py:682-702): lives on the class
  per-frame .npy files    deformed_nodes/%d.npy (warpfield.py:73, 418), visible_nodes/%d.npy (tsdf.py:634): written by
                          WarpField.update_transformations / TSDFVolume.get_visible_nodes

Host-side only (little-endian, native sizes -- what `struct.pack('I' / '=Nf' / '=Ni')` writes on x86-64); numpy
`tofile` / `fromfile` instead of the reference's per-element struct packing.
"""
import os

import numpy as np


def _save(filename, header, body, dtype):
    with open(filename, "wb") as fout:
        np.asarray(header, dtype="<u4").tofile(fout)
        np.ascontiguousarray(body, dtype=dtype).tofile(fout)


def _load(filename, n_header, dtype, cols=None):
    assert os.path.isfile(filename), "File not found: {}".format(filename)
    with open(filename, "rb") as fin:
        header = np.fromfile(fin, dtype="<u4", count=n_header)
        rows = int(header[0])
        ncol = cols if cols is not None else int(header[1])
        data = np.fromfile(fin, dtype=dtype, count=rows * ncol)
    return data.reshape(rows, ncol)


def save_graph_nodes(filename, nodes):                      # utils.py:220-227
    assert len(nodes.shape) == 2
    assert nodes.shape[1] == 3
    _save(filename, [nodes.shape[0]], nodes, "<f4")


def load_graph_nodes(filename):                             # utils.py:205-217
    return _load(filename, 1, "<f4", cols=3)


def save_graph_edges(filename, edges):                      # utils.py:246-253
    assert len(edges.shape) == 2
    _save(filename, [edges.shape[0], edges.shape[1]], edges, "<i4")


def load_graph_edges(filename):                             # utils.py:230-243
    return _load(filename, 2, "<i4")


def save_graph_edges_weights(filename, edges_weights):      # utils.py:270-277
    assert len(edges_weights.shape) == 2
    _save(filename, [edges_weights.shape[0], edges_weights.shape[1]], edges_weights, "<f4")


def load_graph_edges_weights(filename):                     # utils.py:256-268
    return _load(filename, 2, "<f4")


def save_graph_node_deformations(filename, node_deformations):   # utils.py:295-302
    assert len(node_deformations.shape) == 2
    assert node_deformations.shape[1] == 3
    _save(filename, [node_deformations.shape[0]], node_deformations, "<f4")


def load_graph_node_deformations(filename):                 # utils.py:280-292
    return _load(filename, 1, "<f4", cols=3)


def save_graph_clusters(filename, clusters):                # utils.py:314-321
    assert len(clusters.shape) == 2
    _save(filename, [clusters.shape[0], clusters.shape[1]], clusters, "<i4")


def load_graph_clusters(filename):                          # utils.py:300-312 (reads num_nodes values, one column)
    return _load(filename, 2, "<i4", cols=1)
