"""LIBSVM text -> device buffers (utility/io/libsvm_io.hpp:33-926 reads the same format into
Elemental / sparse matrices; python-skylark/skylark/io.py wraps it).  One data point per line:
`label idx:val idx:val ...` with 1-based feature indices."""
import numpy


def read_libsvm(path, n_features=None, device="cuda", sparse=False):
    """Returns (X, y).  X: float64 tensor (points x features, row-major) on `device`, or with
    sparse=True a lib.DeviceCSC of the transposed matrix (features x points, one point per column --
    the layout utility/io/libsvm_io.hpp:433-620 produces) ready for `apply(..., "columnwise")`."""
    import torch
    labels, rows, cols, vals = [], [], [], []
    with open(path) as f:
        for r, line in enumerate(l for l in f if l.strip() and not l.startswith("#")):
            parts = line.split()
            labels.append(float(parts[0]))
            for tok in parts[1:]:
                i, v = tok.split(":")
                rows.append(r); cols.append(int(i) - 1); vals.append(float(v))
    npts = len(labels)
    d = (max(cols) + 1 if cols else 0) if n_features is None else int(n_features)
    y = torch.tensor(labels, dtype=torch.float64, device=device)
    if not sparse:
        X = numpy.zeros((npts, d))
        X[rows, cols] = vals
        return torch.from_numpy(X).to(device), y
    import scipy.sparse
    from .. import lib
    M = scipy.sparse.csc_matrix((vals, (cols, rows)), shape=(d, npts))     # features x points
    M.sort_indices()
    return lib.DeviceCSC(torch.from_numpy(M.indptr.astype(numpy.int32)).to(device),
                         torch.from_numpy(M.indices.astype(numpy.int32)).to(device),
                         torch.from_numpy(M.data.astype(numpy.float64)).to(device), (d, npts)), y
