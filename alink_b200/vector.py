"""Vector (de)serialisation compatible with Alink's VectorUtil
(reference: core/src/main/java/com/alibaba/alink/common/linalg/VectorUtil.java:52-114,244).

Dense vectors are space- or comma-separated numbers ("1 2 3" / "1,2,3"); sparse vectors are
"$size$i:v i:v ..." (the "$size$" header is optional).  Sparse input is densified: adding the
x_i = 0 terms leaves Alink's sequential sums unchanged, so the results are bit-identical.  A column that holds
only sparse vectors can also be handed over as CSR arrays (rows_to_csr -> akm_upload_rows_csr): the expansion then
happens on the device and only the non-zeros cross PCIe.
"""
import numpy as np


def parse_vector(s, size=None):
    """VectorUtil.getVector for a string / sequence / ndarray. Returns a float64 ndarray or None."""
    if s is None:
        return None
    if isinstance(s, np.ndarray):
        return s.astype(np.float64, copy=False)
    if isinstance(s, (list, tuple)):
        return np.asarray(s, dtype=np.float64)
    t = str(s).strip()
    if t == "":
        return np.zeros(0, dtype=np.float64) if size is None else np.zeros(size, dtype=np.float64)
    if ":" in t or t.startswith("$"):
        n = size
        if t.startswith("$"):
            end = t.index("$", 1)
            n = int(t[1:end])
            t = t[end + 1:].strip()
        idx, val = [], []
        for tok in t.replace(",", " ").split():
            i, v = tok.split(":")
            idx.append(int(i))
            val.append(float(v))
        if n is None:
            n = (max(idx) + 1) if idx else 0
        out = np.zeros(n, dtype=np.float64)
        out[idx] = val
        return out
    return np.array([float(x) for x in t.replace(",", " ").split()], dtype=np.float64)


def rows_to_matrix(values, vector_size=None):
    """Stack parsed vectors into an n x d matrix; returns (X, null_mask)."""
    vecs = [parse_vector(v, vector_size) for v in values]
    mask = np.array([v is None for v in vecs], dtype=bool)
    sizes = {v.size for v in vecs if v is not None}
    if not sizes:
        return np.zeros((0, vector_size or 0), dtype=np.float64), mask
    d = vector_size or max(sizes)
    X = np.zeros((len(vecs), d), dtype=np.float64)
    for i, v in enumerate(vecs):
        if v is not None:
            if v.size > d:
                raise ValueError("vector of size %d exceeds vectorSize %d" % (v.size, d))
            X[i, : v.size] = v
    return X, mask


def is_sparse_text(v):
    """True for the textual form of a SparseVector ("$size$i:v ..." or "i:v ...")."""
    if not isinstance(v, str):
        return False
    t = v.strip()
    return t.startswith("$") or ":" in t


def rows_to_csr(values, vector_size=None):
    """Sparse vector strings -> (indptr int64[n+1], indices int32[nnz], data float64[nnz], d).
    d = vector_size, else the largest "$size$" header / index + 1 seen (VectorUtil.java:52-114)."""
    indptr = np.zeros(len(values) + 1, dtype=np.int64)
    idx, val = [], []
    d = 0
    for r, s in enumerate(values):
        t = str(s).strip()
        if t.startswith("$"):
            end = t.index("$", 1)
            d = max(d, int(t[1:end]))
            t = t[end + 1:].strip()
        row = {}
        for tok in t.replace(",", " ").split():
            i, v = tok.split(":")
            row[int(i)] = float(v)                      # a repeated index keeps its last value, like the dense parse
        for i in sorted(row):
            idx.append(i)
            val.append(row[i])
        if row:
            d = max(d, max(row) + 1)
        indptr[r + 1] = len(idx)
    if vector_size is not None:
        if d > vector_size:
            raise ValueError("vector of size %d exceeds vectorSize %d" % (d, vector_size))
        d = vector_size
    return indptr, np.asarray(idx, dtype=np.int32), np.asarray(val, dtype=np.float64), d
