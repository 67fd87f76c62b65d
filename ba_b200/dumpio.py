"""The on-disk interchange of the path: the reference's CSV dumps and a compact binary twin.

The only files the reference's Solve() writes are the ``options.write_reduced_camera_matrix``
dumps (reference src/BundleAdjuster.cpp:600-620): ``s.txt`` (reduced camera matrix, dense, upper
triangle meaningful), ``rhs.txt`` (rhs_p_sc), ``j_pr.txt`` / ``j_l.txt`` (pose and landmark
Jacobians of the projection residuals, already scaled by sqrt(w), densified with
Eigen::LoadDenseFromSparse) and ``r_pr.txt`` (sqrt(w)-scaled residuals), all in Eigen's
``kLongCsvFmt`` (reference include/ba/Utils.h:66): coefficients separated by ", ", rows by "\\n",
no prefix, suffix or trailing newline.  reference matlab/marginalization.py:47-52 reads them back
with ``numpy.loadtxt(delimiter=',')``; ``read_csv`` accepts exactly what that call accepts.

``bagpu_dump`` (include/ba_b200.h) writes ``s.txt`` and ``rhs.txt`` of the last build; the other
three files hold matrices the GPU path never forms (the build is matrix-free).

The binary twin (``*.bad``) exists because the CSV of S is 25 bytes per coefficient and dense:
the header is ``b"BADUMP1\\0"``, then uint32 kind (0 dense, 1 block-CSC), uint32 rows, uint32 cols,
uint32 block rows, uint32 block cols, uint64 nnzb (0 for dense), then for dense ``rows*cols``
float64 row-major, for block-CSC ``colptr[ncols_b+1]`` uint32, ``rowidx[nnzb]`` uint32 (ascending
per column, as reference include/ba/SparseBlockMatrix.h:108-112 stores them) and
``values[nnzb, br, bc]`` float64 row-major blocks.  Little endian throughout.
"""
import os
import struct
import numpy as np

MAGIC = b"BADUMP1\0"
KIND_DENSE, KIND_BLOCK_CSC = 0, 1
DUMP_FILES = ("s", "rhs", "j_pr", "r_pr", "j_l")


# ---- the reference's CSV format ------------------------------------------------------------------
def write_csv(path, a):
    """Eigen ``matrix.format(kLongCsvFmt)``; a 1-D array is a column vector (one value per row)."""
    a = np.asarray(a, dtype=np.float64)
    if a.ndim == 1:
        a = a.reshape(-1, 1)
    assert a.ndim == 2
    with open(path, "w") as f:
        f.write("\n".join(", ".join("%.17g" % v for v in row) for row in a))


def read_csv(path):
    """Inverse of write_csv; vectors come back 1-D, an empty file as an empty vector."""
    with open(path) as f:
        text = f.read()
    rows = [ln for ln in text.split("\n") if ln.strip()]
    if not rows:
        return np.zeros(0)
    a = np.array([[float(t) for t in ln.rstrip(";").split(",")] for ln in rows], dtype=np.float64)
    return a[:, 0].copy() if a.shape[1] == 1 else a


def read_dump(directory):
    """All dump files present in ``directory`` -> dict name -> array (names of DUMP_FILES)."""
    out = {}
    for name in DUMP_FILES:
        p = os.path.join(directory, name + ".txt")
        if os.path.exists(p):
            out[name] = read_csv(p)
    return out


# ---- block-CSC <-> dense (Eigen::LoadDenseFromSparse, reference SparseBlockMatrixOps.h:400-429) ----
def dense_to_block_csc(a, br, bc, pattern=None):
    """Blocks of ``a`` as (colptr, rowidx, values).  Without ``pattern`` every block with a non-zero
    coefficient is kept; with ``pattern=(colptr, rowidx)`` exactly those blocks (explicit zeros stay)."""
    a = np.asarray(a, dtype=np.float64)
    assert a.shape[0] % br == 0 and a.shape[1] % bc == 0
    nrb, ncb = a.shape[0] // br, a.shape[1] // bc
    blocks = a.reshape(nrb, br, ncb, bc).transpose(2, 0, 1, 3)          # [col block, row block, br, bc]
    if pattern is None:
        nz = np.any(blocks != 0, axis=(2, 3))                          # [ncb, nrb]
        colptr = np.concatenate([[0], np.cumsum(nz.sum(axis=1))]).astype(np.uint32)
        cols, rows = np.nonzero(nz)                                    # row-major over (col, row): ascending rows per column
    else:
        colptr = np.asarray(pattern[0], dtype=np.uint32)
        rows = np.asarray(pattern[1], dtype=np.int64)
        assert len(colptr) == ncb + 1 and colptr[-1] == len(rows)
        cols = np.repeat(np.arange(ncb), np.diff(colptr.astype(np.int64)))
    values = blocks[cols, rows] if len(rows) else np.zeros((0, br, bc))
    return colptr, np.asarray(rows, dtype=np.uint32), np.ascontiguousarray(values)


def block_csc_to_dense(nrb, colptr, rowidx, values):
    values = np.asarray(values, dtype=np.float64)
    ncb = len(colptr) - 1
    br, bc = (values.shape[1], values.shape[2]) if values.ndim == 3 and len(values) else (1, 1)
    a = np.zeros((nrb, br, ncb, bc))
    cols = np.repeat(np.arange(ncb), np.diff(np.asarray(colptr, dtype=np.int64)))
    if len(rowidx):
        a[np.asarray(rowidx, dtype=np.int64), :, cols, :] = values
    return a.reshape(nrb * br, ncb * bc)


# ---- the binary twin -----------------------------------------------------------------------------
def write_bin(path, a=None, block_csc=None, shape_blocks=None):
    """Dense: ``write_bin(path, a)``.  Block-CSC: ``write_bin(path, block_csc=(colptr, rowidx, values),
    shape_blocks=(row blocks, col blocks))``."""
    with open(path, "wb") as f:
        f.write(MAGIC)
        if block_csc is None:
            a = np.ascontiguousarray(a, dtype="<f8")
            vec = a.ndim == 1
            a2 = a.reshape(-1, 1) if vec else a
            f.write(struct.pack("<5IQ", KIND_DENSE, a2.shape[0], a2.shape[1], 1, 1, 0))
            f.write(a2.tobytes())
        else:
            colptr, rowidx, values = block_csc
            values = np.ascontiguousarray(values, dtype="<f8")
            nrb, ncb = shape_blocks
            br, bc = (values.shape[1], values.shape[2]) if values.ndim == 3 else (1, 1)
            assert len(colptr) == ncb + 1 and int(colptr[-1]) == len(rowidx) == len(values)
            f.write(struct.pack("<5IQ", KIND_BLOCK_CSC, nrb * br, ncb * bc, br, bc, len(rowidx)))
            f.write(np.ascontiguousarray(colptr, dtype="<u4").tobytes())
            f.write(np.ascontiguousarray(rowidx, dtype="<u4").tobytes())
            f.write(values.tobytes())


def read_bin(path, dense=True):
    """-> dense array (vectors 1-D), or with ``dense=False`` for a block-CSC file
    ``(colptr, rowidx, values, (row blocks, col blocks))``."""
    with open(path, "rb") as f:
        if f.read(8) != MAGIC:
            raise ValueError("%s: not a BADUMP1 file" % path)
        kind, rows, cols, br, bc, nnzb = struct.unpack("<5IQ", f.read(28))
        if kind == KIND_DENSE:
            a = np.frombuffer(f.read(rows * cols * 8), dtype="<f8").reshape(rows, cols).copy()
            return a[:, 0].copy() if cols == 1 else a
        if kind != KIND_BLOCK_CSC:
            raise ValueError("%s: unknown kind %d" % (path, kind))
        nrb, ncb = rows // br, cols // bc
        colptr = np.frombuffer(f.read((ncb + 1) * 4), dtype="<u4").copy()
        rowidx = np.frombuffer(f.read(nnzb * 4), dtype="<u4").copy()
        values = np.frombuffer(f.read(nnzb * br * bc * 8), dtype="<f8").reshape(nnzb, br, bc).copy()
    if dense:
        return block_csc_to_dense(nrb, colptr, rowidx, values)
    return colptr, rowidx, values, (nrb, ncb)


def csv_dump_to_bin(directory, pose_dim=6, lm_dim=1, s_pattern=None):
    """Convert the CSV files of ``directory`` to ``*.bad``: S as upper block-CSC of pose_dim x pose_dim
    blocks (``s_pattern`` = the symbolic pattern from ``get_pattern(PATTERN_S)`` keeps its explicit
    zeros), the Jacobians as block-CSC of 2x6 / 2xlm_dim blocks, the vectors dense."""
    d = read_dump(directory)
    out = {}
    for name, a in d.items():
        p = os.path.join(directory, name + ".bad")
        if a.ndim == 1:
            write_bin(p, a)
        else:
            br, bc = {"s": (pose_dim, pose_dim), "j_pr": (2, 6), "j_l": (2, max(1, lm_dim))}[name]
            src = np.triu(a) if name == "s" else a
            csc = dense_to_block_csc(src, br, bc, s_pattern if name == "s" else None)
            write_bin(p, block_csc=csc, shape_blocks=(a.shape[0] // br, a.shape[1] // bc))
        out[name] = p
    return out
