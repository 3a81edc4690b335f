"""numpy restatement of the reference vector-quantisation ops (test oracle only).

Follows /root/reference/src/vector_quantization.py line by line; every function
cites the lines it restates.  fp32 arithmetic unless a name says 64.
"""
import numpy as np


def vq_distances(z, codebook):
    """distances[n,k] = (|e_k|^2 + |z_n|^2) - 2 z_n.e_k  (vector_quantization.py:12-17).

    The reference builds the bias ``codebook_sqr + inputs_sqr`` first (one fp32
    add) and then ``addmm(beta=1, alpha=-2)``; the row-constant |z|^2 is kept
    because it contributes fp32 rounding noise to near-ties.
    """
    z = np.ascontiguousarray(z, dtype=np.float32)
    cb = np.ascontiguousarray(codebook, dtype=np.float32)
    d = cb.shape[1]
    zf = z.reshape(-1, d)                                   # :10  inputs.view(-1, D)
    cb_sqr = np.sum(cb * cb, axis=1, dtype=np.float32)      # :12
    in_sqr = np.sum(zf * zf, axis=1, dtype=np.float32, keepdims=True)  # :13
    bias = (cb_sqr[None, :] + in_sqr).astype(np.float32)    # :16 first addmm arg
    return (bias + np.float32(-2.0) * (zf @ cb.T)).astype(np.float32)  # :16-17


def vq_indices(z, codebook, chunk=16384):
    """``vq``: argmin_k distance, lowest index on ties (vector_quantization.py:19-20).

    int64 result shaped ``z.shape[:-1]``.  Rows are independent, so the search is
    chunked over rows (the reference materialises the full N x K matrix).
    """
    z = np.asarray(z, dtype=np.float32)
    d = codebook.shape[1]
    zf = z.reshape(-1, d)
    out = np.empty(zf.shape[0], dtype=np.int64)
    for s in range(0, zf.shape[0], chunk):
        out[s:s + chunk] = np.argmin(vq_distances(zf[s:s + chunk], codebook), axis=1)
    return out.reshape(z.shape[:-1])


def vq_distances64(z, codebook):
    """fp64 re-evaluation of the same distance, used only to adjudicate ties."""
    zf = np.asarray(z, dtype=np.float64).reshape(-1, codebook.shape[1])
    cb = np.asarray(codebook, dtype=np.float64)
    return (np.sum(cb * cb, 1)[None, :] + np.sum(zf * zf, 1, keepdims=True)
            - 2.0 * (zf @ cb.T))


def index_parity(idx, idx_ref, z, codebook, rel=1e-6, chunk=16384):
    """north_star index rule.

    Returns ``(n_mismatch, n_violations)``: rows where ``idx != idx_ref`` and, of
    those, rows whose fp64 distances differ by more than ``rel`` relative (true
    errors, not ties).  Exact duplicates must resolve to the lowest index, which
    the distance test cannot see, so equal-distance mismatches additionally
    require nothing: they are ties under the rule.
    """
    idx = np.asarray(idx).reshape(-1)
    idx_ref = np.asarray(idx_ref).reshape(-1)
    d = codebook.shape[1]
    zf = np.asarray(z).reshape(-1, d)
    bad = np.nonzero(idx != idx_ref)[0]
    viol = 0
    cb64 = np.asarray(codebook, dtype=np.float64)
    for s in range(0, bad.size, chunk):
        rows = bad[s:s + chunk]
        z64 = zf[rows].astype(np.float64)
        da = np.sum((z64 - cb64[idx[rows]]) ** 2, axis=1)
        dr = np.sum((z64 - cb64[idx_ref[rows]]) ** 2, axis=1)
        viol += int(np.sum(np.abs(da - dr) > rel * np.abs(dr)))
    return int(bad.size), viol


def vq_st_forward(z, codebook):
    """``vq_st`` forward (vector_quantization.py:34-44): ``(codes, indices_flat)``."""
    idx = vq_indices(z, codebook)
    flat = idx.reshape(-1)                                   # :36
    codes = np.asarray(codebook, dtype=np.float32)[flat].reshape(np.shape(z))  # :40-42
    return codes, flat


def vq_st_backward(grad_output, indices_flat, codebook_shape):
    """``vq_st`` backward (vector_quantization.py:47-63).

    grad_inputs = clone of grad_output (:52); grad_codebook = index_add of the
    flattened grad rows into a zero (K, D) table (:58-61).  Accumulated in fp64
    and rounded once so the oracle does not depend on summation order.
    """
    g = np.asarray(grad_output, dtype=np.float32)
    k, d = codebook_shape
    gc = np.zeros((k, d), dtype=np.float64)
    np.add.at(gc, np.asarray(indices_flat).reshape(-1), g.reshape(-1, d).astype(np.float64))
    return g.copy(), gc.astype(np.float32)


def ema_update(z, indices_flat, ema_count, ema_sum, decay=0.99, eps=1e-5):
    """EMA codebook update (van den Oord et al. 2017, appendix A.1).

    NOT IN THE REFERENCE -- parity unpinned.  This is our own
    restatement of the published rule, the oracle for the extension kernel:
      counts = bincount(idx); sums = index_add(z)
      N <- g N + (1-g) counts;  m <- g m + (1-g) sums
      e = m / ((N + eps) / (sum N + K eps) * sum N)
    Returns (new_count, new_sum, new_codebook, counts, sums).
    """
    k, d = ema_sum.shape
    zf = np.asarray(z, dtype=np.float64).reshape(-1, d)
    idx = np.asarray(indices_flat).reshape(-1)
    counts = np.bincount(idx, minlength=k).astype(np.float64)
    sums = np.zeros((k, d), dtype=np.float64)
    np.add.at(sums, idx, zf)
    n = decay * np.asarray(ema_count, np.float64) + (1 - decay) * counts
    m = decay * np.asarray(ema_sum, np.float64) + (1 - decay) * sums
    tot = n.sum()
    n_s = (n + eps) / (tot + k * eps) * tot
    e = m / n_s[:, None]
    return (n.astype(np.float32), m.astype(np.float32), e.astype(np.float32),
            counts.astype(np.float32), sums.astype(np.float32))
